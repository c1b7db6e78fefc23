"""Seeded weight sets with the reference's state-dict names and shapes.

TEST INFRASTRUCTURE ONLY.  Trained checkpoints are Git-LFS pointers in the reference
(results/test001/checkpoints/epoch.30.ckpt), so every parity run uses random-init
weights.  A 15.7 M-parameter state dict is too large to commit as a fixture, so the
fixtures store only (kind, seed) and both sides regenerate the tensors from a
``torch.Generator`` -- deterministic for a given torch build; ``checksum`` lets a
fixture detect a mismatch.

Names/shapes follow ``Denoiser.state_dict()`` (model/model.py:77-123 and the module
constructors it calls); ``make_golden.py`` asserts the key set equals the live
reference's.

kinds
  'activated'  every matrix ~ N(0, 1/fan_in), every Linear bias ~ N(0, 0.1^2), LayerNorm
               weight ~ 1 + N(0, 0.1^2), LayerNorm bias ~ N(0, 0.1^2), IPA head_weights ~
               softplus^-1(1) + N(0, 0.3^2).  Exercises every term of the forward pass
               (the stock init zeroes 222 of 396 tensors).
  'stock_like' the reference's init *rules* (primitives.py:31-160): 'final' -> 0,
               'gating' -> w 0 / b 1, default/relu -> truncated normal with the
               reference's (accidentally tiny) fan computation, nn.Linear default for
               the plain nn.Linear layers, LayerNorm 1/0, head_weights softplus^-1(1).
               Not bit-identical to a reference construction (that draws from scipy's
               global numpy RNG) -- the fixtures load THIS dict into the reference.
"""
import math

import torch


def schema(cfg):
    """[(name, shape, role)] in state-dict order.  role drives the init."""
    out = []

    def lin(name, o, i, bias=True, init='default'):
        out.append((name + '.weight', (o, i), 'w:' + init))
        if bias:
            out.append((name + '.bias', (o,), 'b:' + init))

    def ln(name, c):
        out.append((name + '.weight', (c,), 'ln_w'))
        out.append((name + '.bias', (c,), 'ln_b'))

    c_s, c_p = cfg.c_s, cfg.c_p
    lin('single_feature_net.linear', c_s, cfg.c_pos_emb + cfg.c_chain_emb + cfg.c_timestep_emb + 23, False, 'nn')
    lin('pair_feature_net.linear_s_p_i', c_p, c_s, False, 'nn')
    lin('pair_feature_net.linear_s_p_j', c_p, c_s, False, 'nn')
    lin('pair_feature_net.linear_relpos', c_p, 2 * cfg.relpos_k + 3, False, 'nn')
    lin('pair_feature_net.linear_template', c_p, cfg.template_dist_n_bin + 6, False, 'nn')
    lin('pair_feature_net.linear_motif_template', c_p, cfg.template_dist_n_bin + 2, False, 'nn')
    for l in range(cfg.n_pair_transform_layer):
        pre = 'pair_transform_net.net.%d' % l
        if cfg.include_mul_update:
            for d in ('tri_mul_out', 'tri_mul_in'):
                q = '%s.%s' % (pre, d)
                lin(q + '.linear_a_p', cfg.c_hidden_mul, c_p)
                lin(q + '.linear_a_g', cfg.c_hidden_mul, c_p, init='gating')
                lin(q + '.linear_b_p', cfg.c_hidden_mul, c_p)
                lin(q + '.linear_b_g', cfg.c_hidden_mul, c_p, init='gating')
                lin(q + '.linear_g', c_p, c_p, init='gating')
                lin(q + '.linear_z', c_p, cfg.c_hidden_mul, init='final')
                ln(q + '.layer_norm_in', c_p)
                ln(q + '.layer_norm_out', cfg.c_hidden_mul)
        if cfg.include_tri_att:
            hc = cfg.c_hidden_tri_att * cfg.n_head_tri
            for d in ('tri_att_start', 'tri_att_end'):
                q = '%s.%s' % (pre, d)
                ln(q + '.layer_norm', c_p)
                lin(q + '.linear', cfg.n_head_tri, c_p, False, 'normal')
                lin(q + '.mha.linear_q', hc, c_p, False, 'glorot')
                lin(q + '.mha.linear_k', hc, c_p, False, 'glorot')
                lin(q + '.mha.linear_v', hc, c_p, False, 'glorot')
                lin(q + '.mha.linear_o', c_p, hc, init='final')
                lin(q + '.mha.linear_g', hc, c_p, init='gating')
        q = pre + '.pair_transition'
        ln(q + '.layer_norm', c_p)
        lin(q + '.linear_1', cfg.pair_transition_n * c_p, c_p, init='relu')
        lin(q + '.linear_2', c_p, cfg.pair_transition_n * c_p, init='final')
    H, C, Pq, Pv = cfg.n_head_ipa, cfg.c_hidden_ipa, cfg.n_qk_point, cfg.n_v_point
    for l in range(cfg.n_structure_layer):
        pre = 'structure_net.net.%d' % l
        out.append((pre + '.ipa.head_weights', (H,), 'head_w'))
        lin(pre + '.ipa.linear_q', H * C, c_s)
        lin(pre + '.ipa.linear_kv', 2 * H * C, c_s)
        lin(pre + '.ipa.linear_q_points', H * Pq * 3, c_s)
        lin(pre + '.ipa.linear_kv_points', H * (Pq + Pv) * 3, c_s)
        lin(pre + '.ipa.linear_b', H, c_p)
        lin(pre + '.ipa.linear_out', c_s, H * (c_p + C + Pv * 4), init='final')
        ln(pre + '.ipa_layer_norm', c_s)
        for t in range(cfg.n_structure_transition_layer):
            q = '%s.transition.layers.%d' % (pre, t)
            lin(q + '.linear_1', c_s, c_s, init='relu')
            lin(q + '.linear_2', c_s, c_s, init='relu')
            lin(q + '.linear_3', c_s, c_s, init='final')
        ln(pre + '.transition.layer_norm', c_s)
        lin(pre + '.bb_update.linear', 6, c_s)
    return out


_TRUNC_STD = 0.87962566103423978   # std of a unit normal truncated to [-2, 2]


def _trunc_normal(shape, std, g):
    t = torch.empty(shape)
    torch.nn.init.trunc_normal_(t, mean=0.0, std=std, a=-2 * std, b=2 * std, generator=g)
    return t


def make_state_dict(cfg, seed=0, kind='activated'):
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for name, shape, role in schema(cfg):
        if kind == 'activated':
            if role.startswith('w:'):
                t = torch.randn(shape, generator=g) / math.sqrt(shape[1])
            elif role.startswith('b:'):
                t = 0.1 * torch.randn(shape, generator=g)
            elif role == 'ln_w':
                t = 1.0 + 0.1 * torch.randn(shape, generator=g)
            elif role == 'ln_b':
                t = 0.1 * torch.randn(shape, generator=g)
            elif role == 'head_w':
                t = 0.541324854612918 + 0.3 * torch.randn(shape, generator=g)
        elif kind == 'stock_like':
            init = role.split(':')[-1]
            if role == 'ln_w':
                t = torch.ones(shape)
            elif role == 'ln_b':
                t = torch.zeros(shape)
            elif role == 'head_w':
                t = torch.full(shape, 0.541324854612918)
            elif role.startswith('b:'):
                t = torch.ones(shape) if init == 'gating' else torch.zeros(shape)
            elif init in ('final', 'gating'):
                t = torch.zeros(shape)
            elif init in ('default', 'relu'):
                # primitives.py:31-61: fan_in = prod(shape[:2]) * shape[0]
                f = shape[0] * shape[1] * shape[0]
                scale = (2.0 if init == 'relu' else 1.0) / max(1, f)
                t = _trunc_normal(shape, math.sqrt(scale) / _TRUNC_STD, g)
            elif init == 'nn':        # torch.nn.Linear default: U(-1/sqrt(fan_in), 1/sqrt(fan_in))
                bound = 1.0 / math.sqrt(shape[1])
                t = (torch.rand(shape, generator=g) * 2 - 1) * bound
            elif init == 'glorot':
                bound = math.sqrt(6.0 / (shape[0] + shape[1]))
                t = (torch.rand(shape, generator=g) * 2 - 1) * bound
            elif init == 'normal':
                t = torch.randn(shape, generator=g) / math.sqrt(shape[1])
        else:
            raise ValueError(kind)
        sd[name] = t.float().contiguous()
    return sd


def checksum(sd):
    """Order-dependent fp64 digest of a state dict (detects RNG drift across boxes)."""
    acc = 0.0
    for i, (k, v) in enumerate(sd.items()):
        acc += float(v.double().sum()) * (1 + (i % 7)) + float(v.double().abs().sum())
    return acc
