"""Drop-in ``Denoiser`` (reference: genie/model/model.py:12-192).

Same keyword constructor, same ``forward(ts, timesteps, features)`` signature, same
``state_dict()`` key names and shapes (396 tensors in the base configuration), so
``Denoiser(**config.model, n_timestep=..., max_n_res=..., max_n_chain=...)``
(genie/diffusion/ddpm.py:26-31) and ``load_state_dict`` of a reference checkpoint keep
working.  The forward pass is one call into the CUDA library (``gb_denoise``); there is
no PyTorch implementation behind it and no CPU fallback.  Forward-only: inputs that
require grad raise (the reference's SMC samplers that backprop through the denoiser are
out of scope).
"""
import ctypes as C
import math

import torch
from torch import nn

from . import _lib
from .affine import T


def _schema(k):
    """[(state-dict name, shape, init rule)] in the reference's registration order."""
    out = []

    def lin(name, o, i, bias=True, init='default'):
        out.append((name + '.weight', (o, i), 'w:' + init))
        if bias:
            out.append((name + '.bias', (o,), 'b:' + init))

    def ln(name, c):
        out.append((name + '.weight', (c,), 'ln_w'))
        out.append((name + '.bias', (c,), 'ln_b'))

    c_s, c_p = k['c_s'], k['c_p']
    lin('single_feature_net.linear', c_s, k['c_pos_emb'] + k['c_chain_emb'] + k['c_timestep_emb'] + 23, False, 'nn')
    lin('pair_feature_net.linear_s_p_i', c_p, c_s, False, 'nn')
    lin('pair_feature_net.linear_s_p_j', c_p, c_s, False, 'nn')
    lin('pair_feature_net.linear_relpos', c_p, 2 * k['relpos_k'] + 3, False, 'nn')
    lin('pair_feature_net.linear_template', c_p, k['template_dist_n_bin'] + 6, False, 'nn')
    lin('pair_feature_net.linear_motif_template', c_p, k['template_dist_n_bin'] + 2, False, 'nn')
    for l in range(k['n_pair_transform_layer']):
        pre = 'pair_transform_net.net.%d' % l
        if k['include_mul_update']:
            for d in ('tri_mul_out', 'tri_mul_in'):
                q = '%s.%s' % (pre, d)
                lin(q + '.linear_a_p', k['c_hidden_mul'], c_p)
                lin(q + '.linear_a_g', k['c_hidden_mul'], c_p, init='gating')
                lin(q + '.linear_b_p', k['c_hidden_mul'], c_p)
                lin(q + '.linear_b_g', k['c_hidden_mul'], c_p, init='gating')
                lin(q + '.linear_g', c_p, c_p, init='gating')
                lin(q + '.linear_z', c_p, k['c_hidden_mul'], init='final')
                ln(q + '.layer_norm_in', c_p)
                ln(q + '.layer_norm_out', k['c_hidden_mul'])
        if k['include_tri_att']:
            hc = k['c_hidden_tri_att'] * k['n_head_tri']
            for d in ('tri_att_start', 'tri_att_end'):
                q = '%s.%s' % (pre, d)
                ln(q + '.layer_norm', c_p)
                lin(q + '.linear', k['n_head_tri'], c_p, False, 'normal')
                lin(q + '.mha.linear_q', hc, c_p, False, 'glorot')
                lin(q + '.mha.linear_k', hc, c_p, False, 'glorot')
                lin(q + '.mha.linear_v', hc, c_p, False, 'glorot')
                lin(q + '.mha.linear_o', c_p, hc, init='final')
                lin(q + '.mha.linear_g', hc, c_p, init='gating')
        q = pre + '.pair_transition'
        ln(q + '.layer_norm', c_p)
        lin(q + '.linear_1', k['pair_transition_n'] * c_p, c_p, init='relu')
        lin(q + '.linear_2', c_p, k['pair_transition_n'] * c_p, init='final')
    H, Ch, Pq, Pv = k['n_head_ipa'], k['c_hidden_ipa'], k['n_qk_point'], k['n_v_point']
    for l in range(k['n_structure_layer']):
        pre = 'structure_net.net.%d' % l
        out.append((pre + '.ipa.head_weights', (H,), 'head_w'))
        lin(pre + '.ipa.linear_q', H * Ch, c_s)
        lin(pre + '.ipa.linear_kv', 2 * H * Ch, c_s)
        lin(pre + '.ipa.linear_q_points', H * Pq * 3, c_s)
        lin(pre + '.ipa.linear_kv_points', H * (Pq + Pv) * 3, c_s)
        lin(pre + '.ipa.linear_b', H, c_p)
        lin(pre + '.ipa.linear_out', c_s, H * (c_p + Ch + Pv * 4), init='final')
        ln(pre + '.ipa_layer_norm', c_s)
        for t in range(k['n_structure_transition_layer']):
            q = '%s.transition.layers.%d' % (pre, t)
            lin(q + '.linear_1', c_s, c_s, init='relu')
            lin(q + '.linear_2', c_s, c_s, init='relu')
            lin(q + '.linear_3', c_s, c_s, init='final')
        ln(pre + '.transition.layer_norm', c_s)
        lin(pre + '.bb_update.linear', 6, c_s)
    return out


def _init_tensor(shape, rule):
    """The reference's initialisers (modules/primitives.py:31-160) re-stated."""
    kind = rule.split(':')[-1]
    if rule == 'ln_w':
        return torch.ones(shape)
    if rule == 'ln_b':
        return torch.zeros(shape)
    if rule == 'head_w':
        return torch.full(shape, 0.541324854612918)            # softplus^-1(1), primitives.py:90-93
    if rule.startswith('b:'):
        return torch.ones(shape) if kind == 'gating' else torch.zeros(shape)
    if kind in ('final', 'gating'):
        return torch.zeros(shape)
    if kind in ('default', 'relu'):                            # truncated normal, the reference's fan rule
        f = shape[0] * shape[1] * shape[0]
        std = math.sqrt((2.0 if kind == 'relu' else 1.0) / max(1, f)) / 0.87962566103423978
        t = torch.empty(shape)
        nn.init.trunc_normal_(t, 0.0, std, -2 * std, 2 * std)
        return t
    t = torch.empty(shape)
    if kind == 'glorot':
        nn.init.xavier_uniform_(t, gain=1)
    elif kind == 'normal':
        nn.init.kaiming_normal_(t, nonlinearity='linear')
    else:                                                      # plain nn.Linear default
        nn.init.kaiming_uniform_(t, a=math.sqrt(5))
    return t


class _Node(nn.Module):
    """Parameter container; the tree of these reproduces the reference's module paths."""


_DIM_KEYS = ['c_s', 'c_p', 'n_timestep', 'rescale', 'c_pos_emb', 'c_chain_emb', 'c_timestep_emb', 'max_n_res',
             'max_n_chain', 'relpos_k', 'template_dist_min', 'template_dist_step', 'template_dist_n_bin',
             'n_pair_transform_layer', 'include_mul_update', 'include_tri_att', 'c_hidden_mul', 'c_hidden_tri_att',
             'n_head_tri', 'pair_transition_n', 'n_structure_layer', 'n_structure_block', 'c_hidden_ipa', 'n_head_ipa',
             'n_qk_point', 'n_v_point', 'n_structure_transition_layer']


class Denoiser(nn.Module):
    """SE(3)-equivariant denoiser: predicts the noise added at timestep t (model.py:12-19)."""

    def __init__(self, c_s, c_p, n_timestep, rescale,
                 c_pos_emb, c_chain_emb, c_timestep_emb, max_n_res, max_n_chain,
                 relpos_k, template_dist_min, template_dist_step, template_dist_n_bin,
                 n_pair_transform_layer, include_mul_update, include_tri_att, c_hidden_mul, c_hidden_tri_att,
                 n_head_tri, tri_dropout, pair_transition_n,
                 n_structure_layer, n_structure_block, c_hidden_ipa, n_head_ipa, n_qk_point, n_v_point, ipa_dropout,
                 n_structure_transition_layer, structure_transition_dropout):
        super().__init__()
        loc = locals()
        self.dims = {k: loc[k] for k in _DIM_KEYS}
        self.rescale = rescale
        self.precision = 'fp32'
        self.return_intermediates = False   # True: also return 's', 'p', 'states' (model.py:186-192)
        for name, shape, rule in _schema(self.dims):
            parts = name.split('.')
            node = self
            for part in parts[:-1]:
                if not hasattr(node, part):
                    node.add_module(part, _Node())
                node = getattr(node, part)
            node.register_parameter(parts[-1], nn.Parameter(_init_tensor(shape, rule), requires_grad=False))
        self._ctx = None
        self._ctx_device = None
        self._dirty = True
        self._feat_key = None
        self._keep = None
        self._shard = None                  # (world, rank, 128-byte NCCL unique id) when row-sharded

    # ---- weight / context life cycle -------------------------------------------------
    def _apply(self, fn, *a, **k):
        self._dirty = True
        return super()._apply(fn, *a, **k)

    def _load_from_state_dict(self, *a, **k):
        self._dirty = True
        return super()._load_from_state_dict(*a, **k)

    def load_state_dict(self, *a, **k):
        self._dirty = True
        return super().load_state_dict(*a, **k)

    def refresh_weights(self):
        """Force a re-upload after in-place edits of parameters."""
        self._dirty = True

    def set_precision(self, precision):
        if precision not in _lib.PRECISIONS:
            raise ValueError('precision must be one of %s' % sorted(_lib.PRECISIONS))
        self.precision = precision
        if self._ctx is not None:
            _lib.check(_lib.load().gb_set_precision(self._ctx, _lib.PRECISIONS[precision]), self._ctx, 'gb_set_precision')

    def shard_rows(self, group=None):
        """Row-shard every following forward over the ranks of ``group`` (default: the world): ONE sample
        (B = 1, N a multiple of 128 * world) is split by rows of the pair representation, each rank on its
        own GPU; all ranks must make the same calls with the same inputs and all receive the full result
        (csrc/shard.cu).  Collective.  Needs precision 'bf16'."""
        import torch.distributed as dist
        from . import shard
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        self._shard = (world, rank, shard.exchange_unique_id(group)) if world > 1 else None
        if self._ctx is not None:           # the communicator is created with the context
            _lib.load().gb_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            if self._ctx is not None:
                _lib.load().gb_destroy(self._ctx)
        except Exception:
            pass

    def context(self, device):
        """The library context bound to ``device`` with the current weights uploaded."""
        lib = _lib.load()
        device = torch.device(device)
        if device.type != 'cuda':
            raise RuntimeError('genie2_b200.Denoiser runs on a CUDA device only (got %s); there is no CPU path' % device)
        idx = device.index if device.index is not None else torch.cuda.current_device()
        if self._ctx is None or self._ctx_device != idx:
            if self._ctx is not None:
                lib.gb_destroy(self._ctx)
                self._ctx = None
            d = _lib.gb_dims()
            for key in _DIM_KEYS:
                setattr(d, key, int(self.dims[key]) if key not in ('rescale', 'template_dist_min', 'template_dist_step')
                        else float(self.dims[key]))
            h = C.c_void_p()
            _lib.check(lib.gb_create(C.byref(h), C.byref(d), idx), None, 'gb_create')
            self._ctx, self._ctx_device = h, idx
            self._dirty, self._feat_key = True, None
            _lib.check(lib.gb_set_precision(h, _lib.PRECISIONS[self.precision]), h, 'gb_set_precision')
            if self._shard is not None:
                world, rank, uid = self._shard
                _lib.check(lib.gb_shard_init(h, world, rank, uid), h, 'gb_shard_init')
        if self._dirty:
            sd = self.state_dict()
            arr = (_lib.gb_tensor * len(sd))()
            keep = []
            for i, (name, t) in enumerate(sd.items()):
                t = t.detach().to(torch.float32).contiguous()
                keep.append(t)
                arr[i].name = name.encode()
                arr[i].data = t.data_ptr()
                arr[i].ndim = t.dim()
                for j, s in enumerate(t.shape):
                    arr[i].shape[j] = s
            if any(t.is_cuda for t in keep):
                torch.cuda.synchronize()
            _lib.check(lib.gb_load_weights(self._ctx, arr, len(sd)), self._ctx, 'gb_load_weights')
            self._dirty, self._feat_key = False, None
        return self._ctx

    def bind_features(self, features, device):
        """gb_set_features once per feature batch (identity-keyed, like the sampler loop reuses one dict)."""
        ctx = self.context(device)
        key = (id(features), features['residue_mask'].data_ptr(), tuple(features['residue_mask'].shape))
        if key == self._feat_key:
            return ctx
        B, N = features['residue_mask'].shape

        def prep(name, dtype, shape):
            t = features[name]
            if tuple(t.shape) != shape:
                raise ValueError('feature %s has shape %s, expected %s' % (name, tuple(t.shape), shape))
            return t.detach().to(dtype).contiguous()

        f = _lib.gb_features()
        f.B, f.N = B, N
        keep = [
            prep('residue_index', torch.int32, (B, N)), prep('chain_index', torch.int32, (B, N)),
            prep('residue_mask', torch.int32, (B, N)), prep('fixed_sequence_mask', torch.uint8, (B, N)),
            prep('fixed_structure_mask', torch.uint8, (B, N, N)), prep('interface_mask', torch.uint8, (B, N)),
            prep('aatype', torch.int32, (B, N, 20)), prep('atom_positions', torch.float32, (B, N, 3)),
        ]
        (f.residue_index, f.chain_index, f.residue_mask, f.fixed_sequence_mask, f.fixed_structure_mask,
         f.interface_mask, f.aatype, f.atom_positions) = [t.data_ptr() for t in keep]
        torch.cuda.synchronize()
        _lib.check(_lib.load().gb_set_features(ctx, C.byref(f)), ctx, 'gb_set_features')
        self._feat_key = key
        self._keep = features
        return ctx

    # ---- forward ----------------------------------------------------------------------
    def forward(self, ts, timesteps, features, quat_follow=None):
        """ts: frames with ``.rots`` [B,N,3,3] / ``.trans`` [B,N,3]; timesteps int [B];
        features: the batched feature dict.  Returns {'z': [B,N,3], 'ts': T, ...}."""
        trans, rots = ts.trans, ts.rots
        if not trans.is_cuda:
            raise RuntimeError('genie2_b200.Denoiser.forward: inputs must be CUDA tensors (no CPU path)')
        if trans.requires_grad or rots.requires_grad:
            raise RuntimeError('genie2_b200.Denoiser is forward-only; inputs must not require grad')
        dev = trans.device
        ctx = self.bind_features(features, dev)
        B, N = features['residue_mask'].shape
        if tuple(trans.shape) != (B, N, 3) or tuple(rots.shape) != (B, N, 3, 3):
            raise ValueError('frames do not match the feature batch (%d, %d)' % (B, N))
        trans_c = trans.detach().to(torch.float32).contiguous()
        rots_c = rots.detach().to(torch.float32).contiguous()
        t_c = timesteps.detach().to(device=dev, dtype=torch.int32).contiguous()
        c_s, c_p = self.dims['c_s'], self.dims['c_p']
        z = torch.empty(B, N, 3, device=dev, dtype=torch.float32)
        o_rots = torch.empty(B, N, 3, 3, device=dev, dtype=torch.float32)
        o_trans = torch.empty(B, N, 3, device=dev, dtype=torch.float32)
        aux = _lib.gb_aux()
        aux.out_rots, aux.out_trans = o_rots.data_ptr(), o_trans.data_ptr()
        out = {}
        if self.return_intermediates:
            n_states = 1 + self.dims['n_structure_layer'] * self.dims['n_structure_block']
            out['s'] = torch.empty(B, N, c_s, device=dev, dtype=torch.float32)
            out['p'] = torch.empty(B, N, N, c_p, device=dev, dtype=torch.float32)
            out['states'] = torch.empty(n_states, B, N, c_s, device=dev, dtype=torch.float32)
            aux.s, aux.p, aux.states = out['s'].data_ptr(), out['p'].data_ptr(), out['states'].data_ptr()
        if quat_follow is not None:
            quat_follow = quat_follow.detach().to(device=dev, dtype=torch.float32).contiguous()
            aux.quat_follow = quat_follow.data_ptr()
        stream = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(_lib.load().gb_denoise(ctx, rots_c.data_ptr(), trans_c.data_ptr(), t_c.data_ptr(), z.data_ptr(),
                                          C.byref(aux), stream), ctx, 'gb_denoise')
        out['z'] = z
        out['ts'] = T(o_rots, o_trans)
        return out

    def collective_count(self):
        return int(_lib.load().gb_collective_count(self._ctx)) if self._ctx is not None else 0

    def peer_barrier_count(self):
        return int(_lib.load().gb_peer_barrier_count(self._ctx)) if self._ctx is not None else 0

    def launch_count(self):
        return int(_lib.load().gb_launch_count(self._ctx)) if self._ctx is not None else 0

    def workspace_bytes(self):
        """Device bytes of the feature + workspace carve of the batch that is bound now."""
        return int(_lib.load().gb_workspace_bytes(self._ctx)) if self._ctx is not None else 0
