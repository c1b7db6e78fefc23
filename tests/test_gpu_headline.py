"""Parity at the BENCHMARKED shape and of the code paths only the benchmark reaches.

VERDICT r01 "what's weak" 1-4: the bench runs L = 256, batch 4 (BASELINE configs[1]) but the largest
whole-forward oracle comparison was N = 64; `ipa_attention_tc_kernel<8>` / `<4>`, the persistent
multi-tile loop of K7 and `gb_sample_host` / `gb_k_ipa` were executed by no test.  Everything here goes
Python -> ctypes -> C ABI and is compared with the CPU oracle (oracle/genie_oracle.py) on the same
seeded inputs; an oracle eval at L = 256, B = 4 takes a few seconds on the box's host cores.

Reference: genie/model/model.py:125-192, modules/invariant_point_attention.py:181-259,
modules/pair_transition.py:48-87, sampler/base.py:216-289.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from common import make_denoiser, feats_to, run_cuda_forward, rel
from oracle import genie_oracle as O, weights as W

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'

# per-precision gates on one denoiser eval (relative L2 vs the fp32 oracle).  s never touches a tensor core.
#   fp32   : CUDA-core kernels; live oracle runs differ by summation order after eight structure layers
#   bf16x3 : split-bf16 tensor-core mode (three bf16 passes per product, ~2^-16): the north star's 1e-3 gate
#   bf16   : stated separately, as the north star asks
GATES = {
    'fp32': {'p': 1e-4, 'states': 1e-3, 'z': 1e-3, 'rots': 1e-3},
    'bf16x3': {'p': 2e-4, 'states': 1e-3, 'z': 1e-3, 'rots': 1e-3},
    'bf16': {'p': 2e-2, 'states': 3e-2, 'z': 3e-2, 'rots': 6e-2},      # measured at C2: p 5.8e-3, states 9.5e-3, z 2.2e-2, rots 4.0e-2
}

_C2 = {}


def _c2_case():
    """BASELINE configs[1]: L = 256, batch 4, full base configuration, activated weights; oracle run once."""
    if not _C2:
        cfg = O.OracleConfig()
        sd = W.make_state_dict(cfg, 11, 'activated')
        feats = O.make_features([[256]] * 4)
        g = torch.Generator().manual_seed(12)
        trans = torch.randn(4, 256, 3, generator=g) * 8
        rots = O.compute_frenet_frames(trans, feats['chain_index'], feats['residue_mask'])
        ts = torch.tensor([1000, 640, 77, 1], dtype=torch.int32)
        with torch.no_grad():
            ref = O.denoiser_forward(sd, cfg, rots, trans, ts, feats, 'canonical')
        _C2.update(cfg=cfg, sd=sd, feats=feats, trans=trans, rots=rots, ts=ts, ref=ref)
    return _C2


@pytest.mark.parametrize('precision', ['fp32', 'bf16x3', 'bf16'])
def test_forward_c2_shape_vs_oracle(precision):
    """Whole Denoiser.forward at the benchmarked shape against the oracle: s, p, states, frames, z."""
    c = _c2_case()
    model = make_denoiser(c['cfg'], c['sd'], precision=precision)
    o, _ = run_cuda_forward(model, c['rots'], c['trans'], c['ts'], c['feats'], None)
    ref, gate = c['ref'], GATES[precision]
    errs = {k: rel(o[k], ref[k]) for k in ('s', 'p', 'states', 'trans', 'rots', 'z')}
    per_layer = [rel(o['states'][i], ref['states'][i]) for i in range(ref['states'].shape[0])]
    print('C2 %s errors' % precision, {k: '%.2e' % v for k, v in errs.items()}, 'states per layer', ['%.1e' % v for v in per_layer])
    assert errs['s'] < 1e-4
    assert errs['p'] < gate['p']
    assert errs['states'] < gate['states'] and max(per_layer) < 2 * gate['states']
    assert errs['trans'] < gate['z'] and errs['rots'] < gate['rots']
    assert errs['z'] < gate['z']


@pytest.mark.parametrize('precision', ['bf16x3', 'bf16'])
@pytest.mark.parametrize('B,N,G', [(4, 256, 8), (2, 200, 4), (1, 96, 2)])
def test_ipa_attention_variants_states_per_layer(precision, B, N, G):
    """Reaches ipa_attention_tc_kernel<G> for every G the dispatcher can choose (tc_struct.cu ia_queries_per_cta:
    the smallest G in {2, 4, 8} whose grid still fits one wave of 148 SMs, capped by shared memory; bf16 at these
    shapes: 8 / 4 / 2, bf16x3 caps G at 4 because its probability images are twice as large) and holds the
    structure-net states of every layer to the oracle, not just z."""
    cfg = O.OracleConfig(n_pair_transform_layer=1, n_structure_layer=3)
    sd = W.make_state_dict(cfg, 13, 'activated')
    lens = [[N]] + [[N - 7 * (i + 1)] for i in range(B - 1)]
    feats = O.make_features(lens)
    g = torch.Generator().manual_seed(N)
    trans = torch.randn(B, N, 3, generator=g) * 7
    rots = O.compute_frenet_frames(trans, feats['chain_index'], feats['residue_mask'])
    ts = torch.randint(1, 1001, (B,), generator=g).int()
    with torch.no_grad():
        ref = O.denoiser_forward(sd, cfg, rots, trans, ts, feats, 'canonical')
    model = make_denoiser(cfg, sd, precision=precision)
    o, _ = run_cuda_forward(model, rots, trans, ts, feats, None)
    gate = GATES[precision]
    for i in range(ref['states'].shape[0]):
        assert rel(o['states'][i], ref['states'][i]) < gate['states'], (i, precision)
    assert rel(o['z'], ref['z']) < gate['z']


@pytest.mark.parametrize('precision', ['bf16x3', 'bf16'])
def test_pair_transition_persistent_loop_c2(precision):
    """K7 alone at B = 4, N = 256: 262,144 pixels = 1024 tile pairs on 148 persistent CTAs, so every CTA wraps its
    weight ring, X ring and TMEM slots ~7 times (the r01 test stopped at 67 pairs: no CTA ever looped)."""
    import genie2_b200._lib as L
    lib = L.load()
    B, N = 4, 256
    cfg = O.OracleConfig(n_pair_transform_layer=1, n_structure_layer=1)
    sd = W.make_state_dict(cfg, 3, 'activated')
    model = make_denoiser(cfg, sd, precision=precision)
    feats = feats_to(O.make_features([[256], [256], [250], [201]]), DEV)
    ctx = model.bind_features(feats, torch.device(DEV))
    g = torch.Generator().manual_seed(5)
    p0 = (torch.randn(B, N, N, 128, generator=g) * 1.5 + 0.3).to(DEV)
    mask = feats['residue_mask'].float()
    pm = (mask[:, :, None] * mask[:, None, :])[..., None]
    p = p0.clone()
    L.check(lib.gb_k_pair_transition(ctx, 0, p.data_ptr(), None), ctx, 'gb_k_pair_transition')
    torch.cuda.synchronize()
    sdd = {k: v.to(DEV) for k, v in sd.items()}
    torch.backends.cuda.matmul.allow_tf32 = False      # the torch restatement below must be a true fp32 product
    pre = 'pair_transform_net.net.0.pair_transition.'
    z = torch.nn.functional.layer_norm(p0, (128,), sdd[pre + 'layer_norm.weight'], sdd[pre + 'layer_norm.bias'])
    hid = torch.relu(z @ sdd[pre + 'linear_1.weight'].T + sdd[pre + 'linear_1.bias'])
    upd = (hid @ sdd[pre + 'linear_2.weight'].T + sdd[pre + 'linear_2.bias']) * pm
    err = rel(p - p0 * pm, upd * pm)
    print('K7 %s at C2: update error %.2e' % (precision, err))
    assert err < (1.2e-2 if precision == 'bf16' else 1e-4)
    masked = pm.expand_as(p) == 0
    assert float(p[masked].abs().max()) == 0.0
    # every tile pair was visited exactly once: no pixel of an unmasked row is left at its input value
    changed = ((p - p0).abs().amax(-1) > 0) | (pm[..., 0] == 0)
    assert bool(changed.all())


def test_k_ipa_entry_point_vs_oracle():
    """gb_k_ipa (fp32 CUDA-core IPA, exported since r01, untested until now) vs O.invariant_point_attention."""
    import genie2_b200._lib as L
    lib = L.load()
    cfg = O.OracleConfig(n_pair_transform_layer=1, n_structure_layer=2)
    sd = W.make_state_dict(cfg, 17, 'activated')
    model = make_denoiser(cfg, sd)
    feats = O.make_features([[70], [61]])
    fd = feats_to(feats, DEV)
    ctx = model.bind_features(fd, torch.device(DEV))
    B, N = feats['residue_mask'].shape
    g = torch.Generator().manual_seed(2)
    s = torch.randn(B, N, cfg.c_s, generator=g)
    p = torch.randn(B, N, N, 128, generator=g)
    trans = torch.randn(B, N, 3, generator=g) * 4
    rots = O.compute_frenet_frames(trans, feats['chain_index'], feats['residue_mask'])
    for layer in (0, 1):
        with torch.no_grad():
            ref = O.invariant_point_attention(sd, cfg, 'structure_net.net.%d.ipa' % layer, s, p, rots, trans, feats['residue_mask'])
        sd_, pd_, rd_, td_ = (t.to(DEV).contiguous() for t in (s, p, rots, trans))
        out = torch.empty(B, N, cfg.c_s, device=DEV)
        L.check(lib.gb_k_ipa(ctx, layer, sd_.data_ptr(), pd_.data_ptr(), rd_.data_ptr(), td_.data_ptr(), out.data_ptr(), None), ctx, 'gb_k_ipa')
        torch.cuda.synchronize()
        valid = feats['residue_mask'].bool()
        assert rel(out.cpu()[valid], ref[valid]) < 1e-4, layer


@pytest.mark.parametrize('precision', ['fp32', 'bf16'])
def test_sample_host_equals_sample_loop_bitwise(precision):
    """gb_sample_host (the call the e2e headline goes through: host buffers, H2D / D2H inside) must equal
    gb_sample_loop on device buffers bit for bit, for one call over the whole range, for one call per step (the
    bench pattern, which switches to graph replay from the second call on), and must reject bad ranges before
    touching the tape."""
    import genie2_b200
    import genie2_b200._lib as L
    lib = L.load()
    T_, B, N = 6, 2, 24
    cfg = O.OracleConfig(n_timestep=T_, n_pair_transform_layer=1, n_structure_layer=2)
    sd = W.make_state_dict(cfg, 8, 'activated')
    gcfg = genie2_b200.Config()
    gcfg.diffusion['n_timestep'] = T_
    gcfg.model['n_pair_transform_layer'] = 1
    gcfg.model['n_structure_layer'] = 2
    genie = genie2_b200.Genie(gcfg)
    genie.model.load_state_dict(sd)
    genie = genie.to(DEV)
    genie.model.set_precision(precision)
    genie.setup_schedule()
    feats = feats_to(O.make_features([[N]] * B), DEV)
    ctx = genie.model.bind_features(feats, torch.device(DEV))
    genie.bind_schedule(ctx)
    tape = torch.randn(T_ + 1, B, N, 3, generator=torch.Generator().manual_seed(4)).contiguous()
    # device path
    from genie2_b200.sampler import compute_frenet_frames
    trans = tape[T_].to(DEV).clone()
    rots = compute_frenet_frames(genie.model, feats, trans)
    tape_d = tape.to(DEV)
    L.check(lib.gb_sample_loop(ctx, trans.data_ptr(), rots.data_ptr(), T_, 1, 0.6, tape_d.data_ptr(), 0, None), ctx, 'gb_sample_loop')
    torch.cuda.synchronize()
    want = trans.cpu().numpy()
    # host path, one call
    x_in = tape[T_].clone().pin_memory()
    x_out = torch.empty(B, N, 3).pin_memory()
    L.check(lib.gb_sample_host(ctx, x_in.data_ptr(), x_out.data_ptr(), T_, 1, 0.6, tape.data_ptr(), 0), ctx, 'gb_sample_host')
    assert np.array_equal(x_out.numpy(), want)
    # host path, one call per step (x_out of a step is x_in of the next)
    cur = tape[T_].clone().pin_memory()
    for step in range(T_, 0, -1):
        L.check(lib.gb_sample_host(ctx, cur.data_ptr(), cur.data_ptr(), step, step, 0.6, tape.data_ptr(), 0), ctx, 'gb_sample_host')
    assert np.array_equal(cur.numpy(), want)
    # bad ranges are rejected up front (the tape copy used to be sized from an unvalidated range)
    for a, b in ((T_ + 1, 1), (3, 0), (2, 3), (3, -1)):
        assert lib.gb_sample_host(ctx, x_in.data_ptr(), x_out.data_ptr(), a, b, 0.6, tape.data_ptr(), 0) != 0
