"""gb_k_ipa_static: the one entry point that allocates nothing (its scratch is a file-scope static __device__ array of the
library, include/genie2_b200.h).  No counterpart in the reference.

Checked through the C ABI against the existing entry point gb_k_ipa on the same inputs, on two calls in a row with
different inputs (the static array keeps the first call's contents: the second call must not see them), that the device's
free memory does not move across the calls, and that a batch beyond the stated maximum is rejected with a clear error.
Collected last on purpose.
"""
import pytest
import torch

from common import make_denoiser, feats_to, rel
from oracle import genie_oracle as O, weights as W

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'
TOL = 1e-4           # the tolerance of the fp32 parity tests (tests/test_gpu_parity.py); same kernels: expected bit-equal


def _inputs(cfg, feats, seed):
    B, N = feats['residue_mask'].shape
    g = torch.Generator().manual_seed(seed)
    s = torch.randn(B, N, cfg.c_s, generator=g)
    p = torch.randn(B, N, N, 128, generator=g)
    trans = torch.randn(B, N, 3, generator=g) * 4
    rots = O.compute_frenet_frames(trans, feats['chain_index'], feats['residue_mask'])
    return [t.to(DEV).contiguous() for t in (s, p, rots, trans)]


def test_ipa_static_matches_ipa_twice_and_allocates_nothing():
    import genie2_b200._lib as L
    lib = L.load()
    cfg = O.OracleConfig(n_pair_transform_layer=1, n_structure_layer=2)
    sd = W.make_state_dict(cfg, 17, 'activated')
    model = make_denoiser(cfg, sd)
    feats = O.make_features([[70], [61]])
    ctx = model.bind_features(feats_to(feats, DEV), torch.device(DEV))
    B, N = feats['residue_mask'].shape
    valid = feats['residue_mask'].bool()
    outs = [torch.empty(B, N, cfg.c_s, device=DEV) for _ in range(4)]
    ins = [_inputs(cfg, feats, 2), _inputs(cfg, feats, 3)]
    # the existing entry point first: it also makes the driver load the kernels (lazy module loading takes device memory
    # at a kernel's first launch, which is not an allocation of the entry point under test)
    for call, layer in ((0, 0), (1, 1)):
        s, p, rots, trans = ins[call]
        L.check(lib.gb_k_ipa(ctx, layer, s.data_ptr(), p.data_ptr(), rots.data_ptr(), trans.data_ptr(),
                             outs[2 + call].data_ptr(), None), ctx, 'gb_k_ipa')
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    for call, layer in ((0, 0), (1, 1)):                 # two calls in a row, different inputs and layers
        s, p, rots, trans = ins[call]
        L.check(lib.gb_k_ipa_static(ctx, layer, s.data_ptr(), p.data_ptr(), rots.data_ptr(), trans.data_ptr(),
                                    outs[call].data_ptr(), None), ctx, 'gb_k_ipa_static')
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(0)[0] == free0, 'gb_k_ipa_static changed the free device memory'
    for call in (0, 1):
        got, want = outs[call].cpu(), outs[2 + call].cpu()
        assert torch.isfinite(got[valid]).all()
        assert rel(got[valid], want[valid]) < TOL, call
        assert torch.equal(got[valid], want[valid]), 'same kernels, same order: expected bit-equal (call %d)' % call
    # and both agree with the oracle (the first call's, as in test_gpu_headline.test_k_ipa_entry_point_vs_oracle)
    s, p, rots, trans = [t.cpu() for t in ins[0]]
    with torch.no_grad():
        ref = O.invariant_point_attention(sd, cfg, 'structure_net.net.0.ipa', s, p, rots, trans, feats['residue_mask'])
    assert rel(outs[0].cpu()[valid], ref[valid]) < TOL


def test_ipa_static_rejects_batches_beyond_its_maximum():
    import genie2_b200._lib as L
    lib = L.load()
    cfg = O.OracleConfig(n_pair_transform_layer=1, n_structure_layer=1)
    sd = W.make_state_dict(cfg, 1, 'activated')
    model = make_denoiser(cfg, sd)
    feats = O.make_features([[300]] * 4)                 # 1200 rows > GB_K_IPA_STATIC_MAX_ROWS = 1024
    ctx = model.bind_features(feats_to(feats, DEV), torch.device(DEV))
    s, p, rots, trans = _inputs(cfg, feats, 5)
    out = torch.empty(4, 300, cfg.c_s, device=DEV)
    rc = lib.gb_k_ipa_static(ctx, 0, s.data_ptr(), p.data_ptr(), rots.data_ptr(), trans.data_ptr(), out.data_ptr(), None)
    assert rc == -2                                      # GB_ERR_UNSUPPORTED
    msg = lib.gb_last_error(ctx).decode()
    assert 'GB_K_IPA_STATIC_MAX_ROWS' in msg and '1200' in msg
    # the ordinary entry point takes the same batch
    L.check(lib.gb_k_ipa(ctx, 0, s.data_ptr(), p.data_ptr(), rots.data_ptr(), trans.data_ptr(), out.data_ptr(), None), ctx, 'gb_k_ipa')
    torch.cuda.synchronize()
    assert torch.isfinite(out).all()
