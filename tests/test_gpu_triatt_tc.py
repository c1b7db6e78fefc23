"""K8 on tensor cores (GB_PREC_BF16 and GB_PREC_BF16X3, csrc/tc_triatt.cu): triangle attention around the starting and the ending node
(modules/triangular_attention.py:67-130, primitives.py:222-281) through gb_k_tri_attention, against the oracle and --
at the size the oracle's [B, I, H, J, J] logits no longer fit on the host -- against the fp32 CUDA-core kernels.
bf16 operands: stated separately from the fp32 gates, like every other bf16 module test (tests/test_gpu_tensorcore.py)."""
import pytest
import torch

from common import O, W, feats_to, make_denoiser, rel

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'
TOL_BF16 = 1.5e-2          # relative L2 of the module's update (q, k, v, the probabilities and o * g are rounded to bf16)
TOL_X3 = 2e-4              # bf16x3: split hi | lo operands, three passes per product (the north star's gate on z is 1e-3)


def _setup(B, N, prec):
    cfg = O.OracleConfig(include_tri_att=True, n_pair_transform_layer=1, n_structure_layer=1)
    sd = W.make_state_dict(cfg, 21, 'activated')
    model = make_denoiser(cfg, sd, precision=prec)
    lens = [[N]] + [[max(3, N - 6 - 2 * i)] for i in range(B - 1)]
    feats = feats_to(O.make_features(lens), DEV)
    ctx = model.bind_features(feats, torch.device(DEV))
    g = torch.Generator().manual_seed(N)
    p0 = torch.randn(B, N, N, 128, generator=g) * 1.2 + 0.2
    return cfg, sd, model, feats, ctx, p0


def _run(model, ctx, p0, starting):
    import genie2_b200._lib as L
    p = p0.to(DEV).contiguous()
    L.check(L.load().gb_k_tri_attention(ctx, 0, starting, p.data_ptr(), None), ctx, 'gb_k_tri_attention')
    torch.cuda.synchronize()
    return p.cpu() - p0


@pytest.mark.parametrize('prec', ['bf16', 'bf16x3'])
@pytest.mark.parametrize('starting', [1, 0])
@pytest.mark.parametrize('B,N', [(2, 21), (1, 70), (2, 130), (1, 280)])
def test_triangle_attention_tc_vs_oracle(B, N, starting, prec):
    cfg, sd, model, feats, ctx, p0 = _setup(B, N, prec)
    rm = feats['residue_mask'].cpu()
    pm = rm.unsqueeze(1) * rm.unsqueeze(2)
    pre = 'pair_transform_net.net.0.tri_att_%s' % ('start' if starting else 'end')
    with torch.no_grad():
        ref = O.triangle_attention(sd, cfg, pre, p0, pm, bool(starting))
    got = _run(model, ctx, p0, starting)
    assert torch.isfinite(got).all()
    valid = pm.bool()
    e_valid, e_all = rel(got[valid], ref[valid]), rel(got, ref)
    print('triatt tc %s B=%d N=%d starting=%d: update rel err valid %.2e, all pixels %.2e' % (prec, B, N, starting, e_valid, e_all))
    assert e_valid < (TOL_BF16 if prec == 'bf16' else TOL_X3)
    launches = model.launch_count()
    assert launches > 0


@pytest.mark.parametrize('prec', ['bf16', 'bf16x3'])
@pytest.mark.parametrize('starting', [1, 0])
def test_triangle_attention_tc_n512_vs_fp32_kernels(starting, prec):
    """N = 512: two (bf16) / four (bf16x3) key blocks per row (the running-maximum rescale of O), four query tiles."""
    cfg, sd, model, feats, ctx, p0 = _setup(1, 512, prec)
    got = _run(model, ctx, p0, starting)
    model32 = make_denoiser(cfg, sd, precision='fp32')
    ctx32 = model32.bind_features(feats, torch.device(DEV))
    want = _run(model32, ctx32, p0, starting)
    e = rel(got, want)
    print('triatt tc N=512 starting=%d: %s tensor-core vs fp32 CUDA-core kernels %.2e' % (starting, prec, e))
    assert torch.isfinite(got).all() and e < (TOL_BF16 if prec == 'bf16' else TOL_X3)
