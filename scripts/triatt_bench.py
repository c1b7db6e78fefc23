"""Triangle attention (includeTriangularAttention True, off in the base configuration) at the C2 shape: per-module time of the
tensor-core kernels (bf16x3 and bf16 modes, csrc/tc_triatt.cu) and of the fp32 CUDA-core kernels through gb_k_tri_attention, and the
whole denoiser eval with the modules on.  Prints one JSON line.  usage: triatt_bench.py [B N]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import torch
from common import make_denoiser, feats_to
from oracle import genie_oracle as O, weights as W
import genie2_b200
import genie2_b200._lib as L


def main():
    B, N = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (4, 256)
    lib = L.load()
    cfg = O.OracleConfig(include_tri_att=True)
    sd = W.make_state_dict(cfg, 0, 'activated')
    out = {'workload': 'L=%d batch %d, base config + includeTriangularAttention True' % (N, B), 'modules_per_eval': 2 * cfg.n_pair_transform_layer}
    feats = feats_to(O.make_features([[N]] * B), 'cuda:0')
    p = torch.randn(B, N, N, 128, device='cuda')
    x = torch.randn(B, N, 3, device='cuda') * 5
    tt = torch.full((B,), 500, dtype=torch.int32, device='cuda')
    # algorithmic work of one module: projections 128 -> 512 + out 128 -> 128 (2 * 640 * 128 per pixel), Q K^T and P V (2 * 2 * N * 128 per pixel)
    M = B * N * N
    flops = M * (2.0 * 640 * 128 + 4.0 * N * 128)
    for prec in ('bf16x3', 'bf16', 'fp32'):
        model = make_denoiser(cfg, sd, precision=prec)
        model.return_intermediates = False
        ctx = model.bind_features(feats, torch.device('cuda:0'))
        rots = genie2_b200.compute_frenet_frames(model, feats, x)
        ts = genie2_b200.T(rots, x)

        def timed(fn, reps=5):
            fn()
            torch.cuda.synchronize()
            e = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
            e[0].record()
            for i in range(reps):
                fn()
                e[i + 1].record()
            torch.cuda.synchronize()
            return min(e[i].elapsed_time(e[i + 1]) for i in range(reps))

        r = {}
        for node, nm in ((1, 'starting'), (0, 'ending')):
            ms = timed(lambda node=node: L.check(lib.gb_k_tri_attention(ctx, 0, node, p.data_ptr(), None), ctx, 'k'))
            r[nm + '_node_ms'] = ms
            r[nm + '_node_tflops'] = flops / (ms * 1e-3) / 1e12
        r['forward_ms'] = timed(lambda: model(ts, tt, feats), 3)
        out[prec] = r
        del model
    print(json.dumps(out))


if __name__ == '__main__':
    main()
