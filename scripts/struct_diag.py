"""bf16 tensor-core forward vs the fp32 CUDA path on the same inputs: per-tensor relative errors
(p, every entry of states, final frames, z).  python scripts/struct_diag.py [B N]..."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import torch
from common import make_denoiser, run_cuda_forward, rel
from oracle import genie_oracle as O, weights as W


def main():
    cases = [(2, 64), (1, 200), (4, 256), (3, 37)]
    cfg = O.OracleConfig()
    sd = W.make_state_dict(cfg, 5, 'activated')
    model = make_denoiser(cfg, sd)
    g = torch.Generator().manual_seed(9)
    for B, N in cases:
        lens = [[N]] + [[max(3, N - 7 - 5 * i)] for i in range(B - 1)]
        feats = O.make_features(lens)
        trans = torch.randn(B, N, 3, generator=g) * 6
        rots = O.compute_frenet_frames(trans, feats['chain_index'], feats['residue_mask'])
        ts = torch.full((B,), 400, dtype=torch.int32)
        model.set_precision('fp32')
        a, _ = run_cuda_forward(model, rots, trans, ts, feats, None)
        model.set_precision('bf16')
        b, _ = run_cuda_forward(model, rots, trans, ts, feats, None)
        st = ' '.join('%.1e' % rel(b['states'][k], a['states'][k]) for k in range(a['states'].shape[0]))
        print('B=%d N=%d  p %.2e | states %s | rots %.2e trans %.2e z %.2e | finite %s' %
              (B, N, rel(b['p'], a['p']), st, rel(b['rots'], a['rots']), rel(b['trans'], a['trans']), rel(b['z'], a['z']),
               bool(torch.isfinite(b['z']).all())), flush=True)


if __name__ == '__main__':
    main()
