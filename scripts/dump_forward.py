"""Run one bf16 forward at (B, N) with seeded inputs and save z (and p) to an .npz: used to compare two builds / two
environment switches bit for bit.  usage: dump_forward.py out.npz [B N]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import numpy as np
import torch
from common import make_denoiser, feats_to
from oracle import genie_oracle as O, weights as W
import genie2_b200


def main():
    out = sys.argv[1]
    B, N = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (4, 256)
    cfg = O.OracleConfig()
    sd = W.make_state_dict(cfg, 0, 'activated')
    model = make_denoiser(cfg, sd, precision='bf16')
    lens = [[N - 7 * i] for i in range(B)]          # ragged: padding masks differ per sample
    feats = feats_to(O.make_features(lens), 'cuda:0')
    model.bind_features(feats, torch.device('cuda:0'))
    g = torch.Generator().manual_seed(3)
    x = (torch.randn(B, N, 3, generator=g) * 5).cuda()
    rots = genie2_b200.compute_frenet_frames(model, feats, x)
    ts = genie2_b200.T(rots, x)
    tt = torch.full((B,), 500, dtype=torch.int32, device='cuda')
    o = model(ts, tt, feats)
    torch.cuda.synchronize()
    np.savez(out, z=o['z'].cpu().numpy(), p=o['p'].cpu().numpy()[:, ::8, ::8])
    print('saved', out, float(o['z'].abs().mean()))


if __name__ == '__main__':
    main()
