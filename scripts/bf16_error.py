"""Relative L2 error of the bf16 tensor-core path against the fp32 CUDA path on the same seeded inputs (z and p).
usage: bf16_error.py [B N] ..."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import torch
from common import make_denoiser, feats_to, rel
from oracle import genie_oracle as O, weights as W
import genie2_b200


def main():
    args = [int(a) for a in sys.argv[1:]] or [1, 1024]
    cfg = O.OracleConfig()
    sd = W.make_state_dict(cfg, 2, 'activated')
    model = make_denoiser(cfg, sd, precision='fp32')
    for B, N in zip(args[0::2], args[1::2]):
        g = torch.Generator().manual_seed(4)
        feats = O.make_features([[N]] * B)
        trans = torch.randn(B, N, 3, generator=g) * 12
        rots = O.compute_frenet_frames(trans, feats['chain_index'], feats['residue_mask'])
        ts = torch.full((B,), 400, dtype=torch.int32)
        f = feats_to(feats, 'cuda:0')
        t = genie2_b200.T(rots.cuda(), trans.cuda())
        model.set_precision('fp32')
        o32 = model(t, ts.cuda(), f)
        z32, p32 = o32['z'].cpu(), o32['p'][:, ::4, ::4].cpu()
        model.set_precision('bf16')
        o16 = model(t, ts.cuda(), f)
        print('B=%d N=%d  PE_TC=%s PIPJ_TC=%s  z %.4e  p %.4e' % (B, N, os.environ.get('GB_PE_TC', '1'), os.environ.get('GB_PIPJ_TC', '1'),
                                                          rel(o16['z'].cpu(), z32), rel(o16['p'][:, ::4, ::4].cpu(), p32)), flush=True)


if __name__ == '__main__':
    main()
