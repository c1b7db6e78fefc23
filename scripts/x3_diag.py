"""Per-layer error of the structure-net states in every precision mode against the CPU oracle (L=64 B=2, base config),
all rows and valid residues only."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import torch
from common import make_denoiser, run_cuda_forward, rel
from oracle import genie_oracle as O, weights as W

cfg = O.OracleConfig()
sd = W.make_state_dict(cfg, 5, 'activated')
g = torch.Generator().manual_seed(9)
feats = O.make_features([[64], [50]])
B, N = feats['residue_mask'].shape
trans = torch.randn(B, N, 3, generator=g) * 6
rots = O.compute_frenet_frames(trans, feats['chain_index'], feats['residue_mask'])
ts = torch.tensor([321, 7], dtype=torch.int32)
m = feats['residue_mask'].bool()
with torch.no_grad():
    ref = O.denoiser_forward(sd, cfg, rots, trans, ts, feats, 'canonical')
for prec in sys.argv[1:] or ['fp32', 'bf16x3', 'bf16']:
    model = make_denoiser(cfg, sd, precision=prec)
    o, _ = run_cuda_forward(model, rots, trans, ts, feats, None)
    print(prec, 'p %.2e z %.2e (valid %.2e)' % (rel(o['p'], ref['p']), rel(o['z'], ref['z']), rel(o['z'][m], ref['z'][m])))
    print('   states all  ', ['%.1e' % rel(o['states'][i], ref['states'][i]) for i in range(9)])
    print('   states valid', ['%.1e' % rel(o['states'][i][m], ref['states'][i][m]) for i in range(9)])
