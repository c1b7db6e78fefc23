"""Print per-tensor relative errors of the CUDA forward vs the golden fixtures (GPU box)."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', 'tests'))
import torch
from common import load_forward_case, make_denoiser, run_cuda_forward, rel
prec = sys.argv[1] if len(sys.argv) > 1 else 'fp32'
for name in ['uncond_base_ragged', 'uncond_stock', 'motif_base', 'multichain'] + (['triatt'] if '--triatt' in sys.argv else []):
    cfg, sd, feats, t = load_forward_case(name)
    model = make_denoiser(cfg, sd, prec)
    o, _ = run_cuda_forward(model, t['rots'], t['trans'], t['timesteps'], feats, t['quat'])
    ref = {'s': t['s'], 'p': t['p'], 'states': t['states'], 'trans': t['out_trans'], 'rots': t['out_rots'], 'z': t['z']}
    print(name, prec, ' '.join('%s=%.2e' % (k, rel(o[k], ref[k])) for k in ref),
          'states/layer', ' '.join('%.1e' % rel(o['states'][i], t['states'][i]) for i in range(t['states'].shape[0])), flush=True)
