"""Shared helpers for the tests: golden-case loading and error metrics."""
import ast
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, 'tests', 'golden')

from oracle import genie_oracle as O, weights as W  # noqa: E402

sys.path.insert(0, GOLDEN)


def rel(a, b):
    a, b = torch.as_tensor(a).double(), torch.as_tensor(b).double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def load_forward_case(name):
    from make_golden import motif_spec
    d = np.load(os.path.join(GOLDEN, 'fwd_%s.npz' % name))
    over = ast.literal_eval(str(d['cfg_over']))
    lengths = ast.literal_eval(str(d['lengths']))
    seed = int(d['seed'])
    cfg = O.OracleConfig(**over)
    sd = W.make_state_dict(cfg, seed, str(d['kind']))
    assert abs(W.checksum(sd) - float(d['weights_checksum'])) <= 1e-6 * abs(float(d['weights_checksum'])), \
        'seeded weights do not reproduce on this box'
    motif = [motif_spec(sum(l), seed * 100 + i) for i, l in enumerate(lengths)] if bool(d['with_motif']) else None
    feats = O.make_features(lengths, motif)
    t = {k: torch.from_numpy(d[k]) for k in ('trans', 'rots', 'timesteps', 'quat', 'z', 's', 'p', 'states',
                                              'out_trans', 'out_rots')}
    return cfg, sd, feats, t


FORWARD_CASES = ['uncond_base_ragged', 'uncond_stock', 'motif_base', 'triatt', 'multichain']


# ------------------------------------------------------------------ CUDA-side helpers (gpu tests)
def make_denoiser(cfg, sd, precision='fp32', device='cuda:0'):
    """genie2_b200.Denoiser holding the oracle's seeded state dict."""
    import genie2_b200
    kw = cfg.model_kwargs()
    m = genie2_b200.Denoiser(**kw)
    m.load_state_dict(sd)
    m = m.to(device)
    m.set_precision(precision)
    m.return_intermediates = True
    return m


def feats_to(feats, device):
    return {k: v.to(device) for k, v in feats.items()}


def run_cuda_forward(model, rots, trans, timesteps, feats, quat_follow=None, device='cuda:0'):
    import genie2_b200
    f = feats_to(feats, device)
    out = model(genie2_b200.T(rots.to(device), trans.to(device)), timesteps.to(device), f,
                quat_follow=None if quat_follow is None else quat_follow.to(device))
    torch.cuda.synchronize()
    return {'z': out['z'].cpu(), 's': out['s'].cpu(), 'p': out['p'].cpu(), 'states': out['states'].cpu(),
            'trans': out['ts'].trans.cpu(), 'rots': out['ts'].rots.cpu()}, f


# ------------------------------------------------------------------ 1000-step reference trajectory (tests/golden/traj_c1k.npz)
def unpack_quat_signs(packed):
    """uint8 [..] -> float [.., 4]: +-1 at the index of the reference quaternion's largest |component| (make_golden.py
    pack_quat_signs).  Used as the 'follow' field: only the sign of its dot product with our quaternion is consumed."""
    packed = torch.as_tensor(packed).long()
    q = torch.zeros(*packed.shape, 4)
    q.scatter_(-1, (packed & 3).unsqueeze(-1), 1.0)
    return q * torch.where((packed & 4) != 0, -1.0, 1.0).unsqueeze(-1)


def centred_rmsd(a, b):
    a = a - a.mean(0, keepdim=True)
    b = b - b.mean(0, keepdim=True)
    return float(((a - b) ** 2).sum(-1).mean().sqrt())
