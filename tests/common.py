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
