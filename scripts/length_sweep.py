"""ms per reverse step (denoiser eval + posterior + frames) over chain lengths at a fixed batch: the shape of BASELINE.json
configs[2] (length sweep L = 50..256, 5 samples per length) and configs[3] (scaffolding, batch 8, L ~ 75).
usage: length_sweep.py [batch=5] [steps=60]     (GB_STEP_GRAPH=0 to time the plain launch loop)"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import genie2_b200
from genie2_b200 import _lib
from oracle import genie_oracle as O, weights as W     # seeded weights / synthetic features only


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 5
    K = int(sys.argv[2]) if len(sys.argv) > 2 else 60
    dev = torch.device('cuda', 0)
    lib = _lib.load()
    genie = genie2_b200.Genie(genie2_b200.Config())
    genie.model.load_state_dict(W.make_state_dict(O.OracleConfig(), 0, 'activated'))
    genie = genie.to(dev)
    genie.setup_schedule()
    genie.model.set_precision('bf16')
    out = {'batch': B, 'steps': K, 'graph': os.environ.get('GB_STEP_GRAPH', '1'), 'ms_per_step': {}}
    for n in (50, 64, 75, 100, 128, 160, 192, 224, 256):
        feats = {k: v.to(dev) for k, v in O.make_features([[n]] * B).items()}
        ctx = genie.model.bind_features(feats, dev)
        genie.bind_schedule(ctx)
        x = torch.randn(B, n, 3, device=dev)
        rots = genie2_b200.compute_frenet_frames(genie.model, feats, x)
        st = torch.cuda.current_stream(dev).cuda_stream
        run = lambda t, k: _lib.check(lib.gb_sample_loop(ctx, x.data_ptr(), rots.data_ptr(), t, t - k + 1, 0.6, None, 5, st), ctx, 'loop')
        run(1000, 10)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run(990, K)
        e1.record()
        torch.cuda.synchronize()
        out['ms_per_step'][n] = e0.elapsed_time(e1) / K
        assert torch.isfinite(x).all()
    print(json.dumps(out))


if __name__ == '__main__':
    main()
