"""Short driver for per-kernel timing under `ncu --metrics gpu__time_duration.sum`: the pair-stack
modules and whole forwards at the C2 shape (B=4, L=256); precision from GB_KBENCH_PREC (default bf16x3, the headline
mode).  Also prints CUDA-event times."""
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
    reps = 3
    lib = L.load()
    cfg = O.OracleConfig()
    sd = W.make_state_dict(cfg, 0, 'activated')
    model = make_denoiser(cfg, sd, precision=os.environ.get('GB_KBENCH_PREC', 'bf16x3'))
    model.return_intermediates = False
    feats = feats_to(O.make_features([[N]] * B), 'cuda:0')
    ctx = model.bind_features(feats, torch.device('cuda:0'))
    p = torch.randn(B, N, N, 128, device='cuda')
    x = torch.randn(B, N, 3, device='cuda') * 5
    rots = genie2_b200.compute_frenet_frames(model, feats, x)
    ts = genie2_b200.T(rots, x)
    tt = torch.full((B,), 500, dtype=torch.int32, device='cuda')

    def timed(name, fn):
        fn()
        torch.cuda.synchronize()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
        e[0].record()
        for i in range(reps):
            fn()
            e[i + 1].record()
        torch.cuda.synchronize()
        print('%-18s %8.1f us' % (name, 1e3 * min(e[i].elapsed_time(e[i + 1]) for i in range(reps))), flush=True)

    for st, nm in ((1, 'K4 proj'), (2, 'K5 contract'), (3, 'K6 out')):
        timed(nm, lambda st=st: L.check(lib.gb_k_trimul_stage(ctx, 0, 1, st, p.data_ptr(), None), ctx, 'k'))
    timed('trimul_out', lambda: L.check(lib.gb_k_trimul(ctx, 0, 1, p.data_ptr(), None), ctx, 'k'))
    timed('trimul_in', lambda: L.check(lib.gb_k_trimul(ctx, 0, 0, p.data_ptr(), None), ctx, 'k'))
    timed('pair_transition', lambda: L.check(lib.gb_k_pair_transition(ctx, 0, p.data_ptr(), None), ctx, 'k'))
    timed('forward', lambda: model(ts, tt, feats))


if __name__ == '__main__':
    main()
