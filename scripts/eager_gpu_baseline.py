"""Reference-style PyTorch eager on the B200 ("the reference on the B200 via PyTorch eager CUDA ... the number
to beat"): the oracle port (a line-by-line restatement of the reference's fp32 eager modules, chunked pair transition
included) run with its tensors on cuda:0.  The reference itself cannot travel to the GPU box (it lives in /root/reference
and needs pytorch_lightning).  Not part of bench.py; prints one JSON line.  usage: eager_gpu_baseline.py [B N steps]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from oracle import genie_oracle as O, weights as W


def main():
    B, N, K = (int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (4, 256, 5)
    dev = torch.device('cuda', 0)
    cfg = O.OracleConfig()
    sd = {k: v.to(dev) for k, v in W.make_state_dict(cfg, 0, 'activated').items()}
    feats = {k: v.to(dev) for k, v in O.make_features([[N]] * B).items()}
    sched = O.setup_schedule(1000)
    if isinstance(sched, dict):
        sched = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in sched.items()}
    torch.set_default_device(dev)          # tensors the oracle creates on the fly (encodings, index ranges) land on the GPU
    x = torch.randn(B, N, 3, device=dev)

    def step(x, t):
        r = O.compute_frenet_frames(x, feats['chain_index'], feats['residue_mask'])
        tt = torch.full((B,), t, dtype=torch.int32, device=dev)
        z = O.denoiser_forward(sd, cfg, r, x, tt, feats, 'canonical')['z']
        _, x2 = O.p_sample(sched, x, z, t, 0.6, torch.randn(B, N, 3, device=dev), feats)
        return x2

    with torch.no_grad():
        x = step(x, 1000)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(K):
            x = step(x, 999 - i)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / K
    print(json.dumps({'what': 'oracle port, PyTorch eager fp32 on cuda:0 (TF32 off)', 'B': B, 'N': N, 'steps': K, 'ms_per_step': dt * 1e3,
                      'evals_per_s': 1.0 / dt, 'finite': bool(torch.isfinite(x).all())}))


if __name__ == '__main__':
    main()
