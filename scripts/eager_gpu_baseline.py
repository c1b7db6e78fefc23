"""The UNMODIFIED reference on the B200 through PyTorch eager ("the reference on the B200 via PyTorch eager
CUDA ... the number to beat"): baseline/_ref (oracle/install_reference.py) with its Genie module moved to cuda:0, timed
through its own UnconditionalSampler._sample with the schedule shortened to the timed step count -- 2 L B host
synchronisations per step in compute_frenet_frames (utils/geo_utils.py:63-83) and the 256-chunk launch storm of the pair
transition (modules/pair_transition.py:77-83) included.  Falls back to the oracle port (un-chunked, closed-form quaternion:
faster than the reference) when baseline/_ref is absent.  Not part of bench.py; prints one JSON line.
usage: eager_gpu_baseline.py [B N steps]"""
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from oracle import genie_oracle as O, weights as W, reference_bridge as R


def main():
    B, N, K = (int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (4, 256, 5)
    dev = torch.device('cuda', 0)
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    sd = W.make_state_dict(O.OracleConfig(), 0, 'activated')
    if R.installed_available():
        params = {'length': N, 'scale': 0.6, 'num_samples': B, 'outdir': tempfile.gettempdir(), 'prefix': 'ref', 'offset': 0}

        def run(n):
            genie, sampler = R.load_installed_sampler(n, sd, dev)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            with torch.no_grad():
                out = sampler._sample(params)
            torch.cuda.synchronize()
            return time.perf_counter() - t0, out

        note = ''
        try:
            run(2)
        except Exception as e:      # torch 2.11 + cuSOLVER reject the 262,144-matrix batched eigh of rot_to_quat on this box
            if 'cusolver' not in str(e).lower():
                raise
            try:
                torch.backends.cuda.preferred_linalg_library('magma')
                run(2)
                note = ' (torch.backends.cuda.preferred_linalg_library("magma"): cuSOLVER syevBatched rejects the batch)'
            except Exception as e2:
                print(json.dumps({'what': 'unmodified reference on cuda:0', 'unavailable': 'torch.linalg.eigh of the [B,N,N,4,4] batch fails on this box: %s / %s'
                                  % (str(e).split(chr(10))[0][:160], str(e2).split(chr(10))[0][:160])}))
                return port(B, N, K, dev, sd)
        dt, out = run(K)
        fin = all(bool(torch.isfinite(torch.as_tensor(o['atom_positions'])).all()) for o in out)
        print(json.dumps({'what': 'unmodified reference (baseline/_ref), PyTorch eager fp32 on cuda:0 (TF32 off), UnconditionalSampler._sample' + note,
                          'B': B, 'N': N, 'steps': K, 'ms_per_step': dt / K * 1e3, 'evals_per_s': K / dt, 'finite': fin}))
        return
    port(B, N, K, dev, sd)


def port(B, N, K, dev, sd):
    cfg = O.OracleConfig()
    sd = {k: v.to(dev) for k, v in sd.items()}
    feats = {k: v.to(dev) for k, v in O.make_features([[N]] * B).items()}
    sched = O.setup_schedule(1000)
    if isinstance(sched, dict):
        sched = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in sched.items()}
    torch.set_default_device(dev)
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
    print(json.dumps({'what': 'oracle port (no chunking, closed-form quaternion), PyTorch eager fp32 on cuda:0 (TF32 off)', 'B': B, 'N': N, 'steps': K,
                      'ms_per_step': dt * 1e3, 'evals_per_s': 1.0 / dt, 'finite': bool(torch.isfinite(x).all())}))


if __name__ == '__main__':
    main()
