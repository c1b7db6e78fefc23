"""Benchmark of the reverse-diffusion sampling hot path (contract: see the task brief).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--precision fp32|bf16x3|bf16] [--impl reference]

A "step" is one reverse-diffusion step over one batch: denoiser eval (Denoiser.forward) + DDPM
posterior update + Frenet frames of x_{t-1}, i.e. one iteration of the loop at
genie/sampler/base.py:239-282.  Workload at every N: BASELINE.json configs[1] -- unconditional,
L=256, batch 4, scale 0.6, base configuration, seeded "activated" random-init weights, synthetic
x_T ~ N(0,1).  With N > 1 every rank runs its own independent batch (the reference's
process-per-GPU farm, utils/multiprocessor.py:59-100; no collective on the data path): weak scaling.

value  = denoiser evals/s summed over ranks, state resident in HBM (gb_sample_loop).
e2e    = the same through the host-buffer entry point gb_sample_host: every step uploads x_t
         [B,N,3] from pinned host memory and downloads x_{t-1}.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

L, B, SCALE, T_STEPS = 256, 4, 0.6, 1000
METRIC = 'denoiser evals/sec at L=256 batch 4'
UNIT = 'evals/s'


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=None)
    ap.add_argument('--warmup', type=int, default=None)
    ap.add_argument('--precision', default=os.environ.get('GB_BENCH_PRECISION', 'auto'))
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--single-mode', action='store_true', help='skip the block of the other tensor-core mode')
    ap.add_argument('--no-sample-1000', action='store_true', help='skip the measured whole 1000-step sample (about 4 s)')
    ap.add_argument('--no-extra-blocks', action='store_true', help='skip the length_sweep (configs[2]) and c5 (configs[4]) blocks')
    ap.add_argument('--sweep-steps', type=int, default=10, help='reverse steps per batch in the length_sweep block')
    ap.add_argument('--workload', default='c2', choices=['c2', 'c5'],
                    help="c2 (default, the contract's line): L=256 batch 4; c5: ONE L=1024 sample, row-sharded over the ranks")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        d = json.load(open(p))
        return {'hbm': d['hbm_gbs'], 'tensor_burst': d['bf16_tflops'],
                'tensor_sustained': d.get('bf16_tflops_sustained', d['bf16_tflops']), 'src': 'measured'}
    return {'hbm': 6650.0, 'tensor_burst': 1590.0, 'tensor_sustained': 1400.0, 'src': 'fallback'}


def flops_per_eval(n, b):
    """algorithmic FLOPs of one denoiser eval (base config)."""
    return b * (2560.0 * n ** 3 + 3378304.0 * n ** 2 + 28022016.0 * n)


# ------------------------------------------------------------------------------- CPU oracle legs
def oracle_stepper(n_threads=None):
    """Time the CPU oracle (port of the reference's fp32 eager path) on ONE sample of the
    workload: a full denoiser eval + posterior step at L=256, batch 1."""
    import torch
    from oracle import genie_oracle as O, weights as W
    # all the host cores this process may use: torchrun exports OMP_NUM_THREADS=1 to its workers, which would time the
    # reference on a single core
    if not n_threads:
        n_threads = host_threads(32)       # beyond 32 threads the [1, 256, 256, 128]-sized eager ops stop scaling
    torch.set_num_threads(n_threads)
    cfg = O.OracleConfig()
    sd = W.make_state_dict(cfg, 0, 'activated')
    feats = O.make_features([[L]])
    g = torch.Generator().manual_seed(1)
    x = torch.randn(1, L, 3, generator=g)
    sched = O.setup_schedule(T_STEPS)

    def step(x, t):
        r = O.compute_frenet_frames(x, feats['chain_index'], feats['residue_mask'])
        tt = torch.full((1,), t, dtype=torch.int32)
        z = O.denoiser_forward(sd, cfg, r, x, tt, feats, 'canonical')['z']
        _, x2 = O.p_sample(sched, x, z, t, SCALE, torch.randn(1, L, 3, generator=g), feats)
        return x2

    return step, x, torch.get_num_threads()


def host_threads(limit=64):
    """All the host cores this process may use (torchrun exports OMP_NUM_THREADS=1 to its workers, which would time the
    reference on a single core)."""
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    return min(n, limit)


def reference_sample_seconds(n_steps, threads, batch=None):
    """The UNMODIFIED reference (baseline/_ref, installed by oracle/install_reference.py) through its own public sampler,
    UnconditionalSampler._sample (sampler/base.py:216-289): the whole workload -- L=256, batch 4, base configuration,
    scale 0.6 -- with the schedule shortened to `n_steps` reverse steps (the work of a step does not depend on the
    schedule length).  Chunked pair transition and the eigh quaternions intact.  Returns the seconds _sample took."""
    import torch
    from oracle import reference_bridge as R, genie_oracle as O, weights as W
    torch.set_num_threads(threads)
    sd = W.make_state_dict(O.OracleConfig(), 0, 'activated')
    genie, sampler = R.load_installed_sampler(n_steps, sd)
    params = {'length': L, 'scale': SCALE, 'num_samples': batch or B, 'outdir': tempfile.gettempdir(), 'prefix': 'ref', 'offset': 0}
    torch.manual_seed(1)
    t0 = time.perf_counter()
    out = sampler._sample(params)
    dt = time.perf_counter() - t0
    if not all(np.isfinite(o['atom_positions']).all() for o in out):
        raise RuntimeError('reference arm produced non-finite coordinates')
    return dt


def reference_available():
    from oracle import reference_bridge as R
    return R.installed_available()


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on the host cores.  With baseline/_ref
    present (kind "reference"): the unmodified reference's sampler on the full batch-4 workload, W warm-up steps in one
    _sample call and K timed steps in a second one.  Without it (kind "port"): the oracle port on ONE of the four
    samples per step, evals/s quoted for batch 4."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    import torch
    K = args.steps if args.steps is not None else 3
    W_ = args.warmup if args.warmup is not None else 1
    threads = host_threads()
    if reference_available():
        # the warm-up call also sizes the timed one: the whole batch if K steps of it fit in ~4 minutes on this box's cores,
        # otherwise one of the four samples per step (rate quoted for batch 4 = 1 / (4 t))
        t_w = reference_sample_seconds(max(W_, 1), threads) / max(W_, 1)
        nb = B if t_w * K <= 240.0 else 1
        dt = reference_sample_seconds(K, threads, nb)
        per_step = dt / K * (B / nb)
        kind = 'reference'
        sample = ('%d reverse steps of %s (L=256, batch %d) through the unmodified reference UnconditionalSampler._sample '
                  'from baseline/_ref, torch CPU fp32, %d threads' % (K, 'the whole workload' if nb == B else 'one sample of four per step; batch-4 rate = 1/(4 t)', nb, threads))
        cfg_sample = ('the whole batch-4 workload' if nb == B else 'one of the 4 samples per step') + ', schedule shortened to the timed step count'
    else:
        step, x, threads = oracle_stepper(threads)
        with torch.no_grad():
            t = T_STEPS
            for _ in range(W_):
                x = step(x, t)
                t -= 1
            t0 = time.perf_counter()
            for _ in range(K):
                x = step(x, t)
                t -= 1
            dt = time.perf_counter() - t0
        per_step = dt / K * B
        kind = 'port'
        sample = '%d reverse steps at L=256 batch 1 (oracle/genie_oracle.py, torch CPU fp32); batch-4 rate = 1/(4 t)' % K
        cfg_sample = 'one reverse step of 1 of the 4 samples per step; batch-4 rate = 1/(4 t)'
    value = 1.0 / per_step
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': K, 'warmup': W_,
        'ms_per_step': per_step * 1e3, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': 'unconditional L=256 batch 4, base config, 1000-step schedule, scale 0.6 (BASELINE configs[1])',
                   'per_gpu_batch': B, 'length': L, 'sample': cfg_sample},
        'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': threads, 'kind': kind, 'sample': sample},
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'seconds_per_1000_step_sample': 1000.0 * per_step,
        'gpu_launches': 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 25 ms during the warm-up and timed steps."""
    Q = ('timestamp,clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile('w+', suffix='.csv', delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(['nvidia-smi', '-i', str(index), '--query-gpu=' + self.Q,
                                       '--format=csv,noheader,nounits', '-lms', '25'], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            pass

    def stop(self, t0=None, t1=None):
        """Median SM clock and throttle reasons of the samples whose nvidia-smi timestamp lies inside [t0, t1]
        (time.time() values bracketing the timed region); all samples if none does (said in 'window')."""
        if self.p is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        rows = [r.strip().split(', ') for r in open(self.f.name) if r.strip()]
        os.unlink(self.f.name)
        import datetime
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        parsed = []
        for r in rows:
            if len(r) < 7:
                continue
            try:
                ts = datetime.datetime.strptime(r[0].strip(), '%Y/%m/%d %H:%M:%S.%f').timestamp()
                parsed.append((ts, float(r[1]), float(r[2]), [n for n, v in zip(names, r[3:7]) if v.strip().lower().startswith('active')]))
            except ValueError:
                continue
        inside = [q for q in parsed if t0 is not None and t1 is not None and t0 - 0.02 <= q[0] <= t1 + 0.02]
        use, window = (inside, 'timed region') if inside else (parsed, 'whole run (no sample fell inside the timed region)')
        sm = sorted(q[1] for q in use)
        reasons = sorted({n for q in use for n in q[3]})
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': use[-1][2] if use else None, 'reasons': reasons,
                'samples': len(sm), 'window': window}


# ------------------------------------------------------------------------------- B200 arm
HEADLINE_PRECISION = 'bf16x3'     # the tensor-core mode that meets the 1e-3 / 0.1 A parity gates (tests/test_gpu_x3.py)


def pick_precision(genie, requested):
    """'auto' = bf16x3: tcgen05 tensor cores with split-bf16 operands (three passes per product, fp32 accumulation), the
    mode held to the reference's fp32 results at 1e-3.  No fall-back: a mode that cannot run is an error."""
    prec = HEADLINE_PRECISION if requested == 'auto' else requested
    genie.model.set_precision(prec)
    probe_forward(genie)
    return prec


_PROBE = {}


def probe_forward(genie):
    import torch
    import genie2_b200
    from oracle import genie_oracle as O   # feature construction only (host side)
    dev = genie.device
    if 'f' not in _PROBE:
        f = O.make_features([[32]])
        _PROBE['f'] = {k: v.to(dev) for k, v in f.items()}
    x = torch.randn(1, 32, 3, device=dev)
    ts = genie2_b200.T(genie2_b200.compute_frenet_frames(genie.model, _PROBE['f'], x), x)
    genie.model(ts, torch.full((1,), 5, dtype=torch.int32, device=dev), _PROBE['f'])
    torch.cuda.synchronize()


def _time_launches(fn, reps):
    import torch
    for _ in range(3):
        fn(0)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    ev[0].record()
    for i in range(reps):
        fn(i)
        ev[i + 1].record()
    torch.cuda.synchronize()
    return sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(reps))[reps // 2]


def kernel_table(genie, ctx, lib, prec, pk, reps=10):
    """CUDA-event times, on the launching stream, of the four pair-stack kernels (95 % of the step) run alone through the
    per-kernel entry points, with the algorithmic bytes / flops of DESIGN.md section 4 (S = B*N^2*128 elements of the pair
    tensor).  Every kernel's inputs + outputs exceed the 126 MB L2, so each timed launch streams from HBM.  Returns the
    roofline object of the kernel with the largest share of the eval and the table of the others."""
    import torch
    from genie2_b200 import _lib
    dev = genie.device
    M = B * L * L
    S = M * 128
    p = torch.randn(B, L, L, 128, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    if prec == 'fp32':
        ms = _time_launches(lambda i: _lib.check(lib.gb_k_pair_transition(ctx, i % 5, p.data_ptr(), stream), ctx, 'k7'), reps)
        achieved = M * 262144.0 / (ms * 1e-3) / 1e12
        return {'kernel': 'pair_transition_simt_kernel (fp32 CUDA cores)', 'bound': 'tensor', 'achieved': achieved, 'peak': pk['tensor_burst'],
                'unit': 'TFLOP/s', 'frac': achieved / pk['tensor_burst'], 'traffic': None, 'ms_per_launch': ms,
                'peak_source': pk['src'] + ' bf16 burst'}
    _lib.check(lib.gb_k_trimul(ctx, 0, 1, p.data_ptr(), stream), ctx, 'gb_k_trimul')
    stage = lambda st: (lambda i: _lib.check(lib.gb_k_trimul_stage(ctx, i % 5, i & 1, st, p.data_ptr(), stream), ctx, 'stage'))
    ms4, ms5, ms6 = _time_launches(stage(1), reps), _time_launches(stage(2), reps), _time_launches(stage(3), reps)
    ms7 = _time_launches(lambda i: _lib.check(lib.gb_k_pair_transition(ctx, i % 5, p.data_ptr(), stream), ctx, 'k7'), reps)
    x3 = prec == 'bf16x3'
    passes = 3.0 if x3 else 1.0
    # name, ms, launches per eval, algorithmic bytes, flops (tensor-core passes included), bound
    rows = [
        ('trimul_proj_x3_kernel x2 (K4: LayerNorm + gated projections; a + g launch, b launch)' if x3 else 'trimul_proj_tc_kernel (K4: LayerNorm + gated projections)',
         ms4, 10, (16.0 if x3 else 10.0) * S, passes * 2.0 * 128 * 640 * M, 'hbm'),
        ('trimul_contract_pers_tc_kernel (K5)', ms5, 10, (12.0 if x3 else 6.0) * S, passes * 2.0 * B * 128 * L ** 3, 'hbm'),
        ('trimul_out_x3_kernel (K6)' if x3 else 'trimul_out_tc_kernel (K6)', ms6, 10, (16.0 if x3 else 12.0) * S, passes * 2.0 * 128 * 128 * M, 'hbm'),
        ('pair_transition_x3_kernel (K7)' if x3 else 'pair_transition_tc_kernel (K7)', ms7, 5, 8.0 * S, passes * M * 262144.0, 'tensor'),
    ]
    table = {}
    for name, ms, cnt, nbytes, flops, bound in rows:
        table[name.split(' ')[0]] = {'ms': ms, 'launches_per_eval': cnt, 'hbm_gbs': nbytes / (ms * 1e-3) / 1e9, 'tflops': flops / (ms * 1e-3) / 1e12,
                                     'frac_hbm': nbytes / (ms * 1e-3) / 1e9 / pk['hbm'], 'frac_tensor': flops / (ms * 1e-3) / 1e12 / pk['tensor_burst']}
    name, ms, cnt, nbytes, flops, bound = max(rows, key=lambda r: r[1] * r[2])
    traffic, tsrc = None, None
    for prof in ('r02_ncu_summary.json',):
        path = os.path.join(ROOT, 'profiles', prof)
        if os.path.exists(path):
            d = json.load(open(path)).get(prec, {}).get(name.split(' ')[0])
            if d:
                traffic, tsrc = d.get('dram_bytes_per_launch'), 'stored constant from profiles/%s (one ncu --set full capture of this kernel, not re-measured by this run)' % prof
    if bound == 'hbm':
        achieved, peak, unit = nbytes / (ms * 1e-3) / 1e9, pk['hbm'], 'GB/s'
    else:
        achieved, peak, unit = flops / (ms * 1e-3) / 1e12, pk['tensor_burst'], 'TFLOP/s'
    return {'kernel': name, 'bound': bound, 'achieved': achieved, 'peak': peak, 'unit': unit, 'frac': achieved / peak, 'traffic': traffic,
            'traffic_source': tsrc, 'ms_per_launch': ms, 'launches_per_eval': cnt, 'algorithmic_bytes_per_launch': nbytes,
            'flops_per_launch': flops, 'peak_source': pk['src'] + (' HBM copy' if bound == 'hbm' else ' bf16 burst (kernel timed alone)'),
            'kernels': table}


def genie_with(genie, mode):
    genie.model.set_precision(mode)
    return genie


def run_b200(args):
    import torch
    import torch.distributed as dist
    import genie2_b200
    from genie2_b200 import _lib
    from oracle import genie_oracle as O, weights as W     # seeded weights / synthetic features (host) + cpu_baseline leg

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise RuntimeError('bench.py needs a CUDA device: the hot path has no CPU fallback')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    K = args.steps if args.steps is not None else 50
    W_ = args.warmup if args.warmup is not None else 5
    W_ = max(W_, 3)
    if K + W_ > T_STEPS:
        raise SystemExit('steps + warmup must be <= %d' % T_STEPS)

    clocks = ClockSampler(local)      # started now: nvidia-smi delivers its first line after a few hundred milliseconds
    lib = _lib.load()
    cfg = O.OracleConfig()
    sd = W.make_state_dict(cfg, 0, 'activated')
    genie = genie2_b200.Genie(genie2_b200.Config())
    genie.model.load_state_dict(sd)
    genie = genie.to(dev)
    genie.setup_schedule()
    prec = pick_precision(genie, args.precision)
    feats = {k: v.to(dev) for k, v in O.make_features([[L]] * B).items()}
    ctx = genie.model.bind_features(feats, dev)
    genie.bind_schedule(ctx)
    g = torch.Generator().manual_seed(1 + rank)
    x_host0 = torch.randn(B, L, 3, generator=g).pin_memory()
    stream = torch.cuda.current_stream(dev).cuda_stream

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def measure(mode):
        """device-resident and host-buffer throughput of one precision mode; max over ranks"""
        genie.model.set_precision(mode)
        x = x_host0.to(dev)
        rots = genie2_b200.compute_frenet_frames(genie.model, feats, x)

        def loop(t_start, n):
            _lib.check(lib.gb_sample_loop(ctx, x.data_ptr(), rots.data_ptr(), t_start, t_start - n + 1, SCALE, None, 1234 + rank, stream),
                       ctx, 'gb_sample_loop')
        # ---- device-resident throughput
        loop(T_STEPS, W_)
        barrier()
        wall0 = time.time()
        l0 = genie.model.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        loop(T_STEPS - W_, K)
        e1.record()
        barrier()
        wall1 = time.time()
        launches = genie.model.launch_count() - l0
        ms = e0.elapsed_time(e1)
        finite = bool(torch.isfinite(x).all())
        # ---- end to end through host buffers: one gb_sample_host call per step
        x_host = x_host0.clone().pin_memory()
        out_host = torch.empty_like(x_host).pin_memory()
        cur, nxt = x_host, out_host
        t = T_STEPS
        for _ in range(3):
            _lib.check(lib.gb_sample_host(ctx, cur.data_ptr(), nxt.data_ptr(), t, t, SCALE, None, 99 + rank), ctx, 'gb_sample_host')
            cur, nxt, t = nxt, cur, t - 1
        barrier()
        t0 = time.perf_counter()
        for _ in range(K):
            _lib.check(lib.gb_sample_host(ctx, cur.data_ptr(), nxt.data_ptr(), t, t, SCALE, None, 99 + rank), ctx, 'gb_sample_host')
            cur, nxt, t = nxt, cur, t - 1
        barrier()
        e2e_s = time.perf_counter() - t0
        times = torch.tensor([ms, e2e_s * 1e3], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(times, op=dist.ReduceOp.MAX)
        return {'ms': float(times[0]), 'e2e_ms': float(times[1]), 'launches': int(launches), 'finite': finite, 'wall': (wall0, wall1)}

    m = measure(prec)
    clk = clocks.stop(*m['wall'])
    other = 'bf16' if prec != 'bf16' else 'bf16x3'
    m2 = measure(other) if not args.single_mode else None
    genie.model.set_precision(prec)
    # ---- blocks of the other multi-GPU configurations, measured after the contract's timed region; every rank takes part
    sweep = c5 = scaf = None
    if not args.no_extra_blocks:
        try:
            sweep = measure_length_sweep(genie, dev, world, rank, barrier, args.sweep_steps, prec)
        except Exception as e:      # never lose the contract line to an extra block
            sweep = {'error': '%s: %s' % (type(e).__name__, e)}
        try:
            scaf = measure_scaffold(genie, dev, rank, barrier, 50, prec)
        except Exception as e:
            scaf = {'error': '%s: %s' % (type(e).__name__, e)}
        genie.model.set_precision(prec)
        ctx = genie.model.bind_features(feats, dev)      # the sweep re-bound the context to its own batches
        genie.bind_schedule(ctx)
        try:
            c5 = measure_c5(5, 3, world, rank, local, barrier)
        except Exception as e:
            c5 = {'error': '%s: %s' % (type(e).__name__, e)}
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    pk = peaks()
    ms, e2e_ms, launches, finite = m['ms'], m['e2e_ms'], m['launches'], m['finite']
    value = world * K / (ms * 1e-3)
    e2e_value = world * K / (e2e_ms * 1e-3)
    roof = kernel_table(genie, ctx, lib, prec, pk)
    fl = flops_per_eval(L, B)
    DT = {'fp32': 'f32 (CUDA cores)', 'bf16x3': 'bf16x3 (tcgen05 bf16 tensor cores, split hi+lo operands, three passes per product, fp32 accumulate and storage: fp32-grade)',
          'tf32': 'bf16x3', 'bf16': 'bf16'}
    GATE = {'fp32': 'z <= 1e-4 vs the reference fp32 path (tests/test_gpu_parity.py)',
            'bf16x3': 'z <= 1e-3 vs the reference fp32 path, 50-step trajectory 0.015 A (tests/test_gpu_x3.py, test_gpu_headline.py)',
            'bf16': 'z <= 3e-2, stated separately (tests/test_gpu_tensorcore.py)'}
    line = {
        'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': K, 'warmup': W_, 'ms_per_step': ms / K,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': DT[prec], 'data': 'synthetic',
        'config': {'workload': 'unconditional L=256 batch 4, base config, 1000-step schedule, scale 0.6 (BASELINE configs[1])',
                   'per_gpu_batch': B, 'length': L, 'precision': prec, 'parity': GATE[prec], 'weights': 'seeded activated random init',
                   'l2': 'no explicit flush: per-step working set (p,a,b,x,g >= 5 x 134 MB) exceeds the 126 MB L2',
                   'parallelism': 'independent batches per GPU, no collective' if world > 1 else 'single GPU'},
        'seconds_per_1000_step_sample': 1000.0 * ms / K * 1e-3,
        'e2e': {'value': e2e_value, 'unit': UNIT, 'h2d_bytes_per_step': B * L * 3 * 4, 'd2h_bytes_per_step': B * L * 3 * 4,
                'api': 'gb_sample_host (one call per step, pinned host buffers, synchronous)'},
        'gpu_launches': int(launches),
        'clocks': clk,
        'roofline': roof,
        'whole_eval': {'algorithmic_tflop_per_eval': fl / 1e12, 'achieved_tflops': fl / (ms / K * 1e-3) / 1e12,
                       'frac_of_tensor_sustained': fl / (ms / K * 1e-3) / 1e12 / pk['tensor_sustained']},
        'finite': finite,
    }
    if sweep is not None:
        line['length_sweep'] = sweep
    if scaf is not None:
        line['scaffold'] = scaf
    if c5 is not None:
        line['c5'] = c5
    if m2 is not None:
        # the other tensor-core mode on the same workload, same K / W, timed right after the headline mode
        r2 = kernel_table(genie_with(genie, other), ctx, lib, other, pk)
        genie.model.set_precision(prec)
        line['%s_mode' % other] = {
            'value': world * K / (m2['ms'] * 1e-3), 'unit': UNIT, 'ms_per_step': m2['ms'] / K, 'dtype': DT[other], 'parity': GATE[other],
            'e2e': {'value': world * K / (m2['e2e_ms'] * 1e-3), 'unit': UNIT}, 'gpu_launches': m2['launches'], 'finite': m2['finite'],
            'roofline': {k: r2[k] for k in ('kernel', 'bound', 'achieved', 'peak', 'unit', 'frac', 'ms_per_launch')}, 'kernels': r2['kernels']}
    # ---- metric (ii) of SURVEY 8(d), measured rather than extrapolated: wall time of ONE whole 1000-step batch through the
    # public sampler API (BaseSampler._sample, base.py:216-289): feature build + upload, 1000 x (denoise + posterior +
    # frames), download of the final coordinates.  Rank 0 only, after the timed region above.
    if not args.no_sample_1000:
        sampler = genie2_b200.UnconditionalSampler(genie)
        params = {'length': L, 'scale': SCALE, 'num_samples': B, 'outdir': tempfile.gettempdir(), 'prefix': 'bench', 'offset': 0,
                  'noise': 'philox', 'seed': 7}
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        outs = sampler._sample(params)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        line['sample_1000'] = {'seconds': dt, 'steps': T_STEPS, 'batch': B, 'length': L,
                               'api': 'UnconditionalSampler._sample (features + 1000 reverse steps + download), in-kernel Philox noise',
                               'finite': bool(all(np.isfinite(o['atom_positions']).all() for o in outs))}
        line['seconds_per_1000_step_sample'] = dt
    if not args.no_cpu_baseline and world == 1:
        if reference_available():
            threads, n = host_threads(), 3
            reference_sample_seconds(1, threads)
            dt = reference_sample_seconds(n, threads) / n
            line['cpu_baseline'] = {'value': 1.0 / dt, 'unit': UNIT, 'cores': threads, 'kind': 'reference',
                                    'sample': '%d reverse steps of the whole workload (L=256, batch 4) through the unmodified reference sampler '
                                              'in baseline/_ref (torch CPU fp32, chunked pair transition, eigh quaternions)' % n}
        else:
            step, xc, threads = oracle_stepper()
            with torch.no_grad():
                xc = step(xc, T_STEPS)
                t0 = time.perf_counter()
                n = 2
                for i in range(n):
                    xc = step(xc, T_STEPS - 1 - i)
                dt = (time.perf_counter() - t0) / n
            line['cpu_baseline'] = {'value': 1.0 / (B * dt), 'unit': UNIT, 'cores': threads, 'kind': 'port',
                                    'sample': '%d reverse steps of one L=256 sample (oracle port, torch CPU fp32); batch-4 rate = 1/(4 t)' % n}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def measure_c5(K, W_, world, rank, local, barrier):
    """BASELINE configs[4]: ONE unconditional L=1024 sample, bf16 mode.  With world > 1 the pair representation is sharded
    by rows over the ranks (csrc/shard.cu: NCCL all-gather / all-to-all of the triangle operands inside gb_denoise); every
    rank holds the same x_t and the same noise, so this is STRONG scaling of one sample.  Returns the result dict on
    rank 0 (None elsewhere).  Collective: every rank must call it."""
    import torch
    import torch.distributed as dist
    import genie2_b200
    from genie2_b200 import _lib, shard
    from oracle import genie_oracle as O, weights as W     # seeded weights / synthetic features (host side only)
    dev = torch.device('cuda', local)
    n = 1024
    lib = _lib.load()
    cfg = O.OracleConfig()
    genie = genie2_b200.Genie(genie2_b200.Config())
    genie.model.load_state_dict(W.make_state_dict(cfg, 0, 'activated'))
    genie = genie.to(dev)
    genie.setup_schedule()
    genie.model.set_precision('bf16')
    if world > 1:
        shard.check_shardable(1, n, world)
        genie.model.shard_rows()
    feats = {k: v.to(dev) for k, v in O.make_features([[n]]).items()}
    ctx = genie.model.bind_features(feats, dev)
    genie.bind_schedule(ctx)
    x = torch.randn(1, n, 3, generator=torch.Generator().manual_seed(1)).to(dev)     # same on every rank
    rots = genie2_b200.compute_frenet_frames(genie.model, feats, x)
    stream = torch.cuda.current_stream(dev).cuda_stream

    def loop(t_start, cnt):
        _lib.check(lib.gb_sample_loop(ctx, x.data_ptr(), rots.data_ptr(), t_start, t_start - cnt + 1, SCALE, None, 1234, stream),
                   ctx, 'gb_sample_loop')

    clocks = ClockSampler(local)
    loop(T_STEPS, W_)
    barrier()
    wall0 = time.time()
    l0, c0, b0 = genie.model.launch_count(), genie.model.collective_count(), genie.model.peer_barrier_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    loop(T_STEPS - W_, K)
    e1.record()
    barrier()
    wall1 = time.time()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    digest = x.double().sum().reshape(1)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        lo, hi = digest.clone(), digest.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        same = bool(lo == hi)
    else:
        same = True
    clk = clocks.stop(wall0, wall1)
    out = None
    if rank == 0:
        ms = float(ms)
        fl = flops_per_eval(n, 1)
        pk = peaks()
        out = {
            'metric': 'denoiser evals/sec at L=1024 batch 1', 'value': K / (ms * 1e-3), 'unit': UNIT, 'n_gpus': world, 'steps': K,
            'warmup': W_, 'ms_per_step': ms / K, 'higher_is_better': True, 'scaling': 'strong', 'vs_baseline': None, 'dtype': 'bf16',
            'data': 'synthetic',
            'config': {'workload': 'unconditional L=1024 batch 1, base config, scale 0.6 (BASELINE configs[4])',
                       'parallelism': ('pair rows sharded over %d GPUs (triangle contraction by channel), operands pulled from peer-mapped '
                                       'memory over NVLink, device-side barriers' % world) if world > 1 else 'single GPU',
                       'weights': 'seeded activated random init'},
            'seconds_per_1000_step_sample': ms / K,
            'gpu_launches': int(genie.model.launch_count() - l0), 'collectives': int(genie.model.collective_count() - c0),
            'peer_barriers': int(genie.model.peer_barrier_count() - b0),
            'recv_bytes_per_rank_per_step': shard.traffic_bytes(n, world), 'ranks_agree_bitwise': same,
            'finite': bool(torch.isfinite(x).all()), 'clocks': clk,
            'whole_eval': {'algorithmic_tflop_per_eval': fl / 1e12, 'achieved_tflops': fl / (ms / K * 1e-3) / 1e12,
                           'frac_of_tensor_sustained': fl / (ms / K * 1e-3) / 1e12 / pk['tensor_sustained'] / world},
        }
    del genie
    torch.cuda.empty_cache()
    return out


def run_c5(args):
    """--workload c5: the c5 block alone, as its own JSON line."""
    import torch
    import torch.distributed as dist
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (('WORLD_SIZE', '1'), ('RANK', '0'), ('LOCAL_RANK', '0')))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    out = measure_c5(args.steps if args.steps is not None else 20, max(args.warmup if args.warmup is not None else 3, 3), world, rank, local, barrier)
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


def measure_scaffold(genie, dev, rank, barrier, steps, prec):
    """BASELINE configs[3]: motif scaffolding with motif pair conditioning, batch 8, scale 0.4 (sample_scaffold.py,
    sampler/scaffold.py:12-168).  The design25 problem files are not in the reference snapshot, so the problem is the
    synthetic two-group one of tests/golden/io_problem.pdb (its parsing and seeded placement are pinned to the reference by
    tests/test_host_io_golden.py); the ragged batch of 8 placements is built by the package's own host path.  Every rank
    runs its own batch (no collective); rank 0 reports."""
    import torch
    import genie2_b200
    from genie2_b200 import _lib, feat_utils as F
    prob = os.path.join(ROOT, 'tests', 'golden', 'io_problem.pdb')
    if not os.path.exists(prob):
        return {'unavailable': 'tests/golden/io_problem.pdb missing'}
    lib = _lib.load()
    genie.model.set_precision(prec)
    np.random.seed(5 + rank)
    feats = F.convert_np_features_to_tensor(F.batchify_np_features([F.create_np_features_from_motif_pdb(prob) for _ in range(8)]), dev)
    nb, n = feats['residue_mask'].shape
    ctx = genie.model.bind_features(feats, dev)
    genie.bind_schedule(ctx)
    x = torch.randn(nb, n, 3, device=dev) * feats['residue_mask'].unsqueeze(-1)
    rots = genie2_b200.compute_frenet_frames(genie.model, feats, x)
    stream = torch.cuda.current_stream(dev).cuda_stream

    def loop(t, k):
        _lib.check(lib.gb_sample_loop(ctx, x.data_ptr(), rots.data_ptr(), t, t - k + 1, 0.4, None, 31 + rank, stream), ctx, 'gb_sample_loop')
    loop(T_STEPS, 5)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    loop(T_STEPS - 5, steps)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    barrier()
    if rank != 0:
        return None
    lens = [int(v) for v in feats['num_residues'].reshape(-1).tolist()] if 'num_residues' in feats else []
    return {'workload': 'motif scaffolding, synthetic two-group problem, batch 8 (BASELINE configs[3]), scale 0.4', 'precision': prec, 'batch': nb,
            'padded_length': n, 'lengths': lens, 'steps': steps, 'ms_per_step': ms, 'evals_per_s_per_gpu': 1e3 / ms,
            'seconds_per_1000_step_batch': ms, 'finite': bool(torch.isfinite(x).all()), 'motif_pair_conditioning': bool(feats['fixed_structure_mask'].any())}


def measure_length_sweep(genie, dev, world, rank, barrier, steps, prec):
    """BASELINE configs[2]: the reference's length-sweep task list (sample_unconditional.py:33-42,101-120: lengths 256..50,
    5 samples per length, batch size 4) through the package's own scheduler -- plan_batches (full batches + ragged packing
    of the left-over samples) and the shared dynamic queue of multiprocessor.py (a TCPStore fetch-add between the ranks; no
    tensor travels, no collective on the data path) -- with `steps` reverse steps per batch instead of 1000.  Wall time
    from a barrier to the barrier after the last rank has finished; every rank must call it."""
    import torch
    import torch.distributed as dist
    import genie2_b200
    from genie2_b200 import _lib
    from genie2_b200.feat_utils import batchify_np_features, convert_np_features_to_tensor, create_empty_np_features
    from genie2_b200.multiprocessor import StoreQueue, TaskQueue
    from genie2_b200.sample_unconditional import plan_batches
    lib = _lib.load()
    genie.model.set_precision(prec)
    lengths = list(range(256, 49, -1))
    units = plan_batches(lengths, 5, 4)
    q = StoreQueue(world, rank, key='gb_sweep') if world > 1 else None
    queue = TaskQueue(units, q.claim if q else iter(range(len(units) + 1)).__next__)
    stream = torch.cuda.current_stream(dev).cuda_stream
    barrier()
    t0 = time.perf_counter()
    l0 = genie.model.launch_count()
    n_units = n_samples = 0
    t_bind = t_feat = t_loop = 0.0
    ok = True
    for u in queue:
        tb = time.perf_counter()
        feats = convert_np_features_to_tensor(batchify_np_features([create_empty_np_features([n]) for n in u['lengths']]), dev)
        t_feat += time.perf_counter() - tb
        ctx = genie.model.bind_features(feats, dev)      # synchronous: features upload, workspace re-carve, static embedding
        genie.bind_schedule(ctx)
        t_bind += time.perf_counter() - tb
        nb = len(u['lengths'])
        x = torch.randn(nb, u['length'], 3, device=dev) * feats['residue_mask'].unsqueeze(-1)
        rots = genie2_b200.compute_frenet_frames(genie.model, feats, x)
        tl = time.perf_counter()
        _lib.check(lib.gb_sample_loop(ctx, x.data_ptr(), rots.data_ptr(), T_STEPS, T_STEPS - steps + 1, SCALE, None, 77 + n_units, stream),
                   ctx, 'gb_sample_loop')
        ok = ok and bool(torch.isfinite(x).all())        # (also the per-batch download a real run does)
        t_loop += time.perf_counter() - tl
        n_units += 1
        n_samples += nb
    torch.cuda.synchronize()
    busy = time.perf_counter() - t0
    launches = genie.model.launch_count() - l0
    if q:
        q.close()
    barrier()
    total = time.perf_counter() - t0
    stats = torch.tensor([busy, float(n_units), float(n_samples), float(launches), 1.0 if ok else 0.0, total, t_bind, t_feat, t_loop], device=dev, dtype=torch.float64)
    if world > 1:
        allst = [torch.zeros_like(stats) for _ in range(world)]
        dist.all_gather(allst, stats)
    else:
        allst = [stats]
    if rank != 0:
        return None
    rows = [[float(v) for v in t.cpu()] for t in allst]
    total = max(r[5] for r in rows)
    return {
        'workload': 'length sweep L=50..256 (207 lengths), 5 samples per length, batch 4 (BASELINE configs[2]); %d reverse steps per batch '
                    'instead of 1000' % steps,
        'precision': prec, 'n_gpus': world, 'steps_per_batch': steps, 'batches': len(units),
        'scheduler': 'plan_batches: 207 full batches + 52 ragged batches of left-over samples; shared dynamic queue, largest first',
        'seconds': total, 'samples': int(sum(r[2] for r in rows)), 'sample_steps_per_s': sum(r[2] for r in rows) * steps / total,
        'per_rank_busy_seconds': [r[0] for r in rows], 'per_rank_batches': [int(r[1]) for r in rows],
        # once per batch whatever the step count (< 1 % of a 1000-step run): host feature build + upload + gb_set_features
        'per_rank_bind_seconds': [r[6] for r in rows], 'per_rank_feature_build_seconds': [r[7] for r in rows],
        'per_rank_sample_loop_seconds': [r[8] for r in rows],

        'gpu_launches': int(sum(r[3] for r in rows)), 'finite': all(r[4] == 1.0 for r in rows),
        'scaling': 'strong (fixed task list)',
    }


if __name__ == '__main__':
    a = parse()
    if a.impl == 'reference':
        run_reference(a)
    elif a.workload == 'c5':
        run_c5(a)
    else:
        run_b200(a)
