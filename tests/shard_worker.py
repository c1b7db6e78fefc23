"""One rank of the row-sharding check / timing (launched by torchrun from test_gpu_shard.py or by hand):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
        tests/shard_worker.py --length 256 --check [--time 20]

Every rank first runs the denoiser on its own GPU unsharded, then row-sharded over all ranks, on the same
seeded inputs, and compares: z (and p, when it fits) must be bit-identical, on every rank.  --time K also
times K evals of each and prints one JSON line on rank 0 (device time, max over ranks)."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from common import O, W, feats_to, make_denoiser  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--length', type=int, default=256)
    ap.add_argument('--check', action='store_true')
    ap.add_argument('--time', type=int, default=0)
    ap.add_argument('--out', default='')
    ap.add_argument('--regrow', type=int, default=0, help='then run the SAME sharded model at this larger length (the workspace is '
                    're-allocated and re-mapped between the ranks) and again at --length (re-carved in place)')
    a = ap.parse_args()
    import genie2_b200
    from genie2_b200 import shard

    rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
    local = int(os.environ.get('LOCAL_RANK', rank))
    torch.cuda.set_device(local)
    dev = 'cuda:%d' % local
    dist.init_process_group('nccl', device_id=torch.device(dev))
    N = a.length
    shard.check_shardable(1, N, world)
    cfg = O.OracleConfig()
    sd = W.make_state_dict(cfg, 2, 'activated')
    ts = torch.tensor([400], dtype=torch.int32, device=dev)

    def inputs(n):
        g = torch.Generator().manual_seed(4)
        feats = O.make_features([[n]])
        trans = torch.randn(1, n, 3, generator=g) * 12
        rots = O.compute_frenet_frames(trans, feats['chain_index'], feats['residue_mask'])
        return feats_to(feats, dev), genie2_b200.T(rots.to(dev), trans.to(dev))

    model = make_denoiser(cfg, sd, precision='bf16', device=dev)
    with_p = a.check and N <= 512
    model.return_intermediates = with_p
    f, t = inputs(N)

    def timed(k):
        for _ in range(3):
            model(t, ts, f)
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            model(t, ts, f)
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / k], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    one = model(t, ts, f)
    one = {k: v.clone() for k, v in one.items() if k in ('z', 'p')}
    ms_one = timed(a.time) if a.time else None
    model.return_intermediates = False
    model.shard_rows()
    model.return_intermediates = with_p
    many = model(t, ts, f)
    torch.cuda.synchronize()
    ok = bool(torch.isfinite(many['z']).all())
    if a.check:
        # both runs use the same kernels on the same tiles in the same k order: bit-identical
        ok = ok and torch.equal(one['z'], many['z'])
        if with_p:
            ok = ok and torch.equal(one['p'], many['p'])
        if not ok:
            dz = float((one['z'] - many['z']).abs().max())
            dp = float((one['p'] - many['p']).abs().max()) if with_p else float('nan')
            print('rank %d: MISMATCH max |dz| = %g, max |dp| = %g (pair stack)' % (rank, dz, dp), flush=True)
    n_coll, n_bar = model.collective_count(), model.peer_barrier_count()
    if a.regrow:
        # a larger sample on the same sharded context, then the first one again: same bits as fresh single-GPU runs
        shard.check_shardable(1, a.regrow, world)
        model.return_intermediates = False
        f2, t2 = inputs(a.regrow)
        big = model(t2, ts, f2)['z'].clone()
        again = model(t, ts, f)['z'].clone()
        fresh = make_denoiser(cfg, sd, precision='bf16', device=dev)
        ok = ok and torch.equal(again, one['z']) and torch.equal(big, fresh(t2, ts, f2)['z'])
        del fresh
        if not ok:
            print('rank %d: MISMATCH after re-growing the workspace' % rank, flush=True)
    model.return_intermediates = False
    ms_many = timed(a.time) if a.time else None
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        line = {'workload': 'C5 row-sharded denoiser eval, L=%d B=1, bf16' % N, 'world': world, 'ok': bool(int(flag)),
                'nccl_collectives_per_eval': n_coll, 'peer_barriers_per_eval': n_bar, 'recv_bytes_per_rank_per_eval': shard.traffic_bytes(N, world),
                'ms_per_eval_one_gpu': ms_one, 'ms_per_eval_sharded': ms_many}
        print(json.dumps(line), flush=True)
        if a.out:
            with open(a.out, 'w') as fh:
                fh.write(json.dumps(line) + '\n')
    dist.destroy_process_group()
    sys.exit(0 if int(flag) else 1)


if __name__ == '__main__':
    main()
