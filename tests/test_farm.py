"""Host-side multi-GPU logic on CPU: the batch plan of the length sweep (ragged packing of the left-over samples), the
static task partition (pure function), and world_size-2 gloo runs of both schedules -- static (every rank derives its
own share with no communication) and dynamic (shared queue: a TCPStore fetch-add, no tensor travels); the only
collective is the test's own all_gather that checks the shares are disjoint and complete."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from genie2_b200.multiprocessor import MultiProcessor, TaskQueue, partition, shard, task_cost
from genie2_b200.sample_unconditional import UnconditionalRunner, build_parser, plan_batches


def _tasks():
    args = build_parser().parse_args(['--name', 'base', '--epoch', '40', '--scale', '0.6', '--outdir', 'o'])
    return UnconditionalRunner().create_tasks(vars(args))


def test_partition_covers_and_balances():
    tasks = _tasks()
    assert [t['length'] for t in tasks] == list(range(256, 49, -1))            # sample_unconditional.py:33-42
    for world in (1, 2, 4, 8):
        bins = partition(tasks, world)
        got = sorted(t['length'] for b in bins for t in b)
        assert got == list(range(50, 257))
        loads = [sum(task_cost(t) for t in b) for b in bins]
        assert max(loads) / min(loads) < 1.02                                   # cost-weighted, not count-weighted
    ref = partition(tasks, 4, sequential_order=True)                           # the reference's contiguous bins
    assert [len(b) for b in ref] == [52, 52, 52, 51] and ref[0][0]['length'] == 256
    assert partition([], 3) == [[], [], []]
    assert partition(tasks[:2], 4)[2:] == [[], []]


def test_plan_batches_packs_leftovers_into_ragged_batches():
    lengths = list(range(256, 49, -1))
    ref = plan_batches(lengths, 5, 4, pack=False)                  # the reference's loop: a batch of 4 and a batch of 1 per length
    assert len(ref) == 414 and [len(u['lengths']) for u in ref[:2]] == [4, 1]
    units = plan_batches(lengths, 5, 4)
    assert len(units) == 207 + 52
    seen = sorted((n, o) for u in units for n, o in zip(u['lengths'], u['offsets']))
    assert seen == sorted((n, o) for n in lengths for o in range(5))         # every (length, sample index) exactly once
    for u in units:
        assert u['length'] == max(u['lengths']) and len(u['lengths']) <= 4
        assert u['length'] - min(u['lengths']) <= 3                            # neighbours only: at most 3 residues of padding
    assert plan_batches([64], 8, 4) == plan_batches([64], 8, 4, pack=False)    # nothing left over: nothing to pack
    # the runner schedules these units, largest first under the dynamic queue
    args = build_parser().parse_args(['--name', 'base', '--epoch', '40', '--scale', '0.6', '--outdir', 'o'])
    r = UnconditionalRunner()
    got = r.schedule_units(r.create_tasks(vars(args)), vars(args))
    assert got == units
    q = TaskQueue(got, iter(range(10 ** 6)).__next__)
    costs = [task_cost(t) for t in q]
    assert len(costs) == 259 and costs == sorted(costs, reverse=True)


class _Recorder(MultiProcessor):
    def create_tasks(self, params):
        return _tasks()

    def create_constants(self, params):
        return {'out': params['out']}

    def execute(self, constants, tasks, device):
        with open(os.path.join(constants['out'], device.replace(':', '_') + '.txt'), 'w') as fh:
            fh.write(' '.join(str(t['length']) for t in tasks))


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    _Recorder().run({'out': out}, num_devices=world, schedule='static')        # torchrun mode: only this rank's share
    mine = [t['length'] for t in shard(_tasks(), world, rank)]
    flags = torch.zeros(257, dtype=torch.int64)
    flags[mine] = 1
    gathered = [torch.zeros_like(flags) for _ in range(world)]
    dist.all_gather(gathered, flags)
    total = torch.stack(gathered).sum(0)
    assert int(total[50:].min()) == 1 and int(total[50:].max()) == 1 and int(total[:50].sum()) == 0
    wrote = open(os.path.join(out, 'cuda_%d.txt' % rank)).read().split()
    assert [int(x) for x in wrote] == mine
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_shares_are_disjoint_and_complete(tmp_path):
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)


def _worker_dynamic(rank, world, port, out):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    _Recorder().run({'out': out}, num_devices=world)                           # default schedule: the shared dynamic queue
    mine = [int(x) for x in open(os.path.join(out, 'cuda_%d.txt' % rank)).read().split()]
    assert mine == sorted(mine, reverse=True)                                  # claimed largest first
    flags = torch.zeros(257, dtype=torch.int64)
    flags[mine] = 1
    gathered = [torch.zeros_like(flags) for _ in range(world)]
    dist.all_gather(gathered, flags)
    total = torch.stack(gathered).sum(0)
    assert int(total[50:].min()) == 1 and int(total[50:].max()) == 1 and int(total[:50].sum()) == 0
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_dynamic_queue_hands_out_every_task_once(tmp_path):
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker_dynamic, args=(2, port, str(tmp_path)), nprocs=2, join=True)


class _SpawnRecorder(_Recorder):
    def execute(self, constants, tasks, device):
        with open(os.path.join(constants['out'], device.replace(':', '_') + '.txt'), 'w') as fh:
            for t in tasks:
                fh.write('%d ' % t['length'])


def test_spawned_workers_share_one_queue(tmp_path):
    _SpawnRecorder().run({'out': str(tmp_path)}, num_devices=2)
    got = []
    for i in range(2):
        got += [int(x) for x in open(os.path.join(str(tmp_path), 'cuda_%d.txt' % i)).read().split()]
    assert sorted(got) == list(range(50, 257))
