"""Host-side multi-GPU logic on CPU: the task partition (pure function) and a world_size-2 gloo run in
which every rank derives its own share with no communication on the data path; the only collective
is the test's own all_gather that checks the shares are disjoint and complete."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from genie2_b200.multiprocessor import MultiProcessor, partition, shard, task_cost
from genie2_b200.sample_unconditional import UnconditionalRunner, build_parser


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
    _Recorder().run({'out': out}, num_devices=world)                           # torchrun mode: only this rank's share
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
