"""Process-per-GPU task farm (reference: genie/utils/multiprocessor.py:8-100).

Same abstract interface (``create_tasks`` / ``create_constants`` / ``execute(constants, tasks,
device)`` / ``run(params, num_devices, sequential_order)``), same data-parallel model: independent
tasks, one OS process per ``cuda:i``, **no collective on the data path**.  Two differences:

* partitioning is cost-weighted longest-processing-time-first instead of shuffled equal-count bins
  (the reference's bins are by task count although a length-L task costs ~F(L)), and it
  is a pure function of (tasks, world) so every rank of a ``torchrun`` launch derives the same
  partition without communicating;
* under ``torchrun`` (``WORLD_SIZE`` set) ``run`` executes only this rank's share on
  ``cuda:LOCAL_RANK`` instead of spawning processes.
"""
import os
from abc import ABC, abstractmethod
from multiprocessing import get_context


def denoiser_flops(n):
    """Algorithmic FLOPs of one denoiser eval of one length-n sample, base config."""
    return 2560.0 * n ** 3 + 3378304.0 * n ** 2 + 28022016.0 * n


def task_cost(task):
    if 'cost' in task:
        return float(task['cost'])
    if 'length' in task:
        return denoiser_flops(int(task['length']))
    return 1.0


def partition(tasks, world, sequential_order=False):
    """Longest-processing-time-first assignment of tasks to ``world`` bins; returns a list of
    ``world`` task lists.  Deterministic; every task appears in exactly one bin.  With
    ``sequential_order`` the bins are contiguous equal-count slices in the given order, exactly the
    reference's ``tasks[binsize*i:binsize*(i+1)]``."""
    world = max(1, int(world))
    if sequential_order:
        binsize = -(-len(tasks) // world)
        return [list(tasks[binsize * i:binsize * (i + 1)]) for i in range(world)]
    order = sorted(range(len(tasks)), key=lambda i: (-task_cost(tasks[i]), i))
    bins, load = [[] for _ in range(world)], [0.0] * world
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        bins[r].append(tasks[i])
        load[r] += task_cost(tasks[i])
    return bins


def shard(tasks, world, rank, sequential_order=False):
    return partition(tasks, world, sequential_order)[rank]


class MultiProcessor(ABC):

    @abstractmethod
    def create_tasks(self, params):
        raise NotImplementedError

    @abstractmethod
    def create_constants(self, params):
        raise NotImplementedError

    @abstractmethod
    def execute(self, constants, tasks, device):
        raise NotImplementedError

    def run(self, params, num_devices, sequential_order=False):
        tasks = self.create_tasks(params)
        constants = self.create_constants(params)
        if 'WORLD_SIZE' in os.environ and int(os.environ['WORLD_SIZE']) > 1:
            world, rank = int(os.environ['WORLD_SIZE']), int(os.environ.get('RANK', '0'))
            local = int(os.environ.get('LOCAL_RANK', str(rank)))
            self.execute(constants, shard(tasks, world, rank, sequential_order), 'cuda:{}'.format(local))
            return
        bins = partition(tasks, num_devices, sequential_order)
        if num_devices <= 1:
            self.execute(constants, bins[0], 'cuda:0')
            return
        ctx = get_context('spawn')          # CUDA contexts do not survive fork
        procs = []
        for i in range(num_devices):
            p = ctx.Process(target=self.execute, args=(constants, bins[i], 'cuda:{}'.format(i)))
            p.start()
            procs.append(p)
        for p in procs:
            p.join()
        bad = [p.exitcode for p in procs if p.exitcode]
        if bad:
            raise RuntimeError('worker process(es) failed with exit codes %s' % bad)
