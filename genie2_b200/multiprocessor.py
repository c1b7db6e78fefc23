"""Process-per-GPU task farm (reference: genie/utils/multiprocessor.py:8-100).

Same abstract interface (``create_tasks`` / ``create_constants`` / ``execute(constants, tasks,
device)`` / ``run(params, num_devices, sequential_order)``), same data-parallel model: independent
tasks, one OS process per ``cuda:i``, **no collective on the data path**.  Three differences:

* the default schedule is a **shared dynamic queue**: the work units are ordered by decreasing cost and every worker
  claims the next unclaimed one through an atomic counter (a ``multiprocessing.Value`` between spawned processes, a
  ``TCPStore`` fetch-add between the ranks of a ``torchrun`` launch), so a GPU that finishes early keeps pulling work
  -- the reference's shuffled equal-count bins leave GPUs idle because a length-L task costs ~F(L).
  ``execute`` still receives "a list of tasks": an iterable that claims on ``next()``;
* ``schedule='static'`` is the communication-free alternative: cost-weighted longest-processing-time-first bins, a pure
  function of (tasks, world), so every rank derives the same partition by itself;
* under ``torchrun`` (``WORLD_SIZE`` set) ``run`` works on ``cuda:LOCAL_RANK`` instead of spawning processes.

A runner may also override ``schedule_units(tasks, params)`` to regroup its tasks into the units that are actually
scheduled (``UnconditionalRunner`` packs the left-over samples of neighbouring lengths into ragged batches).
"""
import os
import time
from abc import ABC, abstractmethod
from multiprocessing import get_context

# seconds-equivalent cost model of one reverse step of a batch (used only to ORDER work): algorithmic flops plus a
# fixed per-step latency (short chains are latency-bound: ~0.7 ms per step against 3.3 ms for L=256 batch 4)
STEP_OVERHEAD_FLOPS = 2.3e11


def denoiser_flops(n):
    """Algorithmic FLOPs of one denoiser eval of one length-n sample, base config."""
    return 2560.0 * n ** 3 + 3378304.0 * n ** 2 + 28022016.0 * n


def task_cost(task):
    if 'cost' in task:
        return float(task['cost'])
    if 'length' in task:
        return denoiser_flops(int(task['length']))
    return 1.0


def batch_cost(n_max, batch):
    """Cost of one reverse step of a (possibly ragged) batch padded to length n_max."""
    return batch * denoiser_flops(int(n_max)) + STEP_OVERHEAD_FLOPS


class TaskQueue:
    """The "list of tasks" handed to ``execute`` under the dynamic schedule: iterating it claims the next unclaimed
    unit (largest cost first) from a counter shared by all workers.  ``claimed`` lists what this worker took."""

    def __init__(self, tasks, claim):
        self.tasks = sorted(tasks, key=lambda t: -task_cost(t))
        self._claim = claim
        self.claimed = []

    def __iter__(self):
        while True:
            i = self._claim()
            if i >= len(self.tasks):
                return
            self.claimed.append(self.tasks[i])
            yield self.tasks[i]

    def __len__(self):
        return len(self.tasks)


def _value_claimer(counter):
    def claim():
        with counter.get_lock():
            i = counter.value
            counter.value += 1
        return i
    return claim


class StoreQueue:
    """Atomic work counter between the ranks of a torchrun launch: a TCPStore served by rank 0 on MASTER_PORT + 29
    (``GB_QUEUE_PORT`` overrides), ``add`` is the fetch-and-increment.  No tensor ever travels through it."""

    def __init__(self, world, rank, key='gb_next'):
        from datetime import timedelta
        from torch.distributed import TCPStore
        host = os.environ.get('MASTER_ADDR', '127.0.0.1')
        port = int(os.environ.get('GB_QUEUE_PORT', int(os.environ.get('MASTER_PORT', '29500')) + 29))
        self.world, self.rank, self.key = world, rank, key
        self.store = TCPStore(host, port, world, is_master=(rank == 0), timeout=timedelta(seconds=600), wait_for_workers=True)

    def claim(self):
        return int(self.store.add(self.key, 1)) - 1

    def close(self):
        """Rank 0 serves the store: it may only leave once every rank has stopped claiming."""
        n = int(self.store.add(self.key + '_done', 1))
        if self.rank == 0:
            while n < self.world:
                time.sleep(0.01)
                n = int(self.store.add(self.key + '_done', 0))


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

    def schedule_units(self, tasks, params):
        """Hook: regroup the tasks into the units that are scheduled (default: the tasks themselves)."""
        return tasks

    def _execute_dynamic(self, constants, tasks, device, counter):
        self.execute(constants, TaskQueue(tasks, _value_claimer(counter)), device)

    def run(self, params, num_devices, sequential_order=False, schedule=None):
        """schedule: 'dynamic' (default: shared queue, largest unit first), 'static' (cost-weighted LPT bins);
        ``sequential_order`` forces the reference's contiguous equal-count bins."""
        tasks = self.schedule_units(self.create_tasks(params), params)
        constants = self.create_constants(params)
        schedule = 'static' if sequential_order else (schedule or params.get('schedule') or 'dynamic')
        if schedule not in ('dynamic', 'static'):
            raise ValueError("schedule must be 'dynamic' or 'static'")
        if 'WORLD_SIZE' in os.environ and int(os.environ['WORLD_SIZE']) > 1:
            world, rank = int(os.environ['WORLD_SIZE']), int(os.environ.get('RANK', '0'))
            local = int(os.environ.get('LOCAL_RANK', str(rank)))
            if schedule == 'static':
                self.execute(constants, shard(tasks, world, rank, sequential_order), 'cuda:{}'.format(local))
                return
            q = StoreQueue(world, rank)
            try:
                self.execute(constants, TaskQueue(tasks, q.claim), 'cuda:{}'.format(local))
            finally:
                q.close()
            return
        if num_devices <= 1:
            self.execute(constants, partition(tasks, 1, sequential_order)[0], 'cuda:0')
            return
        ctx = get_context('spawn')          # CUDA contexts do not survive fork
        procs = []
        bins = partition(tasks, num_devices, sequential_order) if schedule == 'static' else None
        counter = ctx.Value('i', 0) if schedule == 'dynamic' else None
        for i in range(num_devices):
            if schedule == 'static':
                p = ctx.Process(target=self.execute, args=(constants, bins[i], 'cuda:{}'.format(i)))
            else:
                p = ctx.Process(target=self._execute_dynamic, args=(constants, tasks, 'cuda:{}'.format(i), counter))
            p.start()
            procs.append(p)
        for p in procs:
            p.join()
        bad = [p.exitcode for p in procs if p.exitcode]
        if bad:
            raise RuntimeError('worker process(es) failed with exit codes %s' % bad)
