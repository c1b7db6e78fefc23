"""Row sharding of ONE long sample over the GPUs of a box (BASELINE config C5).

No reference counterpart: the reference only farms whole samples (genie/utils/multiprocessor.py:73-100).
Host-side logic only; the exchange itself happens inside the CUDA library (csrc/shard.cu):

* ``row_range(N, world, rank)``    rows of the pair representation a rank owns;
* ``check_shardable(B, N, world)`` the constraints of the kernels (one sample, whole 128-row blocks per rank);
* ``exchange_unique_id(group)``    rank 0 draws the 128-byte NCCL unique id, every rank receives it over the
                                   process group that launched the job (``nccl`` under torchrun, ``gloo`` in tests);
* ``traffic_bytes(N, world, ...)`` bytes each rank receives per denoiser eval (for the bench line).

Exchange pattern (round 2): the triangle contraction is sharded by CHANNEL (x^c = a^c b^cT is independent per channel):
K4 on a rank's rows -> the owner of each 128 / world channel group pulls its blocks of a and b out of the producers' memory
(peer-mapped over NVLink, device-side barrier first) -> contraction of own channels over all rows -> K6 on own rows fetches
its x tiles from the channel owners.  3 (W - 1) / W^2 of an image per rank and update instead of the (W - 1) / W of
round 1's all-gather of b; NCCL only carries the setup (unique id here, CUDA IPC handles inside the library).
"""
import ctypes as C

UNIQUE_ID_BYTES = 128


def row_range(N, world, rank):
    if world < 1 or not 0 <= rank < world:
        raise ValueError('bad rank %d of %d' % (rank, world))
    if N % world:
        raise ValueError('N = %d is not divisible by world = %d' % (N, world))
    rows = N // world
    return rank * rows, (rank + 1) * rows


def check_shardable(B, N, world):
    """Raises ValueError with the reason when (B, N) cannot be row-sharded over ``world`` GPUs."""
    if world == 1:
        return
    if B != 1:
        raise ValueError('row sharding handles one sample per call (B = %d); batches shard by sample instead' % B)
    if N % (128 * world):
        raise ValueError('row sharding over %d GPUs needs N to be a multiple of %d (got %d)' % (world, 128 * world, N))


def traffic_bytes(N, world, n_pair_layer=5, n_struct_layer=8, c_p=128, c_s=384):
    """Bytes each rank reads from its peers per denoiser eval.  Per triangle update (world - 1) of the world^2 blocks of three
    bf16 images (its channel group of a and b from the row owners, its rows of x from the channel owners); per structure
    layer the other ranks' rows of the fp32 single representation."""
    if world == 1:
        return 0
    image = N * N * c_p * 2
    exchanged = n_pair_layer * 2 * 3 * (image // (world * world)) * (world - 1)
    gathered = n_struct_layer * N * c_s * 4 * (world - 1) // world
    return exchanged + gathered


def _draw_unique_id():
    from . import _lib
    buf = (C.c_uint8 * UNIQUE_ID_BYTES)()
    _lib.check(_lib.load().gb_shard_unique_id(buf), None, 'gb_shard_unique_id')
    return bytes(buf)


def exchange_unique_id(group=None, draw=_draw_unique_id):
    """Collective over ``group``: returns the same 128 bytes on every rank."""
    import torch
    import torch.distributed as dist
    rank = dist.get_rank(group)
    payload = draw() if rank == 0 else bytes(UNIQUE_ID_BYTES)
    if len(payload) != UNIQUE_ID_BYTES:
        raise ValueError('unique id must be %d bytes' % UNIQUE_ID_BYTES)
    t = torch.tensor(list(payload), dtype=torch.uint8)
    on_gpu = dist.get_backend(group) == 'nccl'
    if on_gpu:
        t = t.cuda()
    dist.broadcast(t, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    return bytes(t.cpu().tolist())
