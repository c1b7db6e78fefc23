"""Host-side logic of the row sharding (genie2_b200/shard.py) on CPU: row ranges, the shardability rules, the
unique-id exchange over a world_size-2 gloo group, and -- as a numpy model of csrc/tc_common.cuh's ImgLayout --
that per-rank blocks of the operand images, concatenated by an in-place all-gather, address every (row, k)
element exactly once at the offset the contraction kernel reads."""
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from genie2_b200 import shard


def test_row_ranges_and_rules():
    assert [shard.row_range(1024, 8, r) for r in (0, 7)] == [(0, 128), (896, 1024)]
    assert shard.row_range(512, 1, 0) == (0, 512)
    with pytest.raises(ValueError):
        shard.row_range(1000, 3, 0)
    with pytest.raises(ValueError):
        shard.row_range(1024, 2, 2)
    shard.check_shardable(4, 300, 1)                 # world 1: anything goes
    shard.check_shardable(1, 1024, 8)
    with pytest.raises(ValueError):
        shard.check_shardable(2, 1024, 2)            # batches shard by sample, not by row
    with pytest.raises(ValueError):
        shard.check_shardable(1, 768, 4)             # 192 rows per rank: not whole 128-row blocks
    assert shard.traffic_bytes(1024, 1) == 0
    img = 1024 * 1024 * 128 * 2
    assert shard.traffic_bytes(1024, 8) == (5 * 2 * img + 8 * 1024 * 384 * 4 + 5 * img // 8) * 7 // 8


def img_off(Np, B, row_split, k_split, b, ch, kb, row):
    """Python restatement of tc::img_off (element offset of a (b, channel, k-block, row) row of 64)."""
    RL, KBL = Np // row_split, (Np // 64) // k_split
    blk = B * 128 * KBL * RL * 64
    rb, kq = row // RL, kb // KBL
    return (rb * k_split + kq) * blk + (((b * 128 + ch) * KBL + kb % KBL) * RL + row % RL) * 64


@pytest.mark.parametrize('world', [2, 4])
@pytest.mark.parametrize('mode', ['rows', 'kblocks'])
def test_image_layout_is_a_partition(world, mode):
    Np, B, nch = 128 * world, 1, 3                   # 3 of the 128 channels are enough to catch a stride error
    rs, ks = (world, 1) if mode == 'rows' else (1, world)
    total = B * 128 * (Np // 64) * Np * 64
    piece = total // world
    seen = np.zeros(total // 64, dtype=np.int32)
    for rank in range(world):
        rows = range(rank * Np // world, (rank + 1) * Np // world) if mode == 'rows' else range(Np)
        kbs = range(Np // 64) if mode == 'rows' else range(rank * (Np // 64) // world, (rank + 1) * (Np // 64) // world)
        for ch in list(range(nch)) + [127]:
            for kb in kbs:
                for row in rows:
                    off = img_off(Np, B, rs, ks, 0, ch, kb, row)
                    assert off % 64 == 0 and rank * piece <= off < (rank + 1) * piece      # inside this rank's all-gather piece
                    seen[off // 64] += 1
    assert seen.max() == 1
    # unsharded layout is the plain [b][ch][kb][row][64] image
    assert img_off(Np, 1, 1, 1, 0, 5, 1, 7) == ((5 * (Np // 64) + 1) * Np + 7) * 64
    # 128-row operand tiles stay contiguous in both sharded layouts
    for row0 in range(0, Np, 128):
        a = img_off(Np, B, rs, ks, 0, 9, 1, row0)
        assert img_off(Np, B, rs, ks, 0, 9, 1, row0 + 127) == a + 127 * 64


@pytest.mark.parametrize('world', [2, 4])
def test_incoming_a_image_is_a_block_grid(world):
    """Incoming update: a is cut by destination row block AND source k block; rank s writes column s of the
    world x world block grid, the in-place all-to-all moves block (d, s) to rank d, whose contraction then reads
    row d only (csrc/shard.cu, shard_all_to_all)."""
    Np, B = 128 * world, 1
    blk = B * 128 * ((Np // 64) // world) * (Np // world) * 64
    for src in range(world):                                      # writer: k blocks of rank src, every image row
        for kb in range(src * (Np // 64) // world, (src + 1) * (Np // 64) // world):
            for row in (0, Np // world - 1, Np // world, Np - 1):
                off = img_off(Np, B, world, world, 0, 17, kb, row)
                d = row // (Np // world)
                assert (d * world + src) * blk <= off < (d * world + src + 1) * blk
    for dst in range(world):                                      # reader: its own rows, every k block
        lo, hi = dst * world * blk, (dst + 1) * world * blk
        for kb in range(Np // 64):
            for row in range(dst * Np // world, (dst + 1) * Np // world, 37):
                assert lo <= img_off(Np, B, world, world, 0, 3, kb, row) < hi


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    want = bytes((7 * i + 3) % 256 for i in range(shard.UNIQUE_ID_BYTES))

    def draw():
        assert rank == 0, 'only rank 0 draws the id'
        return want

    got = shard.exchange_unique_id(None, draw=draw)
    assert got == want
    lo, hi = shard.row_range(1024, world, rank)
    assert hi - lo == 1024 // world and lo == rank * (1024 // world)
    dist.barrier()
    dist.destroy_process_group()


def test_unique_id_exchange_gloo_world2():
    mp.spawn(_worker, args=(2, _free_port()), nprocs=2, join=True)
