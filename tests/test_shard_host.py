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
    assert shard.traffic_bytes(1024, 8) == 5 * 2 * 3 * (img // 64) * 7 + 8 * 1024 * 384 * 4 * 7 // 8
    assert shard.traffic_bytes(1024, 8) < 0.4 * (5 * 2 * img * 7 // 8)      # well under the all-gather of b alone (round 1)


def img_off(Np, B, row_split, k_split, b, ch, kb, row, cg_split=1):
    """Python restatement of tc::img_off (element offset of a (b, channel, k-block, row) row of 64):
    [channel group][row block][k block] [b][CG][KBL][RL][64]."""
    RL, KBL, CG = Np // row_split, (Np // 64) // k_split, 128 // cg_split
    blk = B * CG * KBL * RL * 64
    rb, kq, cg = row // RL, kb // KBL, ch // CG
    return ((cg * row_split * k_split) + rb * k_split + kq) * blk + (((b * CG + ch % CG) * KBL + kb % KBL) * RL + row % RL) * 64


def x_off(Np, B, world, b, i, jb64, ch):
    """Python restatement of tc::x_off: [row block][channel group][b][local i / 2][jb64][i & 1][CG][64]."""
    RL, CG, njb = Np // world, 128 // world, Np // 64
    blk = B * (RL // 2) * njb * 2 * CG * 64
    rb, il, cg = i // RL, i % RL, ch // CG
    return (rb * world + cg) * blk + ((((b * (RL // 2) + il // 2) * njb + jb64) * 2 + (il & 1)) * CG + ch % CG) * 64


@pytest.mark.parametrize('world', [2, 4, 8])
@pytest.mark.parametrize('mode', ['rows', 'kblocks'])
def test_image_layout_is_a_block_grid_by_channel_group_and_producer(world, mode):
    """Both operand images of a sharded triangle update: the producer (rank s: its rows for the outgoing update, its k-blocks
    for the incoming one) writes block (d, s) for every channel group d; block (d, s) is contiguous, so the pull of csrc/shard.cu
    (rank d copies block (d, s) out of rank s's memory) needs no packing; afterwards rank d reads blocks (d, *) only."""
    Np, B = 128 * world, 1
    rs, ks = (world, 1) if mode == 'rows' else (1, world)
    CG = 128 // world
    total = B * 128 * (Np // 64) * Np * 64
    blk = total // (world * world)
    seen = np.zeros(total // 64, dtype=np.int32)
    for src in range(world):
        rows = range(src * Np // world, (src + 1) * Np // world) if mode == 'rows' else range(Np)
        kbs = range(Np // 64) if mode == 'rows' else range(src * (Np // 64) // world, (src + 1) * (Np // 64) // world)
        for ch in sorted({0, 1, CG - 1, CG, 127}):
            d = ch // CG
            for kb in kbs:
                for row in rows:
                    off = img_off(Np, B, rs, ks, 0, ch, kb, row, world)
                    assert off % 64 == 0 and (d * world + src) * blk <= off < (d * world + src + 1) * blk
                    seen[off // 64] += 1
    assert seen.max() == 1
    # unsharded layout is the plain [b][ch][kb][row][64] image
    assert img_off(Np, 1, 1, 1, 0, 5, 1, 7) == ((5 * (Np // 64) + 1) * Np + 7) * 64
    # the 128-row operand tiles of the contraction stay contiguous
    for row0 in range(0, Np, 128):
        a = img_off(Np, B, rs, ks, 0, 9, 1, row0, world)
        assert img_off(Np, B, rs, ks, 0, 9, 1, row0 + 127, world) == a + 127 * 64


@pytest.mark.parametrize('world', [2, 4, 8])
def test_x_layout_returns_blocks_to_the_row_owners(world):
    """Contraction result: rank d (channel group d) writes block (s, d) for every row block s in its own memory; rank s fetches
    its tiles from there as world x 2 contiguous pieces of CG x 128 bytes."""
    Np, B = 128 * world, 1
    CG = 128 // world
    total = B * Np * (Np // 64) * 128 * 64
    blk = total // (world * world)
    seen = np.zeros(total // 64, dtype=np.int32)
    for i in range(0, Np):
        for jb in range(Np // 64):
            for ch in sorted({0, CG - 1, CG, 127}):
                off = x_off(Np, B, world, 0, i, jb, ch)
                s_, d = i // (Np // world), ch // CG
                assert (s_ * world + d) * blk <= off < (s_ * world + d + 1) * blk
                seen[off // 64] += 1
            a = x_off(Np, B, world, 0, i, jb, CG)                     # a piece: CG consecutive channel rows
            assert x_off(Np, B, world, 0, i, jb, 2 * CG - 1) == a + (CG - 1) * 64 if world > 1 else True
    assert seen.max() == 1
    assert x_off(256, 1, 1, 0, 5, 2, 77) == ((((5 // 2) * 4 + 2) * 2 + 1) * 128 + 77) * 64      # one GPU: the tile images


def pull_pairs(world, me, block):
    """Python restatement of shard_pull_blocks (csrc/shard.cu): (peer, offset) of every block this rank copies out of a peer's
    buffer into the same offset of its own: blocks (me, s), s != me, of the world x world grid, starting on another peer per rank."""
    return [((me + k) % world, (me * world + (me + k) % world) * block) for k in range(1, world)]


@pytest.mark.parametrize('world', [2, 4])
@pytest.mark.parametrize('mode', ['rows', 'kblocks'])
def test_exchange_model_every_rank_ends_up_with_what_its_kernels_read(world, mode):
    """The whole exchange of one triangle update on `world` model ranks with separate buffers (one 64-element row = one number):
    K4 writes its blocks locally, every rank pulls blocks (me, s), K5 finds ITS channels for all rows and all k, writes x blocks
    (s, me) locally, and K6 finds every piece of its rows in the memory of the channel's owner."""
    Np, B, CG = 128 * world, 1, 128 // world
    njb, rs, ks = Np // 64, *((world, 1) if mode == 'rows' else (1, world))
    rows64 = B * 128 * njb * Np                       # rows of 64 elements in an image
    blk = rows64 // (world * world)
    value = lambda ch, kb, row: (ch * njb + kb) * Np + row + 1
    a = [np.zeros(rows64, dtype=np.int64) for _ in range(world)]
    for me in range(world):                           # K4: own rows (outgoing) or own k-blocks (incoming), all channels
        rows = range(me * Np // world, (me + 1) * Np // world) if mode == 'rows' else range(Np)
        kbs = range(njb) if mode == 'rows' else range(me * njb // world, (me + 1) * njb // world)
        for ch in range(0, 128, 3):
            for kb in kbs:
                for row in rows:
                    a[me][img_off(Np, B, rs, ks, 0, ch, kb, row, world) // 64] = value(ch, kb, row)
    snap = [v.copy() for v in a]                      # the pull reads what the producers wrote, not what other pulls wrote
    for me in range(world):
        for peer, off in pull_pairs(world, me, blk):
            a[me][off:off + blk] = snap[peer][off:off + blk]
    for me in range(world):                           # K5: this rank's channel group, every k-block, every row
        for ch in range(me * CG, (me + 1) * CG):
            if ch % 3:
                continue
            for kb in range(njb):
                for row in range(0, Np, 5):
                    assert a[me][img_off(Np, B, rs, ks, 0, ch, kb, row, world) // 64] == value(ch, kb, row)
    xrows = B * Np * njb * 128
    xv = lambda i, jb, ch: (i * njb + jb) * 128 + ch + 1
    x = [np.zeros(xrows, dtype=np.int64) for _ in range(world)]
    for me in range(world):                           # K5 epilogue: own channels, all rows, into its own memory
        for ch in range(me * CG, (me + 1) * CG):
            for i in range(Np):
                for jb in range(njb):
                    x[me][x_off(Np, B, world, 0, i, jb, ch) // 64] = xv(i, jb, ch)
    for me in range(world):                           # K6: own rows; the piece of channel group cg lies in rank cg's memory
        for i in range(me * Np // world, (me + 1) * Np // world, 3):
            for jb in range(njb):
                for cg in range(world):
                    base = x_off(Np, B, world, 0, i, jb, cg * CG) // 64
                    assert list(x[cg][base:base + CG]) == [xv(i, jb, cg * CG + c) for c in range(CG)]


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
