"""Debug aid for the row-sharded triangle update (run under torchrun on 2+ GPUs): one module, stage by stage, sharded
against unsharded on the same GPU, with a map of where the results differ.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 scripts/shard_debug.py
"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
from common import O, W, feats_to, make_denoiser  # noqa: E402


def main():
    import genie2_b200._lib as L
    N = int(os.environ.get('DBG_N', '256'))
    rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
    torch.cuda.set_device(rank)
    dev = 'cuda:%d' % rank
    dist.init_process_group('nccl', device_id=torch.device(dev))
    lib = L.load()
    cfg = O.OracleConfig(n_pair_transform_layer=1, n_structure_layer=1)
    sd = W.make_state_dict(cfg, 3, 'activated')
    model = make_denoiser(cfg, sd, precision='bf16', device=dev)
    feats = feats_to(O.make_features([[N]]), dev)
    g = torch.Generator().manual_seed(3)
    p0 = (torch.randn(1, N, N, 128, generator=g) * 1.5 + 0.3).to(dev)
    Np = (N + 63) // 64 * 64
    njb = Np // 64
    n_img = 128 * Np * Np

    def grab(ctx, which):
        out = torch.empty(n_img, dtype=torch.bfloat16, device=dev)
        assert lib.gb_dbg_tc_copy(ctx, which, out.data_ptr(), n_img * 2) == Np
        return out

    ref = {}
    ctx = model.bind_features(feats, torch.device(dev))
    for outgoing in (1, 0):
        p = p0.clone()
        L.check(lib.gb_k_trimul(ctx, 0, outgoing, p.data_ptr(), None), ctx, 'gb_k_trimul')
        torch.cuda.synchronize()
        ref[outgoing] = (p, grab(ctx, 0), grab(ctx, 1), grab(ctx, 2))
    model.shard_rows()
    ctx = model.bind_features(feats, torch.device(dev))
    RL, CG = N // world, 128 // world
    r0 = rank * RL
    for outgoing in (1, 0):
        p = p0.clone()
        L.check(lib.gb_k_trimul(ctx, 0, outgoing, p.data_ptr(), None), ctx, 'gb_k_trimul sharded')
        torch.cuda.synchronize()
        p_ref, a_ref, b_ref, x_ref = ref[outgoing]
        dp = (p[0, r0:r0 + RL] - p_ref[0, r0:r0 + RL]).abs().amax(-1)            # [RL][N]
        blocks = dp.view(RL // 64, 64, N // 64, 64).amax((1, 3))
        msg = ['rank %d outgoing %d: max |dp| own rows %.4g; per 64x64 block:' % (rank, outgoing, float(dp.max()))]
        msg.append(str(blocks.cpu().numpy().round(4)))
        # x: sharded [rb][cg][il/2][jb][ii][chl][64] -> [i/2][jb][ii][ch][64]
        xs = grab(ctx, 2).view(world, world, RL // 2, njb, 2, CG, 64)
        xr = x_ref.view(Np // 2, njb, 2, 128, 64)[r0 // 2:(r0 + RL) // 2]
        mine = xs[rank].permute(1, 2, 3, 0, 4, 5).reshape(RL // 2, njb, 2, 128, 64)
        dx = (mine.float() - xr.float()).abs()
        msg.append('x own rows: max diff %.4g; by channel group %s; by row-pair block %s' % (
            float(dx.max()), dx.view(RL // 2, njb, 2, world, CG, 64).amax((0, 1, 2, 4, 5)).cpu().numpy().round(4),
            dx.view(RL // 128, 64, -1).amax((1, 2)).cpu().numpy().round(4)))
        # operand images after the all-to-all: this rank's channel group, all rows / k.  Unsharded [ch][kb][Np][64].
        for name, which, r in (('a', 0, a_ref), ('b', 1, b_ref)):
            img = grab(ctx, which)
            full = r.view(128, njb, Np, 64)[rank * CG:(rank + 1) * CG]          # [CG][kb][row][64]
            if outgoing:   # blocks (cg, rb): [CG][njb][RL][64]
                got = img.view(world, world, CG, njb, RL, 64)[rank].permute(1, 2, 0, 3, 4).reshape(CG, njb, Np, 64)
            else:          # blocks (cg, kq): [CG][KBL][Np][64]
                got = img.view(world, world, CG, njb // world, Np, 64)[rank].permute(1, 0, 2, 3, 4).reshape(CG, njb, Np, 64)
            d = (got.float() - full.float()).abs()
            msg.append('%s image, my channel group: max diff %.4g; by source block %s' % (
                name, float(d.max()),
                (d.view(CG, njb, world, RL, 64).amax((0, 1, 3, 4)) if outgoing else d.view(CG, world, -1).amax((0, 2))).cpu().numpy().round(4)))
        for r in range(world):
            if r == rank:
                print('\n'.join(msg), flush=True)
            dist.barrier()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
