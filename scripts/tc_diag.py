"""Per-kernel diagnostics of the tcgen05 (bf16) pair-stack kernels against the fp32 SIMT kernels
and a torch fp32 restatement, one kernel per process so that a trapped kernel does not hide the
others.   python scripts/tc_diag.py            -> runs every case in subprocesses
          python scripts/tc_diag.py CASE B N   -> one case in this process
"""
import ctypes as C
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))


def rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def decode_image(buf, B, Np):
    """[B][128][Np/64][Np][64] swizzled bf16 tile images -> [B,128,Np(row),Np(k)] float."""
    import torch
    t = buf.view(B, 128, Np // 64, Np, 8, 8).float()            # [.., row, chunk position, e]
    r = torch.arange(Np, device=buf.device)
    q = torch.arange(8, device=buf.device)
    pos = (q[None, :] ^ (r[:, None] & 7))                        # position of logical chunk q in row r
    idx = pos[None, None, None, :, :, None].expand(B, 128, Np // 64, Np, 8, 8)
    t = torch.gather(t, 4, idx)                                  # logical chunk order
    return t.reshape(B, 128, Np // 64, Np, 64).permute(0, 1, 3, 2, 4).reshape(B, 128, Np, Np)


def decode_x(buf, B, Np):
    """x tile images [B][Np/2][Np/64][2][128 ch][64 j], chunks swizzled by ch & 7 -> [B,128,Np(i),Np(j)] float."""
    import torch
    t = buf.view(B, Np // 2, Np // 64, 2, 128, 8, 8).float()
    c = torch.arange(128, device=buf.device)
    q = torch.arange(8, device=buf.device)
    pos = q[None, :] ^ (c[:, None] & 7)
    idx = pos[None, None, None, None, :, :, None].expand(B, Np // 2, Np // 64, 2, 128, 8, 8)
    t = torch.gather(t, 5, idx)
    return t.permute(0, 4, 1, 3, 2, 5, 6).reshape(B, 128, Np, Np)


def setup(B, N, seed=3):
    import torch
    from common import make_denoiser, feats_to
    from oracle import genie_oracle as O, weights as W
    cfg = O.OracleConfig(n_pair_transform_layer=1, n_structure_layer=1)
    sd = W.make_state_dict(cfg, seed, 'activated')
    model = make_denoiser(cfg, sd)
    lens = [[N]] + [[max(3, N - 5 - 3 * i)] for i in range(B - 1)]
    feats = feats_to(O.make_features(lens), 'cuda:0')
    ctx = model.bind_features(feats, torch.device('cuda:0'))
    g = torch.Generator().manual_seed(seed)
    p = (torch.randn(B, N, N, 128, generator=g) * 1.5 + 0.3).cuda()
    return model, sd, feats, ctx, p


def run_case(case, B, N):
    import torch
    import genie2_b200._lib as L
    lib = L.load()
    model, sd, feats, ctx, p0 = setup(B, N)
    mask = feats['residue_mask'].float()
    pm = (mask[:, :, None] * mask[:, None, :])[..., None]
    sdc = {k: v.cuda() for k, v in sd.items()}
    ln = torch.nn.functional.layer_norm

    def call(prec, fn):
        model.set_precision(prec)
        p = p0.clone()
        fn(p)
        torch.cuda.synchronize()
        return p

    if case == 'transition':
        f = lambda p: L.check(lib.gb_k_pair_transition(ctx, 0, p.data_ptr(), None), ctx, 'transition')
        ref = call('fp32', f)
        got = call('bf16', f)
        pre = 'pair_transform_net.net.0.pair_transition.'
        z = ln(p0, (128,), sdc[pre + 'layer_norm.weight'], sdc[pre + 'layer_norm.bias'])
        hid = torch.relu(z @ sdc[pre + 'linear_1.weight'].T + sdc[pre + 'linear_1.bias'])
        t = (p0 + (hid @ sdc[pre + 'linear_2.weight'].T + sdc[pre + 'linear_2.bias']) * pm) * pm
        print('transition B=%d N=%d: simt vs torch %.3e | tc vs torch %.3e | tc-delta vs torch-delta %.3e' %
              (B, N, rel(ref, t), rel(got, t), rel(got - p0 * pm, t - p0 * pm)))
        return
    outgoing = 1 if case == 'trimul_out' else 0
    f = lambda p: L.check(lib.gb_k_trimul(ctx, 0, outgoing, p.data_ptr(), None), ctx, 'trimul')
    ref = call('fp32', f)
    got = call('bf16', f)
    pre = 'pair_transform_net.net.0.tri_mul_%s.' % ('out' if outgoing else 'in')
    lin = lambda n, v: v @ sdc[pre + n + '.weight'].T + sdc[pre + n + '.bias']
    z = ln(p0, (128,), sdc[pre + 'layer_norm_in.weight'], sdc[pre + 'layer_norm_in.bias'])
    a = lin('linear_a_p', z) * torch.sigmoid(lin('linear_a_g', z)) * pm
    b = lin('linear_b_p', z) * torch.sigmoid(lin('linear_b_g', z)) * pm
    gate = torch.sigmoid(lin('linear_g', z))
    if outgoing:
        x = torch.einsum('bikc,bjkc->bijc', a, b)
    else:
        x = torch.einsum('bkic,bkjc->bijc', a, b)
    upd = lin('linear_z', ln(x, (128,), sdc[pre + 'layer_norm_out.weight'], sdc[pre + 'layer_norm_out.bias'])) * gate
    t = p0 + upd
    Np = lib.gb_dbg_tc_copy(ctx, 0, None, 0)
    assert Np > 0, Np

    def grab(which, n):
        out = torch.empty(n, dtype=torch.bfloat16, device='cuda')
        rc = lib.gb_dbg_tc_copy(ctx, which, out.data_ptr(), n * 2)
        assert rc == Np, rc
        return out
    ai = decode_image(grab(0, B * 128 * Np * Np), B, Np)[:, :, :N, :N]
    bi = decode_image(grab(1, B * 128 * Np * Np), B, Np)[:, :, :N, :N]
    xi = decode_x(grab(2, B * 128 * Np * Np), B, Np)[:, :, :N, :N]
    gi = grab(3, B * N * N * 128).view(B, N, N, 128).float()
    # operand orientation: outgoing A[row=i][k] = a[i,k]; incoming A[row=i][k] = a[k,i]
    a_ref = a.permute(0, 3, 1, 2) if outgoing else a.permute(0, 3, 2, 1)
    b_ref = b.permute(0, 3, 1, 2) if outgoing else b.permute(0, 3, 2, 1)
    print('%s B=%d N=%d Np=%d: a %.3e  b %.3e  g %.3e  x %.3e | simt vs torch %.3e | tc vs torch %.3e | delta %.3e' %
          (case, B, N, Np, rel(ai, a_ref), rel(bi, b_ref), rel(gi * pm, gate * pm), rel(xi, x.permute(0, 3, 1, 2)),
           rel(ref, t), rel(got, t), rel(got - p0, upd)))


def main():
    if len(sys.argv) >= 4:
        run_case(sys.argv[1], int(sys.argv[2]), int(sys.argv[3]))
        return
    cases = [('transition', 1, 64), ('transition', 2, 50), ('trimul_out', 1, 64), ('trimul_in', 1, 64),
             ('trimul_out', 2, 100), ('trimul_in', 2, 100), ('trimul_out', 1, 256), ('trimul_in', 1, 200),
             ('transition', 1, 256)]
    for c, B, N in cases:
        r = subprocess.run([sys.executable, __file__, c, str(B), str(N)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                           timeout=300)
        out = r.stdout.decode().strip().splitlines()
        print('[%s %d %d] rc=%d  %s' % (c, B, N, r.returncode, out[-1] if out else ''))
        if r.returncode != 0:
            print('\n'.join(out[-8:]))


if __name__ == '__main__':
    main()
