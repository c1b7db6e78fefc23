"""Parity of the CUDA path (through the C ABI) against the oracle and the golden fixtures.

fp32 mode is held to 1e-4 relative L2 on every tensor ``Denoiser.forward`` returns
(north-star gate: 1e-3 on z); the tensor-core modes state their own tolerances in
test_gpu_tensorcore.py.
"""
import ctypes as C
import os

import numpy as np
import pytest
import torch

from common import GOLDEN, load_forward_case, make_denoiser, feats_to, run_cuda_forward, rel
from oracle import genie_oracle as O, weights as W

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'
TOL = 1e-4
# Tests that run the oracle on the box's own CPU (instead of comparing with a recorded fixture) see its fp32 summation
# order change with the CPU and thread count (conftest.py pins 8 threads; a 24-core box still differed by 2e-4 on
# the final frames after eight structure layers).  s and p stay at 1e-4; what comes out of the structure net is held
# to the north-star gate there.
TOL_LIVE_STRUCT = 1e-3


def test_frenet_golden_and_random():
    import genie2_b200._lib as L
    lib = L.load()
    d = np.load(os.path.join(GOLDEN, 'primitives.npz'))
    cases = [(torch.from_numpy(d['fr_coords']), torch.from_numpy(d['fr_chains']), torch.from_numpy(d['fr_mask']),
              torch.from_numpy(d['fr_rots']))]
    g = torch.Generator().manual_seed(3)
    for B, N in ((1, 3), (2, 50), (5, 257), (1, 1024)):
        x = torch.randn(B, N, 3, generator=g) * 5
        ch = torch.zeros(B, N, dtype=torch.int32)
        m = torch.ones(B, N, dtype=torch.int32)
        if N > 20:
            ch[0, N // 2:] = 1
            m[-1, N - 7:] = 0
        cases.append((x, ch, m, O.compute_frenet_frames(x, ch, m)))
    for x, ch, m, ref in cases:
        B, N = m.shape
        xd, cd, md = x.to(DEV).contiguous(), ch.to(DEV).contiguous(), m.to(DEV).contiguous()
        out = torch.empty(B, N, 3, 3, device=DEV)
        L.check(lib.gb_k_frenet(xd.data_ptr(), cd.data_ptr(), md.data_ptr(), B, N, out.data_ptr(), None), None, 'gb_k_frenet')
        torch.cuda.synchronize()
        assert float((out.cpu() - ref).abs().max()) < 2e-5, (B, N)


@pytest.mark.parametrize('name', ['uncond_base_ragged', 'uncond_stock', 'motif_base', 'multichain', 'triatt'])
def test_forward_matches_reference_fixture(name):
    """CUDA forward vs the unmodified reference's recorded outputs (quaternion sign replayed)."""
    cfg, sd, feats, t = load_forward_case(name)
    model = make_denoiser(cfg, sd)
    o, _ = run_cuda_forward(model, t['rots'], t['trans'], t['timesteps'], feats, t['quat'])
    # motif_base is ill-conditioned in the structure net: the reference's OWN fp32 result is
    # 3.6e-3 (z) / 8.2e-3 (rots) away from an fp64 evaluation of the same graph (measured with the
    # oracle in fp64), so two correct fp32 implementations with different summation order differ by
    # ~1e-4 there.  s and p (before the amplification) are still held to 1e-4.
    tol_struct = 5e-4 if name == 'motif_base' else TOL
    assert rel(o['s'], t['s']) < TOL
    assert rel(o['p'], t['p']) < TOL
    assert rel(o['states'], t['states']) < tol_struct
    assert rel(o['trans'], t['out_trans']) < tol_struct
    assert rel(o['rots'], t['out_rots']) < tol_struct
    assert rel(o['z'], t['z']) < tol_struct


def test_forward_canonical_sign_matches_oracle():
    """Production quaternion convention (w >= 0) against the oracle run in the same mode."""
    cfg, sd, feats, t = load_forward_case('uncond_base_ragged')
    model = make_denoiser(cfg, sd)
    o, _ = run_cuda_forward(model, t['rots'], t['trans'], t['timesteps'], feats, None)
    with torch.no_grad():
        ref = O.denoiser_forward(sd, cfg, t['rots'], t['trans'], t['timesteps'], feats, 'canonical')
    for k in ('s', 'p', 'states', 'trans', 'rots', 'z'):
        assert rel(o[k], ref[k]) < (TOL if k in ('s', 'p') else TOL_LIVE_STRUCT), k


def test_forward_base_config_L64_vs_oracle():
    """Config C1 shape (L=64, B=1) plus a ragged B=2 batch, full base config, activated weights."""
    cfg = O.OracleConfig()
    sd = W.make_state_dict(cfg, 5, 'activated')
    model = make_denoiser(cfg, sd)
    g = torch.Generator().manual_seed(9)
    for lengths, steps in (([[64]], [1000]), ([[64], [50]], [321, 7])):
        feats = O.make_features(lengths)
        B, N = feats['residue_mask'].shape
        trans = torch.randn(B, N, 3, generator=g) * 6
        rots = O.compute_frenet_frames(trans, feats['chain_index'], feats['residue_mask'])
        ts = torch.tensor(steps, dtype=torch.int32)
        with torch.no_grad():
            ref = O.denoiser_forward(sd, cfg, rots, trans, ts, feats, 'canonical')
        o, _ = run_cuda_forward(model, rots, trans, ts, feats, None)
        for k in ('s', 'p', 'states', 'trans', 'rots', 'z'):
            assert rel(o[k], ref[k]) < (TOL if k in ('s', 'p') else TOL_LIVE_STRUCT), (k, lengths)


def test_padding_and_batch_invariance_L128():
    """Size-independent properties at a size the oracle is slow on: the result for a sample
    does not depend on what else is in the batch, nor on how far it is padded."""
    cfg = O.OracleConfig()
    sd = W.make_state_dict(cfg, 6, 'activated')
    model = make_denoiser(cfg, sd)
    g = torch.Generator().manual_seed(10)
    x96 = torch.randn(1, 96, 3, generator=g) * 6
    f1 = O.make_features([[96]])
    r1 = O.compute_frenet_frames(x96, f1['chain_index'], f1['residue_mask'])
    ts1 = torch.tensor([400], dtype=torch.int32)
    o1, _ = run_cuda_forward(model, r1, x96, ts1, f1, None)
    f2 = O.make_features([[128], [96]])
    x2 = torch.randn(2, 128, 3, generator=g) * 6
    x2[1, :96] = x96[0]
    x2[1, 96:] = 0
    r2 = O.compute_frenet_frames(x2, f2['chain_index'], f2['residue_mask'])
    o2, _ = run_cuda_forward(model, r2, x2, torch.tensor([77, 400], dtype=torch.int32), f2, None)
    assert rel(o2['z'][1, :96], o1['z'][0]) < 1e-5
    assert rel(o2['p'][1, :96, :96], o1['p'][0]) < 1e-5
    assert float(o2['p'][1, 96:].abs().max()) == 0.0 and float(o2['p'][1, :, 96:].abs().max()) == 0.0


def test_p_sample_matches_oracle():
    import genie2_b200
    from genie2_b200.sampler import _posterior
    cfg = O.OracleConfig(n_pair_transform_layer=1, n_structure_layer=1)
    sd = W.make_state_dict(cfg, 7, 'activated')
    gcfg = genie2_b200.Config()
    gcfg.model['n_pair_transform_layer'] = 1
    gcfg.model['n_structure_layer'] = 1
    genie = genie2_b200.Genie(gcfg)
    genie.model.load_state_dict(sd)
    genie = genie.to(DEV)
    genie.setup_schedule()
    sched = O.setup_schedule(1000)
    feats = O.make_features([[9, 8], [12]])
    fd = feats_to(feats, DEV)
    g = torch.Generator().manual_seed(2)
    B, N = feats['residue_mask'].shape
    for step in (1000, 500, 2, 1):
        x = torch.randn(B, N, 3, generator=g) * 3
        z = torch.randn(B, N, 3, generator=g)
        nz = torch.randn(B, N, 3, generator=g)
        r_ref, x_ref = O.p_sample(sched, x, z, step, 0.6, nz, feats)
        rots0 = O.compute_frenet_frames(x, feats['chain_index'], feats['residue_mask'])
        ts = _posterior(genie, genie2_b200.T(rots0.to(DEV), x.to(DEV)), z.to(DEV), step, fd, 0.6, nz.to(DEV))
        torch.cuda.synchronize()
        assert float((ts.trans.cpu() - x_ref).abs().max()) < 1e-5, step
        assert float((ts.rots.cpu() - r_ref).abs().max()) < 1e-4, step


def test_trajectory_fixture_lockstep():
    """The reference sampler's recorded 6-step trajectory (tests/golden/traj_uncond.npz): same
    tape, reference quaternion signs replayed per step; final C-alpha within 1e-3 A."""
    import genie2_b200
    from genie2_b200.sampler import _posterior, compute_frenet_frames
    d = np.load(os.path.join(GOLDEN, 'traj_uncond.npz'))
    T_, L, B = int(d['T']), int(d['L']), int(d['B'])
    cfg = O.OracleConfig(n_timestep=T_, n_pair_transform_layer=2, n_structure_layer=3)
    sd = W.make_state_dict(cfg, int(d['seed']), 'activated')
    gcfg = genie2_b200.Config()
    gcfg.diffusion['n_timestep'] = T_
    gcfg.model['n_pair_transform_layer'] = 2
    gcfg.model['n_structure_layer'] = 3
    genie = genie2_b200.Genie(gcfg)
    genie.model.load_state_dict(sd)
    genie = genie.to(DEV)
    genie.setup_schedule()
    feats = feats_to(O.make_features([[L]] * B), DEV)
    tape = torch.from_numpy(d['tape']).to(DEV)
    quats = torch.from_numpy(d['quats']).to(DEV)
    trans = tape[T_].clone()
    ts = genie2_b200.T(compute_frenet_frames(genie.model, feats, trans), trans)
    for step in range(T_, 0, -1):
        tt = torch.full((B,), step, dtype=torch.int32, device=DEV)
        z = genie.model(ts, tt, feats, quat_follow=quats[T_ - step])['z']
        ts = _posterior(genie, ts, z, step, feats, float(d['scale']), tape[step])
    torch.cuda.synchronize()
    err = float((ts.trans.cpu() - torch.from_numpy(d['final'])).abs().max())
    assert err < 1e-3, err


@pytest.mark.parametrize('precision,gate', [('fp32', 0.1), ('bf16', None)])
def test_c1_trajectory_fixture(precision, gate):
    """BASELINE.json configs[0] (L = 64, batch 1, 50 steps, full base configuration) against the trajectory the
    unmodified reference sampler produced (tests/golden/traj_c1.npz): same tape, the reference's quaternion signs
    replayed.  fp32 mode must land within the north star's 0.1 A C-alpha RMSD (centred); the trajectory reaches
    |x| ~ 300 A with these random weights, so that is ~3e-4 relative after 50 compounding steps.  bf16 mode reports
    its RMSD (stated separately, as the north star asks) and is held to 5 % of the structure's radius."""
    import genie2_b200
    from genie2_b200.sampler import _posterior, compute_frenet_frames
    d = np.load(os.path.join(GOLDEN, 'traj_c1.npz'))
    T_, L, B = int(d['T']), int(d['L']), int(d['B'])
    sd = W.make_state_dict(O.OracleConfig(n_timestep=T_), int(d['seed']), 'activated')
    gcfg = genie2_b200.Config()
    gcfg.diffusion['n_timestep'] = T_
    genie = genie2_b200.Genie(gcfg)
    genie.model.load_state_dict(sd)
    genie = genie.to(DEV)
    genie.model.set_precision(precision)
    genie.setup_schedule()
    feats = feats_to(O.make_features([[L]] * B), DEV)
    tape = torch.from_numpy(d['tape']).to(DEV)
    quats = torch.from_numpy(d['quats'].astype(np.float32)).to(DEV)
    trans = tape[T_].clone()
    ts = genie2_b200.T(compute_frenet_frames(genie.model, feats, trans), trans)
    for step in range(T_, 0, -1):
        tt = torch.full((B,), step, dtype=torch.int32, device=DEV)
        z = genie.model(ts, tt, feats, quat_follow=quats[T_ - step])['z']
        ts = _posterior(genie, ts, z, step, feats, float(d['scale']), tape[step])
    torch.cuda.synchronize()
    got, ref = ts.trans.cpu(), torch.from_numpy(d['final'])
    assert torch.isfinite(got).all()
    a = got[0] - got[0].mean(0, keepdim=True)
    b = ref[0] - ref[0].mean(0, keepdim=True)
    rmsd = float(((a - b) ** 2).sum(-1).mean().sqrt())
    radius = float((b ** 2).sum(-1).mean().sqrt())
    print('C1 %s: centred C-alpha RMSD %.4f A (structure radius %.1f A)' % (precision, rmsd, radius))
    assert rmsd < (gate if gate is not None else 0.05 * radius), rmsd


def test_sample_loop_equals_stepwise():
    """gb_sample_loop (one call, tape) == the python loop of Denoiser.forward + gb_p_sample, bit for bit."""
    import genie2_b200
    cfg = O.OracleConfig(n_timestep=5, n_pair_transform_layer=1, n_structure_layer=2)
    sd = W.make_state_dict(cfg, 8, 'activated')
    gcfg = genie2_b200.Config()
    gcfg.diffusion['n_timestep'] = 5
    gcfg.model['n_pair_transform_layer'] = 1
    gcfg.model['n_structure_layer'] = 2
    genie = genie2_b200.Genie(gcfg)
    genie.model.load_state_dict(sd)
    genie = genie.to(DEV)
    sampler = genie2_b200.UnconditionalSampler(genie)
    tape = torch.randn(6, 2, 20, 3, generator=torch.Generator().manual_seed(4))
    params = {'length': 20, 'scale': 0.6, 'num_samples': 2, 'outdir': '/tmp/_gb', 'prefix': 'x', 'offset': 0}
    a = sampler._sample(dict(params, noise_tape=tape))
    # stepwise with the same tape
    from genie2_b200.sampler import _posterior, compute_frenet_frames
    feats = feats_to(O.make_features([[20]] * 2), DEV)
    trans = tape[5].to(DEV).clone()
    ts = genie2_b200.T(compute_frenet_frames(genie.model, feats, trans), trans)
    for step in range(5, 0, -1):
        tt = torch.full((2,), step, dtype=torch.int32, device=DEV)
        ts = _posterior(genie, ts, genie.model(ts, tt, feats)['z'], step, feats, 0.6, tape[step].to(DEV))
    torch.cuda.synchronize()
    got = np.stack([f['atom_positions'] for f in a])
    assert np.array_equal(got.astype(np.float32), ts.trans.cpu().numpy())


def test_cli_unconditional_end_to_end(tmp_path):
    """The reference's CLI surface (genie/sample_unconditional.py flags) through the runner, random-init
    weights, 4 diffusion steps: one PDB per (length, sample), C-alpha records only."""
    from genie2_b200 import sample_unconditional as S
    root = tmp_path / 'results' / 'tiny'
    root.mkdir(parents=True)
    (root / 'configuration').write_text('name tiny\nnumTimesteps 4\nnumPairTransformLayers 1\nnumStructureLayers 2\n')
    out = tmp_path / 'out'
    args = S.build_parser().parse_args([
        '--name', 'tiny', '--epoch', '1', '--rootdir', str(tmp_path / 'results'), '--scale', '0.6', '--outdir', str(out),
        '--num_samples', '3', '--batch_size', '2', '--min_length', '20', '--max_length', '22', '--random_init', '0'])
    S.main(args)
    files = sorted(os.listdir(out / 'pdbs'))
    assert files == ['%d_%d.pdb' % (l, i) for l in (20, 21, 22) for i in range(3)]
    lines = [l for l in open(out / 'pdbs' / '22_2.pdb') if l.startswith('ATOM')]
    assert len(lines) == 22 and all(' CA ' in l for l in lines)


@pytest.mark.parametrize('starting', [1, 0])
@pytest.mark.parametrize('B,N', [(2, 21), (1, 70), (1, 280)])
def test_triangle_attention_kernel(B, N, starting):
    """K8 on its own (gb_k_tri_attention) vs the oracle's restatement of
    modules/triangular_attention.py:67-130: starting and ending node, ragged masks, N above one
    thread block of queries."""
    import genie2_b200._lib as L
    lib = L.load()
    cfg = O.OracleConfig(include_tri_att=True, n_pair_transform_layer=1, n_structure_layer=1)
    sd = W.make_state_dict(cfg, 21, 'activated')
    model = make_denoiser(cfg, sd)
    lens = [[N]] + [[max(3, N - 6 - 2 * i)] for i in range(B - 1)]
    feats = feats_to(O.make_features(lens), DEV)
    ctx = model.bind_features(feats, torch.device(DEV))
    g = torch.Generator().manual_seed(N)
    p0 = torch.randn(B, N, N, 128, generator=g) * 1.2 + 0.2
    rm = feats['residue_mask'].cpu()
    pm = rm.unsqueeze(1) * rm.unsqueeze(2)
    pre = 'pair_transform_net.net.0.tri_att_%s' % ('start' if starting else 'end')
    with torch.no_grad():
        ref = O.triangle_attention(sd, cfg, pre, p0, pm, bool(starting))
    p = p0.to(DEV).contiguous()
    L.check(lib.gb_k_tri_attention(ctx, 0, starting, p.data_ptr(), None), ctx, 'gb_k_tri_attention')
    torch.cuda.synchronize()
    assert rel(p.cpu() - p0, ref) < TOL
