"""Length-sweep plumbing on the GPU (BASELINE configs[2]): the workspace is re-carved in place when a smaller batch
follows a larger one (grow-only allocation), and the ragged batches that pack the left-over samples of neighbouring
lengths give every sample the result it gets when it is run alone."""
import os

import numpy as np
import pytest
import torch

from common import O, W, feats_to, make_denoiser, rel

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def _forward(model, lengths, seed):
    import genie2_b200
    feats = feats_to(O.make_features([[n] for n in lengths]), DEV)
    B, N = feats['residue_mask'].shape
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(B, N, 3, generator=g) * 4).to(DEV) * feats['residue_mask'].unsqueeze(-1)
    rots = genie2_b200.compute_frenet_frames(model, feats, x)
    tt = torch.full((B,), 321, dtype=torch.int32, device=DEV)
    out = model(genie2_b200.T(rots, x), tt, feats)
    torch.cuda.synchronize()
    return {k: out[k].clone() for k in ('z', 's', 'p')}, x


@pytest.mark.parametrize('prec', ['fp32', 'bf16x3', 'bf16'])
def test_rebind_smaller_batch_reuses_workspace_bit_identically(prec):
    cfg = O.OracleConfig(n_pair_transform_layer=2, n_structure_layer=2)
    sd = W.make_state_dict(cfg, 5, 'activated')
    fresh = make_denoiser(cfg, sd, precision=prec)
    want, _ = _forward(fresh, [40, 37], 3)
    reused = make_denoiser(cfg, sd, precision=prec)
    _forward(reused, [96, 96, 90], 1)                    # the larger batch allocates and leaves its data behind
    big = reused.workspace_bytes()
    got, _ = _forward(reused, [40, 37], 3)
    assert reused.workspace_bytes() <= big
    for k in want:
        assert torch.equal(got[k], want[k]), (prec, k, rel(got[k], want[k]))
    _forward(reused, [100, 100, 100, 100], 2)            # and growing again still works
    again, _ = _forward(reused, [40, 37], 3)
    for k in want:
        assert torch.equal(again[k], want[k]), (prec, k)


@pytest.mark.parametrize('prec', ['fp32', 'bf16x3'])
def test_ragged_batch_matches_samples_run_alone(prec):
    """Samples of lengths 53, 52, 51, 50 in one padded batch (what plan_batches makes of the left-over samples) against
    each sample run alone: same z on the valid residues."""
    cfg = O.OracleConfig(n_pair_transform_layer=2, n_structure_layer=2)
    sd = W.make_state_dict(cfg, 6, 'activated')
    model = make_denoiser(cfg, sd, precision=prec)
    lengths = [53, 52, 51, 50]
    packed, x = _forward(model, lengths, 9)
    import genie2_b200
    for i, n in enumerate(lengths):
        feats = feats_to(O.make_features([[n]]), DEV)
        xi = x[i:i + 1, :n].contiguous()
        rots = genie2_b200.compute_frenet_frames(model, feats, xi)
        out = model(genie2_b200.T(rots, xi), torch.full((1,), 321, dtype=torch.int32, device=DEV), feats)
        e = rel(packed['z'][i, :n], out['z'][0])
        print(prec, 'length', n, 'packed vs alone', e)
        assert e < (1e-5 if prec == 'fp32' else 1e-4), (prec, n, e)


def test_sampler_writes_ragged_batch_under_reference_names(tmp_path):
    import genie2_b200
    gcfg = genie2_b200.Config()
    gcfg.model['n_pair_transform_layer'] = 1
    gcfg.model['n_structure_layer'] = 1
    gcfg.diffusion['n_timestep'] = 4
    genie = genie2_b200.Genie(gcfg)
    cfg = O.OracleConfig(n_pair_transform_layer=1, n_structure_layer=1, n_timestep=4)
    genie.model.load_state_dict(W.make_state_dict(cfg, 7, 'activated'))
    genie = genie.to(DEV)
    sampler = genie2_b200.UnconditionalSampler(genie)
    sampler.sample({'length': 23, 'lengths': [23, 22, 21], 'offsets': [4, 4, 4], 'scale': 0.6, 'num_samples': 3,
                    'outdir': str(tmp_path), 'prefix': '23', 'offset': 4, 'noise': 'philox', 'seed': 1})
    names = sorted(os.listdir(os.path.join(str(tmp_path), 'pdbs')))
    assert names == ['21_4.pdb', '22_4.pdb', '23_4.pdb']
    for n in (21, 22, 23):
        lines = [l for l in open(os.path.join(str(tmp_path), 'pdbs', '%d_4.pdb' % n)) if l.startswith('ATOM')]
        assert len(lines) == n
