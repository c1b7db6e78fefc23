"""BASELINE.json configs[3]: motif scaffolding with motif pair conditioning, batch 8 (sample_scaffold.py, scaffold.py:12-168).
The design25 problem files are not in the reference snapshot, so the problem is the synthetic two-group one
of tests/golden/io_problem.pdb, whose parsing and seeded placement are pinned to the reference by test_host_io_golden.py.
Here: the features the host path builds for a ragged batch of 8 placements go through the CUDA denoiser and must match the
oracle (fp32, <= 1e-3), bf16 is stated separately, and the CLI writes the sample and motif PDBs the reference would."""
import os
import shutil

import numpy as np
import pytest
import torch

from common import GOLDEN, make_denoiser, run_cuda_forward, rel
from oracle import genie_oracle as O, weights as W

pytestmark = pytest.mark.gpu
PROB = os.path.join(GOLDEN, 'io_problem.pdb')


def _scaffold_batch(n, seed):
    from genie2_b200 import feat_utils as F
    np.random.seed(seed)
    feats = F.convert_np_features_to_tensor(
        F.batchify_np_features([F.create_np_features_from_motif_pdb(PROB) for _ in range(n)]), 'cpu')
    return feats


@pytest.mark.parametrize('precision,tol', [('fp32', 1e-3), ('bf16', 4e-2)])
def test_scaffold_batch8_forward_vs_oracle(precision, tol):
    feats = _scaffold_batch(8, 5)
    B, N = feats['residue_mask'].shape
    assert B == 8 and len(set(feats['num_residues'].tolist())) > 1          # ragged placements
    assert bool(feats['fixed_structure_mask'].any()) and int(feats['fixed_group'].max()) == 2
    cfg = O.OracleConfig()
    sd = W.make_state_dict(cfg, 17, 'activated')
    model = make_denoiser(cfg, sd, precision=precision)
    g = torch.Generator().manual_seed(3)
    trans = torch.randn(B, N, 3, generator=g) * 5 * feats['residue_mask'][..., None]
    rots = O.compute_frenet_frames(trans, feats['chain_index'], feats['residue_mask'])
    ts = torch.tensor([1000, 900, 640, 500, 333, 120, 2, 1], dtype=torch.int32)
    with torch.no_grad():
        ref = O.denoiser_forward(sd, cfg, rots, trans, ts, feats, 'canonical')
    o, _ = run_cuda_forward(model, rots, trans, ts, feats, None)
    errs = {k: rel(o[k], ref[k]) for k in ('s', 'p', 'z')}
    print('scaffold batch 8, %s: %s' % (precision, errs))
    for k, e in errs.items():
        assert e < tol, (k, e)


def test_cli_scaffold_end_to_end(tmp_path):
    import genie2_b200.sample_scaffold as S
    from genie2_b200 import feat_utils as F
    root = tmp_path / 'results' / 'tiny'
    root.mkdir(parents=True)
    (root / 'configuration').write_text('name tiny\nnumTimesteps 4\nnumPairTransformLayers 1\nnumStructureLayers 2\n')
    data = tmp_path / 'design'
    data.mkdir()
    shutil.copy(PROB, data / 'synth.pdb')
    out = tmp_path / 'out'
    args = S.build_parser().parse_args([
        '--name', 'tiny', '--epoch', '1', '--rootdir', str(tmp_path / 'results'), '--scale', '0.4', '--outdir', str(out),
        '--num_samples', '10', '--batch_size', '8', '--motif_name', 'synth', '--datadir', str(data), '--random_init', '0'])
    np.random.seed(0)
    S.main(args)
    pdbs = sorted(os.listdir(out / 'motif=synth' / 'pdbs'))
    assert pdbs == sorted('synth_%d.pdb' % i for i in range(10))
    spec = F.load_motif_spec(PROB)
    for name in pdbs:
        atoms = [l for l in open(out / 'motif=synth' / 'pdbs' / name) if l.startswith('ATOM')]
        assert spec['min_total_length'] <= len(atoms) <= spec['max_total_length']
        # the motif residues carry their group letter in columns 73-76 and the motif's residue names
        motif = [l for l in atoms if l[72] != ' ']
        assert len(motif) == 12 and {l[72] for l in motif} == {'A', 'B'}
        mlines = [l for l in open(out / 'motif=synth' / 'motif_pdbs' / name) if l.startswith('ATOM')]
        assert len(mlines) == 48                                                     # 12 residues x N, CA, C, O
        assert sorted({int(l[22:26]) for l in mlines}) == sorted(int(l[22:26]) for l in motif)
        coords = np.array([[float(l[30:38]), float(l[38:46]), float(l[46:54])] for l in atoms])
        assert np.isfinite(coords).all() and np.abs(coords.mean(0)).max() < 2e-3     # centred
