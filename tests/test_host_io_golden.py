"""Host-side data formats either side of the hot path against golden vectors produced by
the unmodified reference (tests/golden/make_golden_io.py): motif problem parsing, seeded motif placement, feature
dictionaries, pad / batchify / debatchify, and the byte-exact PDB writers.  CPU only."""
import os

import numpy as np
import pytest

from genie2_b200 import feat_utils as F

HERE = os.path.dirname(os.path.abspath(__file__))
PROB = os.path.join(HERE, 'golden', 'io_problem.pdb')


@pytest.fixture(scope='module')
def gold():
    return np.load(os.path.join(HERE, 'golden', 'io_golden.npz'))


def _same(a, b, key):
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape, (key, a.shape, b.shape)
    assert a.dtype.kind == b.dtype.kind, (key, a.dtype, b.dtype)
    assert np.array_equal(a, b), key


def test_problem_header_and_parse(gold):
    spec = F.load_motif_spec(PROB)
    assert spec['name'] == str(gold['spec_name'])
    assert [spec['min_total_length'], spec['max_total_length']] == gold['spec_minmax'].tolist()
    seqs, coords = F.parse_pdb(PROB)
    assert len(seqs) == int(gold['parse_n_chain'])
    for i in range(len(seqs)):
        _same(seqs[i], gold['parse_seq_%d' % i], 'seq')
        _same(coords[i], gold['parse_xyz_%d' % i], 'xyz')


def test_motif_features_follow_the_numpy_seed(gold):
    for sd in gold['seeds'].tolist():
        np.random.seed(sd)
        f = F.create_np_features_from_motif_pdb(PROB)
        keys = [k[len('feat_%d_' % sd):] for k in gold.files if k.startswith('feat_%d_' % sd)]
        assert sorted(keys) == sorted(f.keys())
        for k in keys:
            _same(f[k], gold['feat_%d_%s' % (sd, k)], k)


def test_empty_pad_batchify_debatchify(gold):
    e = F.create_empty_np_features([13, 8])
    for k in e:
        _same(e[k], gold['empty_%s' % k], k)
    feats = []
    for sd in gold['seeds'].tolist()[:3]:
        np.random.seed(sd)
        feats.append(F.create_np_features_from_motif_pdb(PROB))
    cp = lambda f: {k: np.array(v, copy=True) for k, v in f.items()}
    pad0 = F.pad_np_features(cp(feats[0]), 1, 64)
    for k in pad0:
        _same(pad0[k], gold['pad0_%s' % k], k)
    batch = F.batchify_np_features([cp(f) for f in feats])
    for k in batch:
        _same(batch[k], gold['batch_%s' % k], k)
    un = F.debatchify_np_features(batch)
    assert len(un) == int(gold['debatch_n'])
    for i, f in enumerate(un):
        for k in f:
            _same(f[k], gold['debatch_%d_%s' % (i, k)], k)


def test_pdb_writers_are_byte_exact(gold, tmp_path):
    for sd in gold['seeds'].tolist():
        f = {k[len('feat_%d_' % sd):]: gold[k] for k in gold.files if k.startswith('feat_%d_' % sd)}
        f['atom_positions'] = gold['pdb_%d_positions' % sd]
        out = str(tmp_path / 'o.pdb')
        F.save_np_features_to_pdb(f, out)
        assert open(out, 'rb').read() == gold['pdb_%d_bytes' % sd].tobytes(), 'sample pdb, seed %d' % sd
        out = str(tmp_path / 'm.pdb')
        F.save_motif_pdb(PROB, f['fixed_sequence_mask'], out)
        assert open(out, 'rb').read() == gold['motifpdb_%d_bytes' % sd].tobytes(), 'motif pdb, seed %d' % sd


def test_parse_reads_back_what_the_writer_wrote(gold, tmp_path):
    sd = int(gold['seeds'][0])
    f = {k[len('feat_%d_' % sd):]: gold[k] for k in gold.files if k.startswith('feat_%d_' % sd)}
    f['atom_positions'] = gold['pdb_%d_positions' % sd]
    out = str(tmp_path / 'o.pdb')
    F.save_np_features_to_pdb(f, out)
    seqs, coords = F.parse_pdb(out)
    pos = np.asarray(f['atom_positions'], dtype=np.float64)
    pos = np.around(pos - pos.mean(0, keepdims=True), 3)
    assert np.allclose(np.concatenate(coords), pos, atol=5e-4)
    assert np.array_equal(np.concatenate(seqs), np.argmax(f['aatype'], -1))
