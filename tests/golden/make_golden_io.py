"""Golden vectors for the host-side data formats either side of the hot path, made by
running the UNMODIFIED reference functions from /root/reference (read-only import, nothing copied):

    python tests/golden/make_golden_io.py

Writes
  tests/golden/io_problem.pdb    a synthetic two-group motif-scaffolding problem file (REMARK 999 header in the
                                 reference's fixed columns, README.md "Format of Motif Scaffolding Problem Definition",
                                 + full-backbone ATOM lines of two motif segments on chains A and B)
  tests/golden/io_golden.npz     for several numpy seeds: the reference's sample_motif_mask / create_np_features_from_
                                 motif_pdb outputs, its pad/batchify/debatchify results, parse_pdb output, and the exact
                                 bytes of save_np_features_to_pdb and save_motif_pdb
"""
import io
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, '/root/reference')

from genie.utils import feat_utils as RF, motif_utils as RM  # noqa: E402  (the reference, imported where it lies)

RES3 = ['GLY', 'ALA', 'LEU', 'LYS', 'ASP', 'TRP', 'SER', 'GLU', 'VAL', 'HIS', 'PRO', 'THR']


def atom_line(serial, name, res3, chain, resi, xyz, elem):
    return 'ATOM  %5d  %-3s %3s %1s%4d    %8.3f%8.3f%8.3f  1.00  0.00           %1s  \n' % (
        serial, name, res3, chain, resi, xyz[0], xyz[1], xyz[2], elem)


def make_problem(path):
    """Two motif segments in different groups (chain A residues 3-9 -> group A, chain B residues 21-25 -> group B),
    three scaffold segments of variable length, total length 40-60."""
    rng = np.random.RandomState(7)
    out = []
    out.append('REMARK 999 NAME   synthetic_two_group\n')
    out.append('REMARK 999 PDB    XXXX\n')
    out.append('REMARK 999 INPUT  %1s%4d%4d\n' % (' ', 5, 20))
    out.append('REMARK 999 INPUT  %1s%4d%4d %1s\n' % ('A', 3, 9, 'A'))
    out.append('REMARK 999 INPUT  %1s%4d%4d\n' % (' ', 8, 25))
    out.append('REMARK 999 INPUT  %1s%4d%4d %1s\n' % ('B', 21, 25, 'B'))
    out.append('REMARK 999 INPUT  %1s%4d%4d\n' % (' ', 0, 15))
    out.append('REMARK 999 MINIMUM TOTAL LENGTH      %d\n' % 40)
    out.append('REMARK 999 MAXIMUM TOTAL LENGTH      %d\n' % 60)
    serial = 1
    for chain, lo, hi in (('A', 3, 9), ('B', 21, 25)):
        ca = rng.randn(3) * 4
        for resi in range(lo, hi + 1):
            ca = ca + rng.randn(3) * 2.2
            r3 = RES3[rng.randint(len(RES3))]
            for name, elem, off in (('N', 'N', -1.2), ('CA', 'C', 0.0), ('C', 'C', 1.3), ('O', 'O', 2.1)):
                out.append(atom_line(serial, name, r3, chain, resi, ca + off * np.array([0.6, 0.5, -0.4]) + 0.01 * rng.randn(3), elem))
                serial += 1
    out.append('END\n')
    with open(path, 'w') as fh:
        fh.write(''.join(out))


def main():
    prob = os.path.join(HERE, 'io_problem.pdb')
    make_problem(prob)
    g = {}
    spec = RM.load_motif_spec(prob)
    g['spec_name'] = np.array(spec['name'])
    g['spec_minmax'] = np.array([spec['min_total_length'], spec['max_total_length']])
    seqs, coords = RF.parse_pdb(prob)
    g['parse_n_chain'] = np.array(len(seqs))
    for i, (s, c) in enumerate(zip(seqs, coords)):
        g['parse_seq_%d' % i] = np.asarray(s)
        g['parse_xyz_%d' % i] = np.asarray(c)
    seeds = [0, 1, 2, 3, 11]
    g['seeds'] = np.array(seeds)
    feats = []
    with tempfile.TemporaryDirectory() as tmp:
        for sd in seeds:
            np.random.seed(sd)
            f = RF.create_np_features_from_motif_pdb(prob)
            feats.append(f)
            for k, v in f.items():
                g['feat_%d_%s' % (sd, k)] = np.asarray(v)
            # generated-structure PDB: motif coordinates in place, scaffold filled with seeded noise (exercises centring,
            # 3-dp rounding and python's str(float) column formatting incl. values like -6.15 and 1e-3)
            out = dict(f)
            rng = np.random.RandomState(100 + sd)
            pos = np.array(f['atom_positions'], dtype=np.float64)
            pos[~f['fixed_sequence_mask']] = rng.randn(int((~f['fixed_sequence_mask']).sum()), 3) * 9
            pos[0] = [1e-3, -6.15, 12.5]
            out['atom_positions'] = pos.astype(np.float32) if sd % 2 else pos
            p1 = os.path.join(tmp, 'o.pdb')
            RF.save_np_features_to_pdb(out, p1)
            g['pdb_%d_positions' % sd] = out['atom_positions']
            g['pdb_%d_bytes' % sd] = np.frombuffer(open(p1, 'rb').read(), dtype=np.uint8)
            p2 = os.path.join(tmp, 'm.pdb')
            RM.save_motif_pdb(prob, f['fixed_sequence_mask'], p2)
            g['motifpdb_%d_bytes' % sd] = np.frombuffer(open(p2, 'rb').read(), dtype=np.uint8)
    # unconditional features, padding and batching
    e = RF.create_empty_np_features([13, 8])
    for k, v in e.items():
        g['empty_%s' % k] = np.asarray(v)
    max_n_res = 64
    cp = lambda f: {k: np.array(v, copy=True) for k, v in f.items()}     # the reference pads in place
    for k, v in RF.pad_np_features(cp(feats[0]), 1, max_n_res).items():
        g['pad0_%s' % k] = np.asarray(v)
    batch = RF.batchify_np_features([cp(f) for f in feats[:3]])
    for k, v in batch.items():
        g['batch_%s' % k] = np.asarray(v)
    un = RF.debatchify_np_features(batch)
    g['debatch_n'] = np.array(len(un))
    for i, f in enumerate(un):
        for k, v in f.items():
            g['debatch_%d_%s' % (i, k)] = np.asarray(v)
    np.savez_compressed(os.path.join(HERE, 'io_golden.npz'), **g)
    print('wrote', len(g), 'arrays')


if __name__ == '__main__':
    main()
