"""Feature dictionaries at the sampler boundary (host side, once per batch).

Mirrors the data model of the reference (genie/utils/feat_utils.py:17-65,95-130,136-186,
192-340,377-416 and genie/utils/motif_utils.py:4-190): same 12 keys, dtypes, zero padding,
PDB column layout and the python ``str(float)`` coordinate formatting quirk (``-6.15``,
not ``-6.150``).  Everything here is numpy on the host; nothing on the per-step path.
"""
import gzip

import numpy as np
import torch

RESTYPE_1_TO_3 = {
    'A': 'ALA', 'R': 'ARG', 'N': 'ASN', 'D': 'ASP', 'C': 'CYS', 'Q': 'GLN', 'E': 'GLU', 'G': 'GLY', 'H': 'HIS',
    'I': 'ILE', 'L': 'LEU', 'K': 'LYS', 'M': 'MET', 'F': 'PHE', 'P': 'PRO', 'S': 'SER', 'T': 'THR', 'W': 'TRP',
    'Y': 'TYR', 'V': 'VAL',
}
RESTYPE_3_TO_1 = {v: k for k, v in RESTYPE_1_TO_3.items()}
RESTYPES = list(RESTYPE_1_TO_3.keys())
RESTYPE_ORDER = {r: i for i, r in enumerate(RESTYPES)}

_PER_RESIDUE = ('aatype', 'atom_positions', 'residue_mask', 'residue_index', 'chain_index', 'fixed_sequence_mask',
                'fixed_group', 'interface_mask')
_INT_KEYS = ('num_chains', 'num_residues', 'num_residues_per_chain', 'aatype', 'residue_mask', 'residue_index',
             'chain_index', 'fixed_group')
_BOOL_KEYS = ('fixed_sequence_mask', 'fixed_structure_mask', 'interface_mask')


def create_empty_np_features(lengths):
    """Unconditional features for one structure with the given chain lengths (feat_utils.py:17-65)."""
    n = int(np.sum(lengths))
    return {
        'aatype': np.zeros((n, len(RESTYPES)), dtype=int),
        'num_chains': np.array(len(lengths)).astype(int),
        'num_residues': np.array(n).astype(int),
        'num_residues_per_chain': np.array(lengths).astype(int),
        'atom_positions': np.zeros((n, 3), dtype=float),
        'residue_mask': np.ones(n, dtype=int),
        'residue_index': np.concatenate([np.arange(l) for l in lengths]).astype(int),
        'chain_index': np.concatenate([[i] * l for i, l in enumerate(lengths)]).astype(int),
        'fixed_sequence_mask': np.zeros(n, dtype=bool),
        'fixed_structure_mask': np.zeros((n, n), dtype=bool),
        'fixed_group': np.zeros(n, dtype=int),
        'interface_mask': np.zeros(n, dtype=bool),
    }


# ---------------------------------------------------------------- motif problems
def load_motif_spec(filepath):
    """REMARK 999 header of a motif-scaffolding problem file (motif_utils.py:4-57)."""
    structures, name, lo, hi = [], None, None, None
    with open(filepath) as fh:
        for line in fh:
            if line.startswith('REMARK 999 INPUT'):
                if line[18] == ' ':
                    structures.append({'type': 'scaffold', 'min_length': int(line[19:23]), 'max_length': int(line[23:27])})
                else:
                    structures.append({'type': 'motif', 'chain': line[18], 'start_index': int(line[19:23]),
                                       'end_index': int(line[23:27]),
                                       'group': line[28] if len(line) > 28 and line[28] != ' ' else 'A'})
            elif line.startswith('REMARK 999 NAME'):
                name = line[18:]
            elif line.startswith('REMARK 999 MINIMUM TOTAL LENGTH'):
                lo = int(line[37:])
            elif line.startswith('REMARK 999 MAXIMUM TOTAL LENGTH'):
                hi = int(line[37:])
    return {'name': name, 'structures': structures, 'min_total_length': lo, 'max_total_length': hi}


def sample_motif_mask(spec):
    """Draw scaffold segment lengths with ``np.random`` until the total length is admissible,
    then build sequence / structure / group masks (motif_utils.py:59-129)."""
    while True:
        seq, grp = [], []
        for st in spec['structures']:
            if st['type'] == 'scaffold':
                n = np.random.randint(st['min_length'], st['max_length'] + 1)
                seq += [0] * n
                grp += [0] * n
            else:
                n = st['end_index'] - st['start_index'] + 1
                seq += [1] * n
                grp += [ord(st['group']) - ord('A') + 1] * n
        if spec['min_total_length'] <= len(seq) <= spec['max_total_length']:
            break
    grp = np.array(grp).astype(int)
    structure = np.zeros((len(seq), len(seq)))
    for gi in range(1, 1 + int(np.max(grp))):
        m = grp == gi
        structure += m[:, None] * m[None, :]
    return {'sequence': np.array(seq).astype(bool), 'structure': structure.astype(bool), 'group': grp}


def parse_pdb(filepath):
    """Per-chain residue types and C-alpha coordinates (feat_utils.py:377-416)."""
    opener = (lambda p: gzip.open(p, 'rt')) if filepath.endswith('.gz') else (lambda p: open(p, 'r'))
    seqs, coords, cur = [], [], None
    with opener(filepath) as fh:
        for line in fh:
            if line.startswith('ATOM') and line[13:15].strip() == 'CA':
                chain = line[21]
                if cur is None or chain != cur:
                    seqs.append([])
                    coords.append([])
                    cur = chain
                seqs[-1].append(RESTYPE_ORDER[RESTYPE_3_TO_1[line[17:20]]])
                coords[-1].append([float(line[30:38]), float(line[38:46]), float(line[46:54])])
    return seqs, coords


def create_np_features_from_motif_pdb(filepath):
    """One sampled placement of a motif problem (feat_utils.py:95-130)."""
    spec = load_motif_spec(filepath)
    seqs, coords = parse_pdb(filepath)
    aatype = np.eye(len(RESTYPES))[np.concatenate(seqs)]
    pos = np.concatenate(coords)
    mask = sample_motif_mask(spec)
    f = create_empty_np_features([len(mask['sequence'])])
    f['aatype'][mask['sequence']] = aatype
    f['atom_positions'][mask['sequence']] = pos
    f['fixed_sequence_mask'] = mask['sequence']
    f['fixed_structure_mask'] = mask['structure']
    f['fixed_group'] = mask['group']
    return f


# ---------------------------------------------------------------- batching
def pad_np_features(f, max_n_chain, max_n_res):
    """Zero-pad one feature dict to (max_n_chain, max_n_res) (feat_utils.py:192-231)."""
    n_chain, n_res = int(f['num_chains']), int(f['num_residues'])
    for key in f:
        v = f[key]
        if key == 'num_residues_per_chain':
            f[key] = np.concatenate([v, np.zeros(max_n_chain - n_chain, dtype=v.dtype)])
        elif key == 'fixed_structure_mask':
            f[key] = np.pad(v, [(0, max_n_res - n_res)] * 2, 'constant', constant_values=0).astype(v.dtype)
        elif not key.startswith('num'):
            f[key] = np.concatenate([v, np.zeros((max_n_res - n_res, *v.shape[1:]), dtype=v.dtype)])
    return f


def batchify_np_features(list_np_features):
    """Pad to the batch maxima and stack (feat_utils.py:233-268)."""
    max_c = int(np.max([f['num_chains'] for f in list_np_features]))
    max_r = int(np.max([f['num_residues'] for f in list_np_features]))
    padded = [pad_np_features(f, max_c, max_r) for f in list_np_features]
    return {k: np.concatenate([np.expand_dims(f[k], 0) for f in padded]) for k in list_np_features[0].keys()}


def debatchify_np_features(np_features):
    """Inverse of batchify: strip the padding per sample (feat_utils.py:270-302)."""
    out = []
    for i in range(np_features['aatype'].shape[0]):
        nc, nr = int(np_features['num_chains'][i]), int(np_features['num_residues'][i])
        d = {'num_chains': np_features['num_chains'][i], 'num_residues': np_features['num_residues'][i],
             'num_residues_per_chain': np_features['num_residues_per_chain'][i, :nc],
             'fixed_structure_mask': np_features['fixed_structure_mask'][i, :nr, :nr]}
        for k in _PER_RESIDUE:
            d[k] = np_features[k][i, :nr]
        out.append(d)
    return out


def convert_np_features_to_tensor(features, device):
    """numpy -> torch on ``device`` with the reference dtypes (feat_utils.py:304-321)."""
    out = {}
    for k, v in features.items():
        t = torch.as_tensor(np.asarray(v))
        if k in _INT_KEYS:
            t = t.to(torch.int32)
        elif k in _BOOL_KEYS:
            t = t.to(torch.bool)
        else:
            t = t.to(torch.float32)
        out[k] = t.to(device)
    return out


def convert_tensor_features_to_numpy(features):
    """torch -> numpy (int / bool / float64) (feat_utils.py:323-340)."""
    out = {}
    for k, v in features.items():
        a = v.detach().cpu().numpy()
        out[k] = a.astype(int) if k in _INT_KEYS else a.astype(bool) if k in _BOOL_KEYS else a.astype(float)
    return out


# ---------------------------------------------------------------- PDB output
def _put(line, index, sub):
    return line[:index] + sub + line[index + len(sub):]


def format_pdb_lines(np_features):
    """C-alpha ATOM records, centred, 3 dp, fixed columns (feat_utils.py:136-186)."""
    coords = np_features['atom_positions']
    coords = np.around(coords - np.mean(coords, axis=0, keepdims=True), decimals=3)
    lines = []
    for i in range(coords.shape[0]):
        grp = int(np_features['fixed_group'][i])
        line = ' ' * 80
        line = _put(line, 0, 'ATOM')
        line = _put(line, 6, str(i + 1).rjust(5))
        line = _put(line, 13, 'CA')
        line = _put(line, 17, RESTYPE_1_TO_3[RESTYPES[int(np.argmax(np_features['aatype'][i]))]])
        line = _put(line, 21, chr(ord('A') + int(np_features['chain_index'][i])))
        line = _put(line, 22, str(int(np_features['residue_index'][i]) + 1).rjust(4))
        for col, v in zip((30, 38, 46), coords[i]):
            line = _put(line, col, str(v).rjust(8))
        line = _put(line, 72, (' ' if grp == 0 else chr(grp - 1 + ord('A'))).ljust(4))
        line = _put(line, 77, 'C')
        lines.append(line)
    return lines


def save_np_features_to_pdb(np_features, filepath):
    with open(filepath, 'w') as fh:
        for line in format_pdb_lines(np_features):
            fh.write(line + '\n')


def save_motif_pdb(spec_filepath, mask, pdb_filepath):
    """Motif ATOM records of the problem file renumbered to their sampled positions
    (motif_utils.py:131-190)."""
    spec = load_motif_spec(spec_filepath)
    spec_idx = [(st['chain'], i, st['group']) for st in spec['structures'] if st['type'] == 'motif'
                for i in range(st['start_index'], st['end_index'] + 1)]
    pdb_idx = [i + 1 for i, m in enumerate(mask) if m]
    assert len(pdb_idx) == len(spec_idx)
    remap = {'{}_{}'.format(c, i): (pdb_idx[k], g) for k, (c, i, g) in enumerate(spec_idx)}
    with open(spec_filepath) as fh:
        atoms = [l for l in fh if l.startswith('ATOM')]
    out = []
    for line in atoms:
        new_idx, grp = remap['{}_{}'.format(line[21], int(line[22:26]))]
        out.append(line[:21] + 'A' + str(new_idx).rjust(4) + line[26:72] + grp.ljust(4) + line[76:])
    with open(pdb_filepath, 'w') as fh:
        fh.write(''.join(out))
