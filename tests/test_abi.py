"""CPU-side checks of the drop-in boundary: the library builds for sm_100a, loads without a GPU and
exports exactly the symbols include/genie2_b200.h declares; the product path refuses to run
without a CUDA device (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, 'include', 'genie2_b200.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(gb_[a-z0-9_]+)\s*\(', src)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as G
    G.build()
    from genie2_b200 import _lib
    lib = C.CDLL(_lib.LIB_PATH)
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(_lib.SYMBOLS) == names          # the ctypes table binds all of them, nothing else
    assert _lib.load().gb_version() == 200


def test_no_cpu_fallback():
    import genie2_b200
    from oracle import genie_oracle as O
    m = genie2_b200.Denoiser(**O.OracleConfig(n_pair_transform_layer=1, n_structure_layer=1).model_kwargs())
    feats = O.make_features([[8]])
    ts = genie2_b200.T(torch.eye(3).expand(1, 8, 3, 3).contiguous(), torch.zeros(1, 8, 3))
    with pytest.raises(RuntimeError):
        m(ts, torch.ones(1, dtype=torch.int32), feats)
    if not torch.cuda.is_available():
        from genie2_b200 import _lib
        h = C.c_void_p()
        d = _lib.gb_dims()
        assert _lib.load().gb_create(C.byref(h), C.byref(d), 0) != 0


def test_state_dict_schema_matches_reference_names():
    import genie2_b200
    from oracle import genie_oracle as O, weights as W
    cfg = O.OracleConfig()
    m = genie2_b200.Denoiser(**cfg.model_kwargs())
    sd = W.make_state_dict(cfg, 0, 'stock_like')
    assert list(m.state_dict().keys()) == list(sd.keys()) or set(m.state_dict().keys()) == set(sd.keys())
    assert len(sd) == 396
    for k, v in m.state_dict().items():
        assert tuple(v.shape) == tuple(sd[k].shape), k
