"""ctypes binding of the C ABI in ``include/genie2_b200.h``.

The library is the only compute path: if it cannot be loaded, or no B200 is present,
every entry point raises -- there is no CPU / PyTorch fallback.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, 'libgenie2_b200.so')

GB_PREC_FP32, GB_PREC_BF16X3, GB_PREC_BF16 = 0, 1, 2
GB_PREC_TF32 = GB_PREC_BF16X3          # old name of the slot (include/genie2_b200.h)
PRECISIONS = {'fp32': GB_PREC_FP32, 'bf16x3': GB_PREC_BF16X3, 'tf32': GB_PREC_BF16X3, 'bf16': GB_PREC_BF16}


class gb_dims(C.Structure):
    _fields_ = [
        ('c_s', C.c_int32), ('c_p', C.c_int32), ('n_timestep', C.c_int32), ('rescale', C.c_float),
        ('c_pos_emb', C.c_int32), ('c_chain_emb', C.c_int32), ('c_timestep_emb', C.c_int32),
        ('max_n_res', C.c_int32), ('max_n_chain', C.c_int32), ('relpos_k', C.c_int32),
        ('template_dist_min', C.c_float), ('template_dist_step', C.c_float), ('template_dist_n_bin', C.c_int32),
        ('n_pair_transform_layer', C.c_int32), ('include_mul_update', C.c_int32), ('include_tri_att', C.c_int32),
        ('c_hidden_mul', C.c_int32), ('c_hidden_tri_att', C.c_int32), ('n_head_tri', C.c_int32),
        ('pair_transition_n', C.c_int32), ('n_structure_layer', C.c_int32), ('n_structure_block', C.c_int32),
        ('c_hidden_ipa', C.c_int32), ('n_head_ipa', C.c_int32), ('n_qk_point', C.c_int32), ('n_v_point', C.c_int32),
        ('n_structure_transition_layer', C.c_int32),
    ]


class gb_tensor(C.Structure):
    _fields_ = [('name', C.c_char_p), ('data', C.c_void_p), ('ndim', C.c_int32), ('shape', C.c_int64 * 4)]


class gb_features(C.Structure):
    _fields_ = [
        ('B', C.c_int32), ('N', C.c_int32),
        ('residue_index', C.c_void_p), ('chain_index', C.c_void_p), ('residue_mask', C.c_void_p),
        ('fixed_sequence_mask', C.c_void_p), ('fixed_structure_mask', C.c_void_p), ('interface_mask', C.c_void_p),
        ('aatype', C.c_void_p), ('atom_positions', C.c_void_p),
    ]


class gb_aux(C.Structure):
    _fields_ = [('s', C.c_void_p), ('p', C.c_void_p), ('states', C.c_void_p), ('out_rots', C.c_void_p),
                ('out_trans', C.c_void_p), ('quat_follow', C.c_void_p)]


# every symbol include/genie2_b200.h declares: name -> (restype, argtypes)
_P = C.c_void_p
SYMBOLS = {
    'gb_version': (C.c_int, []),
    'gb_last_error': (C.c_char_p, [_P]),
    'gb_create': (C.c_int, [C.POINTER(_P), C.POINTER(gb_dims), C.c_int]),
    'gb_destroy': (None, [_P]),
    'gb_load_weights': (C.c_int, [_P, C.POINTER(gb_tensor), C.c_int]),
    'gb_set_schedule': (C.c_int, [_P, _P, _P, _P, _P]),
    'gb_set_precision': (C.c_int, [_P, C.c_int]),
    'gb_set_features': (C.c_int, [_P, C.POINTER(gb_features)]),
    'gb_frenet': (C.c_int, [_P, _P, _P, _P]),
    'gb_denoise': (C.c_int, [_P, _P, _P, _P, _P, C.POINTER(gb_aux), _P]),
    'gb_p_sample': (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_float, _P, C.c_uint64, _P]),
    'gb_sample_loop': (C.c_int, [_P, _P, _P, C.c_int, C.c_int, C.c_float, _P, C.c_uint64, _P]),
    'gb_sample_host': (C.c_int, [_P, _P, _P, C.c_int, C.c_int, C.c_float, _P, C.c_uint64]),
    'gb_launch_count': (C.c_int64, [_P]),
    'gb_workspace_bytes': (C.c_int64, [_P]),
    'gb_k_frenet': (C.c_int, [_P, _P, _P, C.c_int, C.c_int, _P, _P]),
    'gb_k_trimul': (C.c_int, [_P, C.c_int, C.c_int, _P, _P]),
    'gb_k_trimul_stage': (C.c_int, [_P, C.c_int, C.c_int, C.c_int, _P, _P]),
    'gb_k_pair_transition': (C.c_int, [_P, C.c_int, _P, _P]),
    'gb_k_tri_attention': (C.c_int, [_P, C.c_int, C.c_int, _P, _P]),
    'gb_k_ipa': (C.c_int, [_P, C.c_int, _P, _P, _P, _P, _P, _P]),
    'gb_k_ipa_static': (C.c_int, [_P, C.c_int, _P, _P, _P, _P, _P, _P]),
    'gb_dbg_tc_copy': (C.c_int, [_P, C.c_int, _P, C.c_int64]),
    'gb_shard_unique_id': (C.c_int, [_P]),
    'gb_shard_init': (C.c_int, [_P, C.c_int, C.c_int, _P]),
    'gb_shard_info': (C.c_int, [_P, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    'gb_collective_count': (C.c_int64, [_P]),
    'gb_peer_barrier_count': (C.c_int64, [_P]),
}

_lib = None


def load():
    """dlopen the in-tree library and bind every declared symbol.  Raises if missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            'genie2_b200: %s not found -- build it with `python -m genie2_b200.build` '
            '(there is no CPU or PyTorch fallback for this path)' % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)          # AttributeError if the .so does not export it
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc, ctx=None, what=''):
    if rc != 0:
        msg = load().gb_last_error(ctx)
        raise RuntimeError('genie2_b200 %s failed (%d): %s' % (what, rc, msg.decode() if msg else '?'))
