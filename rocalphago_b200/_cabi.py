"""ctypes binding of include/rocalphago_b200.h -- the only door from Python into the CUDA code.

There is no CPU fallback: if the shared library is missing and cannot be built, or a call
fails, this raises.
"""
import ctypes
import os
import re

from . import build as _build

_lib = None


class RagError(RuntimeError):
    pass


def _declared_symbols():
    """Names of the functions include/rocalphago_b200.h declares (used by tests as well)."""
    hdr = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "include", "rocalphago_b200.h")
    text = open(hdr).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rag_[a-z0-9_]+)\s*\(", text)))


class RagSuperko(ctypes.Structure):
    """rag_superko (include/rocalphago_b200.h)"""
    _fields_ = [(name, ctypes.c_void_p) for name in ("d_zobrist", "d_hash", "d_prev", "d_prev_off", "d_masks", "d_flags")]


def lib():
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB_PATH
    if _build._stale():                 # missing, or built from other sources than the tree holds now (content hash)
        try:
            _build.build_library()
        except Exception as e:  # noqa: BLE001
            if not os.path.exists(path):
                raise RagError("librocalphago_b200.so is not built and nvcc could not build it (%s); "
                               "run `python -c 'import __graft_entry__ as g; g.build()'`" % e)
            import warnings
            warnings.warn("librocalphago_b200.so is older than csrc/ and could not be rebuilt (%s); using it as it is" % e)
    L = ctypes.CDLL(path)
    c = ctypes
    vp, i32, i64, u64, sz = c.c_void_p, c.c_int, c.c_longlong, c.c_uint64, c.c_size_t
    sigs = {
        "rag_last_error": (c.c_char_p, []),
        "rag_abi_version": (i32, []),
        "rag_pack_states": (i32, [i32, i32, vp, vp, vp, vp, vp]),
        "rag_unpack_states": (i32, [i32, i32, vp, vp, vp, vp, vp]),
        "rag_features_workspace_bytes": (sz, [i32]),
        "rag_features_analyze": (i32, [vp, i32, i32, i32, vp, sz, vp]),
        "rag_set_ladder_handover": (i32, [i32]),
        "rag_set_ladder_helpers": (i32, [i32]),
        "rag_set_ladder_watchdog": (i64, [i64]),
        "rag_features_status": (i32, [vp, vp]),
        "rag_features_ladder_cutoffs": (i32, [vp, vp]),
        "rag_features_ladder_longest": (i32, [vp, vp]),
        "rag_features_ladders": (i32, [vp, i32, i32, vp, sz, vp]),
        "rag_rollout_winners": (i32, [vp, i32, i32, i32, c.c_uint64, c.c_float, vp, vp]),
        "rag_features_analyze_superko": (i32, [vp, i32, i32, i32, vp, vp, sz, vp]),
        "rag_features_emit": (i32, [vp, i32, i32, vp, i32, i32, i32, i32, vp, vp]),
        "rag_features": (i32, [vp, i32, i32, vp, i32, i32, i32, i32, vp, vp, sz, vp]),
        "rag_features_static": (i32, [vp, i32, i32, vp, i32, i32, i32, i32, vp, vp]),
        "rag_features_static_max_positions": (i32, []),
        "rag_conv_rows": (i64, [i32, i32]),
        "rag_tree_create": (i32, [c.c_double, vp]),
        "rag_tree_destroy": (i32, [vp]),
        "rag_tree_reset": (i32, [vp]),
        "rag_tree_wave_begin": (i32, [vp, i32, c.c_double]),
        "rag_tree_wave_leaves": (i32, [vp, vp]),
        "rag_tree_wave_expand": (i32, [vp, i32, vp, vp, vp, vp]),
        "rag_tree_wave_step": (i32, [vp, vp]),
        "rag_tree_wave_backup": (i32, [vp, vp]),
        "rag_tree_advance": (i32, [vp, i32]),
        "rag_tree_root": (i32, [vp]),
        "rag_tree_size": (i64, [vp]),
        "rag_tree_node_children": (i32, [vp, i32]),
        "rag_tree_node_stats": (i32, [vp, i32, vp, vp, vp, vp]),
        "rag_tree_node_arrays": (i32, [vp, i32, vp, vp, vp, vp, vp, vp, vp]),
        "rag_select_moves": (i32, [vp, vp, i32, i32, c.c_float, i32, u64, vp, vp]),
        "rag_move_distributions": (i32, [vp, vp, i32, i32, i32, vp, vp, vp, vp]),
        "rag_do_move": (i32, [vp, i32, i32, vp, vp, vp]),
        "rag_random_playouts": (i32, [vp, i32, i32, u64, i32, i32, vp, i32, vp]),
        "rag_playout_quota": (i32, [i32, i32]),
        "rag_net_create": (i32, [vp, vp]),
        "rag_net_destroy": (i32, [vp]),
        "rag_net_num_weight_arrays": (i32, [vp]),
        "rag_net_load_weights": (i32, [vp, vp, i32]),
        "rag_net_set_input_planes": (i32, [vp, vp, i32]),
        "rag_net_workspace_bytes": (sz, [vp, i32]),
        "rag_net_logits_offset": (sz, [vp, i32]),
        "rag_set_conv_pair": (i32, [i32]),
        "rag_conv_stall_report": (i32, [c.c_char_p, i32]),
        "rag_net_set_profiling": (i32, [vp, i32]),
        "rag_net_layer_times": (i32, [vp, vp, i32]),
        "rag_net_input_rows": (i64, [vp, i32]),
        "rag_net_input_channels": (i32, [vp]),
        "rag_net_forward": (i32, [vp, vp, i32, vp, vp, sz, vp]),
        "rag_net_forward_rows": (i32, [vp, vp, i32, vp, vp, sz, vp]),
        "rag_eval_workspace_bytes": (sz, [vp, i32]),
        "rag_eval_batch": (i32, [vp, vp, i32, vp, i32, vp, vp, sz, vp]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(L, name)          # AttributeError here = the library is older than this binding
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def check(rc):
    if rc != 0:
        msg = lib().rag_last_error()
        raise RagError("rocalphago_b200 C ABI error %d: %s" % (rc, msg.decode() if msg else ""))
    return rc
