"""ctypes binding of libgomoku_b200.so (C ABI declared in include/gomoku_b200.h).

There is no CPU fallback: if the CUDA library has not been built, importing this module
raises; if no GPU is visible, creating an engine raises.
"""
import ctypes as C
import os
import subprocess

_PKG = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_PKG)
# GZ_LIB_PATH: load another build of the same library (A/B runs of compile-time variants); never a fallback
LIB_PATH = os.environ.get("GZ_LIB_PATH") or os.path.join(_PKG, "lib", "libgomoku_b200.so")
CSRC = os.path.join(_PKG, "csrc")
HEADER = os.path.join(_ROOT, "include", "gomoku_b200.h")

GZ_OK, GZ_E_INVALID, GZ_E_ILLEGAL, GZ_E_CUDA, GZ_E_NOWEIGHTS, GZ_E_POOL, GZ_E_NOGPU = 0, -1, -2, -3, -4, -5, -6
LAYOUT_TORCH, LAYOUT_THEANO = 0, 1

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC)
                  if f.endswith((".cu", ".cuh", ".h"))) + [HEADER]


def build(force=False, verbose=False):
    """Compile the CUDA library for sm_100a with nvcc (works without a GPU)."""
    srcs = sources()
    if (not force and os.path.exists(LIB_PATH)
            and all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(s) for s in srcs)):
        return LIB_PATH
    os.makedirs(os.path.dirname(LIB_PATH), exist_ok=True)
    nvcc = os.environ.get("NVCC", "nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
        ["-o", LIB_PATH, os.path.join(CSRC, "gz_api.cu")]
    subprocess.check_call(cmd)
    return LIB_PATH


class GzConfig(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("n_in_row", C.c_int32),
                ("n_games", C.c_int32), ("n_playout", C.c_int32), ("nodes_per_game", C.c_int32),
                ("c_puct", C.c_float), ("dirichlet_alpha", C.c_float), ("dirichlet_eps", C.c_float),
                ("trunk", C.c_int32 * 3), ("seed", C.c_uint64), ("game_id_base", C.c_uint64),
                ("device", C.c_int32), ("net_impl", C.c_int32), ("stream", C.c_void_p)]


class GzSelfplayStats(C.Structure):
    _fields_ = [("n_games", C.c_int64), ("n_samples", C.c_int64), ("n_playouts", C.c_int64),
                ("n_leaf_evals", C.c_int64), ("n_plies", C.c_int64), ("pool_overflows", C.c_int64),
                ("n_wins", C.c_int64 * 3), ("max_tree_nodes", C.c_int64)]


class GzCounters(C.Structure):
    _fields_ = [("kernel_launches", C.c_int64), ("select_levels", C.c_int64),
                ("select_children", C.c_int64), ("playouts", C.c_int64), ("leaf_evals", C.c_int64),
                ("expanded_children", C.c_int64), ("rollout_plies", C.c_int64)]


_P = C.c_void_p
_PROTOS = {
    "gz_create": (C.c_int, [C.POINTER(GzConfig), C.POINTER(_P)]),
    "gz_destroy": (None, [_P]),
    "gz_last_error": (C.c_char_p, [_P]),
    "gz_sync": (C.c_int, [_P]),
    "gz_device_count": (C.c_int, []),
    "gz_version": (C.c_char_p, []),
    "gz_set_weights": (C.c_int, [_P, C.POINTER(_P), C.c_int32]),
    "gz_get_weights": (C.c_int, [_P, C.POINTER(_P)]),
    "gz_rules_eval": (C.c_int, [_P, _P, _P, C.c_int64, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "gz_boards_reset": (C.c_int, [_P, _P]),
    "gz_boards_set": (C.c_int, [_P, C.c_int32, _P, C.c_int32, C.c_int32]),
    "gz_boards_do_move": (C.c_int, [_P, _P, _P]),
    "gz_boards_status": (C.c_int, [_P, _P, _P, _P, _P, _P]),
    "gz_boards_legal_mask": (C.c_int, [_P, _P]),
    "gz_boards_planes": (C.c_int, [_P, _P]),
    "gz_boards_cells": (C.c_int, [_P, _P]),
    "gz_net_forward": (C.c_int, [_P, _P, C.c_int64, _P, _P]),
    "gz_net_forward_noalloc": (C.c_int, [_P, _P, C.c_int64, _P, _P]),
    "gz_debug_hash_evaluator": (C.c_int, [_P, C.c_int32, C.c_uint64]),
    "gz_debug_net_layers": (C.c_int, [_P, _P, C.c_int64, _P, _P, _P, _P, _P, _P]),
    "gz_search": (C.c_int, [_P, C.c_int32]),
    "gz_search_vl": (C.c_int, [_P, C.c_int32, C.c_int32, C.POINTER(C.c_int32)]),
    "gz_search_select": (C.c_int, [_P, _P, _P, _P, _P]),
    "gz_search_expand_backup": (C.c_int, [_P, _P, _P]),
    "gz_search_paths": (C.c_int, [_P, _P, _P]),
    "gz_root_visits": (C.c_int, [_P, _P, _P, _P]),
    "gz_update_with_move": (C.c_int, [_P, _P]),
    "gz_set_active": (C.c_int, [_P, _P]),
    "gz_equi_data": (C.c_int, [_P, _P, _P, _P, C.c_int64, _P, _P, _P]),
    "gz_train_step": (C.c_int, [_P, _P, _P, _P, C.c_int64, C.c_float, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "gz_train_backward": (C.c_int, [_P, _P, _P, _P, C.c_int64, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "gz_train_grads": (C.c_int, [_P, _P, C.c_int64, C.POINTER(C.c_int64)]),
    "gz_train_apply": (C.c_int, [_P, _P, C.c_float]),
    "gz_train_forward": (C.c_int, [_P, _P, C.c_int64, _P, _P]),
    "gz_train_reset": (C.c_int, [_P]),
    "gz_root_select_move": (C.c_int, [_P, C.c_float, C.c_int32, _P, _P]),
    "gz_selfplay_begin": (C.c_int, [_P, C.c_float, C.c_int32]),
    "gz_selfplay_step": (C.c_int, [_P, C.c_int32]),
    "gz_selfplay_pending": (C.c_int, [_P, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "gz_selfplay_drain": (C.c_int, [_P, C.c_int64, _P, _P, _P, _P, _P, C.POINTER(C.c_int64)]),
    "gz_selfplay_drain_games": (C.c_int, [_P, C.c_int64, C.c_int64, _P, _P, _P, _P, C.POINTER(C.c_int64)]),
    "gz_selfplay_get_stats": (C.c_int, [_P, C.POINTER(GzSelfplayStats), C.c_int32]),
    "gz_selfplay_last": (C.c_int, [_P, _P, _P]),
    "gz_selfplay_state": (C.c_int, [_P, _P, C.POINTER(C.c_int64)]),
    "gz_pure_search": (C.c_int, [_P, C.c_int32, C.c_uint32]),
    "gz_pure_best_move": (C.c_int, [_P, _P]),
    "gz_rollouts": (C.c_int, [_P, C.c_int32, C.c_uint32, _P, C.POINTER(C.c_int64)]),
    "gz_get_counters": (C.c_int, [_P, C.POINTER(GzCounters), C.c_int32]),
    "gz_profile_enable": (C.c_int, [_P, C.c_int32]),
    "gz_profile_read": (C.c_int, [_P, _P, _P, C.c_int32]),
    "gz_profile_class_name": (C.c_char_p, [C.c_int32]),
}
PROF_CLASSES = 14

_lib = None


def lib():
    """Load the CUDA library; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "alphazero_gomoku_b200: %s is missing. Build it with "
                "`python -c 'import __graft_entry__ as g; g.build()'` (needs nvcc); "
                "there is no CPU fallback." % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in _PROTOS.items():
            fn = getattr(L, name)  # AttributeError if the ABI lost a symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def exported_symbols():
    return sorted(_PROTOS)


class GzError(Exception):
    def __init__(self, code, msg):
        super().__init__("gomoku_b200 error %d: %s" % (code, msg))
        self.code = code
        self.msg = msg


def check(handle, rc):
    """Map a C return code to the reference's exception types."""
    if rc == GZ_OK:
        return
    msg = lib().gz_last_error(handle)
    msg = msg.decode() if msg else ""
    if rc == GZ_E_ILLEGAL:
        raise ValueError("list.remove(x): x not in list (%s)" % (msg or "illegal move"))
    if rc == GZ_E_NOGPU:
        raise RuntimeError("gomoku_b200: %s" % msg)
    if rc == GZ_E_INVALID:
        raise Exception(msg or "invalid argument")
    raise GzError(rc, msg)
