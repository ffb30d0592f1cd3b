"""ctypes binding of liblazac_b200.so -- the C ABI declared in include/lazac_b200.h.

There is no CPU fallback: a missing library or a missing GPU raises.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liblazac_b200.so")

STATUS = {0: "LZG_OK", 1: "LZG_ERR_ARG", 2: "LZG_ERR_NOT_BINARY", 3: "LZG_ERR_RANGE",
          4: "LZG_ERR_CUDA", 5: "LZG_ERR_NO_DEVICE", 6: "LZG_ERR_STATE", 7: "LZG_ERR_NOT_A_MOVE"}

# every symbol include/lazac_b200.h declares: (name, restype, argtypes)
_vp, _i, _i32, _i64p = C.c_void_p, C.c_int, C.c_int32, C.POINTER(C.c_int64)
_i64 = C.c_int64
_i32p = C.POINTER(C.c_int32)
_vpp = C.POINTER(C.c_void_p)
SYMBOLS = [
    ("lzg_init", _i, [_i, _vpp]),
    ("lzg_shutdown", _i, [_vp]),
    ("lzg_last_error", C.c_char_p, [_vp]),
    ("lzg_status_string", C.c_char_p, [_i]),
    ("lzg_version", C.c_char_p, []),
    ("lzg_launch_count", _i, [_vp, _i64p]),
    ("lzg_timer_start", _i, [_vp]),
    ("lzg_timer_stop", _i, [_vp, C.POINTER(C.c_float)]),
    ("lzg_sync", _i, [_vp]),
    ("lzg_stream", _i, [_vp, _vpp]),
    ("lzg_flush_l2", _i, [_vp]),
    ("lzg_host_alloc", _i, [_vp, C.c_uint64, _vpp]),
    ("lzg_host_free", _i, [_vp, _vp]),
    ("lzg_profiles_upload", _i, [_vp, _vp, _i32, _i32, _vpp]),
    ("lzg_profiles_upload_i8", _i, [_vp, _vp, _i32, _i32, _vpp]),
    ("lzg_profiles_free", _i, [_vp]),
    ("lzg_profiles_shape", _i, [_vp, _i32p, _i32p]),
    ("lzg_profiles_distance_path", _i, [_vp, _vp]),
    ("lzg_distance_tc_layout", C.c_int64, [_vp, _vp, _i32, _vp, C.c_int64, _vp]),
    ("lzg_distance_matrix", _i, [_vp, _vp, _vp]),
    ("lzg_distance_matrix_device_async", _i, [_vp, _vp, _vp]),
    ("lzg_distance_matrix_noalloc", _i, [_vp, _vp, _vp]),
    ("lzg_distance_matrix_noalloc_max_bytes", _i, []),
    ("lzg_distance_shard", _i, [_vp, _vp, _i32, _i32, _vp, _i64p]),
    ("lzg_distance_shard_device_async", _i, [_vp, _vp, _i32, _i32, _vp, _i64p]),
    ("lzg_multi_init", _i, [_vp, _i32, _vpp]),
    ("lzg_multi_shutdown", _i, [_vp]),
    ("lzg_multi_size", _i, [_vp]),
    ("lzg_multi_context", _i, [_vp, _i32, _vpp]),
    ("lzg_multi_last_error", C.c_char_p, [_vp]),
    ("lzg_distance_matrix_multi", _i, [_vp, _vp, _i32, _i32, _vp, _i64p]),
    ("lzg_distance_matrix_multi_i8", _i, [_vp, _vp, _i32, _i32, _vp, _i64p]),
    ("lzg_distance_plan", _i, [_i32, _i32, _i32, _i32, _i32, _i32, _i32, _vp, _i64, _vp, _i32, _vp]),
    ("lzg_tree_create", _i, [_vp, _vp, _i32, _vp, _vp, _vp, _vpp]),
    ("lzg_tree_clone", _i, [_vp, _vp, _vpp]),
    ("lzg_tree_free", _i, [_vp]),
    ("lzg_tree_set_topology", _i, [_vp, _vp, _vp, _vp]),
    ("lzg_tree_get_topology", _i, [_vp, _vp, _vp]),
    ("lzg_tree_score", _i, [_vp, _vp, _i32p]),
    ("lzg_nni_best", _i, [_vp, _vp, _i32, _vp, _i32, _i32p, _i32p, _vp]),
    ("lzg_nni_scores_launch", _i, [_vp, _vp, _i32, _vp, _vpp, _i32p]),
    ("lzg_nni_scores_finish", _i, [_vp, _vp, _i32, _i32p, _i32p, _vp]),
    ("lzg_comm_init", _i, [_vp, _i32, _i32, _vp, _vpp]),
    ("lzg_comm_free", _i, [_vp]),
    ("lzg_nni_best_sharded", _i, [_vp, _vp, _vp, _i32, _vp, _i32, _i32p, _i32p, _vp]),
    ("lzg_tree_apply_nni", _i, [_vp, _vp, _i32, _i32, _i32, _i32, _i32p]),
    ("lzg_hill_climb", _i, [_vp, _vp, _i32, _vp, _i32, _i32, _i32, _vp, _vp]),
    ("lzg_hill_climb_sharded", _i, [_vp, _vp, _vp, _i32, _vp, _i32, _i32, _i32, _vp, _vp]),
    ("lzg_tree_last_climb_stamps", _i, [_vp, _i32, _vp, _i32p]),
    ("lzg_tree_last_climb_pass0_ms", _i, [_vp, _vp]),
    ("lzg_tree_get_intervals", _i, [_vp, _vp, _vp, _vp, _vp]),
    ("lzg_tree_get_node_intervals", _i, [_vp, _vp, _i32, _vp, _vp, _vp]),
    ("lzg_tree_ancestral", _i, [_vp, _vp, _vp, _vp]),
    ("lzg_tree_last_candidate_score", _i, [_vp, _i32p]),
    ("lzg_tree_last_nni_work", _i, [_vp, _i64p, _i64p]),
    ("lzg_tree_last_nni_traffic", _i, [_vp, _i64p, _i64p]),
    ("lzg_nj_seed_tree", _i, [_vp, _vp, _i32, _vp, _vp, _vp]),
    ("lzg_nj_seed_tree_device", _i, [_vp, _vp, _i32, _vp, _vp, _vp]),
    ("lzg_nj_seed_tree_profiles", _i, [_vp, _vp, _vp, _vp, _vp]),
]


class ClimbResult(C.Structure):
    """lzg_climb_result (include/lazac_b200.h)."""
    _fields_ = [("n_moves", C.c_int32), ("n_passes", C.c_int32), ("reason", C.c_int32), ("true_score", C.c_int32),
                ("returned_score", C.c_int32), ("candidates", C.c_int64), ("item_units", C.c_int64), ("up_units", C.c_int64),
                ("units_per_full_pass", C.c_int64)]


class LazacGpuError(RuntimeError):
    def __init__(self, status, text):
        self.status = status
        super().__init__("%s: %s" % (STATUS.get(status, status), text))


def build(verbose=False):
    """Compile liblazac_b200.so for sm_100a (nvcc cross-compiles without a GPU)."""
    subprocess.run(["make", "-C", os.path.join(HERE, "csrc"), "-j4"], check=True,
                   stdout=None if verbose else subprocess.DEVNULL)
    return LIB_PATH


_lib = None


def load():
    """dlopen the library and bind every declared symbol.  Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError(
            LIB_PATH + " is not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, res, args in SYMBOLS:
        fn = getattr(lib, name)  # AttributeError if the export is missing
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def ptr(a):
    """Raw pointer of a C-contiguous numpy array (or None)."""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)


def i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)
