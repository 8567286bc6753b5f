"""ctypes binding of libhar_b200.so (include/har_b200.h).  There is no fallback: if the CUDA
library is missing, or a call fails, this raises — the product path never computes on the CPU."""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libhar_b200.so")
_CSRC = os.path.join(os.path.dirname(_HERE), "csrc")
_lib = None

NUM_WINDOW_FEATURES = 43
MAX_CLASSES = 32
MAX_BINS = 256


class HarError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


class FitParams(C.Structure):
    _fields_ = [("num_trees", C.c_int32), ("max_depth", C.c_int32), ("max_bins", C.c_int32),
                ("num_classes", C.c_int32), ("min_instances_per_node", C.c_int32), ("impurity", C.c_int32),
                ("features_per_node", C.c_int32), ("reserved", C.c_int32),
                ("min_info_gain", C.c_double), ("subsampling_rate", C.c_double),
                ("seed", C.c_int64), ("max_memory_bytes", C.c_int64)]


ALLREDUCE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p)


class Dist(C.Structure):
    _fields_ = [("rank", C.c_int32), ("world_size", C.c_int32), ("global_rows", C.c_int64),
                ("partition_offset", C.c_int32), ("total_partitions", C.c_int32),
                ("allreduce_i32", ALLREDUCE_FN), ("ctx", C.c_void_p), ("group_id", C.c_uint64)]


def build(force=False, verbose=False):
    """nvcc build of the library for sm_100a, in-tree (cross-compiles without a GPU)."""
    srcs = [os.path.join(_CSRC, f) for f in os.listdir(_CSRC) if f.endswith((".cu", ".cuh", ".h"))]
    srcs.append(os.path.join(os.path.dirname(os.path.dirname(_HERE)), "include", "har_b200.h"))
    stale = force or not os.path.exists(_SO) or os.path.getmtime(_SO) < max(os.path.getmtime(s) for s in srcs)
    if stale:
        cmd = ["make", "-C", _CSRC, "-j8"] + (["-B"] if force else [])
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if verbose or r.returncode:
            print(r.stdout)
        if r.returncode:
            raise RuntimeError("building libhar_b200.so failed")
    return _SO


_SIGS = {
    "har_last_error": (C.c_char_p, []),
    "har_version": (C.c_int, []),
    "har_device_count": (C.c_int, [C.POINTER(C.c_int32)]),
    "har_set_device": (C.c_int, [C.c_int32]),
    "har_stream_synchronize": (C.c_int, [C.c_void_p]),
    "har_scratch_release": (None, []),
    "har_dist_transport": (C.c_int, [C.POINTER(Dist), C.c_char_p, C.c_int32]),
    "har_dist_release": (None, []),
    "har_debug_launch_count": (C.c_longlong, [C.c_int]),
    "har_debug_comm_stats": (None, [C.POINTER(C.c_longlong), C.POINTER(C.c_longlong), C.POINTER(C.c_int), C.c_int]),
    "har_debug_div_check": (C.c_int, [C.c_longlong, C.c_ulonglong, C.POINTER(C.c_ulonglong), C.c_void_p]),
    "har_debug_smem_atomic_peak": (C.c_int, [C.c_int, C.c_int, C.POINTER(C.c_double), C.c_void_p]),
    "har_debug_profile_enable": (None, [C.c_int]),
    "har_debug_profile_reset": (None, []),
    "har_debug_profile_read": (C.c_int, [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_longlong), C.POINTER(C.c_double)]),
    "har_window_segments": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int64,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int64), C.c_void_p]),
    "har_window_segments_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int64,
                                             C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int64), C.c_void_p]),
    "har_window_features": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p,
                                      C.c_int64, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p]),
    "har_window_transform": (C.c_int, [C.c_void_p] * 6 + [C.c_int64, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.c_void_p]),
    "har_synth_stream": (C.c_int, [C.c_int64, C.c_int64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "har_find_splits": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_void_p, C.c_int32, C.c_int64,
                                  C.c_void_p, C.c_int32, C.POINTER(Dist), C.c_void_p, C.c_void_p, C.c_void_p]),
    "har_bin_features": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_int32, C.c_void_p, C.c_int64, C.c_void_p]),
    "har_bag_weights": (C.c_int, [C.c_int64, C.c_int32, C.c_double, C.c_int64, C.c_void_p, C.c_int32, C.c_int32,
                                  C.c_void_p, C.c_void_p]),
    "har_node_histogram": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32,
                                     C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "har_forest_fit": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                 C.c_int32, C.POINTER(FitParams), C.POINTER(Dist), C.c_void_p, C.c_void_p,
                                 C.POINTER(C.c_void_p)]),
    "har_forest_shape": (C.c_int, [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                   C.POINTER(C.c_int64)]),
    "har_forest_export": (C.c_int, [C.c_void_p] + [C.c_void_p] * 12),
    "har_forest_fit_info": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(C.c_int64),
                                      C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "har_forest_import": (C.c_int, [C.c_int32, C.c_int32, C.c_int32] + [C.c_void_p] * 8 + [C.POINTER(C.c_void_p)]),
    "har_forest_free": (None, [C.c_void_p]),
    "har_forest_predict": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p]),
    "har_parse_wisdm_raw": (C.c_int, [C.c_void_p, C.c_int64, C.POINTER(C.c_char_p), C.c_int32, C.c_int64, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64),
                                      C.c_void_p]),
    "har_count_wisdm_raw": (C.c_int, [C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.c_void_p]),
    "har_csr_to_dense": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_int64,
                                   C.c_void_p]),
}

EXPORTED = sorted(_SIGS)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            raise HarError(-2, "libhar_b200.so is not built (%s); run `python -c 'import __graft_entry__ as g; "
                               "g.build()'` — there is no CPU fallback" % _SO)
        L = C.CDLL(_SO)
        for name, (res, args) in _SIGS.items():
            fn = getattr(L, name)          # AttributeError if the symbol is missing
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise HarError(rc, lib().har_last_error().decode("utf-8", "replace"))
