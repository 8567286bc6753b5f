"""ctypes front-end of the CPU oracle (oracle/har_oracle.c).

TEST INFRASTRUCTURE ONLY — may be imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product path never imports this module.
See the header of har_oracle.c for what is pinned by the reference's golden output and what is not.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None


def _host_tag():
    """The library is built -march=native, so its name carries the host CPU: a copy built on another machine
    (the build box ships its tree to the GPU box) is not loaded there, it is rebuilt."""
    import hashlib
    model, flags = "", ""
    try:
        with open("/proc/cpuinfo") as fh:
            for line in fh:
                if line.startswith("model name") and not model:
                    model = line.split(":", 1)[1].strip()
                elif line.startswith("flags") and not flags:
                    flags = line.split(":", 1)[1].strip()
                if model and flags:
                    break
    except OSError:
        pass
    return hashlib.sha1((model + "|" + flags).encode()).hexdigest()[:12]


_SO = os.path.join(_HERE, "_build", "libhar_oracle-%s.so" % _host_tag())
CFLAGS = None


def build(force=False):
    global CFLAGS
    src = os.path.join(_HERE, "har_oracle.c")
    mk = os.path.join(_HERE, "Makefile")
    stale = force or not os.path.exists(_SO) or os.path.getmtime(_SO) < max(os.path.getmtime(src), os.path.getmtime(mk))
    if stale:
        subprocess.check_call(["make", "-C", _HERE, "-s", "OUT=" + os.path.relpath(_SO, _HERE)] + (["-B"] if force else []))
    if CFLAGS is None:
        CFLAGS = subprocess.run(["make", "-C", _HERE, "-s", "flags"], stdout=subprocess.PIPE, text=True).stdout.strip()
    return _SO


class Params(C.Structure):
    _fields_ = [("num_trees", C.c_int32), ("max_depth", C.c_int32), ("max_bins", C.c_int32),
                ("num_classes", C.c_int32), ("min_instances_per_node", C.c_int32), ("impurity", C.c_int32),
                ("features_per_node", C.c_int32), ("reserved", C.c_int32),
                ("min_info_gain", C.c_double), ("subsampling_rate", C.c_double),
                ("seed", C.c_int64), ("max_memory_bytes", C.c_int64)]


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.orc_fit_forest.restype = C.c_void_p
        _lib.orc_forest_total_nodes.restype = C.c_int64
        _lib.orc_forest_total_nodes.argtypes = [C.c_void_p]
        _lib.orc_forest_num_groups.restype = C.c_int64
        _lib.orc_forest_num_groups.argtypes = [C.c_void_p]
        _lib.orc_forest_free.argtypes = [C.c_void_p]
        _lib.orc_window_segments.restype = C.c_int64
        _lib.orc_split_sample_rows.restype = C.c_int64
        _lib.orc_bin_features.restype = C.c_int64
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _parts(n, part_off):
    if part_off is None:
        part_off = [0, n]
    return np.ascontiguousarray(part_off, dtype=np.int64)


def features_per_node(strategy, F, num_trees, classification=True):
    """DecisionTreeMetadata.buildMetadata featureSubsetStrategy resolution."""
    s = str(strategy).lower()
    if s == "auto":
        s = "all" if num_trees == 1 else ("sqrt" if classification else "onethird")
    if s == "all":
        return F
    if s == "sqrt":
        return int(np.ceil(np.sqrt(F)))
    if s == "log2":
        return max(1, int(np.ceil(np.log(F) / np.log(2))))
    if s == "onethird":
        return int(np.ceil(F / 3.0))
    try:
        k = int(s)
        if k > 0:
            return min(k, F)
    except ValueError:
        pass
    fr = float(s)
    if 0.0 < fr <= 1.0:
        return int(np.ceil(fr * F))
    raise ValueError("Supported values: auto, all, onethird, sqrt, log2, (0.0-1.0], [1-n]. Got %r" % strategy)


class Forest:
    FIELDS = ("tree_off", "feature", "left", "right", "node_id", "arity", "threshold", "left_mask",
              "prediction", "impurity", "gain", "stats", "thresholds", "nsplits")

    def __init__(self, T, Cn, F, **kw):
        self.T, self.C, self.F = T, Cn, F
        for k in self.FIELDS:
            setattr(self, k, kw.get(k))
        self.num_groups = kw.get("num_groups", 0)

    def tree_depth(self, t):
        base = int(self.tree_off[t])

        def d(k):
            if self.feature[base + k] < 0:
                return 0
            return 1 + max(d(int(self.left[base + k])), d(int(self.right[base + k])))
        return d(0)

    def tree_num_nodes(self, t):
        return int(self.tree_off[t + 1] - self.tree_off[t])


def fit_forest(X, y, arity=None, num_classes=None, num_trees=1, max_depth=5, max_bins=32,
               min_instances_per_node=1, min_info_gain=0.0, impurity="gini", feature_subset="auto",
               subsampling_rate=1.0, seed=0, max_memory_mb=256, part_off=None, weights=None):
    X = np.ascontiguousarray(X, dtype=np.float64)
    y = np.ascontiguousarray(y, dtype=np.float64)
    n, F = X.shape
    if num_classes is None:
        num_classes = int(y.max()) + 1 if n else 2
    ar = None if arity is None else np.ascontiguousarray(arity, dtype=np.int32)
    po = _parts(n, part_off)
    p = Params(num_trees, max_depth, max_bins, num_classes, min_instances_per_node,
               0 if impurity == "gini" else 1, features_per_node(feature_subset, F, num_trees), 0,
               float(min_info_gain), float(subsampling_rate), int(seed), int(max_memory_mb) * 1024 * 1024)
    err = C.create_string_buffer(1024)
    w = None if weights is None else np.ascontiguousarray(weights, dtype=np.uint8)
    L = lib()
    h = L.orc_fit_forest(_p(X), C.c_int64(n), C.c_int32(F), C.c_int64(X.strides[0] // 8), _p(y), _p(ar),
                         _p(po), C.c_int32(len(po) - 1), C.byref(p), _p(w), err, C.c_int32(1024))
    if not h:
        raise ValueError(err.value.decode())
    try:
        N = L.orc_forest_total_nodes(h)
        out = dict(tree_off=np.zeros(num_trees + 1, np.int64), feature=np.zeros(N, np.int32),
                   left=np.zeros(N, np.int32), right=np.zeros(N, np.int32), node_id=np.zeros(N, np.int32),
                   arity=np.zeros(N, np.int32), threshold=np.zeros(N, np.float64),
                   left_mask=np.zeros(N, np.uint64), prediction=np.zeros(N, np.float64),
                   impurity=np.zeros(N, np.float64), gain=np.zeros(N, np.float64),
                   stats=np.zeros((N, num_classes), np.float64),
                   thresholds=np.zeros((F, max_bins), np.float64), nsplits=np.zeros(F, np.int32))
        L.orc_forest_export(C.c_void_p(h), *[_p(out[k]) for k in Forest.FIELDS])
        out["num_groups"] = L.orc_forest_num_groups(h)
    finally:
        L.orc_forest_free(h)
    return Forest(num_trees, num_classes, F, **out)


def predict(forest, X, is_forest=None):
    X = np.ascontiguousarray(X, dtype=np.float64)
    n = X.shape[0]
    Cn = forest.C
    raw = np.zeros((n, Cn)); prob = np.zeros((n, Cn)); pred = np.zeros(n)
    if is_forest is None:
        is_forest = forest.T > 1
    lib().orc_predict(C.c_int32(forest.T), C.c_int32(Cn), _p(forest.tree_off), _p(forest.feature),
                      _p(forest.left), _p(forest.right), _p(forest.arity), _p(forest.threshold),
                      _p(forest.left_mask), _p(np.ascontiguousarray(forest.stats)), _p(X), C.c_int64(n),
                      C.c_int64(X.strides[0] // 8 if n else X.shape[1]), C.c_int32(1 if is_forest else 0),
                      _p(raw), _p(prob), _p(pred))
    return raw, prob, pred


def find_splits(X, arity=None, max_bins=32, seed=0, part_off=None):
    X = np.ascontiguousarray(X, dtype=np.float64)
    n, F = X.shape
    ar = None if arity is None else np.ascontiguousarray(arity, dtype=np.int32)
    po = _parts(n, part_off)
    thr = np.zeros((F, max_bins)); ns = np.zeros(F, np.int32)
    lib().orc_find_splits(_p(X), C.c_int64(n), C.c_int32(F), C.c_int64(X.strides[0] // 8), _p(ar),
                          C.c_int32(max_bins), C.c_int64(seed), _p(po), C.c_int32(len(po) - 1), _p(thr), _p(ns))
    return thr, ns


def split_sample_rows(n, max_bins, seed, part_off=None):
    po = _parts(n, part_off)
    out = np.zeros(max(n, 1), np.int64)
    k = lib().orc_split_sample_rows(C.c_int64(n), C.c_int32(min(max_bins, n)), C.c_int64(seed), _p(po),
                                    C.c_int32(len(po) - 1), _p(out))
    return out[:k].copy()


def bin_features(X, thresholds, nsplits, arity=None, ldb=None):
    X = np.ascontiguousarray(X, dtype=np.float64)
    n, F = X.shape
    ldb = ldb or F
    ar = None if arity is None else np.ascontiguousarray(arity, dtype=np.int32)
    bins = np.zeros((n, ldb), np.uint8)
    thresholds = np.ascontiguousarray(thresholds, dtype=np.float64)
    bad = lib().orc_bin_features(_p(X), C.c_int64(n), C.c_int32(F), C.c_int64(X.strides[0] // 8), _p(ar),
                                 _p(thresholds), _p(np.ascontiguousarray(nsplits, np.int32)),
                                 C.c_int32(thresholds.shape[1]), _p(bins), C.c_int64(ldb))
    if bad:
        raise ValueError("invalid categorical value at flat index %d" % (bad - 1))
    return bins


def bag_weights(n, num_trees, rate=1.0, seed=0, part_off=None):
    po = _parts(n, part_off)
    w = np.zeros((num_trees, n), np.uint8)
    lib().orc_bag_weights(C.c_int64(n), C.c_int32(num_trees), C.c_double(rate), C.c_int64(seed), _p(po),
                          C.c_int32(len(po) - 1), _p(w))
    return w


def node_histogram(bins, labels, hist_off, num_classes, weights=None, rows=None):
    bins = np.ascontiguousarray(bins, np.uint8)
    n, ldb = bins.shape
    F = len(hist_off) - 1
    hist_off = np.ascontiguousarray(hist_off, np.int64)
    hist = np.zeros(int(hist_off[-1]), np.int64); parent = np.zeros(num_classes, np.int64)
    rows_a = None if rows is None else np.ascontiguousarray(rows, np.int32)
    lib().orc_node_histogram(_p(bins), C.c_int64(ldb), _p(np.ascontiguousarray(labels, np.uint8)),
                             _p(None if weights is None else np.ascontiguousarray(weights, np.uint8)),
                             _p(rows_a), C.c_int64(n if rows is None else len(rows_a)), C.c_int32(F),
                             C.c_int32(num_classes), _p(hist_off), _p(hist), _p(parent))
    return hist, parent


def window_segments(user, act, ts=None, L=200, gap_ns=0):
    user = np.ascontiguousarray(user, np.int32); act = np.ascontiguousarray(act, np.int32)
    tsa = None if ts is None else np.ascontiguousarray(ts, np.int64)
    n = len(user)
    k = lib().orc_window_segments(_p(user), _p(act), _p(tsa), C.c_int64(n), C.c_int32(L), C.c_int64(gap_ns), None)
    out = np.zeros(k, np.int64)
    lib().orc_window_segments(_p(user), _p(act), _p(tsa), C.c_int64(n), C.c_int32(L), C.c_int64(gap_ns), _p(out))
    return out


def window_features(x, y, z, ts, win_start, L=200):
    x = np.ascontiguousarray(x, np.float64); y = np.ascontiguousarray(y, np.float64)
    z = np.ascontiguousarray(z, np.float64); ts = np.ascontiguousarray(ts, np.int64)
    ws = np.ascontiguousarray(win_start, np.int64)
    out = np.zeros((len(ws), 43))
    lib().orc_window_features(_p(x), _p(y), _p(z), _p(ts), _p(ws), C.c_int64(len(ws)), C.c_int32(L), _p(out))
    return out


# ---- synthetic stream -------------------------------------------------------------
SYNTH_MU = np.array([[0.6, 9.3, 0.4], [0.5, 9.1, 0.7], [0.3, 9.2, 0.9], [0.5, 9.2, 0.5],
                     [1.1, 8.2, 2.1], [0.4, 9.4, 0.6]], dtype=np.float64)
SYNTH_AMP = np.array([[3.0, 4.2, 2.6], [4.4, 5.6, 3.4], [3.1, 4.3, 3.0], [2.9, 4.4, 2.7],
                      [0.6, 0.7, 0.5], [0.8, 0.9, 0.7]], dtype=np.float64)
SYNTH_SIGMA = np.array([[2.2, 2.6, 2.1], [2.6, 3.0, 2.3], [2.3, 2.7, 2.2], [2.4, 2.6, 2.2],
                        [1.1, 1.2, 1.0], [1.2, 1.3, 1.1]], dtype=np.float64)
SYNTH_FREQ = np.array([97, 113, 89, 93, 41, 47], dtype=np.int32)       # table steps per sample
SYNTH_SINE = np.sin(2.0 * np.pi * np.arange(1024) / 1024.0)


def synth_stream(n, first_sample=0, per_user=100000, seed=2018):
    user = np.zeros(n, np.int32); act = np.zeros(n, np.int32); ts = np.zeros(n, np.int64)
    x = np.zeros(n); y = np.zeros(n); z = np.zeros(n)
    lib().orc_synth_stream(C.c_int64(n), C.c_int64(first_sample), C.c_int64(per_user), C.c_uint64(seed),
                           _p(SYNTH_SINE), _p(SYNTH_MU), _p(SYNTH_AMP), _p(SYNTH_SIGMA), _p(SYNTH_FREQ),
                           _p(user), _p(act), _p(ts), _p(x), _p(y), _p(z))
    return user, act, ts, x, y, z


WISDM_ACTIVITIES = ("Walking", "Jogging", "Upstairs", "Downstairs", "Sitting", "Standing")


def parse_wisdm_raw(text, activities=WISDM_ACTIVITIES):
    """strtod / strtoll restatement of har_parse_wisdm_raw -> ((user, act, ts, x, y, z), dropped)."""
    data = bytes(text)
    L = lib()
    L.orc_parse_wisdm_raw.restype = C.c_int64
    names = (C.c_char_p * len(activities))(*[a.encode() for a in activities])
    drop = C.c_int64()
    cap = data.count(b";") + 1
    user = np.zeros(cap, np.int32); act = np.zeros(cap, np.int32); ts = np.zeros(cap, np.int64)
    x = np.zeros(cap); y = np.zeros(cap); z = np.zeros(cap)
    n = L.orc_parse_wisdm_raw(C.c_char_p(data), C.c_int64(len(data)), names, C.c_int32(len(activities)), C.c_int64(cap),
                              _p(user), _p(act), _p(ts), _p(x), _p(y), _p(z), C.byref(drop))
    return (user[:n], act[:n], ts[:n], x[:n], y[:n], z[:n]), int(drop.value)


def num_threads():
    return lib().orc_num_threads()


def set_num_threads(n):
    lib().orc_set_num_threads(C.c_int(int(n)))
