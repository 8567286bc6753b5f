"""Host-side plumbing between Arrow/numpy columns, torch device buffers and the C ABI.

torch is used for device memory, streams and (in dist.py) torch.distributed only; every
computation on the hot path is a kernel of libhar_b200.so.  Functions accept numpy arrays (copied
host->device on the current stream) or CUDA tensors (used in place)."""
import ctypes as C

import numpy as np
import torch

from . import _native as N

WINDOW_FEATURE_NAMES = ([a + str(i) for a in "XYZ" for i in range(10)] +
                        ["XAVG", "YAVG", "ZAVG", "XPEAK", "YPEAK", "ZPEAK", "XABSDEV", "YABSDEV", "ZABSDEV",
                         "XSTDDEV", "YSTDDEV", "ZSTDDEV", "RESULTANT"])

_TORCH = {np.dtype("float64"): torch.float64, np.dtype("int32"): torch.int32, np.dtype("int64"): torch.int64,
          np.dtype("uint8"): torch.uint8}


COUNTERS = {"h2d_bytes": 0, "d2h_bytes": 0}      # host<->device traffic moved by this layer (bench.py's e2e)


def to_host(t):
    """CUDA tensor -> numpy (counted device->host copy)."""
    if t is None:
        return None
    COUNTERS["d2h_bytes"] += t.numel() * t.element_size()
    return t.cpu().numpy()


def require_cuda():
    if not torch.cuda.is_available():
        raise N.HarError(-2, "har_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    N.lib()


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def dev(a, dtype, pinned=False):
    """numpy / tensor -> contiguous CUDA tensor of `dtype` on the current device."""
    td = _TORCH[np.dtype(dtype)]
    if isinstance(a, torch.Tensor):
        if a.is_cuda and a.dtype == td and a.is_contiguous():
            return a
        return a.to(device="cuda", dtype=td).contiguous()
    a = np.ascontiguousarray(a, dtype=dtype)
    t = torch.from_numpy(a)
    COUNTERS["h2d_bytes"] += a.nbytes
    # non_blocking is honoured only for page-locked memory (callers that care keep their columns pinned)
    return t.cuda(non_blocking=True)


def matrix(X):
    """-> (CUDA f64 tensor, n, F, ld) for a row-major matrix; keeps a strided CUDA view in place."""
    if isinstance(X, torch.Tensor) and X.is_cuda and X.dtype == torch.float64 and X.dim() == 2 and X.stride(1) == 1:
        return X, X.shape[0], X.shape[1], (X.stride(0) if X.shape[0] > 1 else max(X.stride(0), X.shape[1]))
    t = dev(X, np.float64)
    if t.dim() != 2:
        raise ValueError("expected a 2-d feature matrix")
    return t, t.shape[0], t.shape[1], t.shape[1]


def ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def hptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def parts(n, part_off):
    po = np.ascontiguousarray([0, n] if part_off is None else part_off, dtype=np.int64)
    if po[0] != 0 or po[-1] != n or np.any(np.diff(po) < 0):
        raise ValueError("part_off must be non-decreasing from 0 to n")
    return po


def comm_stats(reset=False):
    """(all-reduce calls, bytes, 1 if the library's own NCCL communicator carries them) since the last reset."""
    c, b, v = C.c_longlong(), C.c_longlong(), C.c_int()
    N.lib().har_debug_comm_stats(C.byref(c), C.byref(b), C.byref(v), 1 if reset else 0)
    return int(c.value), int(b.value), int(v.value)


def div_check(n, seed=1):
    """Mismatches between the count division of the best-split kernel and __ddiv_rn over n operand pairs."""
    require_cuda()
    m = C.c_ulonglong()
    N.check(N.lib().har_debug_div_check(int(n), int(seed), C.byref(m), stream()))
    return int(m.value)


def release_scratch():
    """Frees the device scratch arena the calling thread's fits keep between calls (har_scratch_release)."""
    N.lib().har_scratch_release()


def launch_count(reset=False):
    return int(N.lib().har_debug_launch_count(1 if reset else 0))


def smem_atomic_peak(mode, words=8192):
    """Measured shared-memory atomic throughput (atomics/s, whole device): mode 0 conflict-free, 1 random."""
    require_cuda()
    v = C.c_double()
    N.check(N.lib().har_debug_smem_atomic_peak(mode, words, C.byref(v), stream()))
    return v.value


PROFILE_IDS = ("node_hist", "best_split", "part_count", "part_scatter", "window_features", "seg_tiles", "predict",
               "bin_features", "find_splits", "bag_weights", "window_span")


def profile(enable=None, reset=False):
    """CUDA-event timing of selected kernels, measured inside the library on the launching stream.
    Returns {kernel: (total_ms, launches, units)} accumulated since the last reset."""
    L = N.lib()
    if enable is not None:
        L.har_debug_profile_enable(1 if enable else 0)
    out = {}
    for i, name in enumerate(PROFILE_IDS):
        ms, ln, un = C.c_double(), C.c_longlong(), C.c_double()
        L.har_debug_profile_read(i, C.byref(ms), C.byref(ln), C.byref(un))
        out[name] = (ms.value, ln.value, un.value)
    if reset:
        L.har_debug_profile_reset()
    return out


# ---- window transform ------------------------------------------------------------------------------
def window_segments(user, activity, ts=None, window_len=200, gap_ns=0):
    require_cuda()
    u, a = dev(user, np.int32), dev(activity, np.int32)
    t = None if ts is None else dev(ts, np.int64)
    n = u.numel()
    cap = max(n // window_len, 1)
    ws = torch.empty(cap, dtype=torch.int64, device="cuda")
    wu = torch.empty(cap, dtype=torch.int32, device="cuda")
    wa = torch.empty(cap, dtype=torch.int32, device="cuda")
    nw = C.c_int64(0)
    N.check(N.lib().har_window_segments(ptr(u), ptr(a), ptr(t), n, window_len, gap_ns, ptr(ws), ptr(wu), ptr(wa),
                                        C.byref(nw), stream()))
    k = nw.value
    return ws[:k], wu[:k], wa[:k]


def window_features(x, y, z, ts, win_start, window_len=200, out=None):
    require_cuda()
    xd, yd, zd, td = dev(x, np.float64), dev(y, np.float64), dev(z, np.float64), dev(ts, np.int64)
    ws = dev(win_start, np.int64)
    nw = ws.numel()
    if out is None:
        out = torch.empty((nw, N.NUM_WINDOW_FEATURES), dtype=torch.float64, device="cuda")
    N.check(N.lib().har_window_features(ptr(xd), ptr(yd), ptr(zd), ptr(td), xd.numel(), ptr(ws), nw, window_len,
                                        ptr(out), out.stride(0) if nw > 1 else N.NUM_WINDOW_FEATURES, stream()))
    return out


def window_transform(user, activity, ts, x, y, z, window_len=200, gap_ns=0):
    """Segmentation + features in one library call (har_window_transform): -> (features [nw, 43] f64, win_user i32,
    win_activity i32, win_start i64), views of buffers sized for n // window_len windows."""
    require_cuda()
    u, a, t = dev(user, np.int32), dev(activity, np.int32), dev(ts, np.int64)
    xd, yd, zd = dev(x, np.float64), dev(y, np.float64), dev(z, np.float64)
    n = u.numel()
    cap = max(n // window_len, 1)
    meta = torch.empty(2 * cap, dtype=torch.int64, device="cuda")           # window starts | users, activities (one allocation)
    ws = meta[:cap]
    ua = meta[cap:].view(torch.int32)
    wu, wa = ua[:cap], ua[cap:]
    out = torch.empty((cap, N.NUM_WINDOW_FEATURES), dtype=torch.float64, device="cuda")
    nw = C.c_int64(0)
    N.check(N.lib().har_window_transform(ptr(u), ptr(a), ptr(t), ptr(xd), ptr(yd), ptr(zd), n, window_len, gap_ns, ptr(ws),
                                         ptr(wu), ptr(wa), ptr(out), N.NUM_WINDOW_FEATURES, C.byref(nw), stream()))
    k = nw.value
    return out[:k], wu[:k], wa[:k], ws[:k]


# ---- tree stages -----------------------------------------------------------------------------------
def resolve_features_per_node(strategy, F, num_trees):
    """DecisionTreeMetadata.buildMetadata's featureSubsetStrategy (classification), SURVEY A.5."""
    s = str(strategy).lower()
    if s == "auto":
        s = "all" if num_trees == 1 else "sqrt"
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
    except ValueError:
        k = None
    if k is not None:
        if k > 0:
            return min(k, F)
    else:
        try:
            fr = float(s)
        except ValueError:
            fr = -1.0
        if 0.0 < fr <= 1.0:
            return int(np.ceil(fr * F))
    raise ValueError("Supported values: auto, all, onethird, sqrt, log2, (0.0-1.0], [1-n]. Got %r" % (strategy,))


def find_splits(X, arity=None, max_bins=32, seed=0, part_off=None, dist=None):
    require_cuda()
    Xd, n, F, ld = matrix(X)
    ar = None if arity is None else np.ascontiguousarray(arity, np.int32)
    po = parts(n, part_off)
    thr = np.zeros((F, max_bins)); ns = np.zeros(F, np.int32)
    N.check(N.lib().har_find_splits(ptr(Xd), n, F, ld, hptr(ar), max_bins, seed, hptr(po), len(po) - 1,
                                    None if dist is None else C.byref(dist.struct), hptr(thr), hptr(ns), stream()))
    return thr, ns


def bin_features(X, thresholds, nsplits, arity=None):
    require_cuda()
    Xd, n, F, ld = matrix(X)
    ar = None if arity is None else np.ascontiguousarray(arity, np.int32)
    thr = np.ascontiguousarray(thresholds, np.float64)
    ldb = (F + 15) & ~15
    bins = torch.zeros((max(n, 1), ldb), dtype=torch.uint8, device="cuda")
    N.check(N.lib().har_bin_features(ptr(Xd), n, F, ld, hptr(ar), hptr(thr), hptr(np.ascontiguousarray(nsplits, np.int32)),
                                     thr.shape[1], ptr(bins), ldb, stream()))
    return bins[:n]


def bag_weights(n, num_trees, rate=1.0, seed=0, part_off=None, partition_offset=0):
    require_cuda()
    po = parts(n, part_off)
    w = torch.empty((num_trees, n), dtype=torch.uint8, device="cuda")
    N.check(N.lib().har_bag_weights(n, num_trees, float(rate), seed, hptr(po), len(po) - 1, partition_offset, ptr(w), stream()))
    return w


def node_histogram(bins, labels, hist_off, num_classes, weights=None, rows=None):
    require_cuda()
    b = dev(bins, np.uint8)
    lab = dev(labels, np.uint8)
    w = None if weights is None else dev(weights, np.uint8)
    r = None if rows is None else dev(rows, np.int32)
    ho = np.ascontiguousarray(hist_off, np.int64)
    F = len(ho) - 1
    hist = torch.zeros(int(ho[-1]), dtype=torch.int32, device="cuda")
    parent = torch.zeros(num_classes, dtype=torch.int32, device="cuda")
    nrows = b.shape[0] if r is None else r.numel()
    N.check(N.lib().har_node_histogram(ptr(b), b.stride(0), ptr(lab), ptr(w), ptr(r), nrows, F, num_classes, hptr(ho),
                                       ptr(hist), ptr(parent), stream()))
    return hist, parent


class Forest:
    """Owns a har_forest_t*; flat arrays are exported once to numpy (see har_forest_export)."""
    FIELDS = ("tree_off", "feature", "left", "right", "node_id", "arity", "threshold", "left_mask",
              "prediction", "impurity", "gain", "stats")

    def __init__(self, handle):
        self._h = C.c_void_p(handle)
        T, Cn, F, tot = C.c_int32(), C.c_int32(), C.c_int32(), C.c_int64()
        N.check(N.lib().har_forest_shape(self._h, C.byref(T), C.byref(Cn), C.byref(F), C.byref(tot)))
        self.T, self.C, self.F, self.total_nodes = T.value, Cn.value, F.value, tot.value
        self._arrays = None          # flat node arrays are copied out of the handle on first use
        self._max_bins = None        # set by fit_forest: the fit by-products below are then read on first use

    def _export(self):
        n = self.total_nodes
        a = {"tree_off": np.zeros(self.T + 1, np.int64), "feature": np.zeros(n, np.int32), "left": np.zeros(n, np.int32),
             "right": np.zeros(n, np.int32), "node_id": np.zeros(n, np.int32), "arity": np.zeros(n, np.int32),
             "threshold": np.zeros(n, np.float64), "left_mask": np.zeros(n, np.uint64), "prediction": np.zeros(n, np.float64),
             "impurity": np.zeros(n, np.float64), "gain": np.zeros(n, np.float64), "stats": np.zeros((n, self.C), np.float64)}
        N.check(N.lib().har_forest_export(self._h, *[hptr(a[k]) for k in self.FIELDS]))
        self._arrays = a

    INFO = ("thresholds", "nsplits", "num_batches", "num_launches")

    def __getattr__(self, name):
        # only reached for attributes not set in __init__: the lazily exported node arrays and fit by-products
        if name in Forest.FIELDS:
            if self.__dict__.get("_arrays") is None:
                self._export()
            return self.__dict__["_arrays"][name]
        if name in Forest.INFO or name == "num_groups":
            mb = self.__dict__.get("_max_bins")
            if mb is None:                                # imported forest: nothing was fitted
                return None if name in ("thresholds", "nsplits") else 0
            if name == "num_groups":
                # Spark's node-group count; the library recounts it from the trees when the fit used its own order
                g = C.c_int64()
                N.check(N.lib().har_forest_fit_info(self._h, None, None, mb, C.byref(g), None, None))
                self.num_groups = g.value
                return g.value
            self._load_fit_info(mb)
            return self.__dict__[name]
        raise AttributeError(name)

    @classmethod
    def from_arrays(cls, T, Cn, F, tree_off, feature, left, right, arity, threshold, left_mask, stats):
        N.lib()
        h = C.c_void_p()
        arrs = [np.ascontiguousarray(tree_off, np.int64), np.ascontiguousarray(feature, np.int32),
                np.ascontiguousarray(left, np.int32), np.ascontiguousarray(right, np.int32),
                np.ascontiguousarray(arity, np.int32), np.ascontiguousarray(threshold, np.float64),
                np.ascontiguousarray(left_mask, np.uint64), np.ascontiguousarray(stats, np.float64)]
        N.check(N.lib().har_forest_import(T, Cn, F, *[hptr(a) for a in arrs], C.byref(h)))
        return cls(h.value)

    def _load_fit_info(self, max_bins):
        self.thresholds = np.zeros((self.F, max_bins)); self.nsplits = np.zeros(self.F, np.int32)
        b, l = C.c_int64(), C.c_int64()
        N.check(N.lib().har_forest_fit_info(self._h, hptr(self.thresholds), hptr(self.nsplits), max_bins,
                                            None, C.byref(b), C.byref(l)))
        self.num_batches, self.num_launches = b.value, l.value

    def tree_num_nodes(self, t):
        return int(self.tree_off[t + 1] - self.tree_off[t])

    def tree_depth(self, t):
        base = int(self.tree_off[t])
        depth = {0: 0}
        best = 0
        for k in range(self.tree_num_nodes(t)):          # pre-order: parents precede children
            if self.feature[base + k] >= 0:
                d = depth[k] + 1
                depth[int(self.left[base + k])] = d
                depth[int(self.right[base + k])] = d
                best = max(best, d)
        return best

    def predict(self, X, is_forest=None, want=("raw", "prob", "pred")):
        require_cuda()
        Xd, n, F, ld = matrix(X)
        if F != self.F:
            raise ValueError("feature matrix has %d columns, the model was trained on %d" % (F, self.F))
        if is_forest is None:
            is_forest = self.T > 1
        raw = torch.empty((n, self.C), dtype=torch.float64, device="cuda") if "raw" in want else None
        prob = torch.empty((n, self.C), dtype=torch.float64, device="cuda") if "prob" in want else None
        pred = torch.empty(n, dtype=torch.float64, device="cuda") if "pred" in want else None
        N.check(N.lib().har_forest_predict(self._h, ptr(Xd), n, ld, 1 if is_forest else 0, ptr(raw), ptr(prob), ptr(pred),
                                           stream()))
        return raw, prob, pred

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            N.lib().har_forest_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def fit_forest(X, y, arity=None, num_classes=None, num_trees=1, max_depth=5, max_bins=32,
               min_instances_per_node=1, min_info_gain=0.0, impurity="gini", feature_subset="auto",
               subsampling_rate=1.0, seed=0, max_memory_mb=256, part_off=None, weights=None, dist=None):
    """RandomForest.run on the GPU (har_forest_fit).  X: [n, F] f64 (numpy or CUDA tensor), y: [n] f64."""
    require_cuda()
    Xd, n, F, ld = matrix(X)
    yd = dev(y, np.float64)
    if yd.numel() != n:
        raise ValueError("labels and features disagree on the number of rows")
    if num_classes is None:
        num_classes = int(yd.max().item()) + 1 if n else 2
    if impurity not in ("gini", "entropy"):
        raise ValueError("impurity must be 'gini' or 'entropy'")
    ar = None if arity is None else np.ascontiguousarray(arity, np.int32)
    po = parts(n, part_off)
    p = N.FitParams(num_trees, max_depth, max_bins, num_classes, min_instances_per_node,
                    0 if impurity == "gini" else 1, resolve_features_per_node(feature_subset, F, num_trees), 0,
                    float(min_info_gain), float(subsampling_rate), int(seed), int(max_memory_mb) * 1024 * 1024)
    w = None if weights is None else dev(weights, np.uint8)
    h = C.c_void_p()
    N.check(N.lib().har_forest_fit(ptr(Xd), n, F, ld, ptr(yd), hptr(ar), hptr(po), len(po) - 1, C.byref(p),
                                   None if dist is None else C.byref(dist.struct), ptr(w), stream(), C.byref(h)))
    fo = Forest(h.value)
    fo._max_bins = max_bins
    return fo


WISDM_ACTIVITIES = ("Walking", "Jogging", "Upstairs", "Downstairs", "Sitting", "Standing")


def parse_wisdm_raw(text, activities=WISDM_ACTIVITIES):
    """Raw WISDM records ("user,activity,timestamp,x,y,z;") -> (user i32, activity i32, ts i64, x, y, z f64) CUDA
    tensors plus the number of dropped records.  `text`: bytes / bytearray / uint8 array / CUDA uint8 tensor."""
    require_cuda()
    if isinstance(text, torch.Tensor):
        t = text.to(device="cuda", dtype=torch.uint8).contiguous()
    else:
        a = np.frombuffer(bytes(text), dtype=np.uint8) if not isinstance(text, np.ndarray) else np.ascontiguousarray(text, np.uint8)
        COUNTERS["h2d_bytes"] += a.nbytes
        t = torch.from_numpy(a.copy()).cuda() if a.size else torch.empty(0, dtype=torch.uint8, device="cuda")
    names = (C.c_char_p * len(activities))(*[s.encode() for s in activities])
    nrec, ndrop, cap = C.c_int64(), C.c_int64(), C.c_int64()
    # one cheap pass counts the terminators (an upper bound on the records); the columns are sized with it, parsed
    # once and trimmed to the valid records
    N.check(N.lib().har_count_wisdm_raw(ptr(t), t.numel(), C.byref(cap), stream()))
    n = cap.value
    user = torch.empty(n, dtype=torch.int32, device="cuda"); act = torch.empty_like(user)
    ts = torch.empty(n, dtype=torch.int64, device="cuda")
    x = torch.empty(n, dtype=torch.float64, device="cuda"); y = torch.empty_like(x); z = torch.empty_like(x)
    if n:
        N.check(N.lib().har_parse_wisdm_raw(ptr(t), t.numel(), names, len(activities), n, ptr(user), ptr(act), ptr(ts), ptr(x),
                                            ptr(y), ptr(z), C.byref(nrec), C.byref(ndrop), stream()))
    k = nrec.value
    user, act, ts, x, y, z = (c[:k] for c in (user, act, ts, x, y, z))
    return (user, act, ts, x, y, z), int(ndrop.value)


def csr_to_dense(indptr, indices, values, n, F):
    require_cuda()
    ip, ix, vv = dev(indptr, np.int64), dev(indices, np.int32), dev(values, np.float64)
    out = torch.empty((n, F), dtype=torch.float64, device="cuda")
    N.check(N.lib().har_csr_to_dense(ptr(ip), ptr(ix), ptr(vv), n, F, ptr(out), F, stream()))
    return out
