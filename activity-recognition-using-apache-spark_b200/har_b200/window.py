"""WisdmWindowFeaturizer — the pyspark.ml Transformer in front of the window-transform kernels.

The reference never runs this stage: it reads the already-windowed WISDM table
(Main/wisdm_main_ver_0.0/data/wisdm_data.csv, Main/main.py:18-20).  This transformer produces that
table's schema (wisdm_data.csv:1: UID, USER, X0..X9, Y0..Y9, Z0..Z9, XAVG..ZAVG, XPEAK..ZPEAK,
X/Y/ZABSDEV, X/Y/ZSTDDEV, RESULTANT, ACTIVITY) from a raw (user, activity, timestamp, x, y, z) frame
on the GPU; definitions in DESIGN.md / SURVEY.md Appendix B.
"""
import numpy as np
import torch
from pyspark.ml import Transformer
from pyspark.ml.linalg import VectorColumn
from pyspark.ml.param import Param, Params, TypeConverters, _make_accessors
from pyspark.sql import DataFrame, DeviceColumn

from . import core

__all__ = ["WisdmWindowFeaturizer"]

_NAMES = ["userCol", "activityCol", "timestampCol", "xCol", "yCol", "zCol", "windowLength", "gapNs", "outputCol",
          "labelCol", "unknownPeakValue"]


class WisdmWindowFeaturizer(Transformer):
    userCol = Param(Params._dummy_parent, "userCol", "user id column (int).", TypeConverters.toString)
    activityCol = Param(Params._dummy_parent, "activityCol", "activity code column (int).", TypeConverters.toString)
    timestampCol = Param(Params._dummy_parent, "timestampCol", "timestamp column (int64 nanoseconds).", TypeConverters.toString)
    xCol = Param(Params._dummy_parent, "xCol", "x acceleration column.", TypeConverters.toString)
    yCol = Param(Params._dummy_parent, "yCol", "y acceleration column.", TypeConverters.toString)
    zCol = Param(Params._dummy_parent, "zCol", "z acceleration column.", TypeConverters.toString)
    windowLength = Param(Params._dummy_parent, "windowLength", "samples per window (10 s at 20 Hz = 200).", TypeConverters.toInt)
    gapNs = Param(Params._dummy_parent, "gapNs", "also break a bout where consecutive timestamps differ by more than "
                  "this many ns (0 = off).", TypeConverters.toInt)
    outputCol = Param(Params._dummy_parent, "outputCol", "optional: also emit the 43 features as one vector column.",
                      TypeConverters.toString)
    labelCol = Param(Params._dummy_parent, "labelCol", "optional: also emit the activity code as a double label column.",
                     TypeConverters.toString)
    unknownPeakValue = Param(Params._dummy_parent, "unknownPeakValue", "value written to outputCol slots whose *PEAK is "
                             "unknown (NaN in the scalar columns, '?' in the reference CSV).", TypeConverters.toFloat)

    def __init__(self, userCol="user", activityCol="activity", timestampCol="timestamp", xCol="x", yCol="y", zCol="z",
                 windowLength=200, gapNs=0, outputCol=None, labelCol=None, unknownPeakValue=0.0):
        super().__init__()
        self._setDefault(userCol="user", activityCol="activity", timestampCol="timestamp", xCol="x", yCol="y", zCol="z",
                         windowLength=200, gapNs=0, unknownPeakValue=0.0)
        kw = dict(userCol=userCol, activityCol=activityCol, timestampCol=timestampCol, xCol=xCol, yCol=yCol, zCol=zCol,
                  windowLength=windowLength, gapNs=gapNs, outputCol=outputCol, labelCol=labelCol,
                  unknownPeakValue=unknownPeakValue)
        self._set(**{k: v for k, v in kw.items() if v is not None})

    def transform_device(self, user, activity, ts, x, y, z):
        """Device-resident variant: CUDA tensors (or host arrays, copied) in, CUDA tensors out:
        (features [nw, 43] f64, win_user i32, win_activity i32, win_start i64)."""
        L = self.getWindowLength()
        u, a = core.dev(user, np.int32, pinned=True), core.dev(activity, np.int32, pinned=True)
        t = core.dev(ts, np.int64, pinned=True)
        xd, yd, zd = core.dev(x, np.float64, pinned=True), core.dev(y, np.float64, pinned=True), core.dev(z, np.float64, pinned=True)
        return core.window_transform(u, a, t, xd, yd, zd, window_len=L, gap_ns=self.getGapNs())

    def _transform(self, dataset):
        c = dataset._cols
        feats, wu, wa, ws = self.transform_device(
            c[dataset._resolve(self.getUserCol())], c[dataset._resolve(self.getActivityCol())],
            c[dataset._resolve(self.getTimestampCol())], c[dataset._resolve(self.getXCol())],
            c[dataset._resolve(self.getYCol())], c[dataset._resolve(self.getZCol())])
        nw = feats.shape[0]
        # the 46 scalar columns of the reference schema stay in HBM (DeviceColumn) until something on the host
        # reads them; transposed once on the device so each column is one contiguous row
        fT = feats.t().contiguous()
        cols, types = {}, {}
        cols["UID"], types["UID"] = np.arange(1, nw + 1, dtype=np.int32), "int"
        cols["USER"], types["USER"] = DeviceColumn(wu, np.int32), "int"
        for j, name in enumerate(core.WINDOW_FEATURE_NAMES):
            cols[name], types[name] = DeviceColumn(fT[j], np.float64), "double"
        cols["ACTIVITY"], types["ACTIVITY"] = DeviceColumn(wa, np.int32), "int"
        meta = {}
        if self.isDefined(self.outputCol):
            # the assembled vector column stays in HBM for the estimator that consumes it; its host copy is lazy
            fv = torch.nan_to_num(feats, nan=self.getUnknownPeakValue())
            cols[self.getOutputCol()], types[self.getOutputCol()] = VectorColumn(feats.shape[1], device=fv), "vector"
            meta[self.getOutputCol()] = {"ml_attr": {"num_attrs": feats.shape[1], "arity": np.zeros(feats.shape[1], np.int32)}}
        if self.isDefined(self.labelCol):
            cols[self.getLabelCol()], types[self.getLabelCol()] = DeviceColumn(wa.to(torch.float64), np.float64), "double"
        return DataFrame(cols, types, meta)


_make_accessors(WisdmWindowFeaturizer, _NAMES)


def read_wisdm_raw(path_or_bytes, activities=core.WISDM_ACTIVITIES):
    """Raw WISDM release text (records "user,activity,timestamp,x,y,z;") -> DataFrame(user, activity, timestamp, x, y,
    z) whose columns stay on the GPU until read; activity is the index into `activities`.  Broken records are
    dropped (their count is in df.dropped_records).  Feed the result to WisdmWindowFeaturizer.transform."""
    if isinstance(path_or_bytes, (bytes, bytearray, memoryview)):
        data = bytes(path_or_bytes)
    else:
        with open(path_or_bytes, "rb") as fh:
            data = fh.read()
    (user, act, ts, x, y, z), dropped = core.parse_wisdm_raw(data, activities)
    cols = {"user": DeviceColumn(user, np.int32), "activity": DeviceColumn(act, np.int32), "timestamp": DeviceColumn(ts, np.int64),
            "x": DeviceColumn(x, np.float64), "y": DeviceColumn(y, np.float64), "z": DeviceColumn(z, np.float64)}
    types = {"user": "int", "activity": "int", "timestamp": "bigint", "x": "double", "y": "double", "z": "double"}
    df = DataFrame(cols, types)
    df.dropped_records = dropped
    return df
