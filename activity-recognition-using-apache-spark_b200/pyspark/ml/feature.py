"""pyspark.ml.feature stages the reference pipeline uses (Main/main.py:51-66) — host-side
restatements of Spark 2.4 behaviour that decide the feature layout the tree kernels train on:

* StringIndexer            (tie order = Scala immutable.HashMap trie order)
* OneHotEncoderEstimator   (dropLast, binary attributes)
* VectorAssembler          (zeros not stored, attribute metadata carried through)

Pinned by result.txt:110-125 (one-hot indices of the first sorted rows).
The window-feature stage (`WisdmWindowFeaturizer`) is re-exported from har_b200.
"""
import numpy as np

from . import Estimator, Model, Transformer
from .linalg import VectorColumn
from .param import Param, Params, TypeConverters, _make_accessors

__all__ = ["StringIndexer", "StringIndexerModel", "OneHotEncoderEstimator", "OneHotEncoderModel",
           "VectorAssembler", "IndexToString", "WisdmWindowFeaturizer"]

_M32 = 0xFFFFFFFF


def _java_string_hash(s):
    h = 0
    data = s.encode("utf-16-be")
    for i in range(0, len(data), 2):
        h = (31 * h + ((data[i] << 8) | data[i + 1])) & _M32
    return h


def _scala_hashmap_improve(h):
    h = (h + (~(h << 9) & _M32)) & _M32
    h ^= h >> 14
    h = (h + (h << 4)) & _M32
    return h ^ (h >> 10)


def scala_trie_order_key(s):
    """Iteration position of key `s` in a Scala 2.11 immutable.HashMap: base-32 digits of the
    improved hash, least-significant digit first."""
    h = _scala_hashmap_improve(_java_string_hash(s))
    return tuple((h >> (5 * k)) & 31 for k in range(7))


def _spark_str(v, dtype):
    """Column value -> the string StringIndexer sees (cast to StringType)."""
    if dtype == "string":
        return v
    if dtype == "double":
        from .linalg import java_double_str
        return java_double_str(v)
    if dtype == "boolean":
        return "true" if v else "false"
    return str(int(v))


class _HasInputOutputCol(Params):
    inputCol = Param(Params._dummy_parent, "inputCol", "input column name.", TypeConverters.toString)
    outputCol = Param(Params._dummy_parent, "outputCol", "output column name.", TypeConverters.toString)
    handleInvalid = Param(Params._dummy_parent, "handleInvalid",
                          "how to handle invalid entries. Options are skip, error, keep.",
                          TypeConverters.toString)


_make_accessors(_HasInputOutputCol, ["inputCol", "outputCol", "handleInvalid"])


class StringIndexer(Estimator, _HasInputOutputCol):
    stringOrderType = Param(Params._dummy_parent, "stringOrderType",
                            "How to order labels of string column. Supported options: frequencyDesc, "
                            "frequencyAsc, alphabetDesc, alphabetAsc.", TypeConverters.toString)

    def __init__(self, inputCol=None, outputCol=None, handleInvalid="error", stringOrderType="frequencyDesc"):
        super().__init__()
        self._setDefault(handleInvalid="error", stringOrderType="frequencyDesc")
        self.setParams(inputCol=inputCol, outputCol=outputCol, handleInvalid=handleInvalid,
                       stringOrderType=stringOrderType)

    def setParams(self, **kw):
        return self._set(**{k: v for k, v in kw.items() if v is not None})

    def _fit(self, dataset):
        name = dataset._resolve(self.getInputCol())
        col, t = dataset._cols[name], dataset._types[name]
        counts = {}
        for v in col:
            if v is None:
                continue
            s = _spark_str(v, t)
            counts[s] = counts.get(s, 0) + 1
        order = self.getOrDefault(self.stringOrderType)
        keys = sorted(counts, key=scala_trie_order_key)          # HashMap iteration order
        if order == "frequencyDesc":
            keys.sort(key=lambda k: -counts[k])                  # stable
        elif order == "frequencyAsc":
            keys.sort(key=lambda k: counts[k])
        elif order == "alphabetDesc":
            keys = sorted(counts, reverse=True)
        elif order == "alphabetAsc":
            keys = sorted(counts)
        else:
            raise ValueError("unsupported stringOrderType %r" % order)
        model = StringIndexerModel(keys)
        return self._copyValues(model)


_make_accessors(StringIndexer, ["stringOrderType"])


class StringIndexerModel(Model, _HasInputOutputCol):
    stringOrderType = StringIndexer.stringOrderType

    def __init__(self, labels=()):
        super().__init__()
        self._setDefault(handleInvalid="error", stringOrderType="frequencyDesc")
        self._labels = list(labels)

    @property
    def labels(self):
        return list(self._labels)

    def _transform(self, dataset):
        name = dataset._resolve(self.getInputCol())
        col, t = dataset._cols[name], dataset._types[name]
        pos = {s: float(i) for i, s in enumerate(self._labels)}
        how = self.getOrDefault(self.handleInvalid)
        out = np.empty(len(col), dtype=np.float64)
        keep = np.ones(len(col), dtype=bool)
        for i, v in enumerate(col):
            s = None if v is None else _spark_str(v, t)
            if s in pos:
                out[i] = pos[s]
            elif how == "keep":
                out[i] = float(len(self._labels))
            elif how == "skip":
                keep[i] = False
            else:
                raise ValueError("Unseen label: %s.  To handle unseen labels, set Param handleInvalid to keep." % s)
        labels = self._labels + (["__unknown"] if how == "keep" else [])
        meta = {"ml_attr": {"type": "nominal", "name": self.getOutputCol(), "vals": labels}, "_nonnull": True}
        df = dataset.withColumn(self.getOutputCol(), out, "double", meta)
        if not keep.all():
            df = df._take_rows(np.nonzero(keep)[0])
        return df


class IndexToString(Transformer, _HasInputOutputCol):
    labels = Param(Params._dummy_parent, "labels", "Optional array of labels", TypeConverters.toListString)

    def __init__(self, inputCol=None, outputCol=None, labels=None):
        super().__init__()
        self._set(**{k: v for k, v in dict(inputCol=inputCol, outputCol=outputCol, labels=labels).items()
                     if v is not None})

    def _transform(self, dataset):
        name = dataset._resolve(self.getInputCol())
        labels = (self.getOrDefault(self.labels) if self.isDefined(self.labels)
                  else dataset._meta[name]["ml_attr"]["vals"])
        out = np.array([labels[int(v)] for v in dataset._cols[name]], dtype=object)
        return dataset.withColumn(self.getOutputCol(), out, "string")


class _OneHotParams(Params):
    inputCols = Param(Params._dummy_parent, "inputCols", "input column names.", TypeConverters.toListString)
    outputCols = Param(Params._dummy_parent, "outputCols", "output column names.", TypeConverters.toListString)
    dropLast = Param(Params._dummy_parent, "dropLast", "whether to drop the last category",
                     TypeConverters.toBoolean)
    handleInvalid = Param(Params._dummy_parent, "handleInvalid", "How to handle invalid data: error or keep.",
                          TypeConverters.toString)


_make_accessors(_OneHotParams, ["inputCols", "outputCols", "dropLast", "handleInvalid"])


class OneHotEncoderEstimator(Estimator, _OneHotParams):
    def __init__(self, inputCols=None, outputCols=None, handleInvalid="error", dropLast=True):
        super().__init__()
        self._setDefault(handleInvalid="error", dropLast=True)
        self._set(**{k: v for k, v in dict(inputCols=inputCols, outputCols=outputCols,
                                           handleInvalid=handleInvalid, dropLast=dropLast).items()
                     if v is not None})

    def _fit(self, dataset):
        sizes = []
        for c in self.getInputCols():
            name = dataset._resolve(c)
            attr = dataset._meta.get(name, {}).get("ml_attr")
            if attr and attr.get("type") == "nominal" and "vals" in attr:
                sizes.append(len(attr["vals"]))
            elif attr and attr.get("type") == "binary":
                sizes.append(2)
            else:
                col = dataset._cols[name]
                sizes.append(int(np.max(col)) + 1 if len(col) else 0)
        return self._copyValues(OneHotEncoderModel(sizes))


class OneHotEncoderModel(Model, _OneHotParams):
    def __init__(self, categorySizes=()):
        super().__init__()
        self._setDefault(handleInvalid="error", dropLast=True)
        self._sizes = list(categorySizes)

    @property
    def categorySizes(self):
        return list(self._sizes)

    def _transform(self, dataset):
        keep_invalid = self.getOrDefault(self.handleInvalid) == "keep"
        drop_last = self.getOrDefault(self.dropLast)
        df = dataset
        for c, out, ncat in zip(self.getInputCols(), self.getOutputCols(), self._sizes):
            name = dataset._resolve(c)
            idx = np.asarray(dataset._cols[name], dtype=np.float64)
            if np.any(idx != np.floor(idx)) or np.any(idx < 0):
                raise ValueError("Values from column %s must be indices, but got %s."
                                 % (c, idx[(idx != np.floor(idx)) | (idx < 0)][0]))
            k = idx.astype(np.int64)
            n_eff = ncat + (1 if keep_invalid else 0)
            if keep_invalid:
                k = np.where(k >= ncat, ncat, k)
            elif np.any(k >= ncat):
                raise ValueError("Unseen value: %s. To handle unseen values, set Param handleInvalid to keep."
                                 % float(idx[k >= ncat][0]))
            size = n_eff - 1 if drop_last else n_eff
            has = k < size
            indptr = np.zeros(len(k) + 1, dtype=np.int64)
            np.cumsum(has, out=indptr[1:])
            vc = VectorColumn(size, indptr, k[has].astype(np.int32), np.ones(int(has.sum()), dtype=np.float64))
            meta = {"ml_attr": {"num_attrs": size, "arity": np.full(size, 2, dtype=np.int32)}}
            df = df.withColumn(out, vc, "vector", meta)
        return df


class VectorAssembler(Transformer):
    inputCols = Param(Params._dummy_parent, "inputCols", "input column names.", TypeConverters.toListString)
    outputCol = Param(Params._dummy_parent, "outputCol", "output column name.", TypeConverters.toString)
    handleInvalid = Param(Params._dummy_parent, "handleInvalid", "How to handle invalid data (NULL "
                          "and NaN values): skip, error or keep.", TypeConverters.toString)

    def __init__(self, inputCols=None, outputCol=None, handleInvalid="error"):
        super().__init__()
        self._setDefault(handleInvalid="error")
        self._set(**{k: v for k, v in dict(inputCols=inputCols, outputCol=outputCol,
                                           handleInvalid=handleInvalid).items() if v is not None})

    def _transform(self, dataset):
        n = dataset.count()
        parts = []          # (offset, kind, payload)
        arity = []
        off = 0
        for c in self.getInputCols():
            name = dataset._resolve(c)
            col, t = dataset._cols[name], dataset._types[name]
            if t == "vector":
                parts.append((off, "vec", col))
                a = dataset._meta.get(name, {}).get("ml_attr", {}).get("arity")
                arity.append(np.zeros(col.size, dtype=np.int32) if a is None else np.asarray(a, np.int32))
                off += col.size
            elif t in ("int", "bigint", "double", "boolean"):
                parts.append((off, "num", np.asarray(col, dtype=np.float64)))
                attr = dataset._meta.get(name, {}).get("ml_attr", {})
                ar = len(attr["vals"]) if attr.get("type") == "nominal" and "vals" in attr else (
                    2 if attr.get("type") == "binary" else 0)
                arity.append(np.array([ar], dtype=np.int32))
                off += 1
            else:
                raise TypeError("Data type %s of column %s is not supported." % (t, c))
        size = off
        # build CSR with zeros dropped (Vectors.sparse(...).compressed keeps explicit zeros out)
        counts = np.zeros(n, dtype=np.int64)
        for o, kind, p in parts:
            if kind == "num":
                counts += (p != 0.0)
            elif p.is_sparse:
                counts += np.diff(p.indptr)
            else:
                counts += (p.dense != 0.0).sum(axis=1)
        indptr = np.zeros(n + 1, dtype=np.int64)
        np.cumsum(counts, out=indptr[1:])
        nnz = int(indptr[-1])
        indices = np.empty(nnz, dtype=np.int32)
        values = np.empty(nnz, dtype=np.float64)
        cursor = indptr[:-1].copy()
        for o, kind, p in parts:
            if kind == "num":
                m = p != 0.0
                pos = cursor[m]
                indices[pos] = o
                values[pos] = p[m]
                cursor[m] += 1
            else:
                if p.is_sparse:
                    rows = np.repeat(np.arange(n, dtype=np.int64), np.diff(p.indptr))
                    within = np.arange(len(rows), dtype=np.int64) - p.indptr[rows]
                    pidx, pval = p.indices, p.values
                else:
                    rows, cols_ = np.nonzero(p.dense)
                    pidx, pval = cols_.astype(np.int32), p.dense[rows, cols_]
                    starts = np.zeros(n + 1, dtype=np.int64)
                    np.cumsum(np.bincount(rows, minlength=n), out=starts[1:])
                    within = np.arange(len(rows), dtype=np.int64) - starts[rows]
                pos = cursor[rows] + within
                indices[pos] = pidx + o
                values[pos] = pval
                cursor += np.bincount(rows, minlength=n)
        # .compressed: sparse iff 1.5 * (nnz + 1) < size, decided per row by Spark; the columnar
        # container is homogeneous, so go sparse only when every row would
        if n and np.all(1.5 * (counts + 1.0) < size):
            vc = VectorColumn(size, indptr, indices, values)
        else:
            vc = VectorColumn(size, dense=VectorColumn(size, indptr, indices, values).to_dense())
        meta = {"ml_attr": {"num_attrs": size,
                            "arity": np.concatenate(arity) if arity else np.zeros(0, np.int32)}}
        return dataset.withColumn(self.getOutputCol(), vc, "vector", meta)


_make_accessors(VectorAssembler, ["inputCols", "outputCol", "handleInvalid"])


def __getattr__(name):
    if name == "WisdmWindowFeaturizer":
        from har_b200.window import WisdmWindowFeaturizer
        return WisdmWindowFeaturizer
    raise AttributeError(name)
