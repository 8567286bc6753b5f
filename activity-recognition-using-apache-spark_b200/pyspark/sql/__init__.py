"""Columnar, Arrow-friendly stand-in for the slice of pyspark.sql that Main/main.py touches
(SURVEY.md Appendix D): SQLContext.read...load (CSV with Spark's inferSchema), DataFrame
select / show / groupBy.count / orderBy / take / dtypes / describe / randomSplit / count /
filter / sample.  It is *not* a SQL engine: columns are numpy arrays (numeric), object arrays
(strings) or VectorColumn (ML vectors), i.e. exactly the buffers handed to the CUDA library.

Behaviour that leaks into the hot path's results is restated bit-for-bit and pinned by the
reference's own output (result.txt:4-18 schema, :105-125 split):
  * CSV type inference         -> ``_infer_type``
  * randomSplit = per-partition sort on all orderable columns + BernoulliCellSampler over one
    XORShiftRandom stream     -> ``DataFrame.randomSplit``
"""
import csv as _csv
import re as _re
from collections import OrderedDict

import numpy as np

from ..ml.linalg import VectorColumn, java_double_str
from .._rng import XORShiftRandom
from .column import Column, col, _ColRef

__all__ = ["SQLContext", "SparkSession", "DataFrame", "Row", "Column", "DeviceColumn"]

_SPARK_TYPE_NAMES = {"int": "integer", "bigint": "long", "double": "double", "string": "string",
                     "vector": "vector", "boolean": "boolean"}

_INT_RE = _re.compile(r"^[+-]?\d+$")
_DBL_RE = _re.compile(r"^\s*[+-]?(\d+\.?\d*([eE][+-]?\d+)?|\.\d+([eE][+-]?\d+)?|NaN|Infinity)[dDfF]?\s*$")


class Row(tuple):
    def __new__(cls, values, fields):
        r = tuple.__new__(cls, values)
        r.__fields__ = list(fields)
        return r

    def __getattr__(self, item):
        if item.startswith("__"):
            raise AttributeError(item)
        try:
            return self[self.__fields__.index(item)]
        except ValueError:
            raise AttributeError(item)

    def __getitem__(self, item):
        if isinstance(item, str):
            return tuple.__getitem__(self, self.__fields__.index(item))
        return tuple.__getitem__(self, item)

    def asDict(self):
        return dict(zip(self.__fields__, self))

    def __repr__(self):
        return "Row(%s)" % ", ".join("%s=%r" % (k, v) for k, v in zip(self.__fields__, self))


def _infer_type(values):
    """Spark 2.4 CSVInferSchema.inferField folded over one column (int -> long -> double -> string)."""
    t = None
    for f in values:
        if f is None or f == "":
            continue
        if t in (None, "int") and _INT_RE.match(f) and -2**31 <= int(f) < 2**31:
            t = "int"
            continue
        if t in (None, "int", "bigint") and _INT_RE.match(f) and -2**63 <= int(f) < 2**63:
            t = "bigint"
            continue
        if t in (None, "int", "bigint", "double") and _DBL_RE.match(f):
            t = "double"
            continue
        if t in (None, "boolean") and f.lower() in ("true", "false"):
            t = "boolean"
            continue
        return "string"
    return t or "string"


def _convert(values, t):
    if t == "int":
        return np.array([int(v) for v in values], dtype=np.int32)
    if t == "bigint":
        return np.array([int(v) for v in values], dtype=np.int64)
    if t == "double":
        return np.array([float(v.strip().rstrip("dDfF").replace("Infinity", "inf")) for v in values],
                        dtype=np.float64)
    if t == "boolean":
        return np.array([v.lower() == "true" for v in values], dtype=bool)
    return np.array(values, dtype=object)


class DataFrameReader:
    def __init__(self, ctx):
        self._ctx = ctx
        self._fmt = "csv"
        self._opts = {}

    def format(self, source):
        self._fmt = source
        return self

    def options(self, **opts):
        self._opts.update({k.lower(): v for k, v in opts.items()})
        return self

    def option(self, key, value):
        self._opts[key.lower()] = value
        return self

    def load(self, path):
        if "csv" not in self._fmt:
            raise ValueError("only CSV sources are supported, got %r" % self._fmt)
        return self.csv(path)

    def csv(self, path, header=None, inferSchema=None):
        hdr = str(self._opts.get("header", header if header is not None else "false")).lower() == "true"
        infer = str(self._opts.get("inferschema", inferSchema if inferSchema is not None else "false")
                    ).lower() == "true"
        with open(path, newline="") as fh:
            rows = [r for r in _csv.reader(fh) if r]
        if hdr:
            names, rows = rows[0], rows[1:]
        else:
            names = ["_c%d" % i for i in range(len(rows[0]))]
        cols = OrderedDict()
        types = OrderedDict()
        for j, name in enumerate(names):
            raw = [r[j] for r in rows]
            t = _infer_type(raw) if infer else "string"
            cols[name] = _convert(raw, t)
            types[name] = t
        return DataFrame(cols, types)


class SQLContext:
    def __init__(self, sparkContext=None, sparkSession=None, jsqlContext=None):
        self._sc = sparkContext

    @property
    def read(self):
        return DataFrameReader(self)

    def createDataFrame(self, data, schema=None):
        return createDataFrame(data, schema)


class SparkSession(SQLContext):
    class _Builder:
        def appName(self, _):
            return self

        def master(self, _):
            return self

        def config(self, *a, **k):
            return self

        def getOrCreate(self):
            return SparkSession()

    builder = _Builder()


def createDataFrame(data, schema=None):
    """Rows (tuples / dicts) or a pandas DataFrame -> DataFrame (minimal)."""
    try:
        import pandas as pd
        if isinstance(data, pd.DataFrame):
            schema = list(data.columns)
            data = [tuple(r) for r in data.itertuples(index=False)]
    except ImportError:  # pragma: no cover
        pass
    data = list(data)
    if data and isinstance(data[0], dict):
        schema = schema or list(data[0].keys())
        data = [tuple(d[k] for k in schema) for d in data]
    if schema is None:
        schema = ["_%d" % (i + 1) for i in range(len(data[0]))]
    from ..ml.linalg import Vector
    cols, types = OrderedDict(), OrderedDict()
    for j, name in enumerate(schema):
        vals = [r[j] for r in data]
        v0 = vals[0] if vals else 0.0
        if isinstance(v0, Vector):
            cols[name], types[name] = VectorColumn.from_vectors(vals), "vector"
        elif isinstance(v0, (bool, np.bool_)):
            cols[name], types[name] = np.array(vals, dtype=bool), "boolean"
        elif isinstance(v0, (int, np.integer)):
            cols[name], types[name] = np.array(vals, dtype=np.int64), "bigint"
        elif isinstance(v0, (float, np.floating)):
            cols[name], types[name] = np.array(vals, dtype=np.float64), "double"
        else:
            cols[name], types[name] = np.array(vals, dtype=object), "string"
    return DataFrame(cols, types)


class DeviceColumn:
    """A scalar column whose values still live in HBM (a 1-d CUDA tensor); it is copied to the host the
    first time anything on the host reads it — the facade's counterpart of Spark's lazy transformations."""

    def __init__(self, tensor, dtype):
        self.tensor, self.dtype = tensor, np.dtype(dtype)

    def __len__(self):
        return int(self.tensor.shape[0])

    def materialize(self):
        from har_b200 import core
        return np.ascontiguousarray(core.to_host(self.tensor.contiguous()), dtype=self.dtype)


class _Cols(OrderedDict):
    """Column store: name -> numpy array / VectorColumn, with DeviceColumn entries materialised on first read."""

    def __init__(self, other=()):
        OrderedDict.__init__(self)
        items = dict.items(other) if isinstance(other, dict) else other      # raw entries: nothing is materialised
        for k, v in items:
            OrderedDict.__setitem__(self, k, v)

    def raw(self, k):
        return OrderedDict.__getitem__(self, k)

    def __getitem__(self, k):
        v = OrderedDict.__getitem__(self, k)
        if isinstance(v, DeviceColumn):
            v = v.materialize()
            OrderedDict.__setitem__(self, k, v)
        return v

    def get(self, k, default=None):
        return self[k] if k in self else default

    def values(self):
        return [self[k] for k in self.keys()]

    def items(self):
        return [(k, self[k]) for k in self.keys()]


def _fmt_cell(v, t):
    if v is None:
        return "null"
    if t == "double":
        return java_double_str(v)
    if t == "boolean":
        return "true" if v else "false"
    return str(v)


class GroupedData:
    def __init__(self, df, names, typed=None):
        self._df, self._names = df, names
        self._typed = typed or names           # Spark keeps the spelling the caller used (result.txt:34 "activity")

    def count(self):
        df = self._df
        keys = list(zip(*[df._cols[n] for n in self._names]))
        counts = OrderedDict()
        for k in keys:
            counts[k] = counts.get(k, 0) + 1
        cols, types = OrderedDict(), OrderedDict()
        for j, (n, shown) in enumerate(zip(self._names, self._typed)):
            vals = [k[j] for k in counts]
            cols[shown] = np.array(vals, dtype=df._cols[n].dtype)
            types[shown] = df._types[n]
        cols["count"] = np.array(list(counts.values()), dtype=np.int64)
        types["count"] = "bigint"
        return DataFrame(cols, types)


class DataFrame:
    def __init__(self, cols, types, meta=None, partitions=None):
        self._cols = _Cols(cols)
        self._types = OrderedDict(types)
        self._meta = dict(meta or {})
        n = len(self._cols.raw(next(iter(self._cols.keys())))) if self._cols else 0
        # row offsets of the Spark partitions this frame stands for (one by default: the
        # reference's 1.2 MB CSV is a single input partition)
        self._partitions = list(partitions) if partitions is not None else [0, n]

    # ---- structure -------------------------------------------------------------------------
    @property
    def columns(self):
        return list(self._cols.keys())

    @property
    def dtypes(self):
        return [(n, self._types[n]) for n in self._cols]

    @property
    def schema(self):
        return self.dtypes

    def _resolve(self, name):
        if name in self._cols:
            return name
        low = name.lower()
        for n in self._cols:                      # Spark resolves column names case-insensitively
            if n.lower() == low:
                return n
        raise KeyError("cannot resolve '`%s`' given input columns: [%s]" % (name, ", ".join(self._cols)))

    def __getitem__(self, item):
        if isinstance(item, str):
            return _ColRef(self._resolve(item))
        if isinstance(item, Column):
            return self.filter(item)
        if isinstance(item, (list, tuple)):
            return self.select(*item)
        raise TypeError("unexpected item type: %s" % type(item))

    def __getattr__(self, name):
        if name.startswith("_"):
            raise AttributeError(name)
        try:
            return _ColRef(self._resolve(name))
        except KeyError:
            raise AttributeError(name)

    def count(self):
        return self._partitions[-1]

    def _take_rows(self, idx, partitions=None):
        idx = np.asarray(idx, dtype=np.int64)
        cols = OrderedDict()
        for n, c in self._cols.items():
            cols[n] = c.take(idx) if isinstance(c, VectorColumn) else c[idx]
        return DataFrame(cols, self._types, self._meta, partitions)

    def select(self, *names):
        if len(names) == 1 and isinstance(names[0], (list, tuple)):
            names = names[0]
        cols, types, meta = OrderedDict(), OrderedDict(), {}
        for n in names:
            if isinstance(n, Column):
                n = n._name
            r = self._resolve(n)
            cols[r], types[r] = self._cols.raw(r), self._types[r]
            if r in self._meta:
                meta[r] = self._meta[r]
        return DataFrame(cols, types, meta, self._partitions)

    def drop(self, *names):
        drop = {self._resolve(n) for n in names if n in self._cols or n.lower() in [c.lower() for c in self._cols]}
        return self.select([n for n in self._cols if n not in drop])

    def withColumn(self, name, data, dtype=None, metadata=None):
        """Facade extension: attach a materialised column (numpy array / VectorColumn)."""
        cols, types, meta = _Cols(self._cols), OrderedDict(self._types), dict(self._meta)
        if isinstance(data, Column):
            data = data._eval(self)
        if dtype is None:
            if isinstance(data, VectorColumn):
                dtype = "vector"
            elif isinstance(data, DeviceColumn):
                dtype = {"f": "double", "i": "bigint"}.get(data.dtype.kind, "double")
            else:
                data = np.asarray(data)
                dtype = {"f": "double", "i": "bigint", "u": "bigint", "b": "boolean"}.get(data.dtype.kind, "string")
                if data.dtype == np.int32:
                    dtype = "int"
        cols[name], types[name] = data, dtype
        if metadata is not None:
            meta[name] = metadata
        else:
            meta.pop(name, None)
        return DataFrame(cols, types, meta, self._partitions)

    def repartition(self, numPartitions):
        """Declare how many Spark partitions the rows stand for (contiguous, equal ranges).
        Only the per-partition random streams (bagging, sampling) depend on it."""
        n = self.count()
        p = max(1, int(numPartitions))
        bounds = [(n * k) // p for k in range(p + 1)]
        return DataFrame(self._cols, self._types, self._meta, bounds)

    def getNumPartitions(self):
        return len(self._partitions) - 1

    # ---- actions ---------------------------------------------------------------------------
    def _row(self, i):
        vals = []
        for n, c in self._cols.items():
            v = c[i]
            if isinstance(v, np.generic):
                v = v.item()
            vals.append(v)
        return Row(vals, self._cols.keys())

    def take(self, n):
        return [self._row(i) for i in range(min(n, self.count()))]

    def head(self, n=None):
        if n is None:
            r = self.take(1)
            return r[0] if r else None
        return self.take(n)

    def first(self):
        return self.head()

    def collect(self):
        return self.take(self.count())

    def limit(self, n):
        return self._take_rows(np.arange(min(n, self.count())))

    def toPandas(self):
        import pandas as pd
        data = OrderedDict()
        for n, c in self._cols.items():
            data[n] = list(c) if isinstance(c, VectorColumn) else c
        return pd.DataFrame(data)

    def toArrow(self):
        """Numeric / string columns as a pyarrow.Table (vector columns as list<double>)."""
        import pyarrow as pa
        arrays, names = [], []
        for n, c in self._cols.items():
            if isinstance(c, VectorColumn):
                arrays.append(pa.array([v.toArray() for v in c]))
            else:
                arrays.append(pa.array(c.tolist() if c.dtype == object else c))
            names.append(n)
        return pa.table(arrays, names=names)

    def printSchema(self):
        print(self._schema_string(), end="")

    def _schema_string(self):
        out = ["root"]
        for n, t in self._types.items():
            nullable = "false" if self._meta.get(n, {}).get("_nonnull") else "true"
            out.append(" |-- %s: %s (nullable = %s)" % (n, _SPARK_TYPE_NAMES.get(t, t), nullable))
        return "\n".join(out) + "\n\n"

    def show(self, n=20, truncate=True, vertical=False):
        print(self._show_string(n, truncate), end="")

    def _show_string(self, n=20, truncate=True):
        trunc = 20 if truncate is True else (0 if truncate is False else int(truncate))
        total = self.count()
        k = min(n, total)
        names = list(self._cols.keys())
        table = [names]
        for i in range(k):
            cells = []
            for name in names:
                c, t = self._cols[name], self._types[name]
                cells.append(_fmt_cell(c[i], t))
            table.append(cells)
        if trunc > 0:
            def cut(s):
                if len(s) > trunc:
                    return s[:trunc] if trunc < 4 else s[: trunc - 3] + "..."
                return s
            table = [[cut(s) for s in row] for row in table]
        widths = [max(3, max(len(row[j]) for row in table)) for j in range(len(names))]
        sep = "+" + "+".join("-" * w for w in widths) + "+\n"
        pad = (lambda s, w: s.rjust(w)) if trunc > 0 or truncate is True else (lambda s, w: s.ljust(w))
        out = [sep, "|" + "|".join(pad(s, w) for s, w in zip(table[0], widths)) + "|\n", sep]
        for row in table[1:]:
            out.append("|" + "|".join(pad(s, w) for s, w in zip(row, widths)) + "|\n")
        out.append(sep)
        if total > k:
            out.append("only showing top %d row%s\n" % (k, "" if k == 1 else "s"))
        out.append("\n")
        return "".join(out)

    def groupBy(self, *names):
        typed = [n._name if isinstance(n, Column) else n for n in names]
        return GroupedData(self, [self._resolve(n) for n in typed], typed)

    groupby = groupBy

    def _sort_key_arrays(self, name):
        c, t = self._cols[name], self._types[name]
        if t == "vector":
            return [_vector_sort_key(c, i) for i in range(len(c))]
        if t == "string":
            return [None if v is None else str(v).encode("utf-8") for v in c]
        return c.tolist()

    def orderBy(self, *cols, **kwargs):
        asc = kwargs.get("ascending", True)
        if len(cols) == 1 and isinstance(cols[0], (list, tuple)):
            cols = cols[0]
        specs = []
        for j, c in enumerate(cols):
            a = asc[j] if isinstance(asc, (list, tuple)) else asc
            if isinstance(c, Column):
                name, a2 = c._sort_spec()
                specs.append((self._resolve(name), a if a2 is None else a2))
            else:
                specs.append((self._resolve(c), bool(a)))
        idx = list(range(self.count()))
        for name, a in reversed(specs):          # stable multi-key sort
            keys = self._sort_key_arrays(name)
            idx.sort(key=lambda i: keys[i], reverse=not a)
        return self._take_rows(idx, self._partitions)

    sort = orderBy

    def filter(self, cond):
        mask = np.asarray(cond._eval(self), dtype=bool)
        return self._take_rows(np.nonzero(mask)[0])

    where = filter

    def describe(self, *names):
        names = [self._resolve(n) for n in names] or [n for n, t in self._types.items()
                                                      if t in ("int", "bigint", "double", "string")]
        cols, types = OrderedDict(), OrderedDict()
        cols["summary"] = np.array(["count", "mean", "stddev", "min", "max"], dtype=object)
        types["summary"] = "string"
        for n in names:
            c, t = self._cols[n], self._types[n]
            if t == "string":
                vals = [v for v in c if v is not None]
                out = [str(len(vals)), None, None, min(vals) if vals else None, max(vals) if vals else None]
            else:
                a = c.astype(np.float64)
                cnt = len(a)
                # Spark's aggregates, row by row in one partition: Average = running sum / count;
                # StddevSamp = CentralMomentAgg's online update (n, avg, m2), sqrt(m2 / (n - 1))
                total, avg, m2, k = 0.0, 0.0, 0.0, 0.0
                for x in a.tolist():
                    total += x
                    k += 1.0
                    delta = x - avg
                    delta_n = delta / k
                    avg += delta_n
                    m2 += delta * (delta - delta_n)
                mean = total / cnt if cnt else float("nan")
                sd = float(np.sqrt(m2 / (cnt - 1))) if cnt > 1 else float("nan")
                mn, mx = (c.min(), c.max()) if cnt else (None, None)
                fm = (lambda v: str(int(v))) if t in ("int", "bigint") else (lambda v: java_double_str(v))
                out = [str(cnt), java_double_str(mean), java_double_str(sd), fm(mn), fm(mx)]
            cols[n] = np.array(out, dtype=object)
            types[n] = "string"
        return DataFrame(cols, types)

    def sample(self, withReplacement=None, fraction=None, seed=None):
        if isinstance(withReplacement, float) and fraction is None:
            withReplacement, fraction = False, withReplacement
        if seed is None:
            import random
            seed = random.getrandbits(63)
        keep = []
        for p in range(len(self._partitions) - 1):
            rng = XORShiftRandom(seed + p)
            for i in range(self._partitions[p], self._partitions[p + 1]):
                if rng.nextDouble() < fraction:
                    keep.append(i)
        return self._take_rows(keep)

    def randomSplit(self, weights, seed=None):
        """Dataset.randomSplit: sort every partition on all orderable columns, then
        split i keeps row r iff lb_i <= x_r < ub_i, x_r the r-th nextDouble of
        XORShiftRandom(seed + partitionIndex)."""
        if any(w < 0.0 for w in weights):
            raise ValueError("Weights must be positive. Found weight value: %s" % min(weights))
        if seed is None:
            import random
            seed = random.getrandbits(63)
        s = float(sum(weights))
        bounds = [0.0]
        for w in weights:
            bounds.append(bounds[-1] + w / s)
        keys = [self._sort_key_arrays(n) for n in self._cols]
        order, draws = [], []
        for p in range(len(self._partitions) - 1):
            rows = list(range(self._partitions[p], self._partitions[p + 1]))
            rows.sort(key=lambda i: tuple(k[i] for k in keys))
            rng = XORShiftRandom(seed + p)
            order.extend(rows)
            draws.extend(rng.nextDouble() for _ in rows)
        order = np.asarray(order, dtype=np.int64)
        draws = np.asarray(draws, dtype=np.float64)
        out = []
        for i in range(len(weights)):
            m = (draws >= bounds[i]) & (draws < bounds[i + 1])
            out.append(self._take_rows(order[m]))
        return out

    def cache(self):
        return self

    persist = cache

    def unpersist(self, blocking=False):
        return self


def _vector_sort_key(c, i):
    """SQL struct order of VectorUDT: (type: byte, size: int, indices: array<int>, values: array<double>),
    nulls first (dense vectors have null size / indices)."""
    if c.is_sparse:
        a, b = c.indptr[i], c.indptr[i + 1]
        return (0, (1, c.size), (1, tuple(c.indices[a:b].tolist())), tuple(c.values[a:b].tolist()))
    return (1, (0, 0), (0, ()), tuple(c.dense[i].tolist()))
