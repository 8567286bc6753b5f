"""pyspark.ml.linalg surface: DenseVector / SparseVector / Vectors, plus the columnar
container (VectorColumn) the facade stores a whole ``features`` column in.

Formatting follows Spark 2.4 (`(3100,[0,934,2335],[1.0,1.0,8.4])`, result.txt:110-114).
"""
import numpy as np

__all__ = ["Vector", "DenseVector", "SparseVector", "Vectors", "VectorColumn", "VectorUDT"]


def java_double_str(x):
    """java.lang.Double.toString for the values a DataFrame prints."""
    x = float(x)
    if x != x:
        return "NaN"
    if x == float("inf"):
        return "Infinity"
    if x == float("-inf"):
        return "-Infinity"
    if x == 0.0:
        return "-0.0" if np.signbit(x) else "0.0"
    a = abs(x)
    r = repr(a)
    if "e" in r:
        mant, exp = r.split("e")
        exp = int(exp)
    else:
        mant, exp = r, 0
    if "." not in mant:
        mant += ".0"
    ip, fp = mant.split(".")
    digits = (ip + fp).lstrip("0")
    # decimal exponent of the first significant digit
    lead = len(ip.lstrip("0"))
    if lead == 0:
        stripped = fp
        lead = -(len(stripped) - len(stripped.lstrip("0")))
    e10 = lead + exp - 1
    digits = digits.rstrip("0") or "0"
    if 1e-3 <= a < 1e7:
        if e10 >= 0:
            ipart = digits[: e10 + 1].ljust(e10 + 1, "0")
            fpart = digits[e10 + 1:] or "0"
        else:
            ipart = "0"
            fpart = "0" * (-e10 - 1) + digits
        s = ipart + "." + fpart
    else:
        s = digits[0] + "." + (digits[1:] or "0") + "E" + str(e10)
    return ("-" if x < 0 else "") + s


class Vector:
    def toArray(self):
        raise NotImplementedError


class DenseVector(Vector):
    def __init__(self, ar):
        self.array = np.asarray(ar, dtype=np.float64).reshape(-1)

    @property
    def values(self):
        return self.array

    @property
    def size(self):
        return len(self.array)

    def toArray(self):
        return self.array

    def __len__(self):
        return len(self.array)

    def __getitem__(self, i):
        return self.array[i]

    def __iter__(self):
        return iter(self.array)

    def __str__(self):
        return "[" + ",".join(java_double_str(v) for v in self.array) + "]"

    def __repr__(self):
        return "DenseVector([%s])" % ", ".join(repr(float(v)) for v in self.array)

    def __eq__(self, other):
        if isinstance(other, Vector):
            a, b = self.toArray(), other.toArray()
            return a.shape == b.shape and bool(np.all(a == b))
        return False

    def __hash__(self):
        return hash(self.array.tobytes())

    def argmax(self):
        return int(np.argmax(self.array)) if len(self.array) else -1


class SparseVector(Vector):
    def __init__(self, size, *args):
        self.size = int(size)
        if len(args) == 1:
            pairs = args[0]
            if isinstance(pairs, dict):
                pairs = sorted(pairs.items())
            else:
                pairs = sorted(pairs)
            self.indices = np.array([p[0] for p in pairs], dtype=np.int32)
            self.values = np.array([p[1] for p in pairs], dtype=np.float64)
        else:
            self.indices = np.asarray(args[0], dtype=np.int32)
            self.values = np.asarray(args[1], dtype=np.float64)
        if len(self.indices) != len(self.values):
            raise ValueError("index and value arrays not the same length")

    def toArray(self):
        a = np.zeros(self.size, dtype=np.float64)
        a[self.indices] = self.values
        return a

    def __len__(self):
        return self.size

    def __getitem__(self, i):
        j = np.searchsorted(self.indices, i)
        if j < len(self.indices) and self.indices[j] == i:
            return self.values[j]
        return 0.0

    def __str__(self):
        return "(%d,[%s],[%s])" % (self.size, ",".join(str(int(i)) for i in self.indices),
                                   ",".join(java_double_str(v) for v in self.values))

    def __repr__(self):
        return "SparseVector(%d, {%s})" % (
            self.size, ", ".join("%d: %s" % (int(i), repr(float(v)))
                                 for i, v in zip(self.indices, self.values)))

    def __eq__(self, other):
        if isinstance(other, Vector):
            a, b = self.toArray(), other.toArray()
            return a.shape == b.shape and bool(np.all(a == b))
        return False

    def __hash__(self):
        return hash((self.size, self.indices.tobytes(), self.values.tobytes()))


class Vectors:
    @staticmethod
    def dense(*elements):
        if len(elements) == 1 and not isinstance(elements[0], (float, int)):
            elements = elements[0]
        return DenseVector(elements)

    @staticmethod
    def sparse(size, *args):
        return SparseVector(size, *args)


class VectorUDT:
    """Only the SQL struct shape matters here (sort order of randomSplit)."""
    @staticmethod
    def simpleString():
        return "vector"


class VectorColumn:
    """A whole column of ML vectors, stored columnar.

    Either CSR (``size``, ``indptr`` int64[n+1], ``indices`` int32[nnz], ``values`` f64[nnz];
    what VectorAssembler's ``.compressed`` gives for the reference's 3100-d rows) or a dense
    row-major ``f64[n, size]`` matrix.  These buffers are what is handed to the CUDA library.
    """

    def __init__(self, size, indptr=None, indices=None, values=None, dense=None, device=None):
        """`device`: an optional CUDA tensor [n, size] holding the same dense values; the host copy is then
        materialised only when something on the host reads it (keeps the column resident in HBM between the
        stages that produce and consume it)."""
        self.size = int(size)
        self._dev = device
        self._dense = None if dense is None else np.ascontiguousarray(dense, dtype=np.float64)
        if dense is None and device is None:
            self.indptr = np.asarray(indptr, dtype=np.int64)
            self.indices = np.asarray(indices, dtype=np.int32)
            self.values = np.asarray(values, dtype=np.float64)
            self._sparse = True
        else:
            self.indptr = self.indices = self.values = None
            self._sparse = False

    @property
    def dense(self):
        if self._dense is None and self._dev is not None:
            from har_b200 import core
            self._dense = core.to_host(self._dev)
        return self._dense

    @property
    def is_sparse(self):
        return self._sparse

    def __len__(self):
        if self.is_sparse:
            return len(self.indptr) - 1
        return int(self._dev.shape[0]) if self._dense is None else self._dense.shape[0]

    def row(self, i):
        if self.is_sparse:
            a, b = self.indptr[i], self.indptr[i + 1]
            return SparseVector(self.size, self.indices[a:b], self.values[a:b])
        return DenseVector(self.dense[i])

    def __getitem__(self, i):
        return self.row(int(i))

    def __iter__(self):
        for i in range(len(self)):
            yield self.row(i)

    def take(self, idx):
        idx = np.asarray(idx, dtype=np.int64)
        if not self.is_sparse:
            return VectorColumn(self.size, dense=self.dense[idx])
        starts, ends = self.indptr[idx], self.indptr[idx + 1]
        lens = ends - starts
        indptr = np.zeros(len(idx) + 1, dtype=np.int64)
        np.cumsum(lens, out=indptr[1:])
        if len(idx):
            gather = np.repeat(starts - indptr[:-1], lens) + np.arange(indptr[-1], dtype=np.int64)
        else:
            gather = np.zeros(0, dtype=np.int64)
        return VectorColumn(self.size, indptr, self.indices[gather], self.values[gather])

    def to_dense(self):
        if not self.is_sparse:
            return self.dense
        n = len(self)
        out = np.zeros((n, self.size), dtype=np.float64)
        rows = np.repeat(np.arange(n, dtype=np.int64), np.diff(self.indptr))
        out[rows, self.indices] = self.values
        return out

    @staticmethod
    def from_vectors(vectors, size=None):
        vectors = list(vectors)
        if size is None:
            size = len(vectors[0]) if vectors else 0
        if all(isinstance(v, DenseVector) for v in vectors) and vectors:
            return VectorColumn(size, dense=np.stack([v.array for v in vectors]))
        indptr = np.zeros(len(vectors) + 1, dtype=np.int64)
        idx, val = [], []
        for k, v in enumerate(vectors):
            if isinstance(v, SparseVector):
                idx.append(v.indices)
                val.append(v.values)
            else:
                a = np.asarray(v.toArray() if isinstance(v, Vector) else v, dtype=np.float64)
                nz = np.nonzero(a)[0]
                idx.append(nz.astype(np.int32))
                val.append(a[nz])
            indptr[k + 1] = indptr[k] + len(idx[-1])
        return VectorColumn(size, indptr,
                            np.concatenate(idx) if idx else np.zeros(0, np.int32),
                            np.concatenate(val) if val else np.zeros(0, np.float64))
