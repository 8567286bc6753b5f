"""Tiny column-expression tree: just what Main/main.py builds (`col("count").desc()`,
`predictions['prediction'] == 0`, `col('label') == col('prediction')`, `~(...)`)."""
import numpy as np

__all__ = ["Column", "col", "lit"]


class Column:
    _name = None

    def _eval(self, df):
        raise NotImplementedError

    def _sort_spec(self):
        raise TypeError("not a sortable column expression")

    def _bin(self, other, op):
        return _BinOp(self, other if isinstance(other, Column) else _Lit(other), op)

    def __eq__(self, other):
        return self._bin(other, np.equal)

    def __ne__(self, other):
        return self._bin(other, np.not_equal)

    def __lt__(self, other):
        return self._bin(other, np.less)

    def __le__(self, other):
        return self._bin(other, np.less_equal)

    def __gt__(self, other):
        return self._bin(other, np.greater)

    def __ge__(self, other):
        return self._bin(other, np.greater_equal)

    def __and__(self, other):
        return self._bin(other, np.logical_and)

    def __or__(self, other):
        return self._bin(other, np.logical_or)

    def __invert__(self):
        return _Not(self)

    __hash__ = None


class _ColRef(Column):
    def __init__(self, name, ascending=None):
        self._name = name
        self._asc = ascending

    def _eval(self, df):
        return df._cols[df._resolve(self._name)]

    def desc(self):
        return _ColRef(self._name, False)

    def asc(self):
        return _ColRef(self._name, True)

    def _sort_spec(self):
        return self._name, self._asc

    def alias(self, name):
        return self


class _Lit(Column):
    def __init__(self, v):
        self._v = v

    def _eval(self, df):
        return self._v


class _BinOp(Column):
    def __init__(self, a, b, op):
        self._a, self._b, self._op = a, b, op

    def _eval(self, df):
        return self._op(self._a._eval(df), self._b._eval(df))


class _Not(Column):
    def __init__(self, a):
        self._a = a

    def _eval(self, df):
        return np.logical_not(self._a._eval(df))


def col(name):
    return _ColRef(name)


column = col


def lit(v):
    return _Lit(v)
