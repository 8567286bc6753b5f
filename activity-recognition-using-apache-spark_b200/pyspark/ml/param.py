"""pyspark.ml.param surface (Param / Params / TypeConverters), Spark 2.4 behaviour.

Mirrors upstream python/pyspark/ml/param/__init__.py (un-vendored).  The
reference uses it at Main/main.py:55-66 (kwargs constructors), :149 and :203-204
(``evaluate(df, {evaluator.metricName: "f1"})`` param-map overrides).
"""
import copy as _copy
import uuid

__all__ = ["Param", "Params", "TypeConverters"]


class TypeConverters:
    @staticmethod
    def identity(v):
        return v

    @staticmethod
    def toInt(v):
        if isinstance(v, bool):
            raise TypeError("Could not convert %r to int" % (v,))
        if isinstance(v, int):
            return v
        if isinstance(v, float) and v == int(v):
            return int(v)
        try:
            import numpy as np
            if isinstance(v, np.integer):
                return int(v)
        except ImportError:  # pragma: no cover
            pass
        raise TypeError("Could not convert %r to int" % (v,))

    @staticmethod
    def toFloat(v):
        if isinstance(v, bool):
            raise TypeError("Could not convert %r to float" % (v,))
        try:
            return float(v)
        except (TypeError, ValueError):
            raise TypeError("Could not convert %r to float" % (v,))

    @staticmethod
    def toString(v):
        if isinstance(v, str):
            return v
        raise TypeError("Could not convert %r to string type" % (v,))

    @staticmethod
    def toBoolean(v):
        if isinstance(v, bool):
            return v
        raise TypeError("Boolean Param requires value of type bool. Found %s." % type(v))

    @staticmethod
    def toListString(v):
        if isinstance(v, (list, tuple)) and all(isinstance(x, str) for x in v):
            return list(v)
        raise TypeError("Could not convert %r to list of strings" % (v,))

    @staticmethod
    def toListFloat(v):
        return [TypeConverters.toFloat(x) for x in v]

    @staticmethod
    def toListInt(v):
        return [TypeConverters.toInt(x) for x in v]


class Param:
    """A param with self-contained documentation."""

    def __init__(self, parent, name, doc, typeConverter=None):
        self.parent = parent.uid if isinstance(parent, Params) else parent
        self.name = str(name)
        self.doc = str(doc)
        self.typeConverter = TypeConverters.identity if typeConverter is None else typeConverter

    def _copy_new_parent(self, parent):
        if self.parent == "undefined":
            param = _copy.copy(self)
            param.parent = parent.uid
            return param
        raise ValueError("Cannot copy from non-dummy parent %s." % parent)

    def __str__(self):
        return str(self.parent) + "__" + self.name

    def __repr__(self):
        return "Param(parent=%r, name=%r, doc=%r)" % (self.parent, self.name, self.doc)

    def __hash__(self):
        return hash(str(self))

    def __eq__(self, other):
        return isinstance(other, Param) and self.parent == other.parent and self.name == other.name


class Params:
    """Components that take parameters; also gives each instance a Spark-2.4 style uid."""

    _dummy_parent = "undefined"

    def __init__(self):
        self._paramMap = {}
        self._defaultParamMap = {}
        self._params = None
        if not hasattr(self, "uid"):
            self.uid = self._randomUID()
        self._copy_params()

    @classmethod
    def _randomUID(cls):
        # pyspark 2.x: ClassName + "_" + uuid4().hex[12:]  (20 hex chars; result.txt:141,231,276)
        return str(cls.__name__ + "_" + uuid.uuid4().hex[12:])

    def _copy_params(self):
        cls = type(self)
        for name in dir(cls):
            attr = getattr(cls, name, None)
            if isinstance(attr, Param):
                setattr(self, name, attr._copy_new_parent(self))

    @property
    def params(self):
        if self._params is None:
            self._params = sorted(
                [getattr(self, x) for x in dir(self)
                 if x != "params" and not isinstance(getattr(type(self), x, None), property)
                 and isinstance(getattr(self, x), Param)],
                key=lambda p: p.name)
        return self._params

    def explainParam(self, param):
        param = self._resolveParam(param)
        values = []
        if self.isDefined(param):
            if param in self._defaultParamMap:
                values.append("default: %s" % self._defaultParamMap[param])
            if param in self._paramMap:
                values.append("current: %s" % self._paramMap[param])
        else:
            values.append("undefined")
        return "%s: %s (%s)" % (param.name, param.doc, ", ".join(values))

    def explainParams(self):
        return "\n".join(self.explainParam(p) for p in self.params)

    def getParam(self, paramName):
        param = getattr(self, paramName, None)
        if isinstance(param, Param):
            return param
        raise ValueError("Cannot find param with name %s." % paramName)

    def isSet(self, param):
        return self._resolveParam(param) in self._paramMap

    def hasDefault(self, param):
        return self._resolveParam(param) in self._defaultParamMap

    def isDefined(self, param):
        return self.isSet(param) or self.hasDefault(param)

    def hasParam(self, paramName):
        if isinstance(paramName, str):
            return isinstance(getattr(self, paramName, None), Param)
        raise TypeError("hasParam(): paramName must be a string")

    def getOrDefault(self, param):
        param = self._resolveParam(param)
        if param in self._paramMap:
            return self._paramMap[param]
        return self._defaultParamMap[param]

    def extractParamMap(self, extra=None):
        paramMap = self._defaultParamMap.copy()
        paramMap.update(self._paramMap)
        if extra:
            paramMap.update(extra)
        return paramMap

    def copy(self, extra=None):
        that = _copy.copy(self)
        that._paramMap = {}
        that._defaultParamMap = {}
        return self._copyValues(that, extra)

    def set(self, param, value):
        self._shouldOwn(param)
        self._paramMap[param] = param.typeConverter(value)

    def _shouldOwn(self, param):
        if not (self.uid == param.parent and self.hasParam(param.name)):
            raise ValueError("Param %r does not belong to %r." % (param, self))

    def _resolveParam(self, param):
        if isinstance(param, Param):
            self._shouldOwn(param)
            return param
        if isinstance(param, str):
            return self.getParam(param)
        raise ValueError("Cannot resolve %r as a param." % (param,))

    def _set(self, **kwargs):
        for name, value in kwargs.items():
            p = getattr(self, name)
            if value is not None:
                try:
                    value = p.typeConverter(value)
                except (TypeError, ValueError) as e:
                    raise TypeError('Invalid param value given for param "%s". %s' % (p.name, e))
            self._paramMap[p] = value
        return self

    def _clear(self, param):
        self._paramMap.pop(param, None)

    def _setDefault(self, **kwargs):
        for name, value in kwargs.items():
            p = getattr(self, name)
            if value is not None:
                value = p.typeConverter(value)
            self._defaultParamMap[p] = value
        return self

    def _copyValues(self, to, extra=None):
        paramMap = self._paramMap.copy()
        if extra is not None:
            paramMap.update(extra)
        for param in self.params:
            if param in self._defaultParamMap and to.hasParam(param.name):
                to._defaultParamMap[to.getParam(param.name)] = self._defaultParamMap[param]
            if param in paramMap and to.hasParam(param.name):
                to._set(**{param.name: paramMap[param]})
        return to

    def _resetUid(self, newUid):
        newUid = str(newUid)
        self.uid = newUid
        newDefault, newMap = {}, {}
        for param in self.params:
            newParam = _copy.copy(param)
            newParam.parent = newUid
            if param in self._defaultParamMap:
                newDefault[newParam] = self._defaultParamMap[param]
            if param in self._paramMap:
                newMap[newParam] = self._paramMap[param]
            param.parent = newUid
        self._defaultParamMap, self._paramMap = newDefault, newMap
        return self


def _make_accessors(cls, names):
    """Generate pyspark-style getX/setX for the named params."""
    for n in names:
        cap = n[0].upper() + n[1:]

        def getter(self, _n=n):
            return self.getOrDefault(getattr(self, _n))

        def setter(self, value, _n=n):
            return self._set(**{_n: value})

        if not hasattr(cls, "get" + cap):
            setattr(cls, "get" + cap, getter)
        if not hasattr(cls, "set" + cap):
            setattr(cls, "set" + cap, setter)
    return cls
