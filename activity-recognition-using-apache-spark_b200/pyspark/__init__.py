"""`pyspark`-shaped facade in front of the B200 library.

Only the surface Main/main.py imports exists: SparkConf / SparkContext stubs here, the
columnar DataFrame in `pyspark.sql`, and the `pyspark.ml` Estimator/Transformer API whose
DecisionTree / RandomForest fit and transform run on the GPU through the C ABI in
include/har_b200.h.  There is no JVM, no py4j and no scheduler behind it.
"""
__version__ = "2.4.0+har_b200"


class SparkConf:
    def __init__(self, loadDefaults=True, _jvm=None, _jconf=None):
        self._conf = {}

    def setAppName(self, value):
        self._conf["spark.app.name"] = value
        return self

    def setMaster(self, value):
        self._conf["spark.master"] = value
        return self

    def set(self, key, value):
        self._conf[key] = value
        return self

    def get(self, key, defaultValue=None):
        return self._conf.get(key, defaultValue)

    def getAll(self):
        return list(self._conf.items())


class SparkContext:
    _active = None

    def __init__(self, master=None, appName=None, conf=None, **kwargs):
        self._conf = conf or SparkConf()
        if master:
            self._conf.setMaster(master)
        if appName:
            self._conf.setAppName(appName)
        SparkContext._active = self

    @classmethod
    def getOrCreate(cls, conf=None):
        return cls._active or cls(conf=conf)

    @property
    def master(self):
        return self._conf.get("spark.master")

    @property
    def appName(self):
        return self._conf.get("spark.app.name")

    def setLogLevel(self, level):
        pass

    def stop(self):
        SparkContext._active = None


__all__ = ["SparkConf", "SparkContext"]
