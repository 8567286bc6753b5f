"""pyspark.ml.classification: DecisionTreeClassifier / RandomForestClassifier and their models
(Spark 2.4 params and defaults), with fit and transform running on the GPU through
har_b200 (include/har_b200.h: har_forest_fit / har_forest_predict).

Reference call sites: Main/main.py:297-305 (DT), :478-486 (RF), and via CrossValidator :389-400, :570-581.
LogisticRegression (main.py:115) is outside the hot path: `_host_lr.py` is a host-only stand-in so main.py completes.
"""
import numpy as np

try:
    import torch
except ImportError:  # pragma: no cover
    torch = None

from . import Estimator, Model
from ..sql import DeviceColumn
from .linalg import DenseVector, VectorColumn
from .param import Param, Params, TypeConverters, _make_accessors

from ._host_lr import LogisticRegression, LogisticRegressionModel  # host-only filler, outside the hot path

__all__ = ["DecisionTreeClassifier", "DecisionTreeClassificationModel",
           "RandomForestClassifier", "RandomForestClassificationModel",
           "LogisticRegression", "LogisticRegressionModel"]


class _TreeClassifierParams(Params):
    featuresCol = Param(Params._dummy_parent, "featuresCol", "features column name.", TypeConverters.toString)
    labelCol = Param(Params._dummy_parent, "labelCol", "label column name.", TypeConverters.toString)
    predictionCol = Param(Params._dummy_parent, "predictionCol", "prediction column name.", TypeConverters.toString)
    probabilityCol = Param(Params._dummy_parent, "probabilityCol", "Column name for predicted class conditional "
                           "probabilities.", TypeConverters.toString)
    rawPredictionCol = Param(Params._dummy_parent, "rawPredictionCol", "raw prediction (a.k.a. confidence) column name.",
                             TypeConverters.toString)
    thresholds = Param(Params._dummy_parent, "thresholds", "Thresholds in multi-class classification to adjust the "
                       "probability of predicting each class.", TypeConverters.toListFloat)
    maxDepth = Param(Params._dummy_parent, "maxDepth", "Maximum depth of the tree. (>= 0) E.g., depth 0 means 1 leaf "
                     "node; depth 1 means 1 internal node + 2 leaf nodes.", TypeConverters.toInt)
    maxBins = Param(Params._dummy_parent, "maxBins", "Max number of bins for discretizing continuous features.  Must "
                    "be >=2 and >= number of categories for any categorical feature.", TypeConverters.toInt)
    minInstancesPerNode = Param(Params._dummy_parent, "minInstancesPerNode", "Minimum number of instances each child "
                                "must have after split.", TypeConverters.toInt)
    minInfoGain = Param(Params._dummy_parent, "minInfoGain", "Minimum information gain for a split to be considered "
                        "at a tree node.", TypeConverters.toFloat)
    maxMemoryInMB = Param(Params._dummy_parent, "maxMemoryInMB", "Maximum memory in MB allocated to histogram "
                          "aggregation.", TypeConverters.toInt)
    cacheNodeIds = Param(Params._dummy_parent, "cacheNodeIds", "If false, the algorithm will pass trees to executors "
                         "to match instances with nodes. (accepted, no effect here)", TypeConverters.toBoolean)
    checkpointInterval = Param(Params._dummy_parent, "checkpointInterval", "set checkpoint interval (>= 1) or disable "
                               "checkpoint (-1). (accepted, no effect here)", TypeConverters.toInt)
    impurity = Param(Params._dummy_parent, "impurity", "Criterion used for information gain calculation "
                     "(case-insensitive). Supported options: entropy, gini", TypeConverters.toString)
    seed = Param(Params._dummy_parent, "seed", "random seed.", TypeConverters.toInt)

    def _tree_defaults(self):
        self._setDefault(featuresCol="features", labelCol="label", predictionCol="prediction",
                         probabilityCol="probability", rawPredictionCol="rawPrediction", maxDepth=5, maxBins=32,
                         minInstancesPerNode=1, minInfoGain=0.0, maxMemoryInMB=256, cacheNodeIds=False,
                         checkpointInterval=10, impurity="gini", seed=hash(type(self).__name__))


_TREE_PARAM_NAMES = ["featuresCol", "labelCol", "predictionCol", "probabilityCol", "rawPredictionCol", "thresholds",
                     "maxDepth", "maxBins", "minInstancesPerNode", "minInfoGain", "maxMemoryInMB", "cacheNodeIds",
                     "checkpointInterval", "impurity", "seed"]
_make_accessors(_TreeClassifierParams, _TREE_PARAM_NAMES)


def _num_classes(dataset, label_name, y):
    """Classifier.getNumClasses: nominal metadata if present, else max label + 1 (<= 100)."""
    attr = dataset._meta.get(label_name, {}).get("ml_attr", {})
    if attr.get("type") == "nominal" and "vals" in attr:
        return len(attr["vals"])
    if attr.get("type") == "binary":
        return 2
    if len(y) == 0:
        raise ValueError("ML algorithm was given empty dataset.")
    mx = float(y.max())
    if mx != int(mx) or mx < 0:
        raise ValueError("Classifier found max label value = %s but requires integers in range [0, ... %d)" % (mx, 2**31 - 1))
    if int(mx) + 1 > 100:
        raise ValueError("Classifier inferred %d from label values in column %s, but this exceeded the max numClasses "
                         "(100) allowed to be inferred from values." % (int(mx) + 1, label_name))
    return int(mx) + 1


def _device_features(col):
    """VectorColumn -> dense row-major f64 CUDA tensor (CSR is expanded on the device)."""
    from har_b200 import core
    if col.is_sparse:
        return core.csr_to_dense(col.indptr, col.indices, col.values, len(col), col.size)
    dev = getattr(col, "_dev", None)
    if dev is not None and dev.is_cuda:
        return dev                                   # still resident from the stage that produced it
    col._dev = core.dev(col.dense, np.float64)       # keep it for the next fit / transform on this column
    return col._dev


class _TreeEstimatorBase(Estimator, _TreeClassifierParams):
    _num_trees_default = 1

    def _fit_forest(self, dataset, num_trees, feature_subset, subsampling_rate):
        from har_b200 import core
        fname = dataset._resolve(self.getFeaturesCol())
        lname = dataset._resolve(self.getLabelCol())
        if dataset._types[fname] != "vector":
            raise TypeError("Column %s must be of type vector but was actually %s." % (fname, dataset._types[fname]))
        col = dataset._cols[fname]
        yraw = dataset._cols.raw(lname)
        if isinstance(yraw, DeviceColumn) and yraw.tensor.is_cuda:
            y = yraw.tensor.to(torch.float64)            # label column still resident in HBM
        else:
            y = np.asarray(dataset._cols[lname], dtype=np.float64)
        C = _num_classes(dataset, lname, y)
        arity = dataset._meta.get(fname, {}).get("ml_attr", {}).get("arity")
        imp = self.getOrDefault(self.impurity).lower()
        if imp not in ("gini", "entropy"):
            raise ValueError("impurity given invalid value %r" % imp)
        X = _device_features(col)
        part_off, dist = dataset._partitions, None
        try:
            import torch.distributed as td
            multi = td.is_available() and td.is_initialized() and td.get_world_size() > 1
        except Exception:
            multi = False
        if multi:
            # one process per GPU, each holding its row shard of the frame: a collective fit
            from har_b200.dist import Dist, exchange_shard_layout
            global_rows, p_off, p_total = exchange_shard_layout(len(y), len(part_off) - 1)
            dist = Dist(global_rows, p_off, p_total)
        forest = core.fit_forest(
            X, y, arity=arity, num_classes=C, num_trees=num_trees, max_depth=self.getMaxDepth(),
            max_bins=self.getMaxBins(), min_instances_per_node=self.getMinInstancesPerNode(),
            min_info_gain=self.getMinInfoGain(), impurity=imp, feature_subset=feature_subset,
            subsampling_rate=subsampling_rate, seed=self.getSeed(), max_memory_mb=self.getMaxMemoryInMB(),
            part_off=part_off, dist=dist)
        return forest, col.size, C


class DecisionTreeClassifier(_TreeEstimatorBase):
    """DecisionTreeClassifier(featuresCol="features", labelCol="label", predictionCol="prediction",
    probabilityCol="probability", rawPredictionCol="rawPrediction", maxDepth=5, maxBins=32,
    minInstancesPerNode=1, minInfoGain=0.0, maxMemoryInMB=256, cacheNodeIds=False, checkpointInterval=10,
    impurity="gini", seed=None) — keyword arguments only, as in pyspark; only what is passed is `set`."""
    _KW = tuple(_TREE_PARAM_NAMES)

    def __init__(self, *args, **kwargs):
        super().__init__()
        self._tree_defaults()
        if args:
            raise TypeError("Method __init__ forces keyword arguments.")
        self.setParams(**kwargs)

    def setParams(self, **kw):
        for k in kw:
            if k not in self._KW:
                raise TypeError("__init__() got an unexpected keyword argument '%s'" % k)
        return self._set(**{k: v for k, v in kw.items() if v is not None})

    def _fit(self, dataset):
        forest, F, C = self._fit_forest(dataset, 1, "all", 1.0)
        model = DecisionTreeClassificationModel(forest, 0, F, C)
        model._resetUid(self.uid)
        return self._copyValues(model)


class _TreeModelBase(Model, _TreeClassifierParams):
    def _append_predictions(self, dataset, forest, is_forest):
        fname = dataset._resolve(self.getFeaturesCol())
        col = dataset._cols[fname]
        X = _device_features(col)
        from har_b200 import core
        from ..sql import DeviceColumn
        raw, prob, pred = forest.predict(X, is_forest=is_forest)
        use_thresholds = self.isDefined(self.thresholds) and self.getOrDefault(self.thresholds) is not None
        if not use_thresholds and getattr(raw, "is_cuda", False):
            # output columns stay in HBM until read (rawPrediction / probability as device-backed vector columns)
            df = dataset
            if self.getRawPredictionCol():
                df = df.withColumn(self.getRawPredictionCol(), VectorColumn(forest.C, device=raw), "vector")
            if self.getProbabilityCol():
                df = df.withColumn(self.getProbabilityCol(), VectorColumn(forest.C, device=prob), "vector")
            if self.getPredictionCol():
                df = df.withColumn(self.getPredictionCol(), DeviceColumn(pred, np.float64), "double")
            return df
        raw, prob, pred = core.to_host(raw), core.to_host(prob), core.to_host(pred)
        if use_thresholds:
            th = np.asarray(self.getOrDefault(self.thresholds), dtype=np.float64)
            if len(th) != forest.C:
                raise ValueError("thresholds must have length equal to numClasses (%d)" % forest.C)
            with np.errstate(divide="ignore", invalid="ignore"):
                scaled = np.where(th[None, :] == 0.0, np.where(prob > 0.0, np.inf, 0.0), prob / th[None, :])
            pred = np.argmax(scaled, axis=1).astype(np.float64)
        df = dataset
        if self.getRawPredictionCol():
            df = df.withColumn(self.getRawPredictionCol(), VectorColumn(forest.C, dense=raw), "vector")
        if self.getProbabilityCol():
            df = df.withColumn(self.getProbabilityCol(), VectorColumn(forest.C, dense=prob), "vector")
        if self.getPredictionCol():
            df = df.withColumn(self.getPredictionCol(), pred, "double")
        return df


def _importances_of_tree(forest, t, out):
    """TreeEnsembleModel.computeFeatureImportance: sum over internal nodes of gain * node count."""
    a, b = int(forest.tree_off[t]), int(forest.tree_off[t + 1])
    for k in range(a, b):
        f = forest.feature[k]
        if f >= 0:
            out[f] += forest.gain[k] * forest.stats[k].sum()


def _normalize(v):
    s = v.sum()
    return v / s if s != 0 else v


def _subtree_string(forest, t, k, indent):
    base = int(forest.tree_off[t])
    i = base + k
    prefix = " " * indent
    if forest.feature[i] < 0:
        return prefix + "Predict: %s\n" % _jd(forest.prediction[i])
    f = forest.feature[i]
    if forest.arity[i] > 0:
        cats = [float(v) for v in range(int(forest.arity[i])) if (int(forest.left_mask[i]) >> v) & 1]
        cs = "{" + ",".join(_jd(c) for c in cats) + "}"
        left, right = "feature %d in %s" % (f, cs), "feature %d not in %s" % (f, cs)
    else:
        th = _jd(forest.threshold[i])
        left, right = "feature %d <= %s" % (f, th), "feature %d > %s" % (f, th)
    return (prefix + "If (%s)\n" % left + _subtree_string(forest, t, int(forest.left[i]), indent + 1) +
            prefix + "Else (%s)\n" % right + _subtree_string(forest, t, int(forest.right[i]), indent + 1))


def _jd(x):
    from .linalg import java_double_str
    return java_double_str(x)


class DecisionTreeClassificationModel(_TreeModelBase):
    def __init__(self, forest=None, tree_index=0, num_features=0, num_classes=0):
        super().__init__()
        self._tree_defaults()
        self._forest, self._t = forest, tree_index
        self._numFeatures, self._numClasses = num_features, num_classes
        self._single = None

    @property
    def numNodes(self):
        return self._forest.tree_num_nodes(self._t)

    @property
    def depth(self):
        return self._forest.tree_depth(self._t)

    @property
    def numFeatures(self):
        return self._numFeatures

    @property
    def numClasses(self):
        return self._numClasses

    @property
    def featureImportances(self):
        imp = np.zeros(self._numFeatures)
        _importances_of_tree(self._forest, self._t, imp)
        imp = _normalize(imp)
        nz = np.nonzero(imp)[0]
        from .linalg import SparseVector
        return SparseVector(self._numFeatures, nz, imp[nz])

    @property
    def toDebugString(self):
        return str(self) + "\n" + _subtree_string(self._forest, self._t, 0, 2)

    def __str__(self):
        return "DecisionTreeClassificationModel (uid=%s) of depth %d with %d nodes" % (self.uid, self.depth, self.numNodes)

    __repr__ = __str__

    def _own_forest(self):
        """A one-tree forest handle (a member tree of a random forest predicts on its own)."""
        fo = self._forest
        if fo.T == 1:
            return fo
        if self._single is None:
            from har_b200.core import Forest
            a, b = int(fo.tree_off[self._t]), int(fo.tree_off[self._t + 1])
            self._single = Forest.from_arrays(1, fo.C, fo.F, [0, b - a], fo.feature[a:b], fo.left[a:b], fo.right[a:b],
                                              fo.arity[a:b], fo.threshold[a:b], fo.left_mask[a:b], fo.stats[a:b])
        return self._single

    def _transform(self, dataset):
        return self._append_predictions(dataset, self._own_forest(), is_forest=False)


class RandomForestClassifier(_TreeEstimatorBase):
    numTrees = Param(Params._dummy_parent, "numTrees", "Number of trees to train (>= 1).", TypeConverters.toInt)
    featureSubsetStrategy = Param(Params._dummy_parent, "featureSubsetStrategy", "The number of features to consider "
                                  "for splits at each tree node. Supported options: auto, all, onethird, sqrt, log2, "
                                  "(0.0-1.0], [1-n].", TypeConverters.toString)
    subsamplingRate = Param(Params._dummy_parent, "subsamplingRate", "Fraction of the training data used for learning "
                            "each decision tree, in range (0, 1].", TypeConverters.toFloat)

    _KW = tuple(_TREE_PARAM_NAMES) + ("numTrees", "featureSubsetStrategy", "subsamplingRate")

    def __init__(self, *args, **kwargs):
        super().__init__()
        self._tree_defaults()
        self._setDefault(numTrees=20, featureSubsetStrategy="auto", subsamplingRate=1.0)
        if args:
            raise TypeError("Method __init__ forces keyword arguments.")
        self.setParams(**kwargs)

    def setParams(self, **kw):
        for k in kw:
            if k not in self._KW:
                raise TypeError("__init__() got an unexpected keyword argument '%s'" % k)
        if "featureSubsetStrategy" in kw and kw["featureSubsetStrategy"] is not None:
            kw["featureSubsetStrategy"] = str(kw["featureSubsetStrategy"])
        return self._set(**{k: v for k, v in kw.items() if v is not None})

    def _fit(self, dataset):
        T = self.getNumTrees()
        if T < 1:
            raise ValueError("numTrees must be >= 1")
        forest, F, C = self._fit_forest(dataset, T, self.getFeatureSubsetStrategy(), self.getSubsamplingRate())
        model = RandomForestClassificationModel(forest, F, C)
        model._resetUid(self.uid)
        return self._copyValues(model)


_make_accessors(RandomForestClassifier, ["numTrees", "featureSubsetStrategy", "subsamplingRate"])


class RandomForestClassificationModel(_TreeModelBase):
    numTrees = RandomForestClassifier.numTrees
    featureSubsetStrategy = RandomForestClassifier.featureSubsetStrategy
    subsamplingRate = RandomForestClassifier.subsamplingRate

    def __init__(self, forest=None, num_features=0, num_classes=0):
        super().__init__()
        self._tree_defaults()
        self._setDefault(numTrees=20, featureSubsetStrategy="auto", subsamplingRate=1.0)
        self._forest = forest
        self._numFeatures, self._numClasses = num_features, num_classes

    @property
    def trees(self):
        out = []
        for t in range(self._forest.T):
            m = DecisionTreeClassificationModel(self._forest, t, self._numFeatures, self._numClasses)
            out.append(m)
        return out

    @property
    def getNumTrees(self):
        return self._forest.T

    @property
    def treeWeights(self):
        return [1.0] * self._forest.T

    @property
    def totalNumNodes(self):
        return int(self._forest.total_nodes)

    @property
    def numFeatures(self):
        return self._numFeatures

    @property
    def numClasses(self):
        return self._numClasses

    @property
    def featureImportances(self):
        total = np.zeros(self._numFeatures)
        for t in range(self._forest.T):
            imp = np.zeros(self._numFeatures)
            _importances_of_tree(self._forest, t, imp)
            total += _normalize(imp)
        total = _normalize(total)
        nz = np.nonzero(total)[0]
        from .linalg import SparseVector
        return SparseVector(self._numFeatures, nz, total[nz])

    @property
    def toDebugString(self):
        s = str(self) + "\n"
        for t in range(self._forest.T):
            s += "  Tree %d (weight 1.0):\n" % t + _subtree_string(self._forest, t, 0, 4)
        return s

    def __str__(self):
        return "RandomForestClassificationModel (uid=%s) with %d trees" % (self.uid, self._forest.T)

    __repr__ = __str__

    def _transform(self, dataset):
        return self._append_predictions(dataset, self._forest, is_forest=True)
