"""pyspark.ml.tuning: ParamGridBuilder / CrossValidator (Spark 2.4 semantics).

Reference call sites: Main/main.py:202-215 (LR grid), :383-395 (DT), :564-576 (RF).  Every fold fit and
the final fit go through Estimator.fit, i.e. through the GPU tree path for DT / RF: one CrossValidator =
numFolds * len(grid) + 1 calls of har_forest_fit.  Fold assignment follows pyspark: a `rand(seed)` column
(XORShiftRandom(seed + partitionIndex).nextDouble() per row) cut into numFolds equal intervals.
"""
import itertools

import numpy as np

from . import Estimator, Model
from .param import Param, Params, TypeConverters, _make_accessors
from .._rng import XORShiftRandom

__all__ = ["ParamGridBuilder", "CrossValidator", "CrossValidatorModel"]


class ParamGridBuilder:
    def __init__(self):
        self._param_grid = {}

    def addGrid(self, param, values):
        self._param_grid[param] = list(values)
        return self

    def baseOn(self, *args):
        if isinstance(args[0], dict):
            self.baseOn(*args[0].items())
        else:
            for (param, value) in args:
                self.addGrid(param, [value])
        return self

    def build(self):
        keys = list(self._param_grid.keys())
        grid_values = [self._param_grid[k] for k in keys]
        return [dict(zip(keys, prod)) for prod in itertools.product(*grid_values)]


class _CVParams(Params):
    estimator = Param(Params._dummy_parent, "estimator", "estimator to be cross-validated")
    estimatorParamMaps = Param(Params._dummy_parent, "estimatorParamMaps", "estimator param maps")
    evaluator = Param(Params._dummy_parent, "evaluator", "evaluator used to select hyper-parameters that maximize the "
                      "validator metric")
    numFolds = Param(Params._dummy_parent, "numFolds", "number of folds for cross validation", TypeConverters.toInt)
    seed = Param(Params._dummy_parent, "seed", "random seed.", TypeConverters.toInt)
    parallelism = Param(Params._dummy_parent, "parallelism", "the number of threads to use when running parallel "
                        "algorithms (>= 1). (accepted; fits run one after another on the GPU)", TypeConverters.toInt)
    collectSubModels = Param(Params._dummy_parent, "collectSubModels", "whether to collect sub-models", TypeConverters.toBoolean)


_make_accessors(_CVParams, ["estimator", "estimatorParamMaps", "evaluator", "numFolds", "seed", "parallelism", "collectSubModels"])


def _rand_column(dataset, seed):
    out = np.empty(dataset.count(), dtype=np.float64)
    po = dataset._partitions
    for p in range(len(po) - 1):
        rng = XORShiftRandom(seed + p)
        for i in range(po[p], po[p + 1]):
            out[i] = rng.nextDouble()
    return out


class CrossValidator(Estimator, _CVParams):
    def __init__(self, estimator=None, estimatorParamMaps=None, evaluator=None, numFolds=3, seed=None, parallelism=1,
                 collectSubModels=False):
        super().__init__()
        self._setDefault(numFolds=3, parallelism=1, collectSubModels=False, seed=hash(type(self).__name__))
        kw = dict(estimator=estimator, estimatorParamMaps=estimatorParamMaps, evaluator=evaluator, numFolds=numFolds,
                  seed=seed, parallelism=parallelism, collectSubModels=collectSubModels)
        self._set(**{k: v for k, v in kw.items() if v is not None})

    def setParams(self, **kw):
        return self._set(**{k: v for k, v in kw.items() if v is not None})

    def _fit(self, dataset):
        est = self.getOrDefault(self.estimator)
        epm = self.getOrDefault(self.estimatorParamMaps)
        eva = self.getOrDefault(self.evaluator)
        n_folds = self.getOrDefault(self.numFolds)
        num_models = len(epm)
        h = 1.0 / n_folds
        rnd = _rand_column(dataset, self.getOrDefault(self.seed))
        metrics = [0.0] * num_models
        for i in range(n_folds):
            lb, ub = i * h, (i + 1) * h
            cond = (rnd >= lb) & (rnd < ub)
            validation = dataset._take_rows(np.nonzero(cond)[0])
            train = dataset._take_rows(np.nonzero(~cond)[0])
            for j, pm in enumerate(epm):
                model = est.fit(train, pm)
                metrics[j] += eva.evaluate(model.transform(validation, pm)) / n_folds
        best = int(np.argmax(metrics)) if eva.isLargerBetter() else int(np.argmin(metrics))
        best_model = est.fit(dataset, epm[best])
        return self._copyValues(CrossValidatorModel(best_model, metrics))

    def copy(self, extra=None):
        extra = extra or {}
        new = Params.copy(self, extra)
        if self.isSet(self.estimator):
            new.setEstimator(self.getEstimator().copy(extra))
        if self.isSet(self.evaluator):
            new.setEvaluator(self.getEvaluator().copy(extra))
        return new


class CrossValidatorModel(Model, _CVParams):
    def __init__(self, bestModel=None, avgMetrics=(), subModels=None):
        super().__init__()
        self.bestModel = bestModel
        self.avgMetrics = list(avgMetrics)
        self.subModels = subModels

    def _transform(self, dataset):
        return self.bestModel.transform(dataset)

    def __str__(self):
        return self.uid

    __repr__ = __str__

    def copy(self, extra=None):
        return CrossValidatorModel(self.bestModel.copy(extra), list(self.avgMetrics), self.subModels)
