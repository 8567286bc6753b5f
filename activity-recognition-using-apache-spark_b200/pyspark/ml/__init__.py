"""pyspark.ml base classes: Estimator / Transformer / Model / Pipeline (Spark 2.4 semantics).

Upstream python/pyspark/ml/{base,pipeline}.py; call sites Main/main.py:68-71 (Pipeline),
:300,:481 (fit), :305,:486 (transform), :395,:576 (fit via CrossValidator).
"""
from .param import Param, Params, TypeConverters

__all__ = ["Estimator", "Transformer", "Model", "Pipeline", "PipelineModel", "Param", "Params"]


class Estimator(Params):
    def _fit(self, dataset):
        raise NotImplementedError

    def fitMultiple(self, dataset, paramMaps):
        estimator = self.copy()

        def gen():
            for i, pm in enumerate(paramMaps):
                yield i, estimator.fit(dataset, pm)
        return gen()

    def fit(self, dataset, params=None):
        if params is None:
            params = dict()
        if isinstance(params, (list, tuple)):
            models = [None] * len(params)
            for index, model in self.fitMultiple(dataset, params):
                models[index] = model
            return models
        if isinstance(params, dict):
            if params:
                return self.copy(params)._fit(dataset)
            return self._fit(dataset)
        raise ValueError("Params must be either a param map or a list/tuple of param maps, "
                         "but got %s." % type(params))


class Transformer(Params):
    def _transform(self, dataset):
        raise NotImplementedError

    def transform(self, dataset, params=None):
        if params is None:
            params = dict()
        if isinstance(params, dict):
            if params:
                return self.copy(params)._transform(dataset)
            return self._transform(dataset)
        raise ValueError("Params must be a param map but got %s." % type(params))


class Model(Transformer):
    pass


class Pipeline(Estimator):
    stages = Param(Params._dummy_parent, "stages", "a list of pipeline stages")

    def __init__(self, stages=None):
        super().__init__()
        if stages is not None:
            self._set(stages=stages)

    def setStages(self, value):
        return self._set(stages=value)

    def getStages(self):
        return self.getOrDefault(self.stages)

    def setParams(self, stages=None):
        if stages is not None:
            self._set(stages=stages)
        return self

    def _fit(self, dataset):
        stages = self.getStages()
        for stage in stages:
            if not isinstance(stage, (Estimator, Transformer)):
                raise TypeError("Cannot recognize a pipeline stage of type %s." % type(stage))
        indexOfLastEstimator = -1
        for i, stage in enumerate(stages):
            if isinstance(stage, Estimator):
                indexOfLastEstimator = i
        transformers = []
        for i, stage in enumerate(stages):
            if i <= indexOfLastEstimator:
                if isinstance(stage, Transformer):
                    transformers.append(stage)
                    dataset = stage.transform(dataset)
                else:
                    model = stage.fit(dataset)
                    transformers.append(model)
                    if i < indexOfLastEstimator:
                        dataset = model.transform(dataset)
            else:
                transformers.append(stage)
        return PipelineModel(transformers)

    def copy(self, extra=None):
        extra = extra or {}
        that = Params.copy(self, extra)
        return that.setStages([stage.copy(extra) for stage in that.getStages()])


class PipelineModel(Model):
    def __init__(self, stages):
        super().__init__()
        self.stages = stages

    def _transform(self, dataset):
        for t in self.stages:
            dataset = t.transform(dataset)
        return dataset

    def copy(self, extra=None):
        return PipelineModel([stage.copy(extra) for stage in self.stages])
