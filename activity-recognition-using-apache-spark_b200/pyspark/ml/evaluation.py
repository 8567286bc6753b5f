"""pyspark.ml.evaluation: the three evaluators Main/main.py applies to the hot path's output columns
(main.py:135-177, 317-366, 498-547).  Host-side restatements of mllib.evaluation.{Multiclass,Binary
Classification,Regression}Metrics; F1 / accuracy are pinned to full precision by
additional_param.csv:3, the others to the 6 digits result.txt prints.
"""
import math

import numpy as np

from .param import Param, Params, TypeConverters, _make_accessors

__all__ = ["Evaluator", "MulticlassClassificationEvaluator", "BinaryClassificationEvaluator", "RegressionEvaluator"]


class Evaluator(Params):
    def _evaluate(self, dataset):
        raise NotImplementedError

    def evaluate(self, dataset, params=None):
        if params is None:
            params = dict()
        if isinstance(params, dict):
            if params:
                return self.copy(params)._evaluate(dataset)
            return self._evaluate(dataset)
        raise ValueError("Params must be a param map but got %s." % type(params))

    def isLargerBetter(self):
        return True


class MulticlassClassificationEvaluator(Evaluator):
    predictionCol = Param(Params._dummy_parent, "predictionCol", "prediction column name.", TypeConverters.toString)
    labelCol = Param(Params._dummy_parent, "labelCol", "label column name.", TypeConverters.toString)
    metricName = Param(Params._dummy_parent, "metricName", "metric name in evaluation "
                       "(f1|weightedPrecision|weightedRecall|accuracy)", TypeConverters.toString)

    def __init__(self, predictionCol="prediction", labelCol="label", metricName="f1"):
        super().__init__()
        self._setDefault(predictionCol="prediction", labelCol="label", metricName="f1")
        self._set(predictionCol=predictionCol, labelCol=labelCol, metricName=metricName)

    def _evaluate(self, dataset):
        pred = np.asarray(dataset._cols[dataset._resolve(self.getPredictionCol())], dtype=np.float64)
        lab = np.asarray(dataset._cols[dataset._resolve(self.getLabelCol())], dtype=np.float64)
        labels = sorted(set(lab.tolist()))                     # classes = distinct ACTUAL labels
        total = len(lab)
        metric = self.getMetricName()
        if metric == "accuracy":
            return float(int((pred == lab).sum())) / total
        acc = 0.0
        for c in labels:
            count = int((lab == c).sum())
            tp = int(((lab == c) & (pred == c)).sum())
            fp = int(((lab != c) & (pred == c)).sum())
            p = 0.0 if tp + fp == 0 else float(tp) / (tp + fp)
            r = float(tp) / count
            if metric == "weightedPrecision":
                v = p
            elif metric == "weightedRecall":
                v = r
            elif metric == "f1":
                v = 0.0 if p + r == 0 else (1 + 1.0) * p * r / (1.0 * p + r)
            else:
                raise ValueError("metricName given invalid value %s" % metric)
            acc += v * float(count) / total
        return acc


_make_accessors(MulticlassClassificationEvaluator, ["predictionCol", "labelCol", "metricName"])


class BinaryClassificationEvaluator(Evaluator):
    rawPredictionCol = Param(Params._dummy_parent, "rawPredictionCol", "raw prediction (a.k.a. confidence) column name.",
                             TypeConverters.toString)
    labelCol = Param(Params._dummy_parent, "labelCol", "label column name.", TypeConverters.toString)
    metricName = Param(Params._dummy_parent, "metricName", "metric name in evaluation (areaUnderROC|areaUnderPR)",
                       TypeConverters.toString)

    def __init__(self, rawPredictionCol="rawPrediction", labelCol="label", metricName="areaUnderROC"):
        super().__init__()
        self._setDefault(rawPredictionCol="rawPrediction", labelCol="label", metricName="areaUnderROC")
        self._set(rawPredictionCol=rawPredictionCol, labelCol=labelCol, metricName=metricName)

    def _evaluate(self, dataset):
        name = dataset._resolve(self.getRawPredictionCol())
        col = dataset._cols[name]
        if dataset._types[name] == "vector":
            score = (col.dense if not col.is_sparse else col.to_dense())[:, 1]     # rawPrediction(1)
        else:
            score = np.asarray(col, dtype=np.float64)
        lab = np.asarray(dataset._cols[dataset._resolve(self.getLabelCol())], dtype=np.float64)
        pos = lab > 0.5
        # one point per distinct score, thresholds descending, cumulative counts
        uniq = np.unique(score)[::-1]
        if len(uniq) > 1000 and len(score) // 1000 >= 2:
            raise NotImplementedError("numBins down-sampling of BinaryClassificationMetrics is not restated")
        P, Nn = int(pos.sum()), int((~pos).sum())
        tps, fps = [], []
        order = np.argsort(-score, kind="stable")
        s_sorted, p_sorted = score[order], pos[order]
        ctp = np.cumsum(p_sorted)
        cfp = np.cumsum(~p_sorted)
        last = np.nonzero(np.append(s_sorted[1:] != s_sorted[:-1], True))[0]
        tps, fps = ctp[last].astype(np.float64), cfp[last].astype(np.float64)
        metric = self.getMetricName()
        if metric == "areaUnderROC":
            xs = np.concatenate([[0.0], fps / Nn if Nn else fps * 0.0, [1.0]])
            ys = np.concatenate([[0.0], tps / P if P else tps * 0.0, [1.0]])
        elif metric == "areaUnderPR":
            rec = tps / P if P else tps * 0.0
            prec = np.where(tps + fps == 0, 1.0, tps / np.maximum(tps + fps, 1.0))
            xs = np.concatenate([[0.0], rec])
            ys = np.concatenate([[prec[0]], prec])
        else:
            raise ValueError("metricName given invalid value %s" % metric)
        area = 0.0
        for i in range(1, len(xs)):                                   # trapezoid, left to right
            area += (xs[i] - xs[i - 1]) * (ys[i] + ys[i - 1]) / 2.0
        return float(area)


_make_accessors(BinaryClassificationEvaluator, ["rawPredictionCol", "labelCol", "metricName"])


class RegressionEvaluator(Evaluator):
    predictionCol = Param(Params._dummy_parent, "predictionCol", "prediction column name.", TypeConverters.toString)
    labelCol = Param(Params._dummy_parent, "labelCol", "label column name.", TypeConverters.toString)
    metricName = Param(Params._dummy_parent, "metricName", "metric name in evaluation (rmse|mse|r2|mae)",
                       TypeConverters.toString)

    def __init__(self, predictionCol="prediction", labelCol="label", metricName="rmse"):
        super().__init__()
        self._setDefault(predictionCol="prediction", labelCol="label", metricName="rmse")
        self._set(predictionCol=predictionCol, labelCol=labelCol, metricName=metricName)

    def _evaluate(self, dataset):
        pred = np.asarray(dataset._cols[dataset._resolve(self.getPredictionCol())], dtype=np.float64)
        lab = np.asarray(dataset._cols[dataset._resolve(self.getLabelCol())], dtype=np.float64)
        n = len(lab)
        err = lab - pred
        ss_err = float(np.dot(err, err))
        metric = self.getMetricName()
        if metric == "rmse":
            return math.sqrt(ss_err / n)
        if metric == "mse":
            return ss_err / n
        if metric == "mae":
            return float(np.abs(err).sum()) / n
        if metric == "r2":
            ss_tot = float(((lab - lab.mean()) ** 2).sum())
            return 1.0 - ss_err / ss_tot
        raise ValueError("metricName given invalid value %s" % metric)

    def isLargerBetter(self):
        return self.getMetricName() == "r2"


_make_accessors(RegressionEvaluator, ["predictionCol", "labelCol", "metricName"])
