"""LogisticRegression — HOST-ONLY, outside the accelerated path.

Main/main.py:115 trains a multinomial LogisticRegression before the tree models; BASELINE.json's
north_star names only the window features and the tree ensembles as the hot path (out
of scope).  This numpy/scipy stand-in exists so that main.py runs to completion against the facade; it
follows MLlib's objective (softmax cross-entropy on std-scaled features, intercepts unpenalised,
regParam * ((1 - a)/2 * ||w||^2 + a * ||w||_1)) but makes no bit-parity claim (breeze L-BFGS / OWL-QN are
not restated; the L1 term is smoothed).
"""
import numpy as np

from . import Estimator, Model
from .linalg import VectorColumn
from .param import Param, Params, TypeConverters, _make_accessors

__all__ = ["LogisticRegression", "LogisticRegressionModel"]

_NAMES = ["featuresCol", "labelCol", "predictionCol", "probabilityCol", "rawPredictionCol", "maxIter", "regParam",
          "elasticNetParam", "tol", "fitIntercept", "standardization", "family"]


class _LRParams(Params):
    featuresCol = Param(Params._dummy_parent, "featuresCol", "features column name.", TypeConverters.toString)
    labelCol = Param(Params._dummy_parent, "labelCol", "label column name.", TypeConverters.toString)
    predictionCol = Param(Params._dummy_parent, "predictionCol", "prediction column name.", TypeConverters.toString)
    probabilityCol = Param(Params._dummy_parent, "probabilityCol", "probability column name.", TypeConverters.toString)
    rawPredictionCol = Param(Params._dummy_parent, "rawPredictionCol", "raw prediction column name.", TypeConverters.toString)
    maxIter = Param(Params._dummy_parent, "maxIter", "max number of iterations (>= 0).", TypeConverters.toInt)
    regParam = Param(Params._dummy_parent, "regParam", "regularization parameter (>= 0).", TypeConverters.toFloat)
    elasticNetParam = Param(Params._dummy_parent, "elasticNetParam", "the ElasticNet mixing parameter, in range [0, 1].",
                            TypeConverters.toFloat)
    tol = Param(Params._dummy_parent, "tol", "the convergence tolerance for iterative algorithms (>= 0).", TypeConverters.toFloat)
    fitIntercept = Param(Params._dummy_parent, "fitIntercept", "whether to fit an intercept term.", TypeConverters.toBoolean)
    standardization = Param(Params._dummy_parent, "standardization", "whether to standardize the training features.",
                            TypeConverters.toBoolean)
    family = Param(Params._dummy_parent, "family", "auto | binomial | multinomial", TypeConverters.toString)

    def _lr_defaults(self):
        self._setDefault(featuresCol="features", labelCol="label", predictionCol="prediction", probabilityCol="probability",
                         rawPredictionCol="rawPrediction", maxIter=100, regParam=0.0, elasticNetParam=0.0, tol=1e-6,
                         fitIntercept=True, standardization=True, family="auto")


_make_accessors(_LRParams, _NAMES)


class LogisticRegression(Estimator, _LRParams):
    def __init__(self, *args, **kwargs):
        super().__init__()
        self._lr_defaults()
        if args:
            raise TypeError("Method __init__ forces keyword arguments.")
        for k in kwargs:
            if k not in _NAMES:
                raise TypeError("__init__() got an unexpected keyword argument '%s'" % k)
        self._set(**{k: v for k, v in kwargs.items() if v is not None})

    def _fit(self, dataset):
        from scipy.optimize import minimize
        col = dataset._cols[dataset._resolve(self.getFeaturesCol())]
        X = col.to_dense()
        y = np.asarray(dataset._cols[dataset._resolve(self.getLabelCol())], dtype=np.int64)
        n, F = X.shape
        K = int(y.max()) + 1
        std = X.std(axis=0, ddof=1)
        inv = np.where(std > 0, 1.0 / np.where(std > 0, std, 1.0), 0.0)
        Xs = X * inv if self.getStandardization() else X
        lam, alpha = self.getRegParam(), self.getElasticNetParam()
        Y = np.zeros((n, K)); Y[np.arange(n), y] = 1.0

        def fg(theta):
            W = theta[:F * K].reshape(F, K)
            b = theta[F * K:]
            Z = Xs @ W + b
            Z -= Z.max(axis=1, keepdims=True)
            P = np.exp(Z); P /= P.sum(axis=1, keepdims=True)
            loss = -np.log(P[np.arange(n), y] + 1e-300).mean()
            G = Xs.T @ (P - Y) / n
            sm = np.sqrt(W * W + 1e-12)
            loss += lam * (0.5 * (1 - alpha) * (W * W).sum() + alpha * sm.sum())
            G += lam * ((1 - alpha) * W + alpha * W / sm)
            gb = (P - Y).mean(axis=0) if self.getFitIntercept() else np.zeros(K)
            return loss, np.concatenate([G.ravel(), gb])

        theta0 = np.zeros(F * K + K)
        if self.getFitIntercept():
            pri = np.bincount(y, minlength=K) / n
            theta0[F * K:] = np.log(np.maximum(pri, 1e-12)) - np.log(np.maximum(pri, 1e-12)).mean()
        res = minimize(fg, theta0, jac=True, method="L-BFGS-B", options={"maxiter": self.getMaxIter(), "ftol": self.getTol()})
        W = res.x[:F * K].reshape(F, K) * (inv[:, None] if self.getStandardization() else 1.0)
        model = LogisticRegressionModel(W, res.x[F * K:])
        model._resetUid(self.uid)
        return self._copyValues(model)


class LogisticRegressionModel(Model, _LRParams):
    def __init__(self, W=None, b=None):
        super().__init__()
        self._lr_defaults()
        self._W, self._b = W, b

    @property
    def coefficientMatrix(self):
        return self._W.T

    @property
    def interceptVector(self):
        return self._b

    def __str__(self):
        return self.uid          # Spark 2.4 prints the uid for this model (result.txt:141)

    __repr__ = __str__

    def _transform(self, dataset):
        col = dataset._cols[dataset._resolve(self.getFeaturesCol())]
        Z = col.to_dense() @ self._W + self._b
        E = np.exp(Z - Z.max(axis=1, keepdims=True))
        P = E / E.sum(axis=1, keepdims=True)
        K = Z.shape[1]
        df = dataset.withColumn(self.getRawPredictionCol(), VectorColumn(K, dense=Z), "vector")
        df = df.withColumn(self.getProbabilityCol(), VectorColumn(K, dense=P), "vector")
        return df.withColumn(self.getPredictionCol(), np.argmax(Z, axis=1).astype(np.float64), "double")
