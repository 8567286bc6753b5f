"""Runs the reference's Main/main.py VERBATIM against the in-tree pyspark facade (CPU container only).

/root/reference exists only in the build container, where there is no GPU; so for this one check the tree
kernels are replaced by the CPU oracle through a test-only monkeypatch (the product never does this).  What
is exercised is everything else main.py touches: SparkConf/SQLContext, the CSV reader, DataFrame ops, the
feature stages, randomSplit, Pipeline, the estimator/model surface, evaluators, CrossValidator."""
import contextlib
import gzip
import io
import os
import runpy
import sys

import numpy as np

REF_MAIN = "/root/reference/Main/main.py"


class _OracleForest:
    """Quacks like har_b200.core.Forest, backed by oracle/oracle.py."""

    def __init__(self, fo):
        self._fo = fo
        self.T, self.C, self.F = fo.T, fo.C, fo.F
        self.total_nodes = int(fo.tree_off[-1])
        for k in ("tree_off", "feature", "left", "right", "node_id", "arity", "threshold", "left_mask", "prediction",
                  "impurity", "gain", "stats"):
            setattr(self, k, getattr(fo, k))

    def tree_num_nodes(self, t):
        return self._fo.tree_num_nodes(t)

    def tree_depth(self, t):
        return self._fo.tree_depth(t)

    def predict(self, X, is_forest=None, want=None):
        import torch
        from oracle import oracle as orc
        raw, prob, pred = orc.predict(self._fo, np.asarray(X), is_forest=is_forest)
        return torch.from_numpy(raw), torch.from_numpy(prob), torch.from_numpy(pred)


@contextlib.contextmanager
def oracle_backend():
    from har_b200 import core
    from oracle import oracle as orc
    import pyspark.ml.classification as cls

    def fit_forest(X, y, arity=None, dist=None, **kw):
        return _OracleForest(orc.fit_forest(np.asarray(X), np.asarray(y), arity=arity, **kw))

    saved = (core.fit_forest, cls._device_features)
    core.fit_forest = fit_forest
    cls._device_features = lambda col: col.to_dense()
    try:
        yield
    finally:
        core.fit_forest, cls._device_features = saved


def run_main_py(workdir, golden_dir):
    """-> (text main.py wrote to result.txt, exception that ended the run or None)."""
    data_dir = os.path.join(workdir, "wisdm_main_ver_0.0", "data")
    os.makedirs(data_dir)
    os.makedirs(os.path.join(workdir, "wisdm_main_ver_0.0", "main_result"))
    with gzip.open(os.path.join(golden_dir, "wisdm_transformed.csv.gz"), "rb") as src:
        open(os.path.join(data_dir, "wisdm_data.csv"), "wb").write(src.read())
    cwd, stdout = os.getcwd(), sys.stdout
    err = None
    os.chdir(workdir)
    try:
        with oracle_backend():
            runpy.run_path(REF_MAIN, init_globals={"master": "local[*]"}, run_name="__main__")
    except BaseException as e:          # main.py ends in IPython / matplotlib calls that do not exist here
        err = e
    finally:
        if sys.stdout is not stdout:
            try:
                sys.stdout.flush()
            except Exception:
                pass
        sys.stdout = stdout
        os.chdir(cwd)
    with open(os.path.join(workdir, "wisdm_main_ver_0.0", "main_result", "result.txt")) as fh:
        return fh.read(), err
