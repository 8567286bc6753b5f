import gzip
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "activity-recognition-using-apache-spark_b200")
for p in (PKG, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        have = torch.cuda.is_available()
    except Exception:
        have = False
    if have:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


def golden_value(g, key):
    return list(g[key].values())[0]


@pytest.fixture(scope="session")
def golden():
    with open(os.path.join(GOLDEN_DIR, "golden.json")) as fh:
        return json.load(fh)


@pytest.fixture(scope="session")
def wisdm_csv(tmp_path_factory):
    """The reference's data file, restored from the committed fixture."""
    path = tmp_path_factory.mktemp("wisdm") / "wisdm_data.csv"
    with gzip.open(os.path.join(GOLDEN_DIR, "wisdm_transformed.csv.gz"), "rb") as src:
        path.write_bytes(src.read())
    return str(path)


def build_wisdm_frames(csv_path):
    from har_b200.pipeline import build_wisdm_frames as f
    return f(csv_path)


@pytest.fixture(scope="session")
def wisdm(wisdm_csv):
    data, df, train, test, model = build_wisdm_frames(wisdm_csv)
    return {"data": data, "df": df, "train": train, "test": test, "pipeline": model}


@pytest.fixture(scope="session")
def wisdm_arrays(wisdm):
    tr, te = wisdm["train"], wisdm["test"]
    return {"Xtr": tr._cols["features"].to_dense(), "ytr": np.asarray(tr._cols["label"]),
            "Xte": te._cols["features"].to_dense(), "yte": np.asarray(te._cols["label"]),
            "arity": np.asarray(tr._meta["features"]["ml_attr"]["arity"], dtype=np.int32)}


def synth_tree_data(n, F=12, C=4, seed=0, n_cat_unordered=0, n_cat_ordered=0, quantize=None):
    """Learnable random classification problem; optional categorical columns (small / large arity)."""
    rng = np.random.RandomState(seed)
    X = rng.randn(n, F)
    if quantize:
        X = np.round(X * quantize) / quantize
    arity = np.zeros(F, dtype=np.int32)
    j = 0
    for _ in range(n_cat_unordered):
        X[:, j] = rng.randint(0, 4, size=n); arity[j] = 4; j += 1
    for _ in range(n_cat_ordered):
        X[:, j] = rng.randint(0, 9, size=n); arity[j] = 9; j += 1
    w = rng.randn(F, C)
    score = X @ w + 0.7 * rng.randn(n, C)
    if n_cat_ordered:
        score[:, 0] += (X[:, j - 1] % 3) * 0.8
    y = np.argmax(score, axis=1).astype(np.float64)
    return X, y, arity


FOREST_FIELDS = ("tree_off", "feature", "left", "right", "node_id", "arity", "threshold", "left_mask",
                 "prediction", "impurity", "gain", "stats")


def assert_forest_equal(a, b, exact_fp=True):
    """Node-by-node comparison of two flat forests (oracle vs GPU)."""
    assert a.T == b.T and a.C == b.C
    for k in FOREST_FIELDS:
        x, y = np.asarray(getattr(a, k)), np.asarray(getattr(b, k))
        assert x.shape == y.shape, "%s: shapes %s vs %s" % (k, x.shape, y.shape)
        if x.dtype.kind == "f" and not exact_fp and k in ("gain", "impurity"):
            np.testing.assert_allclose(x, y, rtol=1e-12, atol=1e-15, err_msg=k)
        elif x.dtype.kind == "f":
            assert np.array_equal(x.view(np.int64), y.view(np.int64)), "%s differs bitwise (max abs diff %g)" % (
                k, float(np.max(np.abs(x - y))))
        else:
            assert np.array_equal(x, y), "%s differs" % k
