"""CPU suite: the host feed (pyspark facade) and the C oracle against the reference's own Spark
output (tests/golden/golden.json <- result.txt / additional_param.csv)."""
import numpy as np
import pytest

from conftest import golden_value as gv
from oracle import oracle as orc


def test_csv_schema_matches_spark_inference(wisdm, golden):
    lines = wisdm["data"]._schema_string().splitlines()
    assert lines[1:16] == gv(golden, "schema_lines")                       # result.txt:4-18


def test_activity_counts_and_label_order(wisdm, golden):
    from pyspark.sql.functions import col
    out = wisdm["data"].groupBy("activity").count().orderBy(col("count").desc())
    got = [[str(a), str(c)] for a, c in zip(out._cols["activity"], out._cols["count"])]
    want = [[c.strip() for c in row] for row in gv(golden, "activity_counts")]
    assert got == want                                                       # result.txt:36-41
    labels = [s for s in wisdm["pipeline"].stages if getattr(s, "getOutputCol", lambda: "")() == "label"][0].labels
    assert labels == ["Walking", "Jogging", "Upstairs", "Downstairs", "Sitting", "Standing"]


def test_string_indexer_tie_order_is_scala_hashmap_order(wisdm):
    xpeak = wisdm["pipeline"].stages[0].labels
    assert xpeak[:12] == ["?", "1100", "1050", "2150", "1450", "1850", "2200", "2000", "1700", "1150", "2450", "1550"]
    assert len(xpeak) == 935
    assert wisdm["df"]._cols["features"].size == 3100


def test_random_split_sizes_and_first_rows(wisdm, golden):
    train, test = wisdm["train"], wisdm["test"]
    assert train.count() == gv(golden, "train_count") and test.count() == gv(golden, "test_count")
    minimized = ["XPEAK", "YPEAK", "ZPEAK", "XABSDEV", "YABSDEV", "ZABSDEV"]
    view = [c for c in test.columns if c not in minimized]
    assert train.select(view)._show_string(5).splitlines()[3:8] == gv(golden, "train_head")   # result.txt:110-114
    assert test.select(view)._show_string(5).splitlines()[3:8] == gv(golden, "test_head")     # result.txt:121-125


def test_xorshift_known_answers():
    from pyspark._rng import XORShiftRandom, xorshift_hash_seed
    assert xorshift_hash_seed(2018) == 0xFE79670DD53901A3
    r = XORShiftRandom(2018)
    got = [r.nextDouble() for _ in range(4)]
    assert got == [0.17860326850261143, 0.2076523294811069, 0.1190947932302171, 0.7710825496303402]


@pytest.mark.parametrize("depth,nodes,correct", [(3, 15, 1187), (5, 39, 1290), (10, 257, 1359)])
def test_oracle_decision_tree_reproduces_spark(wisdm_arrays, golden, depth, nodes, correct):
    a = wisdm_arrays
    fo = orc.fit_forest(a["Xtr"], a["ytr"], arity=a["arity"], num_classes=6, num_trees=1, max_depth=depth)
    raw, prob, pred = orc.predict(fo, a["Xte"])
    assert fo.tree_num_nodes(0) == nodes and fo.tree_depth(0) == depth
    assert int((pred == a["yte"]).sum()) == correct
    if depth == 3:                                                            # Spark-produced goldens
        assert "of depth 3 with 15 nodes" in gv(golden, "dt_model_str")
        assert correct == gv(golden, "dt_correct") and len(pred) == gv(golden, "dt_total")
        assert fo.feature[0] == 3094 and fo.threshold[0] == 5.3
        assert fo.gain[0] == 0.21204147537790785
        np.testing.assert_array_equal(fo.stats[0], [1476, 1146, 426, 361, 212, 172])
        thr = fo.thresholds[3094][:fo.nsplits[3094]]
        assert len(thr) == 31 and thr[0] == 0.045 and thr[1] == 0.10500000000000001 and thr[22] == 5.3


def test_oracle_metrics_match_reference_full_precision(wisdm, wisdm_arrays, golden):
    from pyspark.ml.evaluation import (BinaryClassificationEvaluator, MulticlassClassificationEvaluator,
                                       RegressionEvaluator)
    from pyspark.ml.linalg import VectorColumn
    a = wisdm_arrays
    fo = orc.fit_forest(a["Xtr"], a["ytr"], arity=a["arity"], num_classes=6, num_trees=1, max_depth=3)
    raw, prob, pred = orc.predict(fo, a["Xte"])
    df = (wisdm["test"].withColumn("rawPrediction", VectorColumn(6, dense=raw), "vector")
          .withColumn("probability", VectorColumn(6, dense=prob), "vector").withColumn("prediction", pred, "double"))
    ev = MulticlassClassificationEvaluator(labelCol="label", predictionCol="prediction")
    assert ev.evaluate(df, {ev.metricName: "f1"}) == gv(golden, "dt_f1_full")               # additional_param.csv:3
    assert ev.evaluate(df, {ev.metricName: "accuracy"}) == gv(golden, "dt_accuracy_full")
    assert float("%g" % ev.evaluate(df, {ev.metricName: "weightedPrecision"})) == gv(golden, "dt_weighted_precision")
    assert float("%g" % ev.evaluate(df, {ev.metricName: "weightedRecall"})) == gv(golden, "dt_weighted_recall")
    for name, key in (("rmse", "dt_rmse"), ("r2", "dt_r2"), ("mae", "dt_mae")):
        assert float("%g" % RegressionEvaluator(labelCol="label", predictionCol="prediction", metricName=name).evaluate(df)) == gv(golden, key)
    assert float("%g" % BinaryClassificationEvaluator(labelCol="label", metricName="areaUnderROC").evaluate(df)) == gv(golden, "dt_area_under_roc")
    assert float("%g" % BinaryClassificationEvaluator(labelCol="label", metricName="areaUnderPR").evaluate(df)) == gv(golden, "dt_area_under_pr")
    # probability column of the pure leaf [44,0,0,0,0,0] (result.txt:237-241)
    uid = np.asarray(df._cols["UID"])
    for u in (492, 518, 495, 511, 536):
        rows = np.nonzero((uid == u) & (pred == 0) & (prob[:, 0] == 1.0))[0]
        assert len(rows) >= 1


def test_oracle_random_forest_in_reference_accuracy_band(wisdm_arrays, golden):
    a = wisdm_arrays
    fo = orc.fit_forest(a["Xtr"], a["ytr"], arity=a["arity"], num_classes=6, num_trees=100, max_depth=4,
                        feature_subset="auto", seed=7)
    raw, prob, pred = orc.predict(fo, a["Xte"])
    correct = int((pred == a["yte"]).sum())
    assert abs(correct - gv(golden, "rf_correct")) <= 80       # unseeded in the reference: statistical band only
    assert fo.T == 100 and np.allclose(prob.sum(axis=1), 1.0)


def test_oracle_bagging_is_poisson_one_and_deterministic():
    w1 = orc.bag_weights(20000, 5, 1.0, seed=42)
    w2 = orc.bag_weights(20000, 5, 1.0, seed=42)
    assert np.array_equal(w1, w2)
    assert abs(w1.mean() - 1.0) < 0.02 and abs((w1 == 0).mean() - np.exp(-1)) < 0.01
    assert np.array_equal(orc.bag_weights(100, 1, 1.0, seed=1), np.ones((1, 100), np.uint8))
    # partitions own independent streams: the first partition's draws do not depend on the others
    wp = orc.bag_weights(20000, 5, 1.0, seed=42, part_off=[0, 5000, 20000])
    assert np.array_equal(wp[:, :5000], orc.bag_weights(5000, 5, 1.0, seed=42))


def test_oracle_window_segments_and_feature_invariants():
    u, a, ts, x, y, z = orc.synth_stream(60000, first_sample=0, per_user=20000, seed=2018)
    ws = orc.window_segments(u, a, ts)
    # brute force: consecutive runs of equal (user, activity)
    want, start = [], 0
    for i in range(1, len(u) + 1):
        if i == len(u) or u[i] != u[i - 1] or a[i] != a[i - 1]:
            want += [start + 200 * k for k in range((i - start) // 200)]
            start = i
    assert ws.tolist() == want
    f = orc.window_features(x, y, z, ts, ws)
    assert f.shape == (len(ws), 43)
    np.testing.assert_allclose(f[:, 0:10].sum(1), 1.0, rtol=0, atol=1e-12)      # bins sum to one
    assert np.all(f[:, 36:39] <= f[:, 39:42] + 1e-12)                            # ABSDEV <= STDDEV
    mean_vec = np.sqrt((f[:, 30:33] ** 2).sum(1))
    assert np.all(f[:, 42] >= mean_vec - 1e-9)                                    # RESULTANT >= |mean vector|
    k = 7
    s = ws[k]
    np.testing.assert_allclose(f[k, 30], x[s:s + 200].mean(), rtol=1e-12)
    np.testing.assert_allclose(f[k, 40], y[s:s + 200].std(), rtol=1e-12)
    pk = f[:, 33:36]
    assert np.all(np.isnan(pk) | ((pk >= 50.0) & (pk <= 10000.0)))


def test_oracle_errors_mirror_spark_messages():
    X = np.random.RandomState(0).rand(50, 3)
    y = np.zeros(50)
    with pytest.raises(ValueError, match="maxDepth <= 30"):
        orc.fit_forest(X, y, num_classes=2, max_depth=31)
    y2 = y.copy(); y2[3] = 2.0
    with pytest.raises(ValueError, match="invalid label"):
        orc.fit_forest(X, y2, num_classes=2)
    Xc = X.copy(); Xc[:, 0] = 40
    with pytest.raises(ValueError, match="requires maxBins"):
        orc.fit_forest(Xc, y, arity=[41, 0, 0], num_classes=2, max_bins=32)


def test_oracle_raw_wisdm_ingest_rules():
    """The strtod/strtoll restatement of the raw-text ingest: what is kept, what is dropped."""
    from oracle import oracle as orc
    txt = (b"33,Jogging,49105962326000,-0.6946377,12.680544,0.50395286;\n 17,Walking,1,1e-3,2.5E2,-.5;"
           b"11,Walking,1867172313000,4.4,4.4,;\nbad;;  \n5,Flying,1,1,1,1;7,Sitting,9,1,2,3")
    (u, a, ts, x, y, z), dropped = orc.parse_wisdm_raw(txt)
    assert u.tolist() == [33, 17, 7] and a.tolist() == [1, 0, 4] and ts.tolist() == [49105962326000, 1, 9]
    assert x.tolist() == [-0.6946377, 0.001, 1.0] and y.tolist() == [12.680544, 250.0, 2.0] and z.tolist() == [0.50395286, -0.5, 3.0]
    assert dropped == 3
