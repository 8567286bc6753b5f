"""The feature pipeline of the reference script, Main/main.py:18-80, as one function: load the transformed WISDM
table, drop the histogram columns, index + one-hot the three PEAK columns, index the label, assemble the 3100-d
feature vector, split 70/30 with seed 2018.  Used by the tests and by bench.py's WISDM timings."""


def build_wisdm_frames(csv_path):
    """Main/main.py:18-80 verbatim in spirit: load, drop, index, one-hot, assemble, split(seed=2018)."""
    from pyspark.ml import Pipeline
    from pyspark.ml.feature import OneHotEncoderEstimator, StringIndexer, VectorAssembler
    from pyspark.sql import SQLContext
    data = SQLContext(None).read.format("com.databricks.spark.csv").options(header="true", inferschema="true").load(csv_path)
    drop_list = ["USER"] + [a + str(i) for a in "XYZ" for i in range(10)]
    data = data.select([c for c in data.columns if c not in drop_list])
    cols = data.columns
    stages = []
    for c in ["XPEAK", "YPEAK", "ZPEAK"]:
        si = StringIndexer(inputCol=c, outputCol=c + "Index")
        stages += [si, OneHotEncoderEstimator(inputCols=[si.getOutputCol()], outputCols=[c + "classVec"])]
    stages += [StringIndexer(inputCol="ACTIVITY", outputCol="label")]
    numeric = ["XAVG", "YAVG", "ZAVG", "XABSDEV", "YABSDEV", "ZABSDEV", "XSTDDEV", "YSTDDEV", "ZSTDDEV", "RESULTANT"]
    stages += [VectorAssembler(inputCols=[c + "classVec" for c in ["XPEAK", "YPEAK", "ZPEAK"]] + numeric, outputCol="features")]
    model = Pipeline(stages=stages).fit(data)
    df = model.transform(data).select(["label", "features"] + cols)
    train, test = df.randomSplit([0.7, 0.3], seed=2018)
    return data, df, train, test, model
