"""The extra ``pyspark.ml`` stand-ins (``spark/feature.py``, ``spark/ml_evaluation.py``, DataFrame filter / orderBy /
groupBy): results against NumPy / scikit-learn, persistence of fitted stages, a pipeline that ends in ElephasEstimator."""
import numpy as np
import pytest
from sklearn import metrics as sk

from elephas_b200 import spark as S
from elephas_b200.spark.ml import PipelineModel
from elephas_b200.spark.sql import Row

pytestmark = pytest.mark.skipif(S.HAVE_PYSPARK, reason="stand-ins are only used without PySpark")


def _frame(n=60, seed=0):
    rng = np.random.default_rng(seed)
    session = S.SparkSession.builder.getOrCreate()
    cats = np.array(['a', 'b', 'c'])[rng.choice(3, n, p=[0.5, 0.3, 0.2])]
    rows = [Row(x=float(rng.normal()), v=S.Vectors.dense(rng.normal(size=2) * [1.0, 10.0]), cat=str(c)) for c in cats]
    return session.createDataFrame(rows)


def test_assembler_scalers_encoder_normalizer():
    df = _frame()
    x = np.array([r.x for r in df.collect()])
    v = np.array([r.v.toArray() for r in df.collect()])
    out = S.VectorAssembler(inputCols=['x', 'v'], outputCol='f').transform(df)
    f = np.array([r.f.toArray() for r in out.collect()])
    np.testing.assert_allclose(f, np.column_stack([x, v]))
    scaled = S.MinMaxScaler(inputCol='f', outputCol='s', min=-1.0, max=1.0).fit(out).transform(out)
    s = np.array([r.s.toArray() for r in scaled.collect()])
    np.testing.assert_allclose(s.min(axis=0), -1.0), np.testing.assert_allclose(s.max(axis=0), 1.0)
    np.testing.assert_allclose(s, 2 * (f - f.min(0)) / (f.max(0) - f.min(0)) - 1)
    unit = S.Normalizer(inputCol='f', outputCol='u', p=1.0).transform(out)
    np.testing.assert_allclose(np.abs(np.array([r.u.toArray() for r in unit.collect()])).sum(1), 1.0)
    idx = S.StringIndexer(inputCol='cat', outputCol='ci').fit(df)
    hot = S.OneHotEncoder(inputCol='ci', outputCol='oh').fit(idx.transform(df)).transform(idx.transform(df))
    rows = hot.collect()
    assert all(len(r.oh.toArray()) == 2 for r in rows)                       # three categories, the last one dropped
    for r in rows:
        want = np.zeros(2)
        if r.ci < 2:
            want[int(r.ci)] = 1.0
        np.testing.assert_array_equal(r.oh.toArray(), want)
    full = S.OneHotEncoder(inputCol='ci', outputCol='oh', dropLast=False).fit(idx.transform(df)).transform(idx.transform(df))
    assert all(r.oh.toArray().sum() == 1.0 and len(r.oh.toArray()) == 3 for r in full.collect())
    back = S.IndexToString(inputCol='ci', outputCol='name', labels=idx.labels).transform(idx.transform(df))
    assert all(r.name == r.cat for r in back.collect())
    nan = S.SparkSession.builder.getOrCreate().createDataFrame([Row(a=1.0, b=2.0), Row(a=float('nan'), b=3.0)])
    with pytest.raises(ValueError):
        S.VectorAssembler(inputCols=['a', 'b'], outputCol='f').transform(nan)
    assert S.VectorAssembler(inputCols=['a', 'b'], outputCol='f', handleInvalid='skip').transform(nan).count() == 1


def test_fitted_stages_survive_save_and_load(tmp_path):
    df = _frame()
    pipe = S.Pipeline(stages=[S.StringIndexer(inputCol='cat', outputCol='ci'),
                              S.VectorAssembler(inputCols=['x', 'v'], outputCol='f'),
                              S.StandardScaler(inputCol='f', outputCol='z', withMean=True),
                              S.MinMaxScaler(inputCol='z', outputCol='m'),
                              S.OneHotEncoder(inputCol='ci', outputCol='oh')])
    fitted = pipe.fit(df)
    want = fitted.transform(df).collect()
    path = str(tmp_path / 'fitted')
    fitted.save(path)
    again = PipelineModel.load(path).transform(df).collect()
    for a, b in zip(want, again):
        assert a.ci == b.ci
        for col in ('z', 'm', 'oh'):
            np.testing.assert_allclose(getattr(a, col).toArray(), getattr(b, col).toArray())


def test_dataframe_filter_order_group():
    df = _frame(40, seed=1)
    rows = df.collect()
    assert df.filter(df.x > 0).count() == sum(r.x > 0 for r in rows)
    assert df.where(S.functions.col('cat') == 'a').count() == sum(r.cat == 'a' for r in rows)
    assert df.filter(df['cat'].isin('a', 'b')).count() == sum(r.cat in 'ab' for r in rows)
    assert df.filter(lambda r: r.x < -0.5).count() == sum(r.x < -0.5 for r in rows)
    ordered = [r.x for r in df.orderBy('x', ascending=False).collect()]
    assert ordered == sorted((r.x for r in rows), reverse=True)
    two = [(r.cat, r.x) for r in df.sort('cat', 'x').collect()]
    assert two == sorted((r.cat, r.x) for r in rows)
    counts = {r['cat']: r['count'] for r in df.groupBy('cat').count().collect()}
    assert counts == {c: sum(r.cat == c for r in rows) for c in set(r.cat for r in rows)}
    means = {r[0]: r[1] for r in df.groupBy('cat').avg('x').collect()}
    for c, m in means.items():
        assert m == pytest.approx(np.mean([r.x for r in rows if r.cat == c]))
    holes = S.SparkSession.builder.getOrCreate().createDataFrame([Row(a=1.0, b=None), Row(a=2.0, b='k')])
    assert holes.dropna().count() == 1 and holes.dropna(subset=['a']).count() == 2
    with pytest.raises(ValueError):
        df['nope']


def test_evaluators_match_scikit_learn():
    rng = np.random.default_rng(2)
    session = S.SparkSession.builder.getOrCreate()
    label = rng.integers(0, 3, 200)
    pred = np.where(rng.random(200) < 0.7, label, rng.integers(0, 3, 200))
    df = session.createDataFrame([Row(prediction=float(p), label=float(l)) for p, l in zip(pred, label)])
    ev = S.MulticlassClassificationEvaluator()
    assert ev.evaluate(df) == pytest.approx(sk.f1_score(label, pred, average='weighted'))
    assert ev.setMetricName('accuracy').evaluate(df) == pytest.approx(sk.accuracy_score(label, pred))
    assert S.MulticlassClassificationEvaluator(metricName='weightedRecall').evaluate(df) == \
        pytest.approx(sk.recall_score(label, pred, average='weighted'))
    assert ev.isLargerBetter()
    with pytest.raises(ValueError):
        S.MulticlassClassificationEvaluator(metricName='auc')
    y = rng.normal(size=100)
    yhat = y + 0.3 * rng.normal(size=100)
    reg = session.createDataFrame([Row(prediction=float(a), label=float(b)) for a, b in zip(yhat, y)])
    assert S.RegressionEvaluator().evaluate(reg) == pytest.approx(np.sqrt(sk.mean_squared_error(y, yhat)))
    assert S.RegressionEvaluator(metricName='r2').evaluate(reg) == pytest.approx(sk.r2_score(y, yhat))
    assert S.RegressionEvaluator(metricName='mae').evaluate(reg) == pytest.approx(sk.mean_absolute_error(y, yhat))
    assert not S.RegressionEvaluator().isLargerBetter()
    z = (rng.random(300) < 0.4).astype(float)
    score = np.clip(0.3 * z + 0.7 * rng.random(300), 0, 1)
    bdf = session.createDataFrame([Row(rawPrediction=S.Vectors.dense([1 - s, s]), label=float(t)) for s, t in zip(score, z)])
    assert S.BinaryClassificationEvaluator().evaluate(bdf) == pytest.approx(sk.roc_auc_score(z, score), abs=1e-9)
    pr = S.BinaryClassificationEvaluator(metricName='areaUnderPR').evaluate(bdf)
    assert pr == pytest.approx(sk.average_precision_score(z, score), abs=0.02)


def test_a_pipeline_of_stand_ins_ends_in_the_estimator():
    from elephas_b200.ml_model import ElephasEstimator
    from elephas_b200.models import Dense, Sequential, optimizers
    rng = np.random.default_rng(3)
    session = S.SparkSession.builder.getOrCreate()
    a, b = rng.normal(size=400), rng.normal(size=400)
    cat = np.where(a + b > 0, 'pos', 'neg')
    df = session.createDataFrame([Row(a=float(x), b=float(y), cat=str(c)) for x, y, c in zip(a, b, cat)])
    model = Sequential([Dense(8, activation='relu', input_shape=(2,)), Dense(2, activation='softmax')])
    est = ElephasEstimator(featuresCol='features', labelCol='label')
    est.set_keras_model_config(model.to_json())
    est.set_categorical_labels(True).set_nb_classes(2).set_num_workers(1).set_epochs(12).set_batch_size(32)
    est.set_validation_split(0.0).set_verbosity(0).set_mode('synchronous').set_loss('categorical_crossentropy')
    est.set_metrics(['acc']).set_optimizer_config(optimizers.serialize(optimizers.Adam(learning_rate=0.05)))
    pipe = S.Pipeline(stages=[S.StringIndexer(inputCol='cat', outputCol='label'),
                              S.VectorAssembler(inputCols=['a', 'b'], outputCol='raw'),
                              S.MinMaxScaler(inputCol='raw', outputCol='features'), est])
    out = pipe.fit(df).transform(df)
    scored = out.select(S.functions.array_argmax('prediction').alias('prediction'), 'label')
    scored = scored.withColumn('prediction', S.functions.col('prediction').cast(S.DoubleType()))
    assert S.MulticlassClassificationEvaluator(metricName='accuracy').evaluate(scored) > 0.9


def test_reader_and_pandas_frames(tmp_path):
    import pandas as pd
    session = S.SparkSession.builder.getOrCreate()
    path = tmp_path / 'data.csv'
    path.write_text("id,score,name\n1,0.5,ann\n2,,bob\n3,2.25,cy\n")
    df = session.read.csv(str(path), header=True, inferSchema=True)
    assert df.columns == ['id', 'score', 'name']
    rows = df.collect()
    assert rows[0].id == 1 and rows[0].score == 0.5 and rows[1].score is None and rows[2].name == 'cy'
    assert df.dropna().count() == 2
    same = session.read.format('csv').option('header', 'true').option('inferSchema', 'true').load(str(path))
    assert same.collect() == rows
    text = session.read.csv(str(path))
    assert text.columns == ['_c0', '_c1', '_c2'] and text.first()[0] == 'id'
    lines = tmp_path / 'data.json'
    lines.write_text('{"a": 1, "b": "x"}\n{"a": 2}\n')
    js = session.read.json(str(lines))
    assert js.columns == ['a', 'b'] and js.collect()[1].b is None
    frame = pd.DataFrame({'u': [1.5, 2.5], 'w': ['p', 'q']})
    back = session.createDataFrame(frame)
    assert back.columns == ['u', 'w'] and back.collect()[1].w == 'q' and isinstance(back.first().u, float)
    assert back.toPandas().equals(frame)
