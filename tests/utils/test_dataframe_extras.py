"""DataFrame helpers the reference's notebooks use on real PySpark frames (``limit``, ``randomSplit``)."""
from elephas_b200.spark import SparkSession


def _frame(n=100):
    spark = SparkSession.builder.appName('t').getOrCreate()
    return spark.createDataFrame([(i, float(i) * 0.5) for i in range(n)], ['a', 'b'])


def test_limit_keeps_order_and_schema():
    df = _frame()
    small = df.limit(7)
    assert small.count() == 7 and small.columns == ['a', 'b']
    assert [r.a for r in small.collect()] == list(range(7))
    assert df.limit(1000).count() == 100


def test_random_split_partitions_every_row_once():
    df = _frame(200)
    train, test = df.randomSplit([0.8, 0.2], seed=3)
    assert train.count() + test.count() == 200
    assert sorted([r.a for r in train.collect()] + [r.a for r in test.collect()]) == list(range(200))
    assert 120 < train.count() < 195
    again, _ = df.randomSplit([0.8, 0.2], seed=3)
    assert [r.a for r in again.collect()] == [r.a for r in train.collect()]


def test_rows_share_one_class_per_schema_and_survive_pickle():
    import pickle
    from elephas_b200.spark import Row
    df = _frame(5)
    rows = df.collect()
    assert type(rows[0]) is type(rows[1]) and isinstance(rows[0], Row)
    assert rows[2].a == 2 and rows[2]['b'] == 1.0 and rows[2][0] == 2 and rows[2].asDict() == {'a': 2, 'b': 1.0}
    back = pickle.loads(pickle.dumps(rows[3]))
    assert back == rows[3] and back.b == 1.5 and isinstance(back, Row)
    kw = Row(x=1, y='q')
    assert kw.x == 1 and pickle.loads(pickle.dumps(kw)).y == 'q'
    wider = df.withColumn('c', __import__('elephas_b200.spark', fromlist=['functions']).functions.col('a'))
    assert wider.columns == ['a', 'b', 'c'] and wider.collect()[4].c == 4 and type(wider.collect()[0]) is not type(rows[0])


def test_df_to_simple_rdd_block_path_equals_row_path():
    import numpy as np
    from elephas_b200.ml import adapter
    from elephas_b200.spark import SparkSession, Vectors
    spark = SparkSession.builder.appName('t').getOrCreate()
    rng = np.random.default_rng(0)
    feats = rng.normal(size=(50, 6))
    labels = rng.integers(0, 4, 50).astype(float)
    df = spark.createDataFrame([(Vectors.dense(f), float(l), 'x') for f, l in zip(feats, labels)],
                               ['scaled', 'idx', 'junk']).repartition(3)
    for categorical, nb in ((True, 4), (True, None), (False, None)):
        fast = adapter.df_to_simple_rdd(df, categorical=categorical, nb_classes=nb, features_col='scaled', label_col='idx')
        slow_lp = df.select('scaled', 'idx').rdd.map(lambda r: adapter.LabeledPoint(r[1], adapter._as_mllib_vector(r[0])))
        slow = adapter.lp_to_simple_rdd(slow_lp, categorical, nb)
        a, b = fast.collect(), slow.collect()
        assert len(a) == len(b) == 50 and fast.getNumPartitions() == 3
        for (fa, la), (fb, lb) in zip(a, b):
            assert np.allclose(fa, fb) and np.allclose(la, lb)


def test_standard_scaler_block_path_matches_formula_and_keeps_other_columns():
    import numpy as np
    from elephas_b200.spark import SparkSession, StandardScaler, Vectors
    spark = SparkSession.builder.appName('t').getOrCreate()
    rng = np.random.default_rng(1)
    feats = rng.normal(loc=3.0, scale=2.0, size=(40, 5))
    feats[:, 2] = 7.0  # constant column: std 0 -> scaled to 0 (Spark's behaviour)
    df = spark.createDataFrame([(Vectors.dense(f), i) for i, f in enumerate(feats)], ['features', 'id']).repartition(2)
    model = StandardScaler(inputCol='features', outputCol='scaled', withMean=True, withStd=True).fit(df)
    out = model.transform(df)
    assert out.columns == ['features', 'id', 'scaled']
    got = np.stack([r.scaled.toArray() for r in out.collect()])
    std = feats.std(axis=0, ddof=1)
    want = (feats - feats.mean(axis=0)) * np.where(std > 0, 1.0 / np.where(std > 0, std, 1.0), 0.0)
    assert np.allclose(got, want) and [r.id for r in out.collect()] == list(range(40))


def test_repartition_of_a_single_block_shares_memory():
    import numpy as np
    from elephas_b200.spark import SparkContext
    from elephas_b200.utils.rdd_utils import to_simple_rdd
    x, y = np.arange(40, dtype=np.float32).reshape(10, 4), np.arange(10, dtype=np.float32)
    rdd = to_simple_rdd(SparkContext.getOrCreate(), x, y)
    one = rdd.repartition(1) if rdd.getNumPartitions() != 1 else rdd
    parts = one.repartition(3).partitions()
    assert [len(p) for p in parts] == [4, 3, 3]
    assert np.shares_memory(np.asarray(parts[1].x), np.asarray(one.partitions()[0].x))
    assert np.array_equal(np.concatenate([np.asarray(p.x) for p in parts]), x)


def test_pipeline_stages_exchange_column_blocks_and_build_rows_lazily():
    """StringIndexer -> StandardScaler -> ElephasTransformer pass whole columns as ndarrays ("blocks"); Row objects for
    the appended columns only come into being when a consumer asks for rows, and then hold exactly the block values."""
    import numpy as np
    from elephas.ml_model import ElephasTransformer
    from elephas_b200.models import Dense, Sequential
    from elephas_b200.spark import PipelineModel, SparkSession, StandardScaler, StringIndexer, Vectors
    from elephas_b200.utils.model_utils import ModelType
    rng = np.random.default_rng(5)
    feats = rng.normal(size=(257, 6))
    cats = ['c%d' % (i % 3) for i in range(257)]
    spark = SparkSession.builder.getOrCreate()
    df = spark.createDataFrame([(Vectors.dense(f), c) for f, c in zip(feats, cats)], ['features', 'category'])
    model = Sequential([Dense(5, input_dim=6, activation='relu'), Dense(3, activation='softmax')])
    stages = [StringIndexer(inputCol='category', outputCol='idx').fit(df),
              StandardScaler(inputCol='features', outputCol='scaled', withMean=True, withStd=True).fit(df),
              ElephasTransformer(keras_model_config=model.to_json(), weights=model.get_weights(), featuresCol='scaled',
                                 labelCol='idx', outputCol='prediction', model_type=ModelType.CLASSIFICATION)]
    out = PipelineModel(stages).transform(df)
    assert out.columns == ['features', 'category', 'idx', 'scaled', 'prediction'] and out.count() == 257
    assert out._rdd is None  # nothing has been turned into rows yet
    scaled = (feats - feats.mean(0)) / feats.std(0, ddof=1)
    assert np.allclose(out.column_block('scaled'), scaled)
    expect = model.predict(scaled.astype(np.float32))
    assert np.allclose(out.column_block('prediction'), expect, atol=1e-6)
    labels = sorted(set(cats), key=lambda c: (-cats.count(c), c))
    assert np.array_equal(out.column_block('idx'), np.array([float(labels.index(c)) for c in cats]))
    rows = out.collect()  # now the rows exist, in the original order, with plain Python values
    assert len(rows) == 257 and rows[7].category == cats[7] and isinstance(rows[7].prediction, list)
    assert np.allclose(rows[7].prediction, expect[7], atol=1e-6) and np.allclose(rows[7].scaled.toArray(), scaled[7])
    assert rows[7].idx == float(labels.index(cats[7]))
    # a row-level consumer downstream still works on the lazily built frame
    assert out.select('prediction').count() == 257
    import pytest
    with pytest.raises(ValueError):
        out.with_block_column('prediction', np.zeros((257, 3)), None)
    with pytest.raises(ValueError):
        out.with_block_column('other', np.zeros((5, 3)), None)
