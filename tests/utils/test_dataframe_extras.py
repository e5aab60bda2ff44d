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
