import numpy as np

from elephas.ml import adapter


def test_to_data_frame(spark_context):
    df = adapter.to_data_frame(spark_context, np.ones((2, 10)), np.asarray([[2.0], [1.0]]))
    assert df.count() == 2 and df.columns == ['features', 'label']


def test_to_data_frame_cat(spark_context):
    labels = np.asarray([[0, 0, 1.0], [0, 1.0, 0]])
    df = adapter.to_data_frame(spark_context, np.ones((2, 10)), labels, categorical=True)
    assert df.count() == 2 and df.first().label == 2.0


def test_from_data_frame(spark_context):
    features, labels = np.ones((2, 10)), np.asarray([[2.0], [1.0]]).reshape((2,))
    df = adapter.to_data_frame(spark_context, features, labels)
    x, y = adapter.from_data_frame(df)
    assert features.shape == x.shape and labels.shape == y.shape


def test_from_data_frame_cat(spark_context):
    features, labels = np.ones((2, 10)), np.asarray([[0, 0, 1.0], [0, 1.0, 0]])
    df = adapter.to_data_frame(spark_context, features, labels, categorical=True)
    x, y = adapter.from_data_frame(df, True, 3)
    assert features.shape == x.shape and labels.shape == y.shape


def test_df_to_simple_rdd(spark_context):
    df = adapter.to_data_frame(spark_context, np.ones((2, 10)), np.asarray([[2.0], [1.0]]).reshape((2,)))
    rdd = adapter.df_to_simple_rdd(df, False)
    assert rdd.count() == 2 and rdd.first()[0].shape == (10,) and rdd.first()[1] == 2.0


def test_df_to_simple_rdd_custom_columns(spark_context):
    df = adapter.to_data_frame(spark_context, np.ones((2, 10)), np.asarray([0.0, 1.0]))
    df = df.withColumnRenamed('features', 'f').withColumnRenamed('label', 'l')
    rdd = adapter.df_to_simple_rdd(df, True, 2, features_col='f', label_col='l')
    assert rdd.first()[1].shape == (2,)
