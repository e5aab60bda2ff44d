import numpy as np

from elephas.utils import rdd_utils


def test_to_simple_rdd(spark_context):
    features, labels = np.ones((5, 10)), np.ones((5,))
    rdd = rdd_utils.to_simple_rdd(spark_context, features, labels)
    assert rdd.count() == 5
    first = rdd.first()
    assert first[0].shape == (10,) and first[1] == 1.0


def test_to_labeled_rdd_categorical(spark_context):
    features = np.ones((2, 10))
    labels = np.asarray([[0, 0, 1.0], [0, 1.0, 0]])
    lp_rdd = rdd_utils.to_labeled_point(spark_context, features, labels, True)
    assert lp_rdd.count() == 2
    first = lp_rdd.first()
    assert first.features.toArray().shape == (10,) and first.label == 2.0


def test_to_labeled_rdd_not_categorical(spark_context):
    lp_rdd = rdd_utils.to_labeled_point(spark_context, np.ones((2, 10)), np.asarray([[2.0], [1.0]]), False)
    assert lp_rdd.first().label == 2.0


def test_from_labeled_rdd(spark_context):
    features, labels = np.ones((2, 10)), np.asarray([[2.0], [1.0]]).reshape((2,))
    lp_rdd = rdd_utils.to_labeled_point(spark_context, features, labels, False)
    x, y = rdd_utils.from_labeled_point(lp_rdd, False, None)
    assert x.shape == features.shape and y.shape == labels.shape


def test_from_labeled_rdd_categorical(spark_context):
    features = np.ones((2, 10))
    labels = np.asarray([[0, 0, 1.0], [0, 1.0, 0]])
    lp_rdd = rdd_utils.to_labeled_point(spark_context, features, labels, True)
    x, y = rdd_utils.from_labeled_point(lp_rdd, True, 3)
    assert x.shape == features.shape and np.array_equal(y, labels)


def test_encode_label():
    assert np.array_equal(rdd_utils.encode_label(3, 10), np.eye(10)[3])


def test_lp_to_simple_rdd_categorical(spark_context):
    features = np.ones((2, 10))
    labels = np.asarray([[0, 0, 1.0], [0, 1.0, 0]])
    lp_rdd = rdd_utils.to_labeled_point(spark_context, features, labels, True)
    rdd = rdd_utils.lp_to_simple_rdd(lp_rdd, categorical=True, nb_classes=3)
    assert rdd.first()[0].shape == (10,) and rdd.first()[1].shape == (3,)


def test_lp_to_simple_rdd_not_categorical(spark_context):
    lp_rdd = rdd_utils.to_labeled_point(spark_context, np.ones((2, 10)), np.asarray([[2.0], [1.0]]).reshape((2,)), False)
    rdd = rdd_utils.lp_to_simple_rdd(lp_rdd, categorical=False, nb_classes=3)
    assert rdd.first()[0].shape == (10,) and rdd.first()[1] == 2.0


def test_lp_to_simple_rdd_infers_nb_classes(spark_context):
    labels = np.eye(5)[[0, 4, 2]]
    lp_rdd = rdd_utils.to_labeled_point(spark_context, np.ones((3, 4)), labels, True)
    rdd = rdd_utils.lp_to_simple_rdd(lp_rdd, categorical=True)
    assert rdd.first()[1].shape == (5,)
