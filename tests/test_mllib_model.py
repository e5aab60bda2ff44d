import numpy as np

from elephas.spark_model import SparkMLlibModel, load_spark_model
from elephas.utils.rdd_utils import to_labeled_point
from elephas_b200.models import RMSprop
from elephas_b200.spark import Matrices, Vectors


def test_serialization(tmp_path, classification_model):
    classification_model.compile(RMSprop(), "categorical_crossentropy", ['acc'])
    spark_model = SparkMLlibModel(model=classification_model, frequency='epoch', mode='synchronous', num_workers=2)
    path = str(tmp_path / "test.h5")
    spark_model.save(path)
    loaded = load_spark_model(path)
    assert loaded.master_network.to_json() == classification_model.to_json()
    assert loaded.num_workers == 2


def test_mllib_model(spark_context, classification_model, mnist_data):
    x_train, y_train, x_test, y_test = mnist_data
    classification_model.compile(RMSprop(), "categorical_crossentropy", ['acc'])
    lp_rdd = to_labeled_point(spark_context, x_train[:400], y_train[:400], categorical=True)
    spark_model = SparkMLlibModel(model=classification_model, frequency='epoch', mode='synchronous', num_workers=2)
    spark_model.fit(lp_rdd, epochs=2, batch_size=32, verbose=0, validation_split=0.1, categorical=True, nb_classes=10)
    score = spark_model.master_network.evaluate(x_test, y_test)
    assert score[1] > 0.5
    probs = spark_model.predict(Matrices.dense(2, 784, x_test[:2].ravel(order='F')))
    assert probs.toArray().shape == (2, 10)
    assert np.allclose(probs.toArray(), spark_model.master_network.predict(x_test[:2]), atol=1e-6)
    assert len(spark_model.predict(Vectors.dense(x_test[0]))) == 10
