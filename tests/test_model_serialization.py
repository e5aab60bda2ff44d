import numpy as np

from elephas.spark_model import SparkMLlibModel, SparkModel, load_spark_model


def test_sequential_serialization(tmp_path, classification_model):
    classification_model.compile(optimizer="sgd", loss="categorical_crossentropy", metrics=["acc"])
    spark_model = SparkModel(classification_model, frequency='epoch', mode='synchronous')
    path = str(tmp_path / "elephas_sequential.h5")
    spark_model.save(path)
    loaded = load_spark_model(path)
    assert isinstance(loaded, SparkModel)
    assert loaded.master_network.to_json() == classification_model.to_json()
    for a, b in zip(loaded.master_network.get_weights(), classification_model.get_weights()):
        assert np.array_equal(a, b)
    assert loaded.mode == 'synchronous' and loaded.master_optimizer['class_name'] == 'SGD'


def test_model_serialization(tmp_path, classification_model_functional):
    classification_model_functional.compile(optimizer="sgd", loss="categorical_crossentropy", metrics=["acc"])
    spark_model = SparkModel(classification_model_functional, frequency='epoch', mode='synchronous', foo="bar")
    path = str(tmp_path / "elephas_model.h5")
    spark_model.save(path)
    assert spark_model.get_config()['foo'] == 'bar'
    loaded = load_spark_model(path)
    assert loaded.get_config()['foo'] == 'bar'
    assert loaded.master_network.to_json() == classification_model_functional.to_json()


def test_mllib_model_class_round_trips(tmp_path, classification_model):
    classification_model.compile("rmsprop", "categorical_crossentropy", ['acc'])
    path = str(tmp_path / "mllib.keras")
    SparkMLlibModel(classification_model, mode='synchronous').save(path, overwrite=True)
    assert isinstance(load_spark_model(path), SparkMLlibModel)


def test_save_refuses_other_suffix(tmp_path, classification_model):
    import pytest
    classification_model.compile("sgd", "mse")
    with pytest.raises(AssertionError):
        SparkModel(classification_model, mode='synchronous').save(str(tmp_path / "m.bin"))
