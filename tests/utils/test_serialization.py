import numpy as np

from elephas.utils import serialization


def test_model_to_dict(classification_model):
    d = serialization.model_to_dict(classification_model)
    assert list(d.keys()) == ['model', 'weights']


def test_dict_to_model(classification_model):
    d = serialization.model_to_dict(classification_model)
    recovered = serialization.dict_to_model(d)
    assert recovered.to_json() == classification_model.to_json()
    for a, b in zip(recovered.get_weights(), classification_model.get_weights()):
        assert np.array_equal(a, b)
