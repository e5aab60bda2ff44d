"""Initializer vocabulary of the layer configs (``models/layers.py::_initialize``): names and configs as Keras writes them."""
import json
import math

import numpy as np
import pytest

from elephas_b200.models import Dense, Sequential, model_from_json


def _kernel(init, shape=(300, 200)):
    m = Sequential([Dense(shape[1], input_shape=(shape[0],), kernel_initializer=init)])
    return m.get_weights()[0], m


@pytest.mark.parametrize('init,std', [
    ({'class_name': 'VarianceScaling', 'config': {'scale': 2.0, 'mode': 'fan_in', 'distribution': 'truncated_normal', 'seed': 1}},
     math.sqrt(2.0 / 300)),
    ({'class_name': 'VarianceScaling', 'config': {'scale': 1.0, 'mode': 'fan_avg', 'distribution': 'uniform', 'seed': 1}},
     math.sqrt(1.0 / 250)),
    ({'class_name': 'VarianceScaling', 'config': {'scale': 1.0, 'mode': 'fan_out', 'distribution': 'untruncated_normal', 'seed': 1}},
     math.sqrt(1.0 / 200)),
    ({'class_name': 'TruncatedNormal', 'config': {'mean': 0.0, 'stddev': 0.1, 'seed': 1}}, 0.1 * 0.8796),
    ('he_normal', math.sqrt(2.0 / 300)), ('glorot_uniform', math.sqrt(2.0 / 500)), ('lecun_normal', math.sqrt(1.0 / 300)),
])
def test_random_initializers_have_the_keras_spread(init, std):
    k, _ = _kernel(init)
    assert k.shape == (300, 200)
    assert abs(float(k.mean())) < 0.01
    assert float(k.std()) == pytest.approx(std, rel=0.05)


def test_seeded_initializers_repeat_and_unseeded_ones_do_not():
    seeded = {'class_name': 'GlorotUniform', 'config': {'seed': 7}}
    np.testing.assert_array_equal(_kernel(seeded)[0], _kernel(seeded)[0])
    assert not np.array_equal(_kernel('glorot_uniform')[0], _kernel('glorot_uniform')[0])


def test_structured_initializers():
    k, _ = _kernel('orthogonal', (64, 64))
    np.testing.assert_allclose(k @ k.T, np.eye(64), atol=1e-4)
    k, _ = _kernel({'class_name': 'Orthogonal', 'config': {'gain': 2.0, 'seed': 3}}, (32, 64))
    np.testing.assert_allclose(k @ k.T, 4.0 * np.eye(32), atol=1e-3)
    k, _ = _kernel('identity', (5, 5))
    np.testing.assert_array_equal(k, np.eye(5, dtype=np.float32))
    k, m = _kernel({'class_name': 'Constant', 'config': {'value': 0.25}}, (4, 3))
    np.testing.assert_array_equal(k, np.full((4, 3), 0.25, np.float32))
    cfg = json.loads(m.to_json())['config']['layers'][-1]['config']['kernel_initializer']
    assert cfg == {'class_name': 'Constant', 'config': {'value': 0.25}}
    assert model_from_json(m.to_json()).get_weights()[0].tolist() == k.tolist()
    with pytest.raises(ValueError):
        _kernel('sideways')
    with pytest.raises(ValueError):
        _kernel({'class_name': 'VarianceScaling', 'config': {'distribution': 'cauchy'}})


def test_the_newer_activation_names():
    import torch

    from elephas_b200.models import activations
    x = torch.linspace(-8, 8, 33)
    for name, ref in (('relu6', torch.nn.functional.relu6), ('mish', torch.nn.functional.mish),
                      ('hard_silu', torch.nn.functional.hardswish), ('hard_swish', torch.nn.functional.hardswish),
                      ('leaky_relu', lambda t: torch.nn.functional.leaky_relu(t, 0.2)), ('silu', torch.nn.functional.silu)):
        got_name, fn = activations.get(name)
        assert got_name == name
        np.testing.assert_allclose(fn(x).numpy(), ref(x).numpy(), rtol=1e-5, atol=1e-6)
    m = Sequential([Dense(4, activation='mish', input_shape=(3,)), Dense(2, activation='log_softmax')])
    back = model_from_json(m.to_json())
    assert [l.activation_name for l in back.layers[1:]] == ['mish', 'log_softmax']
    with pytest.raises(ValueError):
        activations.get('squareplus')
