"""Model JSON in the Keras 3 dialect (``module`` / ``registered_name`` entries, ``DTypePolicy`` dictionaries, ``batch_shape``,
inbound nodes as ``{"args", "kwargs"}``) loads like the Keras 2.10 dialect the reference's TF version writes."""
import json

import numpy as np

from elephas_b200.models import model_from_json

_POLICY = {"module": "keras", "class_name": "DTypePolicy", "config": {"name": "float32"}, "registered_name": None}


def _init(name, **cfg):
    return {"module": "keras.initializers", "class_name": name, "config": cfg, "registered_name": None}


def _layer(cls, name, config, inbound=None, build=None):
    spec = {"module": "keras.layers", "class_name": cls, "config": dict(config, name=name), "registered_name": None}
    if cls != 'InputLayer':
        spec["config"].update(trainable=True, dtype=_POLICY)
        spec["build_config"] = {"input_shape": build or [None, 4]}
    if inbound is not None:
        spec["name"] = name
        spec["inbound_nodes"] = inbound
    return spec


def _kt(name, shape):
    return {"class_name": "__keras_tensor__", "config": {"shape": shape, "dtype": "float32", "keras_history": [name, 0, 0]}}


def _dense(units, activation, **extra):
    return dict(units=units, activation=activation, use_bias=True, kernel_initializer=_init("GlorotUniform", seed=None),
                bias_initializer=_init("Zeros"), kernel_regularizer=None, bias_regularizer=None, kernel_constraint=None,
                bias_constraint=None, **extra)


def test_sequential_in_the_keras3_dialect():
    reg = {"module": "keras.regularizers", "class_name": "L2", "config": {"l2": 0.01}, "registered_name": None}
    spec = {"module": "keras", "class_name": "Sequential", "registered_name": None,
            "config": {"name": "sequential", "trainable": True, "dtype": _POLICY, "build_input_shape": [None, 4], "layers": [
                _layer("InputLayer", "input_layer", {"batch_shape": [None, 4], "dtype": "float32", "sparse": False}),
                _layer("Dense", "dense", dict(_dense(8, "relu", lora_rank=None), kernel_regularizer=reg)),
                _layer("Dropout", "dropout", {"rate": 0.25, "seed": None, "noise_shape": None}),
                _layer("Dense", "dense_1", _dense(3, "softmax"))]},
            "build_config": {"input_shape": [None, 4]},
            "compile_config": {"optimizer": "sgd", "loss": "categorical_crossentropy"}}
    m = model_from_json(json.dumps(spec))
    assert [type(l).__name__ for l in m.layers] == ['InputLayer', 'Dense', 'Dropout', 'Dense']
    assert m.input_shape == (None, 4) and m.output_shape == (None, 3) and m.has_weight_options()
    assert m.predict(np.zeros((2, 4), np.float32)).shape == (2, 3)
    again = model_from_json(m.to_json())      # and from there on it is the 2.10 dialect
    assert json.loads(again.to_json())['config']['layers'][1]['config']['kernel_regularizer'] == {'class_name': 'L2', 'config': {'l2': 0.01}}


def test_functional_graph_with_merge_and_attention_kwargs():
    a, b = _kt("left", [None, 5, 4]), _kt("right", [None, 5, 4])
    spec = {"module": "keras", "class_name": "Functional", "registered_name": None,
            "config": {"name": "functional", "trainable": True, "layers": [
                _layer("InputLayer", "left", {"batch_shape": [None, 5, 4], "dtype": "float32", "sparse": False, "optional": False}, inbound=[]),
                _layer("InputLayer", "right", {"batch_shape": [None, 5, 4], "dtype": "float32", "sparse": False}, inbound=[]),
                _layer("Add", "add", {}, inbound=[{"args": [[a, b]], "kwargs": {}}]),
                _layer("MultiHeadAttention", "mha", {"num_heads": 2, "key_dim": 3, "value_dim": 3, "dropout": 0.0, "use_bias": True,
                                                   "output_shape": None, "attention_axes": None,
                                                   "kernel_initializer": _init("GlorotUniform", seed=None),
                                                   "bias_initializer": _init("Zeros"), "seed": None},
                       inbound=[{"args": [_kt("add", [None, 5, 4])], "kwargs": {"value": b, "query_mask": None}}]),
                _layer("Bidirectional", "bi", {"merge_mode": "concat", "layer": _layer("LSTM", "lstm", {
                    "units": 3, "activation": "tanh", "recurrent_activation": "sigmoid", "use_bias": True, "seed": None,
                    "return_sequences": False, "return_state": False, "go_backwards": False, "stateful": False, "unroll": False,
                    "zero_output_for_mask": False, "dropout": 0.0, "recurrent_dropout": 0.0, "unit_forget_bias": True})},
                       inbound=[{"args": [_kt("mha", [None, 5, 4])], "kwargs": {"mask": None}}]),
                _layer("Dense", "out", _dense(2, "softmax"), inbound=[{"args": [_kt("bi", [None, 6])], "kwargs": {}}])],
                "input_layers": [["left", 0, 0], ["right", 0, 0]], "output_layers": [["out", 0, 0]]}}
    m = model_from_json(json.dumps(spec))
    assert m.output_shape == (None, 2) and len(m._inputs) == 2
    mha = m.get_layer('mha')
    assert [s.layer.name for s in mha.inbound[0]] == ['add', 'right']         # query from the sum, value from 'right'
    x = np.random.default_rng(0).standard_normal((3, 5, 4)).astype(np.float32)
    p = m.predict([x, x[::-1].copy()])
    assert p.shape == (3, 2) and np.allclose(p.sum(1), 1.0, atol=1e-5)
