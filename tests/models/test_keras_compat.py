"""Keras fidelity without Keras: JSON schema, weight order/layout, optimizer formulas (hand-written NumPy
oracles), loss/metric definitions, fit semantics."""
import json
import math

import numpy as np
import pytest
import torch

from elephas_b200.models import (SGD, Adam, Dense, RMSprop, Sequential, losses, metrics, model_from_json, optimizers,
                                 to_categorical, zoo)


def test_sequential_json_schema(classification_model):
    spec = json.loads(classification_model.to_json())
    assert spec['class_name'] == 'Sequential' and spec['keras_version'] and spec['backend'] == 'tensorflow'
    layers = spec['config']['layers']
    assert [l['class_name'] for l in layers] == ['InputLayer', 'Dense', 'Activation', 'Dropout', 'Dense', 'Activation',
                                                 'Dropout', 'Dense', 'Activation']
    dense = layers[1]['config']
    assert dense['units'] == 128 and dense['activation'] == 'linear' and dense['batch_input_shape'] == [None, 784]
    assert dense['kernel_initializer']['class_name'] == 'GlorotUniform'
    assert layers[3]['config']['rate'] == 0.2


def test_functional_json_schema(classification_model_functional):
    spec = json.loads(classification_model_functional.to_json())
    assert spec['class_name'] == 'Functional'
    cfg = spec['config']
    assert cfg['input_layers'][0][1:] == [0, 0] and len(cfg['output_layers']) == 1
    dense = [l for l in cfg['layers'] if l['class_name'] == 'Dense'][0]
    assert dense['inbound_nodes'][0][0][0] == cfg['input_layers'][0][0]
    rebuilt = model_from_json(classification_model_functional.to_json())
    assert rebuilt.to_json() == classification_model_functional.to_json()


def test_parses_keras_written_json():
    """A config as written by real Keras 2.10 (extra keys, InputLayer first) must load."""
    keras_json = json.dumps({"class_name": "Sequential", "config": {"name": "sequential", "layers": [
        {"class_name": "InputLayer", "config": {"batch_input_shape": [None, 13], "dtype": "float32", "sparse": False,
                                                "ragged": False, "name": "dense_input"}},
        {"class_name": "Dense", "config": {"name": "dense", "trainable": True, "batch_input_shape": [None, 13],
                                           "dtype": "float32", "units": 64, "activation": "relu", "use_bias": True,
                                           "kernel_initializer": {"class_name": "GlorotUniform", "config": {"seed": None}},
                                           "bias_initializer": {"class_name": "Zeros", "config": {}},
                                           "kernel_regularizer": None, "bias_regularizer": None,
                                           "activity_regularizer": None, "kernel_constraint": None,
                                           "bias_constraint": None}},
        {"class_name": "Dense", "config": {"name": "dense_1", "trainable": True, "dtype": "float32", "units": 1,
                                           "activation": "linear", "use_bias": True,
                                           "kernel_initializer": {"class_name": "GlorotUniform", "config": {"seed": None}},
                                           "bias_initializer": {"class_name": "Zeros", "config": {}},
                                           "kernel_regularizer": None, "bias_regularizer": None,
                                           "activity_regularizer": None, "kernel_constraint": None,
                                           "bias_constraint": None}}]},
        "keras_version": "2.10.0", "backend": "tensorflow"})
    m = model_from_json(keras_json)
    assert [w.shape for w in m.get_weights()] == [(13, 64), (64,), (64, 1), (1,)]


def test_weight_layout_is_keras(regression_model):
    w = regression_model.get_weights()
    assert [a.shape for a in w] == [(13, 64), (64,), (64, 64), (64,), (64, 1), (1,)]
    x = np.random.rand(5, 13).astype(np.float32)
    h = np.maximum(x @ w[0] + w[1], 0)
    h = np.maximum(h @ w[2] + w[3], 0)
    assert np.allclose(regression_model.predict(x), h @ w[4] + w[5], atol=1e-5)
    new = [a + 1 for a in w]
    regression_model.set_weights(new)
    assert all(np.allclose(a, b) for a, b in zip(regression_model.get_weights(), new))


def test_glorot_uniform_bounds():
    m = Sequential([Dense(50, input_dim=30)])
    k = m.get_weights()[0]
    limit = math.sqrt(6.0 / 80)
    assert np.abs(k).max() <= limit and np.abs(k).max() > 0.8 * limit and np.all(m.get_weights()[1] == 0)


def test_optimizer_serialisation_round_trip():
    sgd = optimizers.SGD(learning_rate=0.01, decay=1e-6, momentum=0.9, nesterov=True)
    conf = optimizers.serialize(sgd)
    assert conf == {'class_name': 'SGD', 'config': {'name': 'SGD', 'learning_rate': 0.01, 'decay': 1e-6,
                                                    'momentum': 0.9, 'nesterov': True}}
    assert optimizers.serialize(optimizers.deserialize(conf)) == conf
    assert optimizers.get('adam').get_config()['epsilon'] == 1e-7
    assert optimizers.get('rmsprop').get_config()['rho'] == 0.9
    assert isinstance(optimizers.get({'class_name': 'Adam', 'config': {'learning_rate': 0.5}}), Adam)
    assert SGD(lr=0.3).get_config()['learning_rate'] == 0.3


def _run(opt, grads_seq, w0):
    w = torch.tensor(w0, dtype=torch.float64)
    for g in grads_seq:
        opt.apply_gradients([w], [torch.tensor(g, dtype=torch.float64)])
    return w.numpy()


def test_sgd_momentum_nesterov_decay_oracle():
    rng = np.random.default_rng(0)
    w0, gs = rng.normal(size=5), [rng.normal(size=5) for _ in range(4)]
    lr, mom, decay = 0.1, 0.9, 0.01
    w, v = w0.copy(), np.zeros(5)
    for t, g in enumerate(gs):
        lr_t = lr / (1 + decay * t)
        v = mom * v - lr_t * g
        w = w + mom * v - lr_t * g
    assert np.allclose(_run(SGD(learning_rate=lr, momentum=mom, nesterov=True, decay=decay), gs, w0), w)


def test_rmsprop_oracle():
    rng = np.random.default_rng(1)
    w0, gs = rng.normal(size=5), [rng.normal(size=5) for _ in range(4)]
    w, rms = w0.copy(), np.zeros(5)
    for g in gs:
        rms = 0.9 * rms + 0.1 * g * g
        w = w - 0.001 * g / (np.sqrt(rms) + 1e-7)
    assert np.allclose(_run(RMSprop(), gs, w0), w)


def test_adam_oracle():
    rng = np.random.default_rng(2)
    w0, gs = rng.normal(size=5), [rng.normal(size=5) for _ in range(5)]
    w, m, v = w0.copy(), np.zeros(5), np.zeros(5)
    for t, g in enumerate(gs, start=1):
        m = 0.9 * m + 0.1 * g
        v = 0.999 * v + 0.001 * g * g
        lr_t = 0.001 * math.sqrt(1 - 0.999 ** t) / (1 - 0.9 ** t)
        w = w - lr_t * m / (np.sqrt(v) + 1e-7)
    assert np.allclose(_run(Adam(), gs, w0), w)


def test_losses_against_numpy():
    rng = np.random.default_rng(3)
    p = rng.random((6, 4)); p /= p.sum(1, keepdims=True)
    y = np.eye(4)[rng.integers(0, 4, 6)]
    tp, ty = torch.tensor(p), torch.tensor(y)
    assert np.allclose(losses.categorical_crossentropy(ty, tp).numpy(), -(y * np.log(np.clip(p, 1e-7, 1 - 1e-7))).sum(1))
    b = rng.random((6, 1)); yb = rng.integers(0, 2, (6, 1)).astype(float)
    bc = np.clip(b, 1e-7, 1 - 1e-7)
    assert np.allclose(losses.binary_crossentropy(torch.tensor(yb), torch.tensor(b)).numpy(),
                       -(yb * np.log(bc) + (1 - yb) * np.log(1 - bc)).mean(1))
    assert np.allclose(losses.mean_squared_error(ty, tp).numpy(), ((p - y) ** 2).mean(1))
    assert np.allclose(losses.mean_absolute_error(ty, tp).numpy(), np.abs(p - y).mean(1))
    yt = rng.random((6, 1)) + 1
    assert np.allclose(losses.mean_absolute_percentage_error(torch.tensor(yt), torch.tensor(b)).numpy(),
                       100 * np.abs((yt - b) / yt).mean(1))
    # 1-D labels against an (n, 1) output (the Boston case)
    assert losses.mean_squared_error(torch.tensor(yt[:, 0]), torch.tensor(b)).shape == (6,)


def test_acc_resolution():
    assert metrics.resolve('acc', 'categorical_crossentropy', 10)[1] is metrics.categorical_accuracy
    assert metrics.resolve('acc', 'binary_crossentropy', 1)[1] is metrics.binary_accuracy
    assert metrics.resolve('accuracy', 'sparse_categorical_crossentropy', 10)[1] is metrics.sparse_categorical_accuracy
    assert metrics.resolve('mae', 'mse', 1)[0] == 'mae'


def test_fit_semantics(mnist_data):
    x, y, _, _ = mnist_data
    m = zoo.mnist_mlp()
    m.compile(SGD(lr=0.1), 'categorical_crossentropy', ['acc'])
    h = m.fit(x, y, batch_size=64, epochs=3, validation_split=0.1, verbose=0)
    assert set(h.history) == {'loss', 'acc', 'val_loss', 'val_acc'} and len(h.history['loss']) == 3
    assert h.history['loss'][-1] < h.history['loss'][0]
    assert m.optimizer.iterations == 3 * math.ceil(900 / 64)  # last partial batch kept, 10% held out
    out = m.train_on_batch(x[:32], y[:32])
    assert isinstance(out, list) and len(out) == 2
    assert isinstance(m.evaluate(x[:50], y[:50]), list)


def test_dropout_only_in_training():
    m = zoo.mnist_mlp(dropout=0.5)
    x = np.random.rand(8, 784).astype(np.float32)
    assert np.array_equal(m.predict(x), m.predict(x))
    t = torch.from_numpy(x)
    assert not torch.equal(m(t, training=True), m(t, training=True))


def test_requires_compile(classification_model):
    with pytest.raises(RuntimeError):
        classification_model.fit(np.zeros((4, 784)), np.zeros((4, 10)))


def test_wide_mlp_param_count_formula():
    m = zoo.wide_mlp(input_dim=32, hidden=32, depth=8, out_dim=1)
    assert m.count_params() == 32 * 32 * 8 + 32 * 8 + 32 + 1
    assert 4096 * 4096 * 8 + 4096 * 8 + 4096 + 1 == 134254593


def test_resnet50_shape():
    m = zoo.resnet50(input_shape=(64, 64, 3), classes=10)
    out = m.predict(np.random.rand(2, 64, 64, 3).astype(np.float32))
    assert out.shape == (2, 10) and np.allclose(out.sum(1), 1, atol=1e-4)
    assert zoo.resnet50().count_params() == 25636712  # keras.applications.ResNet50


def test_batchnorm_fusion_plan_computes_what_the_separate_layers_compute(monkeypatch):
    """The executor folds ``BN -> ReLU`` and ``BN -> Add -> ReLU`` (the tail of a residual block) into the
    BatchNormalization call when nothing else reads the intermediates. Same numbers as layer by layer, in training
    and in inference mode; a BN whose output IS read elsewhere stays unfused."""
    import torch
    from elephas_b200.models import Model
    from elephas_b200.models.layers import Activation, Add, BatchNormalization, Conv2D, GlobalAveragePooling2D, Input
    torch.manual_seed(4)
    inp = Input(shape=(8, 8, 8))
    short = BatchNormalization(name='bn0')(Conv2D(16, 1, name='c0')(inp))          # read by the Add only: plain BN
    y = Activation('relu', name='r1')(BatchNormalization(name='bn1')(Conv2D(16, 3, padding='same', name='c1')(inp)))
    y = BatchNormalization(name='bn2')(Conv2D(16, 1, name='c2')(y))
    out = Activation('relu', name='r2')(Add(name='add')([short, y]))
    side = BatchNormalization(name='bn3')(Conv2D(16, 1, name='c3')(out))            # read twice: must not be fused
    both = Add(name='add2')([Activation('relu', name='r3')(side), side])
    model = Model(inputs=inp, outputs=Dense(3, name='head')(GlobalAveragePooling2D()(both)))
    fused, alias = model._fusion_plan()
    names = {id(l): l.name for l in model.layers}
    assert {names[k[0]]: v[0] for k, v in fused.items()} == {'bn1': True, 'bn2': True}
    assert {names[k[0]] for k in alias} == {'r1', 'add', 'r2'}
    assert [names[v[1][0]] for k, v in fused.items() if names[k[0]] == 'bn2'] == ['bn0']
    x = torch.randn(5, 8, 8, 8)
    results = {}
    for flag in ('1', '0'):
        monkeypatch.setenv('ELEPHAS_B200_FUSE_BN', flag)
        state = [w.detach().clone() for w in model.weights]
        results[flag] = (model(x, training=True).detach().clone(), [w.detach().clone() for w in model.weights])
        results[flag + 'eval'] = model(x, training=False).detach().clone()
        for w, s in zip(model.weights, state):
            w.data.copy_(s)  # the training pass moved the running statistics
    assert torch.allclose(results['1'][0], results['0'][0], atol=1e-6)
    assert all(torch.allclose(a, b, atol=1e-6) for a, b in zip(results['1'][1], results['0'][1]))
    assert torch.allclose(results['1eval'], results['0eval'], atol=1e-6)


def test_zero_padding_in_front_of_a_max_pool_is_folded_into_the_pool(monkeypatch):
    """``ZeroPadding2D -> MaxPooling2D('valid')`` (ResNet's pool1): the executor hands the padding to the pool and skips
    the padded copy when nothing else reads it - same numbers and gradients; a padded tensor that IS read elsewhere, or a
    pool with padding='same', stays as written."""
    import torch
    from elephas_b200.models import Model
    from elephas_b200.models.layers import Add, GlobalAveragePooling2D, Input, MaxPooling2D, ZeroPadding2D
    inp = Input(shape=(9, 7, 8))
    a = MaxPooling2D(3, strides=2, name='pool_a')(ZeroPadding2D(1, name='pad_a')(inp))          # folded
    padded = ZeroPadding2D(((0, 2), (1, 1)), name='pad_b')(inp)                                  # read twice: kept
    b = MaxPooling2D(3, strides=2, name='pool_b')(padded)
    c = MaxPooling2D(3, strides=2, name='pool_c')(padded)
    d = MaxPooling2D(3, strides=2, padding='same', name='pool_d')(ZeroPadding2D(1, name='pad_d')(inp))  # 'same': kept
    model = Model(inputs=inp, outputs=[GlobalAveragePooling2D()(Add()([a, b, c])), GlobalAveragePooling2D()(d)])
    pools, skipped = model._pool_pad_plan()
    names = {id(l): l.name for l in model.layers}
    assert {names[k[0]] for k in pools} == {'pool_a'} and {names[k[0]] for k in skipped} == {'pad_a'}
    x = (torch.randn(3, 9, 7, 8) - 0.5).requires_grad_(True)
    outs = {}
    for flag in ('1', '0'):
        monkeypatch.setenv('ELEPHAS_B200_FUSE_BN', flag)
        y1, y2 = model(x, training=True)
        (g,) = torch.autograd.grad(y1.sum() + 2 * y2.sum(), x)
        outs[flag] = (y1.detach(), y2.detach(), g)
    for u, v in zip(outs['1'], outs['0']):
        assert torch.equal(u, v)
