"""Regularizers and constraints of the Keras-compatible layers (``models/regularizers.py``, ``models/constraints.py``): the
layer configs a Keras model JSON carries them in, the penalty in the minimised and reported loss, the projection after
every optimizer step."""
import json

import numpy as np
import pytest
import torch

from elephas_b200.models import (LSTM, Bidirectional, Conv2D, Dense, Flatten, Sequential, constraints, model_from_json,
                                 regularizers)
from elephas_b200.models.optimizers import SGD


def _xy(n=64, d=6, k=3, seed=0):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((n, d)).astype(np.float32)
    y = np.eye(k, dtype=np.float32)[rng.integers(0, k, n)]
    return x, y


def test_regularizer_values_and_identifiers():
    w = torch.tensor([[1.0, -2.0], [0.5, 0.0]])
    assert float(regularizers.l1(0.1)(w)) == pytest.approx(0.35)
    assert float(regularizers.l2(0.1)(w)) == pytest.approx(0.525)
    assert float(regularizers.l1_l2(0.1, 0.1)(w)) == pytest.approx(0.875)
    assert regularizers.get('l2').l2 == pytest.approx(0.01)
    assert regularizers.get({'class_name': 'L1L2', 'config': {'l1': 0.5, 'l2': 0.25}}).get_config() == {'l1': 0.5, 'l2': 0.25}
    assert regularizers.get(None) is None and regularizers.serialize(None) is None
    with pytest.raises(ValueError):
        regularizers.get('l3')
    eye = torch.eye(3)
    assert float(regularizers.OrthogonalRegularizer(0.1)(eye)) == pytest.approx(0.0)
    assert float(regularizers.OrthogonalRegularizer(0.1)(torch.ones(3, 3))) > 0.0


def test_constraint_values_follow_the_keras_formulas():
    w = torch.tensor([[3.0, 0.3], [4.0, -0.4]])            # columns have norms 5 and 0.5
    out = constraints.MaxNorm(2.0, axis=0)(w)
    np.testing.assert_allclose(out.norm(dim=0).numpy(), [2.0, 0.5], rtol=1e-5)
    np.testing.assert_allclose(constraints.UnitNorm(axis=0)(w).norm(dim=0).numpy(), [1.0, 1.0], rtol=1e-5)
    np.testing.assert_allclose(constraints.NonNeg()(w).numpy(), [[3.0, 0.3], [4.0, 0.0]], rtol=1e-6)
    out = constraints.MinMaxNorm(1.0, 2.0, rate=1.0, axis=0)(w)
    np.testing.assert_allclose(out.norm(dim=0).numpy(), [2.0, 1.0], rtol=1e-5)
    half = constraints.MinMaxNorm(1.0, 2.0, rate=0.5, axis=0)(w)
    np.testing.assert_allclose(half.norm(dim=0).numpy(), [3.5, 0.75], rtol=1e-5)
    assert isinstance(constraints.get('max_norm'), constraints.MaxNorm)
    assert constraints.get({'class_name': 'MaxNorm', 'config': {'max_value': 3, 'axis': [0, 1, 2]}}).axis == [0, 1, 2]
    with pytest.raises(ValueError):
        constraints.get('square')


def test_configs_round_trip_through_json():
    m = Sequential([Dense(8, activation='relu', input_shape=(6,), kernel_regularizer=regularizers.l2(1e-3),
                          bias_regularizer='l1', activity_regularizer=regularizers.l1(1e-4),
                          kernel_constraint=constraints.MaxNorm(3.0), bias_constraint='non_neg'),
                    Dense(3, activation='softmax')])
    cfg = json.loads(m.to_json())['config']['layers']
    dense = [l for l in cfg if l['class_name'] == 'Dense'][0]['config']
    assert dense['kernel_regularizer'] == {'class_name': 'L2', 'config': {'l2': 1e-3}}
    assert dense['bias_regularizer'] == {'class_name': 'L1', 'config': {'l1': 0.01}}
    assert dense['kernel_constraint'] == {'class_name': 'MaxNorm', 'config': {'max_value': 3.0, 'axis': 0}}
    assert dense['bias_constraint'] == {'class_name': 'NonNeg', 'config': {}}
    plain = [l for l in cfg if l['class_name'] == 'Dense'][1]['config']
    assert plain['kernel_regularizer'] is None and plain['kernel_constraint'] is None
    back = model_from_json(m.to_json())
    assert json.loads(back.to_json()) == json.loads(m.to_json())
    assert back.has_weight_options() and not Sequential([Dense(2, input_shape=(3,))]).has_weight_options()


def test_l2_penalty_is_in_the_loss_and_in_the_gradient():
    x, y = _xy()
    lam = 0.05

    def build(reg):
        torch.manual_seed(3)
        m = Sequential([Dense(5, activation='tanh', input_shape=(6,), kernel_regularizer=reg),
                        Dense(3, activation='softmax', kernel_regularizer=reg, bias_regularizer=reg)])
        m.compile(SGD(learning_rate=0.1), 'categorical_crossentropy')
        return m

    m = build(regularizers.l2(lam))
    ref = build(None)
    ref.set_weights(m.get_weights())
    w0 = [torch.tensor(w) for w in m.get_weights()]
    loss = m.train_on_batch(x, y)
    loss = loss[0] if isinstance(loss, list) else loss
    # oracle: plain autograd on the same weights in the Keras layout
    ws = [w.clone().requires_grad_(True) for w in w0]
    h = torch.tanh(torch.tensor(x) @ ws[0] + ws[1])
    p = torch.softmax(h @ ws[2] + ws[3], dim=1)
    ce = -(torch.tensor(y) * torch.log(p)).sum(1).mean()
    total = ce + lam * ((ws[0] ** 2).sum() + (ws[2] ** 2).sum() + (ws[3] ** 2).sum())
    total.backward()
    assert loss == pytest.approx(float(total.detach()), rel=1e-5)
    for got, w, wg in zip(m.get_weights(), w0, ws):
        np.testing.assert_allclose(got, (w - 0.1 * wg.grad).numpy(), rtol=1e-4, atol=1e-6)
    plain = ref.train_on_batch(x, y)
    plain = plain[0] if isinstance(plain, list) else plain
    assert plain == pytest.approx(float(ce.detach()), rel=1e-5) and plain < loss


def test_activity_penalty_is_divided_by_the_batch():
    x, y = _xy(32)
    m = Sequential([Dense(4, activation='relu', input_shape=(6,), activity_regularizer=regularizers.l1(0.1)),
                    Dense(3, activation='softmax')])
    m.compile(SGD(learning_rate=0.0), 'categorical_crossentropy')
    ws = [torch.tensor(w) for w in m.get_weights()]
    h = torch.relu(torch.tensor(x) @ ws[0] + ws[1])
    p = torch.softmax(h @ ws[2] + ws[3], dim=1)
    expect = float(-(torch.tensor(y) * torch.log(p)).sum(1).mean() + 0.1 * h.abs().sum() / 32)
    got = m.train_on_batch(x, y)
    got = got[0] if isinstance(got, list) else got
    assert got == pytest.approx(expect, rel=1e-5)
    # evaluation reports data loss + penalties too (keras.Model.test_step passes model.losses to the compiled loss)
    assert m.evaluate(x, y, batch_size=32) == pytest.approx(expect, rel=1e-5)


def test_constraints_hold_after_every_step_in_the_keras_layout():
    x, y = _xy(128)
    m = Sequential([Dense(16, activation='relu', input_shape=(6,), kernel_constraint=constraints.MaxNorm(0.5, axis=0),
                          bias_constraint=constraints.NonNeg()),
                    Dense(3, activation='softmax', kernel_constraint=constraints.UnitNorm(axis=0))])
    m.compile(SGD(learning_rate=0.5), 'categorical_crossentropy')
    m.fit(x, y, batch_size=16, epochs=2)
    k0, b0, k1, _ = m.get_weights()
    assert k0.shape == (6, 16) and np.all(np.linalg.norm(k0, axis=0) <= 0.5 + 1e-5)
    assert np.all(b0 >= 0.0)
    np.testing.assert_allclose(np.linalg.norm(k1, axis=0), np.ones(3), rtol=1e-4)


def test_conv_and_recurrent_layers_take_the_options():
    conv = Sequential([Conv2D(4, 3, input_shape=(8, 8, 2), kernel_regularizer='l2',
                              kernel_constraint=constraints.MaxNorm(1.0, axis=[0, 1, 2])),
                       Flatten(), Dense(3, activation='softmax')])
    conv.compile(SGD(learning_rate=0.2), 'categorical_crossentropy')
    rng = np.random.default_rng(0)
    x = rng.standard_normal((32, 8, 8, 2)).astype(np.float32)
    y = np.eye(3, dtype=np.float32)[rng.integers(0, 3, 32)]
    conv.fit(x, y, batch_size=8, epochs=1)
    k = conv.get_weights()[0]
    assert k.shape == (3, 3, 2, 4)
    assert np.all(np.sqrt((k ** 2).sum(axis=(0, 1, 2))) <= 1.0 + 1e-5)

    rnn = Sequential([Bidirectional(LSTM(4, kernel_regularizer='l2', recurrent_regularizer=regularizers.l1(0.01),
                                         recurrent_constraint='max_norm'), input_shape=(5, 3)),
                      Dense(2, activation='softmax')])
    rnn.compile(SGD(learning_rate=0.1), 'categorical_crossentropy')
    penalties, cons, activity = rnn._weight_options()
    assert len(penalties) == 4 and len(cons) == 2 and not activity   # forward + backward copy of the LSTM
    spec = json.loads(rnn.to_json())['config']['layers']
    inner = [l for l in spec if l['class_name'] == 'Bidirectional'][0]['config']['layer']['config']
    assert inner['recurrent_regularizer'] == {'class_name': 'L1', 'config': {'l1': 0.01}}
    xs = rng.standard_normal((16, 5, 3)).astype(np.float32)
    ys = np.eye(2, dtype=np.float32)[rng.integers(0, 2, 16)]
    with_penalty = rnn.train_on_batch(xs, ys)
    with_penalty = with_penalty[0] if isinstance(with_penalty, list) else with_penalty
    assert with_penalty == pytest.approx(rnn.evaluate(xs, ys, batch_size=16), rel=0.2)
    back = model_from_json(rnn.to_json())
    assert len(back._weight_options()[0]) == 4
