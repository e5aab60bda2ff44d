"""Learning-rate schedules (``models/schedules.py``): values against the closed forms of ``tf.keras.optimizers.schedules``,
the optimizer config round trip, and the rate an optimizer really applies."""
import json
import math

import numpy as np
import pytest
import torch

from elephas_b200.models import Dense, Sequential, optimizers, schedules


def test_values_match_the_closed_forms():
    e = schedules.ExponentialDecay(0.1, decay_steps=10, decay_rate=0.5)
    assert e(0) == pytest.approx(0.1) and e(10) == pytest.approx(0.05) and e(5) == pytest.approx(0.1 * 0.5 ** 0.5)
    s = schedules.ExponentialDecay(0.1, 10, 0.5, staircase=True)
    assert s(9) == pytest.approx(0.1) and s(19) == pytest.approx(0.05)
    i = schedules.InverseTimeDecay(0.1, 10, 0.5)
    assert i(20) == pytest.approx(0.1 / 2.0)
    p = schedules.PiecewiseConstantDecay([100, 200], [1.0, 0.5, 0.1])
    assert [p(0), p(100), p(101), p(200), p(201)] == [1.0, 1.0, 0.5, 0.5, 0.1]
    with pytest.raises(ValueError):
        schedules.PiecewiseConstantDecay([1], [1.0])
    poly = schedules.PolynomialDecay(0.1, 100, end_learning_rate=0.01, power=2.0)
    assert poly(50) == pytest.approx(0.09 * 0.25 + 0.01) and poly(1000) == pytest.approx(0.01)
    cyc = schedules.PolynomialDecay(0.1, 100, end_learning_rate=0.01, cycle=True)
    assert cyc(150) == pytest.approx(0.09 * (1 - 150 / 200) + 0.01)
    c = schedules.CosineDecay(0.1, 100, alpha=0.1)
    assert c(0) == pytest.approx(0.1) and c(100) == pytest.approx(0.01) and c(50) == pytest.approx(0.1 * (0.9 * 0.5 + 0.1))
    r = schedules.CosineDecayRestarts(0.1, 10, t_mul=2.0, m_mul=0.5)
    assert r(0) == pytest.approx(0.1) and r(10) == pytest.approx(0.05) and r(30) == pytest.approx(0.025)
    assert r(20) == pytest.approx(0.05 * 0.5 * (1 + math.cos(math.pi * 0.5)))
    flat = schedules.CosineDecayRestarts(0.1, 10, t_mul=1.0)
    assert flat(25) == pytest.approx(0.1 * 0.5 * (1 + math.cos(math.pi * 0.5)))


def test_optimizer_config_round_trip_and_applied_rate():
    sched = schedules.ExponentialDecay(0.5, decay_steps=1, decay_rate=0.5)
    opt = optimizers.SGD(learning_rate=sched)
    cfg = optimizers.serialize(opt)
    assert cfg['config']['learning_rate'] == {'class_name': 'ExponentialDecay', 'config': sched.get_config()}
    json.dumps(cfg)                                           # what travels to the workers is plain JSON
    back = optimizers.deserialize(cfg)
    assert isinstance(back.schedule, schedules.ExponentialDecay) and back.schedule.get_config() == sched.get_config()
    w = torch.zeros(1, requires_grad=True)
    for expected in (0.5, 0.25, 0.125):                        # gradient 1: the step IS the rate
        before = float(w.detach())
        back.apply_gradients([w], [torch.ones(1)])
        assert before - float(w.detach()) == pytest.approx(expected)
    with pytest.raises(ValueError):
        optimizers.SGD(learning_rate={'class_name': 'Sideways', 'config': {}})


def test_a_model_with_a_schedule_trains_and_keeps_it_through_json():
    from elephas_b200.models import model_from_json
    rng = np.random.default_rng(0)
    x = rng.standard_normal((128, 5)).astype(np.float32)
    y = (x @ rng.standard_normal((5, 1))).astype(np.float32)
    m = Sequential([Dense(8, activation='tanh', input_shape=(5,)), Dense(1)])
    m.compile(optimizers.Adam(learning_rate=schedules.CosineDecay(0.05, 200)), 'mse')
    h = m.fit(x, y, batch_size=16, epochs=8)
    assert h.history['loss'][-1] < 0.5 * h.history['loss'][0]
    assert m.optimizer.iterations == 64
    from elephas_b200.ops.engine import _optimizer_spec
    assert _optimizer_spec(m.optimizer) is None and _optimizer_spec(optimizers.Adam(0.01)) is not None
    clone = model_from_json(m.to_json())
    clone.compile(optimizers.deserialize(optimizers.serialize(m.optimizer)), 'mse')
    assert clone.optimizer.schedule(100) == pytest.approx(0.025)


def test_adamw_and_the_newer_optimizer_config_keys(tmp_path):
    from elephas_b200.models import load_model
    w = torch.full((3,), 2.0, requires_grad=True)
    ref = torch.full((3,), 2.0, requires_grad=True)
    ours = optimizers.AdamW(learning_rate=0.1, weight_decay=0.05)
    theirs = torch.optim.AdamW([ref], lr=0.1, weight_decay=0.05, eps=1e-7)
    for step in range(5):
        g = torch.tensor([0.5, -1.0, 2.0]) * (step + 1)
        ours.apply_gradients([w], [g])
        ref.grad = g.clone()
        theirs.step()
    np.testing.assert_allclose(w.detach().numpy(), ref.detach().numpy(), rtol=1e-5)
    assert optimizers.AdamW().weight_decay == 0.004 and optimizers.Adam().weight_decay is None
    assert 'weight_decay' not in optimizers.Adam().get_config()
    # a Keras 2.11+ / Keras 3 optimizer config: new keys, no `decay`
    cfg = {'class_name': 'Adam', 'config': {'name': 'Adam', 'weight_decay': None, 'clipnorm': None, 'global_clipnorm': None,
                                            'clipvalue': None, 'use_ema': False, 'ema_momentum': 0.99,
                                            'ema_overwrite_frequency': None, 'jit_compile': True, 'is_legacy_optimizer': False,
                                            'learning_rate': 0.001, 'beta_1': 0.9, 'beta_2': 0.999, 'epsilon': 1e-07,
                                            'amsgrad': False, 'loss_scale_factor': None, 'gradient_accumulation_steps': None}}
    opt = optimizers.deserialize(cfg)
    assert opt.learning_rate == 0.001 and opt.weight_decay is None
    with pytest.raises(ValueError):
        optimizers.deserialize({'class_name': 'SGD', 'config': {'use_ema': True}})
    from elephas_b200.ops.engine import _optimizer_spec
    assert _optimizer_spec(optimizers.SGD(0.1, weight_decay=1e-4)) is None
    # through a model file
    m = Sequential([Dense(4, activation='relu', input_shape=(5,)), Dense(1)])
    m.compile(optimizers.AdamW(0.01), 'mse')
    rng = np.random.default_rng(0)
    x, y = rng.standard_normal((64, 5)).astype(np.float32), rng.standard_normal((64, 1)).astype(np.float32)
    m.fit(x, y, batch_size=16, epochs=2)
    path = str(tmp_path / 'adamw.h5')
    m.save(path)
    back = load_model(path)
    assert type(back.optimizer).__name__ == 'AdamW' and back.optimizer.weight_decay == 0.004
    assert back.optimizer.iterations == 8
    a = m.train_on_batch(x[:16], y[:16])
    b = back.train_on_batch(x[:16], y[:16])
    assert a == pytest.approx(b, rel=1e-6)
    np.testing.assert_allclose(back.get_weights()[0], m.get_weights()[0], rtol=1e-5, atol=1e-7)


def test_gradient_accumulation_applies_the_mean_every_k_calls():
    w = torch.zeros(2, requires_grad=True)
    opt = optimizers.SGD(0.5, gradient_accumulation_steps=3)
    grads = [torch.tensor([1.0, 2.0]), torch.tensor([3.0, 0.0]), torch.tensor([2.0, 4.0]), torch.tensor([1.0, 1.0])]
    seen = []
    for g in grads:
        opt.apply_gradients([w], [g])
        seen.append(w.detach().tolist())
    assert seen == [[0.0, 0.0], [0.0, 0.0], [-1.0, -1.0], [-1.0, -1.0]] and opt.iterations == 1
    back = optimizers.deserialize(optimizers.serialize(opt))
    assert back.gradient_accumulation_steps == 3
    with pytest.raises(ValueError):
        optimizers.SGD(gradient_accumulation_steps=1)
    from elephas_b200.ops.engine import _optimizer_spec
    assert _optimizer_spec(opt) is None
    # k batches of size b with accumulation == one batch of size k * b without (plain SGD, mean-reduced loss)
    rng = np.random.default_rng(0)
    x, y = rng.standard_normal((48, 5)).astype(np.float32), rng.standard_normal((48, 1)).astype(np.float32)
    big = Sequential([Dense(4, activation='tanh', input_shape=(5,)), Dense(1)])
    small = Sequential([Dense(4, activation='tanh', input_shape=(5,)), Dense(1)])
    small.set_weights(big.get_weights())
    big.compile(optimizers.SGD(0.1), 'mse')
    small.compile(optimizers.SGD(0.1, gradient_accumulation_steps=3), 'mse')
    big.fit(x, y, batch_size=48, epochs=1, shuffle=False)
    small.fit(x, y, batch_size=16, epochs=1, shuffle=False)
    for a, b in zip(big.get_weights(), small.get_weights()):
        np.testing.assert_allclose(a, b, rtol=1e-5, atol=1e-6)
