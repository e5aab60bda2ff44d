"""Functional models with several inputs: ``fit([xa, xb], y)`` / ``evaluate`` / ``predict`` / ``train_on_batch``."""
import numpy as np
import pytest
import torch

from elephas_b200.models import Concatenate, Dense, Embedding, Flatten, Input, Model, model_from_json
from elephas_b200.models.optimizers import SGD


def _two_tower():
    a = Input(shape=(4,), name='dense_in')
    b = Input(shape=(3,), name='ids')
    ea = Dense(6, activation='relu')(a)
    eb = Flatten()(Embedding(10, 2)(b))
    out = Dense(2, activation='softmax')(Concatenate()([ea, eb]))
    return Model(inputs=[a, b], outputs=out)


def _data(n=300, seed=0):
    rng = np.random.default_rng(seed)
    xa = rng.standard_normal((n, 4)).astype(np.float32)
    xb = rng.integers(0, 10, (n, 3)).astype(np.float32)
    label = (xa[:, 0] + 2.0 * (xb[:, 0] > 4) - 1.0 > 0).astype(int)   # needs both inputs
    return xa, xb, np.eye(2, dtype=np.float32)[label]


def test_fit_evaluate_predict_with_two_inputs():
    xa, xb, y = _data()
    torch.manual_seed(0)
    m = _two_tower()
    m.compile(SGD(learning_rate=0.3), 'categorical_crossentropy', ['acc'])
    h = m.fit([xa, xb], y, batch_size=32, epochs=25, validation_split=0.2)
    assert h.history['loss'][-1] < h.history['loss'][0] and len(h.history['val_acc']) == 25
    loss, acc = m.evaluate([xa, xb], y, batch_size=64)
    assert acc > 0.8
    p = m.predict([xa, xb])
    assert p.shape == (300, 2) and float((p.argmax(1) == y.argmax(1)).mean()) == pytest.approx(acc, abs=1e-6)
    # the same rows in other batchings, by name, one batch at a time
    np.testing.assert_allclose(m.predict([xa[:7], xb[:7]]), p[:7], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(m.predict({'ids': xb[:7], 'dense_in': xa[:7]}), p[:7], rtol=1e-6, atol=1e-7)
    before = m.test_on_batch([xa[:32], xb[:32]], y[:32], return_dict=True)['loss']
    for _ in range(5):
        m.train_on_batch([xa[:32], xb[:32]], y[:32])
    assert m.test_on_batch([xa[:32], xb[:32]], y[:32], return_dict=True)['loss'] < before
    back = model_from_json(m.to_json())
    back.set_weights(m.get_weights())
    np.testing.assert_allclose(back.predict([xa, xb]), m.predict([xa, xb]), rtol=1e-6, atol=1e-7)


def test_wrong_input_lists_are_rejected_and_single_input_lists_accepted():
    xa, xb, y = _data(16)
    m = _two_tower()
    m.compile('sgd', 'categorical_crossentropy')
    with pytest.raises(ValueError):
        m.fit(xa, y)
    with pytest.raises(ValueError):
        m.predict([xa, xb[:5]])
    from elephas_b200.models import Sequential
    s = Sequential([Dense(2, input_shape=(4,))])
    np.testing.assert_array_equal(s.predict([xa]), s.predict(xa))
    assert s.predict([[0.0, 1.0, 2.0, 3.0]]).shape == (1, 2)    # a list of rows is still a batch of rows
