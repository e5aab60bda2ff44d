"""Functional models with several outputs: one loss per output, ``loss_weights``, per-output metrics, list predictions."""
import numpy as np
import pytest
import torch

from elephas_b200.models import Dense, Input, Model, load_model, model_from_json
from elephas_b200.models.optimizers import SGD


def _model():
    inp = Input(shape=(5,), name='features')
    trunk = Dense(8, activation='tanh', name='trunk')(inp)
    cls = Dense(3, activation='softmax', name='cls')(trunk)
    reg = Dense(1, name='reg')(trunk)
    return Model(inputs=inp, outputs=[cls, reg])


def _data(n=240, seed=0):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((n, 5)).astype(np.float32)
    yc = np.eye(3, dtype=np.float32)[(x[:, :3]).argmax(1)]
    yr = (x @ rng.standard_normal((5, 1))).astype(np.float32)
    return x, yc, yr


def test_weighted_sum_of_losses_and_the_reported_names():
    x, yc, yr = _data(32)
    m = _model()
    m.compile(SGD(learning_rate=0.1), loss={'cls': 'categorical_crossentropy', 'reg': 'mse'},
              loss_weights={'cls': 1.0, 'reg': 0.25}, metrics={'cls': ['acc'], 'reg': ['mae']})
    assert m.metrics_names == ['loss', 'cls_loss', 'reg_loss', 'cls_acc', 'reg_mae']
    w0 = [torch.tensor(w, requires_grad=True) for w in m.get_weights()]
    h = torch.tanh(torch.tensor(x) @ w0[0] + w0[1])
    p = torch.softmax(h @ w0[2] + w0[3], dim=1)
    r = h @ w0[4] + w0[5]
    ce = -(torch.tensor(yc) * torch.log(p)).sum(1).mean()
    mse = ((r - torch.tensor(yr)) ** 2).mean()
    total = ce + 0.25 * mse
    total.backward()
    stats = m.train_on_batch(x, [yc, yr], return_dict=True)
    assert stats['loss'] == pytest.approx(float(total.detach()), rel=1e-5)
    assert stats['cls_loss'] == pytest.approx(float(ce.detach()), rel=1e-5)
    assert stats['reg_loss'] == pytest.approx(float(mse.detach()), rel=1e-5)
    for got, w in zip(m.get_weights(), w0):
        np.testing.assert_allclose(got, (w - 0.1 * w.grad).detach().numpy(), rtol=1e-4, atol=1e-6)


def test_fit_evaluate_predict_and_files(tmp_path):
    x, yc, yr = _data()
    torch.manual_seed(0)
    m = _model()
    m.compile(SGD(learning_rate=0.1), loss=['categorical_crossentropy', 'mse'], metrics=['mae'])
    assert m.metrics_names == ['loss', 'cls_loss', 'reg_loss', 'cls_mae', 'reg_mae']
    h = m.fit(x, {'reg': yr, 'cls': yc}, batch_size=24, epochs=15, validation_split=0.2)
    assert h.history['loss'][-1] < 0.6 * h.history['loss'][0] and 'val_reg_loss' in h.history
    ev = m.evaluate(x, [yc, yr], batch_size=64, return_dict=True)
    assert ev['loss'] == pytest.approx(ev['cls_loss'] + ev['reg_loss'], rel=1e-5)
    pc, pr = m.predict(x)
    assert pc.shape == (240, 3) and pr.shape == (240, 1)
    assert float(np.mean((pr - yr) ** 2)) == pytest.approx(ev['reg_loss'], rel=1e-4)
    np.testing.assert_allclose(m.predict(x[:5])[0], pc[:5], rtol=1e-6, atol=1e-7)
    back = model_from_json(m.to_json())
    back.set_weights(m.get_weights())
    np.testing.assert_allclose(back.predict(x)[1], pr, rtol=1e-6, atol=1e-7)
    path = str(tmp_path / 'two_heads.h5')
    m.save(path)
    again = load_model(path)
    assert again.metrics_names == m.metrics_names
    assert again.evaluate(x, [yc, yr], batch_size=64)[0] == pytest.approx(ev['loss'], rel=1e-5)


def test_bad_arguments_are_rejected():
    x, yc, yr = _data(16)
    m = _model()
    with pytest.raises(ValueError):
        m.compile('sgd', loss={'cls': 'mse', 'nope': 'mse'})
    with pytest.raises(ValueError):
        m.compile('sgd', loss=['mse'])
    with pytest.raises(ValueError):
        m.compile('sgd', loss={'cls': 'mse'})
    m.compile('sgd', loss='mse')
    with pytest.raises(ValueError):
        m.fit(x, yc)
    with pytest.raises(ValueError):
        m.fit(x, [yc, yr], class_weight={0: 2.0})
    m.fit(x, [yc, yr], sample_weight=np.ones(16), epochs=1)
