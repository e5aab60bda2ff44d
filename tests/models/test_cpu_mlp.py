"""CPU training step of Dense chains without the autograd tape (``elephas_b200/ops/cpu_mlp.py``) against the tape."""
import numpy as np
import pytest
import torch

from elephas_b200.models import SGD, Activation, Adam, Dense, Dropout, RMSprop, Sequential
from elephas_b200.ops import cpu_mlp


def _model(kind, optimizer):
    torch.manual_seed(5)
    if kind == 'classifier':
        m = Sequential([Dense(24, input_dim=13), Activation('relu'), Dense(17, activation='tanh'), Dense(5),
                        Activation('softmax')])
        m.compile(optimizer, 'categorical_crossentropy', ['acc'])
    elif kind == 'multilabel':
        m = Sequential([Dense(16, input_dim=13, activation='relu'), Dense(2, activation='sigmoid')])
        m.compile(optimizer, 'binary_crossentropy', ['binary_accuracy'])
    else:
        m = Sequential([Dense(16, input_dim=13, activation='sigmoid'), Dense(8, activation='relu'),
                        Dense(2, activation='linear' if kind != 'bounded_regressor' else 'sigmoid')])
        m.compile(optimizer, 'mae' if kind == 'robust_regressor' else 'mse', ['mae'])
    return m


def _data(kind):
    rng = np.random.default_rng(0)
    x = rng.normal(size=(96, 13)).astype('float32')
    if kind == 'classifier':
        y = np.eye(5, dtype='float32')[rng.integers(0, 5, 96)]
    elif kind == 'multilabel':
        y = rng.integers(0, 2, (96, 2)).astype('float32')
    else:
        y = rng.random((96, 2)).astype('float32')
    return x, y


@pytest.mark.parametrize('kind', ['classifier', 'regressor', 'bounded_regressor', 'robust_regressor', 'multilabel'])
@pytest.mark.parametrize('make_opt', [lambda: SGD(learning_rate=0.1, momentum=0.9), lambda: Adam(learning_rate=0.01),
                                      lambda: RMSprop(learning_rate=0.01)])
def test_hand_written_step_equals_the_autograd_step(monkeypatch, kind, make_opt):
    x, y = _data(kind)
    results = {}
    for flag in ('1', '0'):
        monkeypatch.setenv('ELEPHAS_B200_CPU_ENGINE', flag)
        model = _model(kind, make_opt())
        assert (model._cpu_step() is not None) == (flag == '1')
        stats = [model.train_on_batch(x[i:i + 32], y[i:i + 32]) for i in (0, 32, 64, 0)]
        results[flag] = (np.asarray(stats, dtype=np.float64), [w.copy() for w in model.get_weights()])
    assert np.allclose(results['1'][0], results['0'][0], atol=1e-5, rtol=1e-5)   # loss and metric of every step
    for a, b in zip(results['1'][1], results['0'][1]):
        assert np.allclose(a, b, atol=2e-6, rtol=1e-5), np.abs(a - b).max()


def test_dropout_chain_trains_and_other_models_stay_on_the_tape():
    torch.manual_seed(1)
    rng = np.random.default_rng(1)
    x = rng.normal(size=(512, 20)).astype('float32')
    y = np.eye(2, dtype='float32')[(x[:, :3].sum(1) > 0).astype(int)]
    m = Sequential([Dense(32, input_dim=20), Activation('relu'), Dropout(0.25), Dense(2), Activation('softmax')])
    m.compile(SGD(learning_rate=0.2), 'categorical_crossentropy', ['acc'])
    assert isinstance(m._cpu_step(), cpu_mlp.CpuMlpStep)
    before = m.evaluate(x, y)
    m.fit(x, y, batch_size=32, epochs=6, verbose=0)
    after = m.evaluate(x, y)
    assert after[0] < before[0] and after[1] > 0.9
    # a loss / activation the hand-written step does not cover, a custom activation, gradient clipping: the tape
    other = Sequential([Dense(4, input_dim=20, activation='elu'), Dense(1)])
    other.compile(SGD(learning_rate=0.1), 'mse')
    assert other._cpu_step() is None
    mape = Sequential([Dense(4, input_dim=20, activation='relu'), Dense(1)])
    mape.compile(SGD(learning_rate=0.1), 'mean_absolute_percentage_error')
    assert mape._cpu_step() is None
    clipped = Sequential([Dense(4, input_dim=20, activation='relu'), Dense(1)])
    clipped.compile(SGD(learning_rate=0.1, clipnorm=1.0), 'mse')
    assert clipped._cpu_step() is None


def test_penalties_and_constraints_stay_on_the_cpu_step(monkeypatch):
    """L1 / L2 penalties (closed-form gradient) and constraints do not cost a Dense chain its hand-written CPU step; the
    result equals the autograd path. Activity penalties do go to autograd."""
    import numpy as np
    import torch
    from elephas_b200.models import Dense, Sequential, constraints, regularizers
    from elephas_b200.models.optimizers import Adam

    def build(**first):
        torch.manual_seed(1)
        m = Sequential([Dense(16, activation='relu', input_shape=(8,), kernel_regularizer=regularizers.l1_l2(1e-3, 1e-2),
                              bias_regularizer='l2', kernel_constraint=constraints.MaxNorm(0.8), **first),
                        Dense(3, activation='softmax', kernel_regularizer='l2', bias_constraint='non_neg')])
        m.compile(Adam(0.05), 'categorical_crossentropy', ['acc'])
        return m

    rng = np.random.default_rng(0)
    x = rng.standard_normal((64, 8)).astype('float32')
    y = np.eye(3, dtype='float32')[rng.integers(0, 3, 64)]
    fast = build()
    assert fast._cpu_step() is not None
    assert build(activity_regularizer='l1')._cpu_step() is None
    monkeypatch.setenv('ELEPHAS_B200_CPU_ENGINE', '0')
    tape = build()
    assert tape._cpu_step() is None
    for _ in range(5):
        a, b = fast.train_on_batch(x, y), tape.train_on_batch(x, y)
        assert a[0] == pytest.approx(b[0], rel=1e-5) and a[1] == b[1]
    for wa, wb in zip(fast.get_weights(), tape.get_weights()):
        np.testing.assert_allclose(wa, wb, rtol=1e-4, atol=1e-6)
    assert np.all(np.linalg.norm(fast.get_weights()[0], axis=0) <= 0.8 + 1e-5) and np.all(fast.get_weights()[3] >= 0)
