"""GPU forms of the Keras-surface routes added at the end of round 2 (penalties / constraints inside the captured torch
step, graphs re-recorded when a callback assigns a learning rate, callbacks on the native engine).

Written after that round's GPU time was spent, so they are OPT-IN until they have run once on a B200:
``ELEPHAS_B200_PENDING_GPU_TESTS=1 python -m pytest tests/ops/test_keras_surface_gpu.py -m gpu`` (see
``scripts/gpu_round3_a.sh``). Each compares the GPU path with the eager CPU path of the same model.
"""
import os

import numpy as np
import pytest
import torch

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(os.environ.get('ELEPHAS_B200_PENDING_GPU_TESTS') != '1',
                                 reason="pending their first run on a B200 (set ELEPHAS_B200_PENDING_GPU_TESTS=1)")]

from elephas_b200.models import Dense, Sequential, callbacks, constraints, regularizers  # noqa: E402
from elephas_b200.models.optimizers import SGD  # noqa: E402


def _data(n=256, seed=0):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((n, 16)).astype(np.float32)
    y = np.eye(4, dtype=np.float32)[(x[:, :4]).argmax(1)]
    return x, y


def _regularised():
    torch.manual_seed(5)
    m = Sequential([Dense(32, activation='relu', input_shape=(16,), kernel_regularizer=regularizers.l2(1e-2),
                          kernel_constraint=constraints.MaxNorm(1.0)),
                    Dense(4, activation='softmax', bias_regularizer=regularizers.l1(1e-3))])
    m.compile(SGD(learning_rate=0.1), 'categorical_crossentropy', ['acc'])
    return m


def test_penalties_and_constraints_inside_the_captured_step_match_the_cpu():
    x, y = _data()
    cpu, gpu = _regularised(), _regularised()
    gpu.set_weights(cpu.get_weights())
    gpu.to('cuda')
    assert gpu._native_engine() is None and gpu._torch_engine() is not None     # routed to the captured torch step
    for lo in range(0, 256, 64):
        a = cpu.train_on_batch(x[lo:lo + 64], y[lo:lo + 64])
        b = gpu.train_on_batch(x[lo:lo + 64], y[lo:lo + 64])
        assert a[0] == pytest.approx(b[0], rel=2e-3)
    for wa, wb in zip(cpu.get_weights(), gpu.get_weights()):
        np.testing.assert_allclose(wa, wb, rtol=2e-3, atol=2e-4)
    assert np.all(np.linalg.norm(gpu.get_weights()[0], axis=0) <= 1.0 + 1e-4)


def test_a_learning_rate_assigned_by_a_callback_reaches_the_native_engine():
    x, y = _data()

    def build():
        torch.manual_seed(6)
        m = Sequential([Dense(32, activation='relu', input_shape=(16,)), Dense(4, activation='softmax')])
        m.compile(SGD(learning_rate=0.2), 'categorical_crossentropy', ['acc'])
        return m

    cpu, gpu = build(), build()
    gpu.set_weights(cpu.get_weights())
    gpu.to('cuda')
    assert gpu._native_engine() is not None
    schedule = lambda epoch, lr: 0.2 if epoch == 0 else 0.0   # noqa: E731 - the second epoch must not move the weights
    kw = dict(batch_size=64, epochs=2, shuffle=False)
    hc = cpu.fit(x, y, callbacks=[callbacks.LearningRateScheduler(schedule)], **kw)
    after_first = []
    hg = gpu.fit(x, y, callbacks=[callbacks.LearningRateScheduler(schedule),
                                  callbacks.LambdaCallback(on_epoch_end=lambda e, logs: after_first.append(gpu.get_weights()))],
                 **kw)
    assert hg.history['lr'] == pytest.approx([0.2, 0.0]) and hc.history['lr'] == pytest.approx([0.2, 0.0])
    for a, b in zip(after_first[0], after_first[1]):
        np.testing.assert_array_equal(a, b)                    # lr 0 really reached the recorded step
    for wa, wb in zip(cpu.get_weights(), gpu.get_weights()):
        np.testing.assert_allclose(wa, wb, rtol=2e-2, atol=2e-3)   # bf16 GEMMs against fp32


def test_early_stopping_and_checkpoints_on_the_native_engine(tmp_path):
    from elephas_b200.models import load_model
    x, y = _data()
    torch.manual_seed(7)
    m = Sequential([Dense(32, activation='relu', input_shape=(16,)), Dense(4, activation='softmax')])
    m.compile(SGD(learning_rate=0.2), 'categorical_crossentropy', ['acc'])
    m.to('cuda')
    path = str(tmp_path / 'best.h5')

    class Flat(callbacks.Callback):
        def on_epoch_end(self, epoch, logs=None):
            logs['val_loss'] = 1.0

    h = m.fit(x, y, batch_size=64, epochs=10, callbacks=[Flat(), callbacks.EarlyStopping(patience=2),
                                                        callbacks.ModelCheckpoint(path)])
    assert len(h.history['loss']) == 3
    back = load_model(path)
    np.testing.assert_allclose(back.predict(x[:8]), m.predict(x[:8]), rtol=2e-2, atol=2e-3)


def test_frozen_layers_stay_frozen_on_the_gpu():
    x, y = _data()
    torch.manual_seed(8)
    m = Sequential([Dense(32, activation='relu', input_shape=(16,), name='trunk'), Dense(4, activation='softmax', name='head')])
    m.get_layer('trunk').trainable = False
    m.compile(SGD(learning_rate=0.2), 'categorical_crossentropy')
    before = m.get_weights()
    m.to('cuda')
    assert m._native_engine() is None          # the fused step would update the whole flat buffer
    m.fit(x, y, batch_size=64, epochs=2)
    after = m.get_weights()
    np.testing.assert_array_equal(after[0], before[0])
    np.testing.assert_array_equal(after[1], before[1])
    assert not np.array_equal(after[2], before[2])
    assert m.predict(x).shape == (256, 4)


@pytest.mark.parametrize('family', ['lstm', 'transformer'])
def test_sequence_models_train_in_the_captured_step(family):
    from elephas_b200.models import zoo
    from elephas_b200.models.optimizers import Adam
    rng = np.random.default_rng(0)
    tokens = rng.integers(1, 30, (256, 12)).astype(np.float32)
    tokens[:, 9:] *= (rng.random((256, 3)) < 0.5)
    y = np.eye(2, dtype=np.float32)[(np.sum(tokens < 10, axis=1) >= 4).astype(int)]
    torch.manual_seed(0)
    m = zoo.lstm_classifier(vocab=30, maxlen=12, embed=16, units=16, bidirectional=True, dropout=0.0) if family == 'lstm' \
        else zoo.transformer_classifier(vocab=30, maxlen=12, embed=16, heads=2, ff=32, blocks=1, dropout=0.0)
    m.compile(Adam(learning_rate=0.02), 'categorical_crossentropy', ['acc'])
    cpu_out = m.predict(tokens[:16])
    m.to('cuda')
    np.testing.assert_allclose(m.predict(tokens[:16]), cpu_out, rtol=1e-3, atol=1e-4)
    h = m.fit(tokens, y, batch_size=64, epochs=8)
    assert m._torch_step is not None                       # the step was captured, not run eagerly
    assert h.history['loss'][-1] < 0.8 * h.history['loss'][0]
