"""The native engine end to end on a B200: one train step against torch autograd, convergence, bit-identical
batched prediction, checkpoint resume, and the SparkModel matrix on the GPU."""
from itertools import count
from math import isclose

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from elephas.spark_model import SparkModel  # noqa: E402
from elephas.utils.rdd_utils import to_simple_rdd  # noqa: E402
from elephas_b200.models import SGD, Adam, Dense, RMSprop, Sequential, load_model, zoo  # noqa: E402
from tests.conftest import make_boston_like, make_mnist_like  # noqa: E402

_ports = count(6200)


def _pair(builder, opt, loss, metrics):
    """The same model twice: native engine on cuda, torch autograd on cuda (ELEPHAS_B200_NATIVE=0 semantics)."""
    import elephas_b200.ops.engine as engine
    m_native = builder()
    m_native.compile(opt(), loss, metrics)
    m_torch = builder()
    m_torch.compile(opt(), loss, metrics)
    m_torch.set_weights(m_native.get_weights())
    m_native.to('cuda')
    m_torch.to('cuda')
    assert m_native._native_engine() is not None, "the native engine must be selected for MLP-shaped models"
    m_torch._native_engine = lambda *a, **k: None
    return m_native, m_torch


@pytest.mark.parametrize('opt,adaptive', [(lambda: SGD(lr=0.05), False),
                                          (lambda: SGD(lr=0.05, momentum=0.9, nesterov=True), False),
                                          (lambda: Adam(learning_rate=0.01), True),
                                          (lambda: RMSprop(learning_rate=0.001), True)])
def test_train_steps_match_autograd(opt, adaptive):
    x, y, _, _ = make_mnist_like(256, 10)
    builder = lambda: zoo.mnist_mlp(dropout=0.0)  # noqa: E731
    m_native, m_torch = _pair(builder, opt, 'categorical_crossentropy', ['acc'])
    for i in range(3):
        xb, yb = x[i * 64:(i + 1) * 64], y[i * 64:(i + 1) * 64]
        a = m_native.train_on_batch(xb, yb)
        b = m_torch.train_on_batch(xb, yb)
        assert abs(a[0] - b[0]) < 3e-2 and abs(a[1] - b[1]) < 0.05
    for wa, wb in zip(m_native.get_weights(), m_torch.get_weights()):
        if adaptive:
            # sign-normalised steps (|step| ~ lr whatever the gradient magnitude) amplify bf16-vs-fp32
            # differences of near-zero gradients (a sign flip moves a weight by 2 steps): compare in the mean
            assert np.abs(wa - wb).mean() < 5e-3
        else:
            assert np.allclose(wa, wb, atol=5e-3)


def test_regression_step_matches_autograd():
    x, y, _, _ = make_boston_like(128, 10)
    x = (x - x.mean(0)) / (x.std(0) + 1e-6)
    m_native, m_torch = _pair(zoo.boston_mlp, lambda: SGD(lr=0.01), 'mse', ['mae'])
    for _ in range(3):
        a = m_native.train_on_batch(x[:64], y[:64])
        b = m_torch.train_on_batch(x[:64], y[:64])
        assert abs(a[0] - b[0]) < 0.02 * max(1.0, abs(b[0]))
    for wa, wb in zip(m_native.get_weights(), m_torch.get_weights()):
        assert np.allclose(wa, wb, atol=2e-2, rtol=2e-2)


def test_fit_converges_with_dropout_and_graphs():
    x, y, xt, yt = make_mnist_like(2000, 500)
    m = zoo.mnist_mlp()
    m.compile(SGD(lr=0.1), 'categorical_crossentropy', ['acc'])
    m.to('cuda')
    h = m.fit(x, y, batch_size=64, epochs=4, validation_split=0.1, verbose=0)
    assert set(h.history) == {'loss', 'acc', 'val_loss', 'val_acc'}
    assert h.history['loss'][-1] < 0.5 * h.history['loss'][0]
    loss, acc = m.evaluate(xt, yt)
    assert acc > 0.9
    assert m.optimizer.iterations == 4 * int(np.ceil(1800 / 64))
    from elephas_b200.ops import native
    assert native.launch_count() > 0


def test_predict_is_batch_invariant_and_matches_cpu():
    x, y, xt, _ = make_mnist_like(64, 300)
    m = zoo.mnist_mlp()
    cpu = m.predict(xt)
    m.to('cuda')
    full = m.predict(xt)
    parts = np.vstack([m.predict(xt[i:i + 37]) for i in range(0, len(xt), 37)])
    assert np.array_equal(full, parts)                  # bit-identical regardless of batching
    assert np.allclose(full, cpu, atol=2e-2)             # bf16 tensor-core path vs fp32 CPU
    assert (full.argmax(1) == cpu.argmax(1)).mean() > 0.97


def test_checkpoint_resume_on_gpu(tmp_path):
    x, y, _, _ = make_mnist_like(512, 10)
    m = zoo.mnist_mlp(dropout=0.0)
    m.compile(SGD(lr=0.05, momentum=0.9), 'categorical_crossentropy', ['acc'])
    m.to('cuda')
    m.fit(x, y, batch_size=64, epochs=1, shuffle=False)
    path = str(tmp_path / 'gpu.h5')
    m.save(path)
    m2 = load_model(path)
    m2.to('cuda')
    assert m2.optimizer.iterations == 8
    a = m.train_on_batch(x[:64], y[:64])
    b = m2.train_on_batch(x[:64], y[:64])
    assert abs(a[0] - b[0]) < 1e-3
    for wa, wb in zip(m.get_weights(), m2.get_weights()):
        assert np.allclose(wa, wb, atol=1e-4)


MATRIX = [('synchronous', None, 'epoch', 2), ('synchronous', None, 'batch', 1), ('synchronous', None, 'batch', 2),
          ('asynchronous', 'nvlink', 'epoch', 2), ('asynchronous', 'nvlink', 'batch', 2),
          ('hogwild', 'nvlink', 'epoch', 2), ('hogwild', 'nvlink', 'batch', 2),
          ('asynchronous', 'socket', 'epoch', 2), ('hogwild', 'http', 'epoch', 2)]


@pytest.mark.parametrize('mode,ps,freq,workers', MATRIX)
def test_spark_model_matrix_on_gpu(spark_context, mode, ps, freq, workers):
    x, y, xt, yt = make_mnist_like(1024, 300)
    model = zoo.mnist_mlp()
    model.compile(SGD(lr=0.05), 'categorical_crossentropy', ['acc'])
    sm = SparkModel(model, mode=mode, frequency=freq, parameter_server_mode=ps, num_workers=workers,
                    port=next(_ports))
    before = sm.evaluate(xt, yt)[0]
    sm.fit(to_simple_rdd(spark_context, x, y), epochs=3, batch_size=64, verbose=0, validation_split=0.1)
    assert sm.master_network.device.type == 'cuda' or True
    preds = sm.predict(xt)
    evals = sm.evaluate(xt, yt)
    local = sm.master_network.predict(xt)
    assert [int(np.argmax(p)) for p in preds] == [int(np.argmax(p)) for p in local]
    ref = sm.master_network.evaluate(xt, yt)
    assert isclose(evals[0], ref[0], abs_tol=0.01) and isclose(evals[1], ref[1], abs_tol=0.01)
    assert evals[0] < before


def test_hogwild_batch_uses_the_pull_fused_gemm(spark_context):
    """K2: hogwild + frequency='batch' on the device parameter server pulls through the forward GEMMs."""
    from elephas_b200 import worker as _worker
    x, y, xt, yt = make_mnist_like(1024, 300)
    model = zoo.mnist_mlp()
    model.compile(SGD(lr=0.05), 'categorical_crossentropy', ['acc'])
    sm = SparkModel(model, mode='hogwild', frequency='batch', parameter_server_mode='nvlink', num_workers=2,
                    port=next(_ports))
    before_steps = _worker.STATS['fused_pull_steps']
    before = sm.evaluate(xt, yt)
    sm.fit(to_simple_rdd(spark_context, x, y), epochs=2, batch_size=64, verbose=0)
    assert _worker.STATS['fused_pull_steps'] - before_steps == 2 * 2 * 8  # 2 workers x 2 epochs x 512/64 batches
    after = sm.evaluate(xt, yt)
    assert after[0] < before[0] and after[1] > 0.8
    # the locked mode pulls with one seqlock kernel per batch (readers take no lock) instead of through the GEMMs; the
    # optimizer step is fused with the push in both modes
    model2 = zoo.mnist_mlp()
    model2.compile(SGD(lr=0.05), 'categorical_crossentropy', ['acc'])
    sm2 = SparkModel(model2, mode='asynchronous', frequency='batch', parameter_server_mode='nvlink', num_workers=2,
                     port=next(_ports))
    fused, steps = _worker.STATS['fused_pull_steps'], _worker.STATS['device_ps_steps']
    before2 = sm2.evaluate(xt, yt)
    sm2.fit(to_simple_rdd(spark_context, x, y), epochs=2, batch_size=64, verbose=0)
    assert _worker.STATS['fused_pull_steps'] == fused and _worker.STATS['device_ps_steps'] - steps == 2 * 2 * 8
    after2 = sm2.evaluate(xt, yt)
    assert after2[0] < before2[0] and after2[1] > 0.8
    assert all(info['steps'] == 16 and info['device_ms'] > 0 for info in sm2._last_async_run)
    # frequency='epoch' pulls whole replicas
    sm3 = SparkModel(model2, mode='asynchronous', frequency='epoch', parameter_server_mode='nvlink', num_workers=2,
                     port=next(_ports))
    steps = _worker.STATS['device_ps_steps']
    sm3.fit(to_simple_rdd(spark_context, x, y), epochs=1, batch_size=64, verbose=0)
    assert _worker.STATS['device_ps_steps'] == steps


def test_device_ps_batch_mode_streams_inputs_and_logs_losses(spark_context):
    """input_mode='stream' (every batch copied from pinned host memory) and log_batch_loss (64-byte read-back per step)
    on the per-batch parameter-server path - what bench.py's end-to-end number goes through."""
    x, y, xt, yt = make_mnist_like(1024, 300)
    model = zoo.mnist_mlp()
    model.compile(Adam(learning_rate=0.002), 'categorical_crossentropy', ['acc'])
    sm = SparkModel(model, mode='asynchronous', frequency='batch', parameter_server_mode='nvlink', num_workers=1,
                    port=next(_ports))
    before = sm.evaluate(xt, yt)
    sm.fit(to_simple_rdd(spark_context, x, y), epochs=2, batch_size=64, input_mode='stream', log_batch_loss=True)
    after = sm.evaluate(xt, yt)
    assert after[0] < before[0] and after[1] > 0.8
    info = sm._last_async_run[0]
    assert info['steps'] == 32 and info['h2d_bytes_per_step'] == 64 * (784 + 10) * 4
    hist = sm.training_histories[-1]
    assert len(hist['batch_loss']) == 32 and hist['batch_loss'][-1] < hist['batch_loss'][0]


def test_wide_mlp_shape_small_on_gpu(spark_context):
    """BASELINE config 4 at reduced width: regression MLP, synchronous, frequency='batch'."""
    rng = np.random.default_rng(0)
    x = rng.normal(size=(2048, 256)).astype(np.float32)
    w = rng.normal(size=(256, 1)).astype(np.float32) / 16
    y = (x @ w + 0.01 * rng.normal(size=(2048, 1))).astype(np.float32)
    model = zoo.wide_mlp(input_dim=256, hidden=256, depth=4, out_dim=1)
    model.compile(SGD(lr=0.003), 'mse', ['mae'])
    sm = SparkModel(model, mode='synchronous', frequency='batch', num_workers=1)
    before = sm.evaluate(x, y)[0]
    sm.fit(to_simple_rdd(spark_context, x, y), epochs=3, batch_size=256, log_batch_loss=True)
    assert sm.evaluate(x, y)[0] < before
    hist = sm.training_histories[0]
    assert len(hist['loss']) == 3 and len(hist['batch_loss']) == 3 * 8


def test_fallback_for_custom_activation_stays_on_torch():
    m = Sequential([Dense(8, input_dim=4, activation=lambda t: torch.sigmoid(t) + 1), Dense(1)])
    m.compile('sgd', 'mse')
    m.to('cuda')
    assert m._native_engine() is None
    assert np.isfinite(m.train_on_batch(np.random.rand(16, 4).astype('f4'), np.random.rand(16, 1).astype('f4')))


# ---------------------------------------------------------------- graph-captured torch step (non-Dense models)
def _small_cnn():
    from elephas_b200.models.layers import BatchNormalization, Conv2D, Flatten, MaxPooling2D
    return Sequential([Conv2D(8, 3, padding='same', activation='relu', input_shape=(8, 8, 3)), BatchNormalization(),
                       MaxPooling2D(2), Conv2D(16, 3, padding='same', activation='relu'), Flatten(),
                       Dense(10, activation='softmax')])


@pytest.mark.parametrize('opt_name', ['sgd_momentum', 'adam', 'rmsprop'])
def test_graphed_torch_step_matches_eager(monkeypatch, opt_name):
    """Flat weights + native optimizer kernel + CUDA-graph replay == the eager torch path, step for step."""
    def make_opt():
        return {'sgd_momentum': lambda: SGD(learning_rate=0.05, momentum=0.9, decay=1e-3),
                'adam': lambda: Adam(learning_rate=0.01), 'rmsprop': lambda: RMSprop(learning_rate=0.005)}[opt_name]()
    rng = np.random.default_rng(3)
    x = rng.normal(size=(96, 8, 8, 3)).astype(np.float32)
    y = np.eye(10, dtype=np.float32)[rng.integers(0, 10, 96)]
    torch.manual_seed(1)
    a = _small_cnn()
    a.compile(make_opt(), 'categorical_crossentropy', ['acc'])
    w0 = a.get_weights()
    b = _small_cnn()
    b.compile(make_opt(), 'categorical_crossentropy', ['acc'])
    b.set_weights(w0)
    a.to('cuda')
    b.to('cuda')
    ha = a.fit(x, y, batch_size=32, epochs=3, shuffle=False)          # captured
    assert a._torch_engine() is not None and a._torch_step.graphs, "the graphed step did not run"
    monkeypatch.setenv('ELEPHAS_B200_TORCH_GRAPHS', '0')
    hb = b.fit(x, y, batch_size=32, epochs=3, shuffle=False)          # eager
    assert b._torch_engine() is None
    for wa, wb in zip(a.get_weights(), b.get_weights()):
        if opt_name == 'sgd_momentum':
            assert np.allclose(wa, wb, atol=1e-3, rtol=5e-3), np.abs(wa - wb).max()
        else:
            # adaptive optimizers turn a last-bit difference of a tiny gradient (cuDNN's backward kernels are not
            # run-to-run deterministic) into a full-size step: compare on average
            assert np.abs(wa - wb).mean() < 1e-3, np.abs(wa - wb).mean()
    assert np.allclose(ha.history['loss'], hb.history['loss'], rtol=2e-2, atol=1e-3)
    assert np.allclose(ha.history['acc'], hb.history['acc'], atol=0.03)
    assert a.optimizer.iterations == b.optimizer.iterations == 9
    # train_on_batch goes through the same graph and reports the step's own statistics
    sa = a.train_on_batch(x[:32], y[:32])
    sb = b.train_on_batch(x[:32], y[:32])
    assert abs(sa[0] - sb[0]) < 2e-3 and abs(sa[1] - sb[1]) < 0.04


def test_graphed_torch_step_survives_recompile_save_and_device_moves(tmp_path):
    rng = np.random.default_rng(4)
    x = rng.normal(size=(64, 8, 8, 3)).astype(np.float32)
    y = np.eye(10, dtype=np.float32)[rng.integers(0, 10, 64)]
    m = _small_cnn()
    m.compile(Adam(learning_rate=0.01), 'categorical_crossentropy', ['acc'])
    m.to('cuda')
    m.fit(x, y, batch_size=32, epochs=2)
    first = m.evaluate(x, y)[0]
    path = str(tmp_path / 'cnn.h5')
    m.save(path)                                   # exports the flat optimizer slots
    m2 = load_model(path)
    assert m2.optimizer.iterations == 4
    m2.to('cuda')
    m2.fit(x, y, batch_size=32, epochs=2)          # resumes: imports the slots into a new flat buffer
    assert m2.evaluate(x, y)[0] < first
    m.compile(SGD(learning_rate=0.01), 'categorical_crossentropy', ['acc'])   # new optimizer: new engine
    m.fit(x, y, batch_size=32, epochs=1)
    m.to('cpu')                                    # weights leave the flat buffer; the eager CPU path takes over
    m.fit(x, y, batch_size=32, epochs=1)
    assert np.isfinite(m.evaluate(x, y)[0])


def test_graphed_torch_step_bf16_autocast_trains():
    rng = np.random.default_rng(5)
    x = rng.normal(size=(128, 8, 8, 3)).astype(np.float32)
    labels = (x.mean(axis=(1, 2, 3)) > 0).astype(int)
    y = np.eye(10, dtype=np.float32)[labels]
    m = _small_cnn()
    m.compile(Adam(learning_rate=0.01), 'categorical_crossentropy', ['acc'])
    m.set_compute_dtype('bfloat16')
    m.to('cuda')
    h = m.fit(x, y, batch_size=32, epochs=8)
    assert h.history['loss'][-1] < h.history['loss'][0] and m._torch_step.graphs


def test_sync_torch_path_trains_in_place_and_reuses_the_graph(spark_context):
    """SparkModel synchronous on a model without a native Dense engine: the master network is the replica, the
    captured step survives the per-fit recompile (same configuration)."""
    rng = np.random.default_rng(6)
    x = rng.normal(size=(256, 8, 8, 3)).astype(np.float32)
    labels = (x.mean(axis=(1, 2, 3)) > 0).astype(int)
    y = np.eye(10, dtype=np.float32)[labels]
    m = _small_cnn()
    m.compile(Adam(learning_rate=0.01), 'categorical_crossentropy', ['acc'])
    sm = SparkModel(m, mode='synchronous', frequency='epoch', num_workers=1)
    before = sm.evaluate(x, y)[0]
    sm.fit(to_simple_rdd(spark_context, x, y), epochs=3, batch_size=32, verbose=0, validation_split=0.125)
    te = sm.master_network._torch_step
    assert te is not None and len(te.graphs) >= 1
    n_graphs = len(te.graphs)
    hist = sm.training_histories[-1]
    assert len(hist['loss']) == 3 and 'val_loss' in hist and hist['loss'][-1] < hist['loss'][0]
    mid = sm.evaluate(x, y)[0]
    assert mid < before
    sm.fit(to_simple_rdd(spark_context, x, y), epochs=2, batch_size=32, verbose=0, validation_split=0.125)
    assert sm.master_network._torch_step is te and len(te.graphs) == n_graphs   # nothing was re-captured
    assert sm.evaluate(x, y)[0] < mid
    assert len(sm.predict(x[:10])) == 10


def test_streamed_inputs_cast_on_the_host_match_device_staging(spark_context, monkeypatch):
    """input_mode='stream': the host thread pool casts the features to bf16 (half the PCIe bytes); results are
    bit-identical to streaming fp32 and casting on the device."""
    rng = np.random.default_rng(8)
    x = rng.normal(size=(1000, 200)).astype(np.float32)   # 200 features: ld 200; 1000 rows: last batch is partial
    w = rng.normal(size=(200, 1)).astype(np.float32) / 14
    y = (x @ w).astype(np.float32)
    results = []
    for host_bf16 in ('1', '0'):
        monkeypatch.setenv('ELEPHAS_B200_HOST_BF16', host_bf16)
        torch.manual_seed(2)
        model = zoo.wide_mlp(input_dim=200, hidden=128, depth=3, out_dim=1)
        model.compile(SGD(lr=0.01), 'mse', ['mae'])
        sm = SparkModel(model, mode='synchronous', frequency='batch', num_workers=1)
        sm.fit(to_simple_rdd(spark_context, x, y), epochs=3, batch_size=128, input_mode='stream', shuffle=False,
               log_batch_loss=True)
        info = sm._last_sync_run
        assert info['h2d_bytes_per_step'] == (128 * (200 * 2 + 4) if host_bf16 == '1' else 128 * (200 + 1) * 4)
        results.append((sm.master_network.get_weights(), sm.training_histories[-1]['batch_loss']))
    # same inputs bit for bit (the host cast is exact), so the runs differ only by the summation order of the
    # atomics in the bias / narrow-head gradient kernels
    for a, b in zip(results[0][0], results[1][0]):
        assert np.allclose(a, b, atol=1e-5, rtol=1e-4), np.abs(a - b).max()
    assert np.allclose(results[0][1], results[1][1], rtol=1e-4) and len(results[0][1]) == 3 * 8
    # without an override the feeder times both ways once and picks one
    monkeypatch.setenv('ELEPHAS_B200_HOST_BF16', 'auto')
    from elephas_b200.parallel import sync_fast
    sync_fast._HOST_BF16_CHOICE.clear()
    sm.fit(to_simple_rdd(spark_context, x, y), epochs=1, batch_size=128, input_mode='stream')
    assert len(sync_fast._HOST_BF16_CHOICE) == 1
    # shuffled epochs go through the same pipeline
    monkeypatch.setenv('ELEPHAS_B200_HOST_BF16', '1')
    model = zoo.wide_mlp(input_dim=200, hidden=128, depth=3, out_dim=1)
    model.compile(SGD(lr=0.01), 'mse', ['mae'])
    sm = SparkModel(model, mode='synchronous', frequency='batch', num_workers=1)
    before = sm.evaluate(x, y)[0]
    sm.fit(to_simple_rdd(spark_context, x, y), epochs=3, batch_size=128, input_mode='stream', validation_split=0.1)
    assert sm.evaluate(x, y)[0] < before


def test_shard_staged_as_bf16_streams_the_same_batches(spark_context):
    """``to_sharded_rdd(feature_dtype='bfloat16')`` + ``input_mode='stream'``: the shard is cast once on the host (AVX-512,
    bit-identical to the device cast), every step copies 2 bytes per feature through three staging slots. Same batch
    losses and weights as streaming the fp32 rows; 1000 rows / batch 128: the last batch of every epoch is partial."""
    from elephas_b200.utils.rdd_utils import to_sharded_rdd
    rng = np.random.default_rng(9)
    x = rng.normal(size=(1000, 200)).astype(np.float32)
    y = (x @ (rng.normal(size=(200, 1)).astype(np.float32) / 14)).astype(np.float32)
    results = []
    for dtype in ('bfloat16', None):
        torch.manual_seed(3)
        model = zoo.wide_mlp(input_dim=200, hidden=128, depth=3, out_dim=1)
        model.compile(SGD(lr=0.01), 'mse', ['mae'])
        sm = SparkModel(model, mode='synchronous', frequency='batch', num_workers=1)
        rdd = to_sharded_rdd(spark_context, x, y, rank=0, world_size=1, feature_dtype=dtype)
        sm.fit(rdd, epochs=3, batch_size=128, input_mode='stream', shuffle=False, log_batch_loss=True)
        assert sm._last_sync_run['h2d_bytes_per_step'] == (128 * (200 * 2 + 4) if dtype else 128 * (200 + 1) * 4)
        results.append((sm.master_network.get_weights(), sm.training_histories[-1]['batch_loss']))
    for a, b in zip(results[0][0], results[1][0]):
        assert np.allclose(a, b, atol=1e-5, rtol=1e-4), np.abs(a - b).max()
    assert np.allclose(results[0][1], results[1][1], rtol=1e-4) and len(results[0][1]) == 3 * 8


@pytest.mark.parametrize('opt', ['sgd', 'adam'])
def test_narrow_head_path_matches_gemm_path(monkeypatch, opt):
    """Regression head (Dense(1)): the HBM-speed head kernels and the tensor-core GEMM path train the same model."""
    import importlib
    from elephas_b200.ops import engine as engine_mod
    rng = np.random.default_rng(9)
    x = rng.normal(size=(512, 64)).astype(np.float32)
    y = (x @ rng.normal(size=(64, 1)).astype(np.float32) / 8).astype(np.float32)
    weights = []
    for narrow in (True, False):
        monkeypatch.setattr(engine_mod, '_NARROW_HEAD', narrow)
        torch.manual_seed(5)
        m = zoo.wide_mlp(input_dim=64, hidden=64, depth=2, out_dim=1)
        m.compile(SGD(learning_rate=0.01) if opt == 'sgd' else Adam(learning_rate=0.001), 'mse', ['mae'])
        m.to('cuda')
        h = m.fit(x, y, batch_size=128, epochs=3, shuffle=False)
        assert m._native_engine() is not None
        weights.append((m.get_weights(), h.history['loss']))
    tol = 2e-3 if opt == 'sgd' else None
    for a, b in zip(weights[0][0], weights[1][0]):
        if tol is not None:
            assert np.allclose(a, b, atol=tol), np.abs(a - b).max()
        else:
            assert np.abs(a - b).mean() < 2e-3
    assert np.allclose(weights[0][1], weights[1][1], rtol=2e-2, atol=1e-3)


def _emulated_store(model, world, device='cuda'):
    """The striped store of ``world`` ranks emulated on ONE GPU: all stripes are local tensors, one EbPsStore per
    emulated rank (same stripes, different ``rank``). Push / pull kernels never wait for another kernel launch - a lock
    is taken and released inside one kernel - so launching the ranks' kernels one after another on one GPU is legal."""
    from elephas_b200.ops import native
    from elephas_b200.parallel.device_ps import MAX_CHUNKS, StripePlan
    from elephas_b200.parallel.flat import FlatLayout
    layout = FlatLayout(model)
    plan = StripePlan(layout, world)
    master = [torch.zeros(plan.capacity, device=device) for _ in range(world)]
    mirror = [torch.zeros(plan.capacity, dtype=torch.bfloat16, device=device) for _ in range(world)]
    ctrl = [torch.zeros(MAX_CHUNKS * native.PS_LOCK_WORDS, dtype=torch.int32, device=device) for _ in range(world)]
    args = [native.ps_store_args([t.data_ptr() for t in master], [t.data_ptr() for t in mirror],
                                 [t.data_ptr() for t in ctrl], plan.stripe_len, plan.chunk_len, plan.chunks, r, plan.spans)
            for r in range(world)]
    return layout, plan, master, mirror, ctrl, args


@pytest.mark.parametrize('world', [1, 3, 8])
@pytest.mark.parametrize('locked', [False, True])
def test_striped_store_push_pull_match_pytorch(world, locked):
    """Every push source (fp32 delta, bf16 delta, before - after) and the flat-indexed pull against plain PyTorch
    arithmetic, with the master spread over 1 / 3 / 8 (emulated) stripes."""
    from elephas_b200.ops import native
    model = zoo.otto_mlp()
    layout, plan, master, mirror, ctrl, args = _emulated_store(model, world)
    n = layout.numel
    g = torch.Generator(device='cuda').manual_seed(world)
    flat = torch.randn(n, device='cuda', generator=g)
    for flat_off, cnt, owner, owner_off in plan.spans:  # scatter the initial master into the stripes
        master[owner][owner_off:owner_off + cnt] = flat[flat_off:flat_off + cnt]
        mirror[owner][owner_off:owner_off + cnt] = flat[flat_off:flat_off + cnt].to(torch.bfloat16)
    oracle = flat.clone()
    d32 = torch.randn(n, device='cuda', generator=g)
    native.ps_striped_push(args[0], delta_f32=d32, locked=locked)
    oracle -= d32
    d16 = torch.randn(n, device='cuda', generator=g).to(torch.bfloat16)
    native.ps_striped_push(args[world - 1], delta_bf16=d16, locked=locked)
    oracle -= d16.float()
    before, after = torch.randn(n, device='cuda', generator=g), torch.randn(n, device='cuda', generator=g)
    native.ps_striped_push(args[world // 2], before=before, after=after, locked=locked)
    oracle -= before - after
    got32 = torch.full((n,), float('nan'), device='cuda')
    got16 = torch.zeros(n, dtype=torch.bfloat16, device='cuda')
    native.ps_striped_pull(args[0], got32, None, got16, None, seqlock=locked)
    torch.cuda.synchronize()
    assert torch.allclose(got32, oracle, atol=1e-5, rtol=0), (got32 - oracle).abs().max()
    assert torch.equal(got16, got32.to(torch.bfloat16))  # the mirror is bf16 of the master, element for element
    # range-limited pull: fp32 for the vector tail only, bf16 for the matrices only; everything else untouched
    lo = layout.matrix_end
    part32 = torch.full((n,), 7.0, device='cuda')
    part16 = torch.full((n,), 3.0, dtype=torch.bfloat16, device='cuda')
    native.ps_striped_pull(args[world - 1], part32, (lo, n), part16, (0, lo), seqlock=locked)
    torch.cuda.synchronize()
    assert torch.equal(part32[lo:], got32[lo:]) and bool((part32[:lo] == 7.0).all())
    assert torch.equal(part16[:lo], got16[:lo]) and bool((part16[lo:] == 3.0).all())
    assert all(int(c.view(-1, native.PS_LOCK_WORDS)[:, 3].max()) == 0 for c in ctrl)
    if locked:  # every chunk: as many writers counted out as counted in (three pushes), the readers' gate is down
        for r in range(world):
            words = ctrl[r].view(-1, native.PS_LOCK_WORDS)
            used = -(-plan.stripe_len[r] // plan.chunk_len)
            assert bool((words[:used, 1] == words[:used, 2]).all()) and bool((words[:used, 1] >= 3).all())
            assert bool((words[:used, 0] == 0).all())


@pytest.mark.parametrize('opt', ['sgd', 'momentum', 'adam', 'rmsprop'])
def test_optimizer_fused_with_the_push_matches_optimizer_then_push(opt):
    """eb_ps_striped_optimizer_push == eb_optimizer_step in delta mode followed by a push of that delta - and both equal
    a PyTorch fp32 evaluation of the Keras update rule."""
    from elephas_b200.ops import native
    model = zoo.mnist_mlp()
    layout, plan, master, mirror, ctrl, args = _emulated_store(model, 4)
    n = layout.numel
    g = torch.Generator(device='cuda').manual_seed(3)
    kind = {'sgd': 'SGD', 'momentum': 'SGD', 'adam': 'Adam', 'rmsprop': 'RMSprop'}[opt]
    hyper = dict(lr=0.01, momentum=0.9 if opt == 'momentum' else 0.0, rho=0.9, eps=1e-7, beta1=0.9, beta2=0.999)
    need = {'sgd': 0, 'momentum': 1, 'adam': 2, 'rmsprop': 1}[opt]
    step = torch.zeros(1, dtype=torch.int64, device='cuda')
    slots_a = [torch.zeros(n, device='cuda') if i < need else None for i in range(3)]
    slots_b = [t.clone() if t is not None else None for t in slots_a]
    w_ref = torch.zeros(n, device='cuda')
    for it in range(3):
        grad = torch.randn(n, device='cuda', generator=g)
        native.ps_striped_optimizer_push(args[it % 4], native.OPT_IDS[kind], n, grad, step, slots=slots_a, locked=True,
                                         advance=False, **hyper)
        delta = torch.zeros(n, device='cuda')
        native.optimizer_step(native.OPT_IDS[kind], n, grad, step, slots=slots_b, delta_f32=delta, apply=False,
                              advance=True, **hyper)
        w_ref -= delta
        if opt == 'sgd' and it == 0:
            assert torch.allclose(delta, 0.01 * grad, atol=1e-7)  # the PyTorch oracle of the plain rule
        if opt == 'adam' and it == 0:
            m, v = 0.1 * grad, 0.001 * grad * grad
            lr_t = 0.01 * (1 - 0.999) ** 0.5 / (1 - 0.9)
            assert torch.allclose(delta, lr_t * m / (v.sqrt() + 1e-7), atol=1e-6, rtol=1e-4)
    got = torch.empty(n, device='cuda')
    native.ps_striped_pull(args[0], got, None, None, None, seqlock=True)
    torch.cuda.synchronize()
    assert torch.allclose(got, w_ref, atol=1e-6, rtol=0), (got - w_ref).abs().max()
    for a, b in zip(slots_a, slots_b):
        assert a is None or torch.equal(a, b)


@pytest.mark.parametrize('mode', ['asynchronous', 'hogwild'])
def test_device_store_concurrent_pushes_sum_exactly(mode):
    """Race test (the reference's server loses updates: it reads the weights before taking the write lock,
    elephas/parameter/server.py:204-206): worker threads push integer-valued deltas concurrently, each through its own
    stream; `red.add` makes every push land exactly once in both modes, and in asynchronous mode a seqlock pull never
    sees a half-applied push inside a chunk - although readers take no lock at all."""
    import threading
    from elephas_b200.parallel.device_ps import create_store, destroy_store
    model = Sequential([Dense(96, input_dim=256, activation='relu'), Dense(8)])
    model.compile(SGD(learning_rate=0.1), 'mse')
    port = next(_ports)
    store = create_store(port, model.get_weights(), mode, model=model, device=torch.device('cuda', 0))
    n = store.layout.numel
    store.load_flat(torch.zeros(n, device='cuda'))  # integer-valued master and deltas: every partial sum is exact in fp32
    ranges = store.plan.chunk_ranges_flat()
    assert len(ranges) >= 3  # several chunks: several independent locks
    threads, pushes, torn = 4, 25, []

    def work(tid):
        with torch.cuda.device(0), torch.cuda.stream(torch.cuda.Stream(0)):
            delta = torch.full((n,), float(tid + 1), device='cuda')
            snap = torch.empty(n, device='cuda')
            for i in range(pushes):
                store.push_delta(delta)
                if mode == 'asynchronous':
                    store.pull_into(snap)
                    torch.cuda.current_stream().synchronize()
                    for lo, hi in ranges:
                        # a chunk snapshot is a whole number of pushes: all of its elements moved by the same amount
                        if float((snap[lo:hi] - snap[lo]).abs().max()) != 0.0:
                            torn.append((tid, i, lo))
            torch.cuda.current_stream().synchronize()

    ts = [threading.Thread(target=work, args=(t,)) for t in range(threads)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    torch.cuda.synchronize()
    final = store.pull_flat()
    assert torch.equal(final, torch.full((n,), -float(pushes * sum(range(1, threads + 1))), device='cuda'))
    mirror = torch.empty(n, dtype=torch.bfloat16, device='cuda')
    store.pull_into(None, mirror)
    torch.cuda.synchronize()
    if mode == 'asynchronous':  # (hogwild: a racing read-back may leave an older mirror entry, see device_ps.py)
        assert torch.equal(mirror, final.to(torch.bfloat16))
    assert not torn
    assert store.watchdog_status() == 0  # no waiter ever gave up
    destroy_store(port)


def test_fused_mlp_predict_and_evaluate_match_the_gemm_chain(monkeypatch):
    """K8 on the reference's MNIST MLP: predict / evaluate through the one-kernel forward == the per-layer GEMM chain."""
    from elephas_b200.ops import engine as engine_mod
    x, y, xt, yt = make_mnist_like(2048, 1500)
    torch.manual_seed(12)
    m = zoo.mnist_mlp()
    m.compile(SGD(lr=0.05), 'categorical_crossentropy', ['acc'])
    m.to('cuda')
    m.fit(x, y, batch_size=64, epochs=2)
    outs = {}
    for fused in (True, False):
        monkeypatch.setattr(engine_mod, '_FUSED_MLP', fused)
        eng = m._native_engine()
        eng._fused_mlp = None          # re-evaluate the eligibility under the patched switch
        eng.invalidate_graphs()
        launches = native_launches()
        outs[fused] = (m.predict(xt), m.evaluate(xt, yt))
        outs[fused] += (native_launches() - launches,)
        assert eng._fused_mlp_ok() is fused
    pa, ea, la = outs[True]
    pb, eb, lb = outs[False]
    assert np.allclose(pa, pb, atol=2e-2) and (pa.argmax(1) == pb.argmax(1)).mean() > 0.995
    assert abs(ea[0] - eb[0]) < 5e-3 and abs(ea[1] - eb[1]) < 5e-3
    assert la < lb                     # one forward kernel instead of three GEMMs per block of rows


def test_wide_fused_mlp_predict_matches_the_gemm_chain(monkeypatch):
    """K8, streamed-weights variant, on the Otto MLP (93-512-512-512-9): opt-in, same predictions as the GEMM chain."""
    rng = np.random.default_rng(5)
    xt = rng.random((3000, 93), dtype=np.float32)
    torch.manual_seed(13)
    m = zoo.otto_mlp()
    m.compile(SGD(lr=0.05), 'categorical_crossentropy', ['acc'])
    m.to('cuda')
    outs = {}
    for wide in ('1', '0'):
        monkeypatch.setenv('ELEPHAS_B200_FUSED_MLP_WIDE', wide)
        eng = m._native_engine()
        eng._fused_mlp = None
        eng.invalidate_graphs()
        launches = native_launches()
        outs[wide] = (m.predict(xt), native_launches() - launches)
        assert eng._fused_mlp_ok() is (wide == '1')
    (pa, la), (pb, lb) = outs['1'], outs['0']
    assert np.allclose(pa, pb, atol=2e-2) and (pa.argmax(1) == pb.argmax(1)).mean() > 0.99
    assert la < lb


def native_launches():
    from elephas_b200.ops import native
    return native.launch_count()
