"""The NumPy restatement of the reference's algorithm (baseline/cpu_restatement.py, written without this package) as an
oracle: the package's synchronous mode, run on CPU ranks from the same initial weights on the same ordered data, must
end with the same master weights; its parameter-server modes must conserve the sum of the pushed deltas."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), 'baseline'))
import cpu_restatement as ref  # noqa: E402

from elephas.spark_model import SparkModel  # noqa: E402
from elephas.utils.rdd_utils import to_simple_rdd  # noqa: E402
from elephas_b200.models import SGD, Dense, Sequential  # noqa: E402


def _pair(dims, output, seed):
    """The restated MLP and the package's model with identical weights."""
    oracle = ref.Mlp(dims, output, seed=seed)
    model = Sequential()
    for i, (a, b) in enumerate(zip(dims[:-1], dims[1:])):
        last = i == len(dims) - 2
        act = (output if output == 'softmax' else 'linear') if last else 'relu'
        model.add(Dense(b, input_dim=a, activation=act) if i == 0 else Dense(b, activation=act))
    model.compile(SGD(learning_rate=0.1), 'categorical_crossentropy' if output == 'softmax' else 'mse')
    model.set_weights(oracle.get_weights())
    return oracle, model


@pytest.mark.parametrize('workers,epochs', [(2, 1), (3, 2)])
def test_synchronous_mode_equals_the_restated_reference(spark_context, workers, epochs):
    x, y = ref.synthetic_mnist(96 * workers + 17, seed=4)
    x, y = x[:, :40].copy(), y
    oracle, model = _pair([40, 16, 10], 'softmax', seed=5)
    ref.sync_fit(oracle, ref.partitions(x, y, workers), epochs, 32, 0.1)
    sm = SparkModel(model, mode='synchronous', frequency='epoch', num_workers=workers, sync_rounds='fit')
    sm.fit(to_simple_rdd(spark_context, x, y), epochs=epochs, batch_size=32, verbose=0, shuffle=False)
    for got, want in zip(sm.master_network.get_weights(), oracle.get_weights()):
        assert np.allclose(got, want, atol=2e-5), np.abs(got - want).max()
    assert abs(ref.Mlp.loss(oracle, x, y) - sm.master_network.evaluate(x, y)) < 1e-4


def test_regression_head_matches_too(spark_context):
    rng = np.random.default_rng(6)
    x = rng.normal(size=(150, 13)).astype(np.float32)
    y = (x[:, :1] * 0.5 - x[:, 1:2]).astype(np.float32)
    oracle, model = _pair([13, 8, 1], 'linear', seed=7)
    ref.sync_fit(oracle, ref.partitions(x, y, 2), 2, 16, 0.1)
    sm = SparkModel(model, mode='synchronous', num_workers=2, sync_rounds='fit')
    sm.fit(to_simple_rdd(spark_context, x, y), epochs=2, batch_size=16, verbose=0, shuffle=False)
    for got, want in zip(sm.master_network.get_weights(), oracle.get_weights()):
        assert np.allclose(got, want, atol=2e-5), np.abs(got - want).max()


def test_restated_parameter_server_loses_no_update_and_single_worker_async_is_sequential_sgd():
    """One worker against the restated server is plain SGD (pull, step, push of exactly that step), in both modes; with
    several workers the locked server applies every delta exactly once (its weights are start - sum of deltas)."""
    x, y = ref.synthetic_mnist(200, seed=8)
    x = x[:, :30].copy()
    for mode in ('asynchronous', 'hogwild'):
        a, b = ref.Mlp([30, 12, 10], seed=9), ref.Mlp([30, 12, 10], seed=9)
        ref.async_fit(a, [(x, y)], 2, 25, 0.1, mode, 'batch')
        b.fit(x, y, 2, 25, 0.1)
        for got, want in zip(a.get_weights(), b.get_weights()):
            assert np.allclose(got, want, atol=1e-6)
    server = ref.ParameterServer([np.zeros(1000, dtype=np.float64)], 'asynchronous')
    import threading
    ts = [threading.Thread(target=lambda k=k: [server.update([np.full(1000, float(k + 1))]) for _ in range(50)])
          for k in range(4)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert np.array_equal(server.get()[0], np.full(1000, -50.0 * 10))
