"""The reference's acceptance matrix (tests/integration/test_end_to_end.py:18-144): every training mode x
parameter-server transport x {default partitioning, 2 workers}, plus the device-resident 'nvlink' server.
Invariants asserted: predict(ndarray) == predict(rdd) == master_network.predict, and distributed evaluate
within 0.01 of single-process evaluate."""
from itertools import count
from math import isclose

import numpy as np
import pytest

from elephas.spark_model import SparkModel
from elephas.utils.rdd_utils import to_simple_rdd
from elephas_b200.models import SGD

_ports = count(3001)

MATRIX = [('synchronous', None, None), ('synchronous', None, 2),
          ('asynchronous', 'http', None), ('asynchronous', 'http', 2),
          ('asynchronous', 'socket', None), ('asynchronous', 'socket', 2),
          ('hogwild', 'http', None), ('hogwild', 'http', 2),
          ('hogwild', 'socket', None), ('hogwild', 'socket', 2),
          ('asynchronous', 'nvlink', 2), ('hogwild', 'nvlink', 2)]


@pytest.mark.parametrize('mode,parameter_server_mode,num_workers', MATRIX)
def test_training_classification(spark_context, mode, parameter_server_mode, num_workers, mnist_data,
                                 classification_model):
    batch_size, epochs = 64, 4
    x_train, y_train, x_test, y_test = mnist_data
    classification_model.compile(SGD(lr=0.1), 'categorical_crossentropy', ['acc'])
    rdd = to_simple_rdd(spark_context, x_train, y_train)
    spark_model = SparkModel(classification_model, frequency='epoch', num_workers=num_workers, mode=mode,
                             parameter_server_mode=parameter_server_mode, port=next(_ports))
    before = spark_model.master_network.evaluate(x_test, y_test)[0]
    spark_model.fit(rdd, epochs=epochs, batch_size=batch_size, verbose=0, validation_split=0.1)
    predictions = spark_model.predict(x_test)
    evals = spark_model.evaluate(x_test, y_test)
    assert len(predictions) == len(x_test)
    test_rdd = spark_context.parallelize(x_test)
    assert [np.argmax(x) for x in predictions] == [np.argmax(x) for x in spark_model.predict(test_rdd)]
    local = spark_model.master_network.predict(x_test)
    assert [np.argmax(x) for x in predictions] == [np.argmax(x) for x in local]
    ref = spark_model.master_network.evaluate(x_test, y_test)
    assert isclose(evals[0], ref[0], abs_tol=0.01)
    assert isclose(evals[1], ref[1], abs_tol=0.01)
    assert evals[0] < before  # training made progress (the reference never checks this)


@pytest.mark.parametrize('mode,parameter_server_mode,num_workers', MATRIX[:10])
def test_training_regression(spark_context, mode, parameter_server_mode, num_workers, boston_housing_dataset,
                             regression_model):
    x_train, y_train, x_test, y_test = boston_housing_dataset
    rdd = to_simple_rdd(spark_context, x_train, y_train)
    regression_model.compile(SGD(lr=0.0000001), 'mse', ['mae', 'mean_absolute_percentage_error'])
    spark_model = SparkModel(regression_model, frequency='epoch', mode=mode, num_workers=num_workers,
                             parameter_server_mode=parameter_server_mode, port=next(_ports))
    spark_model.fit(rdd, epochs=3, batch_size=64, verbose=0, validation_split=0.1)
    predictions = spark_model.predict(x_test)
    evals = spark_model.evaluate(x_test, y_test)
    test_rdd = spark_context.parallelize(x_test)
    assert all(np.isclose(x, y, 0.01) for x, y in zip(predictions, spark_model.predict(test_rdd)))
    assert all(np.isclose(x, y, 0.01) for x, y in zip(predictions, spark_model.master_network.predict(x_test)))
    ref = spark_model.master_network.evaluate(x_test, y_test)
    for i in range(3):
        assert isclose(evals[i], ref[i], rel_tol=1e-3, abs_tol=0.01)


def test_training_regression_no_metrics(spark_context, boston_housing_dataset, regression_model):
    x_train, y_train, x_test, y_test = boston_housing_dataset
    rdd = to_simple_rdd(spark_context, x_train, y_train)
    regression_model.compile(SGD(lr=0.0000001), 'mse')
    spark_model = SparkModel(regression_model, frequency='epoch', mode='synchronous', port=next(_ports))
    spark_model.fit(rdd, epochs=1, batch_size=64, verbose=0, validation_split=0.1)
    result = spark_model.evaluate(x_test, y_test)
    assert isinstance(result, float)  # scalar when there are no metrics (spark_model.py:308)
    assert isclose(result, spark_model.master_network.evaluate(x_test, y_test), rel_tol=1e-3, abs_tol=0.01)


@pytest.mark.parametrize('mode,ps', [('synchronous', None), ('asynchronous', 'socket'), ('hogwild', 'nvlink')])
def test_batch_frequency(spark_context, mode, ps, mnist_data, classification_model):
    x_train, y_train, x_test, y_test = mnist_data
    classification_model.compile(SGD(lr=0.05), 'categorical_crossentropy', ['acc'])
    spark_model = SparkModel(classification_model, frequency='batch', mode=mode, parameter_server_mode=ps,
                             num_workers=2, port=next(_ports))
    before = spark_model.master_network.evaluate(x_test, y_test)[0]
    spark_model.fit(to_simple_rdd(spark_context, x_train[:512], y_train[:512]), epochs=1, batch_size=64)
    assert spark_model.evaluate(x_test, y_test)[0] < before


def test_sync_average_matches_numpy_oracle(spark_context, mnist_data, classification_model):
    """sync_rounds='fit' reproduces the reference exactly: w_new = mean_i(w_i) over independently trained
    replicas (spark_model.py:221-227). Dropout is removed so replicas are deterministic."""
    from elephas.worker import SparkWorker
    from elephas_b200.models import Dense, Sequential, optimizers
    x_train, y_train, _, _ = mnist_data
    model = Sequential([Dense(16, input_dim=784, activation='relu'), Dense(10, activation='softmax')])
    model.compile(SGD(lr=0.05), 'categorical_crossentropy')
    w0 = model.get_weights()
    cfg = dict(epochs=2, batch_size=50, shuffle=False)
    sm = SparkModel(model, mode='synchronous', num_workers=2, sync_rounds='fit')
    sm.fit(to_simple_rdd(spark_context, x_train[:400], y_train[:400]), **cfg)
    replicas = []
    for lo, hi in ((0, 200), (200, 400)):
        w = SparkWorker(model.to_json(), w0, cfg, optimizers.serialize(SGD(lr=0.05)), 'categorical_crossentropy', [], {})
        w.begin(list(zip(x_train[lo:hi], y_train[lo:hi])))
        delta, _ = w.train_round(None, epochs=2)
        replicas.append([a - d for a, d in zip(w0, delta)])
    for got, parts in zip(sm.master_network.get_weights(), zip(*replicas)):
        assert np.allclose(got, np.mean(parts, axis=0), atol=1e-5)


def test_small_partition_contributes_zero_delta(spark_context, classification_model):
    """rows <= batch_size: the worker trains nothing and its zero delta still counts in N (worker.py:41)."""
    rng = np.random.default_rng(0)
    x = rng.random((40, 784), dtype=np.float32)
    y = np.eye(10, dtype=np.float32)[rng.integers(0, 10, 40)]
    classification_model.compile(SGD(lr=0.1), 'categorical_crossentropy')
    w0 = classification_model.get_weights()
    sm = SparkModel(classification_model, mode='synchronous', num_workers=2)
    sm.fit(to_simple_rdd(spark_context, x, y), epochs=1, batch_size=64)
    assert all(np.array_equal(a, b) for a, b in zip(w0, sm.master_network.get_weights()))
    assert sm.training_histories == [None, None]


def test_unknown_mode(spark_context, classification_model, mnist_data):
    classification_model.compile('sgd', 'categorical_crossentropy')
    sm = SparkModel(classification_model, mode='synchronous')
    sm.mode = 'telepathy'
    with pytest.raises(ValueError):
        sm.fit(to_simple_rdd(spark_context, mnist_data[0][:10], mnist_data[1][:10]), epochs=1, batch_size=2)


def test_uncompiled_model_is_rejected(classification_model):
    with pytest.raises(Exception):
        SparkModel(classification_model)


@pytest.mark.parametrize('mode,frequency', [('asynchronous', 'batch'), ('hogwild', 'epoch')])
def test_cpu_workers_move_flat_tensors_against_the_in_process_store(spark_context, mode, frequency):
    """``parameter_server_mode='nvlink'`` without a GPU: the store is a flat tensor in this process and the workers pull /
    push flat tensors in the model's own layout (no list of arrays, no pickling) - same training result as the socket
    server within the tolerance of asynchronous updates, and every round goes through the fast path."""
    import numpy as np
    from elephas_b200 import worker as worker_mod
    from elephas_b200.models import SGD, Dense, Sequential
    from elephas_b200.parallel import get_runtime
    if get_runtime().device.type != 'cpu':
        pytest.skip('CPU form of the device path')
    import torch
    torch.manual_seed(17)  # reproducible initial weights
    rng = np.random.default_rng(3)
    x = rng.normal(size=(512, 20)).astype('float32')
    labels = (x[:, :4].sum(1) > 0).astype(int)
    y = np.eye(2, dtype='float32')[labels]
    model = Sequential([Dense(19, input_dim=20, activation='relu'), Dense(2, activation='softmax')])   # 19: padded rows
    model.compile(SGD(learning_rate=0.2), 'categorical_crossentropy', ['acc'])
    sm = SparkModel(model, mode=mode, frequency=frequency, parameter_server_mode='nvlink', num_workers=2)
    before = sm.evaluate(x, y)
    rounds = worker_mod.STATS['host_store_rounds']
    sm.fit(to_simple_rdd(spark_context, x, y), epochs=4, batch_size=32, verbose=0)
    after = sm.evaluate(x, y)
    expected = 2 * 4 * (8 if frequency == 'batch' else 1)   # workers x epochs x (batches of a 256-row partition | 1)
    assert worker_mod.STATS['host_store_rounds'] - rounds == expected
    assert after[0] < before[0] and after[1] > 0.85
