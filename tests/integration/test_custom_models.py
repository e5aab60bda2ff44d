"""A user-defined activation has to reach every worker replica (and the parameter server's copy of the model) through
``custom_objects`` and actually be the function that runs there. The reference only checks that predict / evaluate
return something (tests/integration/test_custom_models.py:37-38); here the result is compared with a hand computation
that uses the custom function, and the model must NOT be taken over by the native Dense engine, which only knows the
built-in activations."""
from itertools import count

import numpy as np
import pytest
import torch

from elephas.spark_model import SparkModel
from elephas.utils import to_simple_rdd
from elephas_b200.models import SGD, Dense, Sequential

_ports = count(4810)
CALLS = {'n': 0}


def shifted_sigmoid(x):
    CALLS['n'] += 1
    return torch.sigmoid(x) + 1


def _by_hand(weights, x):
    """The two-layer network evaluated with NumPy and the custom activation written out."""
    k1, b1, k2, b2 = [np.asarray(w, dtype=np.float64) for w in weights]
    hidden = 1.0 / (1.0 + np.exp(-(x.reshape(-1, 1) @ k1 + b1))) + 1.0
    return 1.0 / (1.0 + np.exp(-(hidden @ k2 + b2)))


@pytest.mark.parametrize('mode,workers', [('synchronous', 2), ('asynchronous', 2), ('hogwild', 3)])
def test_custom_activation_trains_and_predicts_on_every_worker(mode, workers, spark_context):
    rng = np.random.default_rng(3)
    model = Sequential()
    model.add(Dense(4, input_dim=1, activation=shifted_sigmoid))
    model.add(Dense(1, activation='sigmoid'))
    model.compile(SGD(learning_rate=0.1), 'binary_crossentropy', ['acc'])
    assert model._native_engine() is None  # a custom activation keeps the model on the generic (torch) path

    x_train = rng.random(240).astype(np.float32)
    y_train = (x_train > 0.5).astype(np.float32)
    x_test = rng.random(31).astype(np.float32)
    y_test = (x_test > 0.5).astype(np.float32)
    before = model.get_weights()
    spark_model = SparkModel(model, frequency='epoch', mode=mode, num_workers=workers,
                             custom_objects={'shifted_sigmoid': shifted_sigmoid}, port=next(_ports))
    loss_before = spark_model.evaluate(x_test, y_test)[0]
    calls = CALLS['n']
    spark_model.fit(to_simple_rdd(spark_context, x_train, y_train), epochs=3, batch_size=16, verbose=0,
                    validation_split=0.1)
    assert CALLS['n'] > calls  # the replicas built from JSON resolved the name to OUR function and ran it
    after = spark_model.master_network.get_weights()
    assert any(not np.allclose(a, b) for a, b in zip(before, after))  # training moved the master weights

    preds = np.asarray(spark_model.predict(x_test)).reshape(-1)
    assert preds.shape == (31,)
    # the distributed prediction is the network with the +1-shifted sigmoid, not with a plain one
    assert np.allclose(preds, _by_hand(after, x_test).reshape(-1), atol=1e-5)
    loss, acc = spark_model.evaluate(x_test, y_test)
    expect = _by_hand(after, x_test).reshape(-1)
    by_hand_loss = float(np.mean(-(y_test * np.log(expect) + (1 - y_test) * np.log(1 - expect))))
    assert abs(loss - by_hand_loss) < 1e-4 and 0.0 <= acc <= 1.0
    if mode == 'synchronous':  # (a parameter server ADDS the workers' deltas: no monotonicity promise after one fit)
        assert loss < loss_before


@pytest.mark.parametrize('mode,ps', [('synchronous', None), ('asynchronous', 'socket'), ('hogwild', 'nvlink')])
def test_recurrent_model_trains_through_every_mode(spark_context, mode, ps):
    """A sequence model (Embedding -> LSTM -> Dense) travels to the workers as JSON like any Keras model the reference
    accepts (elephas/worker.py:27-34), trains in all three modes and predicts the same through SparkModel and locally."""
    import numpy as np
    import torch
    from elephas.spark_model import SparkModel
    from elephas.utils.rdd_utils import to_simple_rdd
    from elephas_b200.models import LSTM, Activation, Adam, Dense, Embedding, Sequential
    torch.manual_seed(11)
    rng = np.random.default_rng(11)
    tokens = rng.integers(1, 12, (192, 8))
    y = np.eye(2, dtype='float32')[(tokens[:, 0] > 6).astype(int)]      # the first token decides the class
    model = Sequential([Embedding(12, 6, input_length=8), LSTM(8, go_backwards=True), Dense(2), Activation('softmax')])
    model.compile(Adam(learning_rate=0.02), 'categorical_crossentropy', ['acc'])
    kwargs = {'parameter_server_mode': ps, 'port': 4000 + int(rng.integers(100, 900))} if ps else {}
    sm = SparkModel(model, mode=mode, frequency='epoch', num_workers=2, **kwargs)
    before = sm.evaluate(tokens, y)
    sm.fit(to_simple_rdd(spark_context, tokens, y), epochs=10, batch_size=16, verbose=0)
    after = sm.evaluate(tokens, y)
    assert after[0] < before[0] and after[1] > 0.8
    assert np.allclose(sm.predict(tokens[:10]), sm.master_network.predict(tokens[:10]), atol=1e-6)
