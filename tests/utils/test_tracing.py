"""Tracing hooks (new: the reference has no timers or profiler ranges)."""
import numpy as np

from elephas_b200.models import SGD, Dense, Sequential
from elephas_b200.spark import SparkContext
from elephas_b200.spark_model import SparkModel
from elephas_b200.utils import tracing
from elephas_b200.utils.rdd_utils import to_simple_rdd


def test_spans_accumulate_and_report():
    tracing.reset()
    tracing.enable()
    try:
        for _ in range(3):
            with tracing.span('phase/a'):
                pass
        with tracing.device_span('phase/b'):
            pass
        tracing.add_host('phase/c', 0.25)
        s = tracing.summary()
        assert s['host']['phase/a']['calls'] == 3
        assert abs(s['host']['phase/c']['total_ms'] - 250.0) < 1e-6
        assert 'phase/b' in s['host'] or 'phase/b' in s['device']
        text = tracing.report()
        assert 'phase/a' in text and 'phase/c' in text
    finally:
        tracing.enable(False)
        tracing.reset()


def test_disabled_tracing_records_nothing():
    tracing.reset()
    tracing.enable(False)
    with tracing.span('x'):
        pass
    tracing.add_host('y', 1.0)
    assert tracing.summary() == {'host': {}, 'device': {}}


def test_fit_is_traced():
    tracing.reset()
    tracing.enable()
    try:
        rng = np.random.RandomState(0)
        x = rng.rand(64, 8).astype('float32')
        y = rng.rand(64, 1).astype('float32')
        m = Sequential([Dense(4, input_dim=8, activation='relu'), Dense(1)])
        m.compile(SGD(learning_rate=0.01), 'mse')
        sm = SparkModel(m, mode='synchronous', frequency='epoch', num_workers=2)
        sm.fit(to_simple_rdd(SparkContext.getOrCreate(), x, y), epochs=1, batch_size=8)
        host = tracing.summary()['host']
        assert host['fit']['calls'] == 1 and 'fit/materialize_partitions' in host
    finally:
        tracing.enable(False)
        tracing.reset()
