"""Shared fixtures. The reference's fixtures download MNIST / Boston housing through Keras
(``tests/conftest.py:43-63``); there is no network here, so the data is synthetic with the same
shapes and dtypes, and learnable (class-dependent means / a noisy linear target)."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault('SPARK_LOCAL_IP', '127.0.0.1')

from elephas_b200.models import Activation, Dense, Dropout, Input, Model, Sequential, to_categorical  # noqa: E402
from elephas_b200.spark import SparkContext, SparkSession  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200)")


def pytest_collection_modifyitems(config, items):
    import torch
    # test_zz_no_alloc.py runs after everything else (directories such as tests/utils sort behind it otherwise)
    last = [it for it in items if it.fspath.basename == 'test_zz_no_alloc.py']
    if last:
        items[:] = [it for it in items if it.fspath.basename != 'test_zz_no_alloc.py'] + last
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if 'gpu' in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope='session')
def spark_context():
    sc = SparkContext.getOrCreate()
    sc.defaultParallelism = 2
    yield sc


@pytest.fixture(scope='session')
def spark_session(spark_context):
    return SparkSession.builder.getOrCreate()


@pytest.fixture
def classification_model():
    model = Sequential()
    model.add(Dense(128, input_dim=784))
    model.add(Activation('relu'))
    model.add(Dropout(0.2))
    model.add(Dense(128))
    model.add(Activation('relu'))
    model.add(Dropout(0.2))
    model.add(Dense(10))
    model.add(Activation('softmax'))
    return model


@pytest.fixture
def regression_model():
    model = Sequential()
    model.add(Dense(64, activation='relu', input_shape=(13,)))
    model.add(Dense(64, activation='relu'))
    model.add(Dense(1, activation='linear'))
    return model


@pytest.fixture
def classification_model_functional():
    input_layer = Input(shape=(784,))
    hidden = Dense(128, activation='relu')(input_layer)
    dropout = Dropout(0.2)(hidden)
    hidden2 = Dense(128, activation='relu')(dropout)
    dropout2 = Dropout(0.2)(hidden2)
    output = Dense(10, activation='softmax')(dropout2)
    return Model(inputs=input_layer, outputs=output)


def make_mnist_like(n_train=1000, n_test=300, seed=0):
    """MNIST-shaped and MNIST-like in statistics: sparse non-negative 'strokes' per class (mean pixel
    value ~0.13), pixel noise, values in [0, 1]."""
    rng = np.random.default_rng(seed)
    centers = (rng.random((10, 784)) < 0.18) * (0.5 + 0.5 * rng.random((10, 784)))

    def draw(n):
        labels = rng.integers(0, 10, n)
        keep = rng.random((n, 784)) < 0.85
        x = centers[labels] * keep + 0.08 * rng.random((n, 784)) * (rng.random((n, 784)) < 0.3)
        return np.clip(x, 0.0, 1.0).astype(np.float32), to_categorical(labels, 10)

    x_train, y_train = draw(n_train)
    x_test, y_test = draw(n_test)
    return x_train, y_train, x_test, y_test


def make_boston_like(n_train=404, n_test=102, seed=1):
    rng = np.random.default_rng(seed)
    w = rng.normal(size=13)

    def draw(n):
        x = rng.random((n, 13)) * np.array([80, 100, 25, 1, 1, 8, 100, 12, 24, 700, 22, 400, 35])
        y = 22.0 + (x / x.max(axis=0)) @ w * 5.0 + rng.normal(scale=0.5, size=n)
        return x.astype(np.float32), y.astype(np.float32)

    x_train, y_train = draw(n_train)
    x_test, y_test = draw(n_test)
    return x_train, y_train, x_test, y_test


@pytest.fixture(scope='session')
def mnist_data():
    return make_mnist_like()


@pytest.fixture(scope='session')
def boston_housing_dataset():
    return make_boston_like()


@pytest.fixture(autouse=True)
def _seeded():
    """CPU tests start from the same torch / NumPy seeds: weight initialisation and dropout masks are reproducible
    (tests that want other seeds set them themselves)."""
    import torch
    if not torch.cuda.is_available():
        # CPU runs (the per-round check) are fully reproducible; GPU runs keep fresh seeds, which is what the numerics
        # tolerances there were exercised with
        torch.manual_seed(20240607)
        np.random.seed(20240607)
    yield
