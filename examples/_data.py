"""Synthetic stand-ins for the datasets the reference's examples download (MNIST, Boston housing, Otto):
same shapes and value ranges, learnable structure, no network access needed."""
import os
import sys

import numpy as np

# run from a checkout without installing the package
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault('SPARK_LOCAL_IP', '127.0.0.1')


def mnist_like(n_train=6000, n_test=1000, seed=0):
    rng = np.random.default_rng(seed)
    centers = (rng.random((10, 784)) < 0.18) * (0.5 + 0.5 * rng.random((10, 784)))

    def draw(n):
        labels = rng.integers(0, 10, n)
        x = centers[labels] * (rng.random((n, 784)) < 0.85) + 0.08 * rng.random((n, 784)) * (rng.random((n, 784)) < 0.3)
        return np.clip(x, 0, 1).astype('float32'), np.eye(10, dtype='float32')[labels]

    return draw(n_train) + draw(n_test)


def boston_like(n_train=404, n_test=102, seed=1):
    rng = np.random.default_rng(seed)
    w = rng.normal(size=13)
    scale = np.array([80, 100, 25, 1, 1, 8, 100, 12, 24, 700, 22, 400, 35])

    def draw(n):
        x = rng.random((n, 13)) * scale
        y = 22.0 + (x / scale) @ w * 5.0 + rng.normal(scale=0.5, size=n)
        return x.astype('float32'), y.astype('float32')

    return draw(n_train) + draw(n_test)


def otto_like(n=6000, seed=2):
    rng = np.random.default_rng(seed)
    centers = rng.gamma(1.0, 2.0, size=(9, 93))
    labels = rng.integers(0, 9, n)
    feats = rng.poisson(centers[labels]).astype('float32')
    return feats, ['Class_%d' % (l + 1) for l in labels]
