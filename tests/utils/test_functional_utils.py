import numpy as np
import pytest
import torch

from elephas.utils import functional_utils


def test_add_params():
    res = functional_utils.add_params([np.ones((5, 5)), np.ones(3)], [np.ones((5, 5)), 2 * np.ones(3)])
    assert np.array_equal(res[0], 2 * np.ones((5, 5)))
    assert np.array_equal(res[1], 3 * np.ones(3))


def test_subtract_params():
    res = functional_utils.subtract_params([np.ones((5, 5)), np.ones(3)], [np.ones((5, 5)), 3 * np.ones(3)])
    assert np.array_equal(res[0], np.zeros((5, 5)))
    assert np.array_equal(res[1], -2 * np.ones(3))


def test_get_neutral():
    res = functional_utils.get_neutral([np.ones((3, 4)), np.ones(2)])
    assert res[0].shape == (3, 4) and res[0].sum() == 0
    assert res[1].shape == (2,) and res[1].sum() == 0


def test_divide_by():
    res = functional_utils.divide_by([np.ones((5, 5)), 4 * np.ones(2)], 2)
    assert np.array_equal(res[0], 0.5 * np.ones((5, 5)))
    assert np.array_equal(res[1], 2 * np.ones(2))


def test_length_mismatch_is_an_error():
    with pytest.raises(ValueError):
        functional_utils.add_params([np.ones(2)], [np.ones(2), np.ones(2)])


def test_torch_tensors_stay_tensors():
    a, b = [torch.ones(3)], [torch.full((3,), 2.0)]
    assert torch.equal(functional_utils.subtract_params(a, b)[0], -torch.ones(3))
    assert torch.equal(functional_utils.get_neutral(a)[0], torch.zeros(3))


def test_sync_average_identity():
    """w0 - mean_i(w0 - w_i) == mean_i(w_i): the arithmetic of spark_model.py:221-227."""
    rng = np.random.default_rng(0)
    w0 = [rng.normal(size=(4, 3)), rng.normal(size=3)]
    workers = [[w + rng.normal(size=w.shape) for w in w0] for _ in range(4)]
    new = w0
    for wi in workers:
        delta = functional_utils.subtract_params(w0, wi)
        new = functional_utils.subtract_params(new, functional_utils.divide_by(delta, 4))
    for got, parts in zip(new, zip(*workers)):
        assert np.allclose(got, np.mean(parts, axis=0))
