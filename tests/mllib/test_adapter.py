import numpy as np
import pytest

from elephas.mllib.adapter import from_matrix, from_vector, to_matrix, to_vector


def test_to_matrix():
    assert to_matrix(np.ones((4, 2))).numRows == 4


def test_from_matrix():
    x = np.arange(8, dtype=float).reshape(4, 2)
    assert np.array_equal(from_matrix(to_matrix(x)), x)  # layout-correct round trip (D6)


def test_to_matrix_rejects_non_2d():
    with pytest.raises(Exception):
        to_matrix(np.ones(3))


def test_vector_round_trip():
    x = np.arange(3, dtype=float)
    assert len(to_vector(x)) == 3 and np.array_equal(from_vector(to_vector(x)), x)
    with pytest.raises(Exception):
        to_vector(np.ones((2, 2)))
