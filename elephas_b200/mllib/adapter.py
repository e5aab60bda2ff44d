"""NumPy <-> MLlib ``Matrix`` / ``Vector`` conversions.

Reference: ``elephas/mllib/adapter.py:5-35``. One deliberate difference: the
reference hands a row-major ``ravel()`` to the column-major ``Matrices.dense`` (``adapter.py:14-16``),
which silently transposes non-square data; here ``to_matrix`` emits column-major values, so
``from_matrix(to_matrix(a)) == a``.
"""
import numpy as np

from ..spark import Matrices, Matrix, Vector, Vectors


def from_matrix(matrix: Matrix) -> np.ndarray:
    """MLlib Matrix -> 2-D NumPy array."""
    return matrix.toArray()


def to_matrix(np_array: np.ndarray) -> Matrix:
    """2-D NumPy array -> MLlib dense Matrix."""
    np_array = np.asarray(np_array)
    if np_array.ndim != 2:
        raise Exception("An MLLib Matrix can only be created from a two-dimensional "
                        "numpy array, got {}".format(np_array.ndim))
    return Matrices.dense(np_array.shape[0], np_array.shape[1], np_array.ravel(order='F'))


def from_vector(vector: Vector) -> np.ndarray:
    """MLlib Vector -> 1-D NumPy array."""
    return vector.toArray()


def to_vector(np_array: np.ndarray) -> Vector:
    """1-D NumPy array -> MLlib dense Vector."""
    np_array = np.asarray(np_array)
    if np_array.ndim != 1:
        raise Exception("An MLLib Vector can only be created from a one-dimensional "
                        "numpy array, got {}".format(np_array.ndim))
    return Vectors.dense(np_array)
