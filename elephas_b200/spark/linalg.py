"""Dense vectors / matrices with the ``pyspark.mllib.linalg`` / ``pyspark.ml.linalg`` surface that
Elephas touches: ``Vectors.dense``, ``Vectors.fromML``, ``Matrices.dense`` (column-major, like Spark),
``toArray``. Used when PySpark is not installed."""
import numpy as np


class Vector:
    pass


class DenseVector(Vector):
    def __init__(self, values):
        self.values = np.asarray(values, dtype=np.float64).reshape(-1)

    @classmethod
    def _wrap(cls, values):
        """A vector around an existing 1-D float64 array (no copy, no checks): bulk producers use this."""
        v = cls.__new__(cls)
        v.values = values
        return v

    def toArray(self):
        return self.values

    def __len__(self):
        return self.values.shape[0]

    def __getitem__(self, i):
        return self.values[i]

    def __iter__(self):
        return iter(self.values)

    def __array__(self, dtype=None, copy=None):
        return self.values.astype(dtype) if dtype is not None else self.values

    def __eq__(self, other):
        return isinstance(other, DenseVector) and np.array_equal(self.values, other.values)

    def __hash__(self):
        return hash(self.values.tobytes())

    def __repr__(self):
        return "DenseVector(%s)" % (self.values.tolist(),)

    def asML(self):
        return self


class Vectors:
    @staticmethod
    def dense(*values):
        if len(values) == 1 and not np.isscalar(values[0]):
            values = values[0]
        return DenseVector(values)

    @staticmethod
    def fromML(vec):
        return vec if isinstance(vec, DenseVector) else DenseVector(vec.toArray())


class Matrix:
    pass


class DenseMatrix(Matrix):
    """``values`` are stored column-major, exactly like Spark's DenseMatrix."""

    def __init__(self, numRows, numCols, values, isTransposed=False):
        self.numRows, self.numCols = int(numRows), int(numCols)
        self.values = np.asarray(values, dtype=np.float64).reshape(-1)
        self.isTransposed = isTransposed
        if self.values.size != self.numRows * self.numCols:
            raise ValueError("DenseMatrix needs numRows*numCols values")

    def toArray(self):
        if self.isTransposed:
            return self.values.reshape(self.numRows, self.numCols)
        return self.values.reshape(self.numCols, self.numRows).T

    def __repr__(self):
        return "DenseMatrix(%d, %d)" % (self.numRows, self.numCols)


class Matrices:
    @staticmethod
    def dense(numRows, numCols, values):
        return DenseMatrix(numRows, numCols, values)
