"""More of ``pyspark.ml.feature`` for pipelines around ``ElephasEstimator`` when PySpark itself is not installed:
``VectorAssembler``, ``MinMaxScaler``, ``OneHotEncoder``, ``Normalizer``, ``IndexToString``.

Same parameter names and results as Spark's stages (dense vectors where Spark would choose sparse ones). The stages work
on whole columns: they read a column as one matrix (``DataFrame.column_block``) and append their output as one matrix
(``with_block_column``), so a pipeline that ends in ``ElephasEstimator`` never builds per-row Python objects - the
reference's pipelines (``examples/ml_pipeline_otto.py``) pay for those at every stage.
"""
import numpy as np

from .ml import (DefaultParamsReadable, DefaultParamsWritable, Estimator, HasInputCol, HasOutputCol, Model, Param,
                 Transformer, _WithData, keyword_only)
from .sql import DoubleType, StringType, VectorUDT


def _matrix(dataset, name):
    block = np.asarray(dataset.column_block(name), dtype=np.float64)
    return block.reshape(len(block), -1)


class VectorAssembler(Transformer, HasOutputCol, DefaultParamsReadable, DefaultParamsWritable):
    """Concatenates numeric and vector columns into one vector column."""

    @keyword_only
    def __init__(self, inputCols=None, outputCol=None, handleInvalid='error'):
        super().__init__()
        self.inputCols = Param(self, 'inputCols', 'input column names')
        self.handleInvalid = Param(self, 'handleInvalid', "how to handle NaN: 'error', 'skip' or 'keep'")
        self._setDefault(handleInvalid='error')
        self._set(**{k: v for k, v in self._input_kwargs.items() if v is not None})

    def setInputCols(self, value):
        return self._set(inputCols=list(value))

    def getInputCols(self):
        return self.getOrDefault(self.inputCols)

    def _transform(self, dataset):
        out = np.hstack([_matrix(dataset, c) for c in self.getInputCols()])
        mode = self.getOrDefault(self.handleInvalid)
        bad = np.isnan(out).any(axis=1)
        if bad.any():
            if mode == 'error':
                raise ValueError("VectorAssembler: NaN in the input columns (handleInvalid='error')")
            if mode == 'skip':
                keep = np.flatnonzero(~bad).tolist()
                rows = dataset.collect()
                from .context import RDD
                from .sql import DataFrame
                dataset = DataFrame(RDD([[rows[i] for i in keep]], dataset.rdd.context), dataset.schema, dataset.sparkSession)
                out = out[~bad]
        return dataset.with_block_column(self.getOutputCol(), out, VectorUDT())


class MinMaxScalerModel(_WithData, Model, HasInputCol, HasOutputCol, DefaultParamsReadable, DefaultParamsWritable):
    _state = ('originalMin', 'originalMax', 'lo', 'hi')

    def __init__(self, originalMin=None, originalMax=None, lo=0.0, hi=1.0, inputCol=None, outputCol=None):
        super().__init__()
        self.originalMin, self.originalMax, self.lo, self.hi = originalMin, originalMax, lo, hi
        self._set(inputCol=inputCol, outputCol=outputCol)

    def _transform(self, dataset):
        x = _matrix(dataset, self.getInputCol())
        span = self.originalMax - self.originalMin
        unit = np.where(span > 0, (x - self.originalMin) / np.where(span > 0, span, 1.0), 0.5)  # constant features -> middle
        return dataset.with_block_column(self.getOutputCol(), unit * (self.hi - self.lo) + self.lo, VectorUDT())


class MinMaxScaler(Estimator, HasInputCol, HasOutputCol, DefaultParamsReadable, DefaultParamsWritable):
    @keyword_only
    def __init__(self, min=0.0, max=1.0, inputCol=None, outputCol=None):  # noqa: A002 - Spark's parameter names
        super().__init__()
        self.min = Param(self, 'min', 'lower bound of the output range')
        self.max = Param(self, 'max', 'upper bound of the output range')
        self._setDefault(min=0.0, max=1.0)
        self._set(**{k: v for k, v in self._input_kwargs.items() if v is not None})

    def _fit(self, dataset):
        x = _matrix(dataset, self.getInputCol())
        return MinMaxScalerModel(x.min(axis=0), x.max(axis=0), float(self.getOrDefault(self.min)),
                                 float(self.getOrDefault(self.max)), self.getInputCol(), self.getOutputCol())


class OneHotEncoderModel(_WithData, Model, HasInputCol, HasOutputCol, DefaultParamsReadable, DefaultParamsWritable):
    _state = ('categorySize', 'dropLast')

    def __init__(self, categorySize=0, dropLast=True, inputCol=None, outputCol=None):
        super().__init__()
        self.categorySize, self.dropLast = int(categorySize), bool(dropLast)
        self._set(inputCol=inputCol, outputCol=outputCol)

    def _transform(self, dataset):
        idx = _matrix(dataset, self.getInputCol())[:, 0].astype(np.int64)
        if idx.size and (idx.min() < 0 or idx.max() >= self.categorySize):
            raise ValueError("OneHotEncoder: index outside of [0, %d)" % self.categorySize)
        width = self.categorySize - 1 if self.dropLast else self.categorySize
        out = np.zeros((len(idx), width), dtype=np.float64)
        keep = idx < width
        out[np.flatnonzero(keep), idx[keep]] = 1.0
        return dataset.with_block_column(self.getOutputCol(), out, VectorUDT())


class OneHotEncoder(Estimator, HasInputCol, HasOutputCol, DefaultParamsReadable, DefaultParamsWritable):
    """Category indices (``StringIndexer`` output) -> indicator vectors; ``dropLast`` omits the last category as Spark does."""

    @keyword_only
    def __init__(self, inputCol=None, outputCol=None, dropLast=True):
        super().__init__()
        self.dropLast = Param(self, 'dropLast', 'whether to drop the last category')
        self._setDefault(dropLast=True)
        self._set(**{k: v for k, v in self._input_kwargs.items() if v is not None})

    def _fit(self, dataset):
        idx = _matrix(dataset, self.getInputCol())[:, 0]
        size = int(idx.max()) + 1 if idx.size else 0
        return OneHotEncoderModel(size, self.getOrDefault(self.dropLast), self.getInputCol(), self.getOutputCol())


class Normalizer(Transformer, HasInputCol, HasOutputCol, DefaultParamsReadable, DefaultParamsWritable):
    """Scales every row vector to unit ``p``-norm."""

    @keyword_only
    def __init__(self, p=2.0, inputCol=None, outputCol=None):
        super().__init__()
        self.p = Param(self, 'p', 'the p norm value')
        self._setDefault(p=2.0)
        self._set(**{k: v for k, v in self._input_kwargs.items() if v is not None})

    def _transform(self, dataset):
        x = _matrix(dataset, self.getInputCol())
        norms = np.linalg.norm(x, ord=float(self.getOrDefault(self.p)), axis=1, keepdims=True)
        return dataset.with_block_column(self.getOutputCol(), x / np.where(norms > 0, norms, 1.0), VectorUDT())


class IndexToString(Transformer, HasInputCol, HasOutputCol, DefaultParamsReadable, DefaultParamsWritable):
    """Maps predicted indices back to the labels of a ``StringIndexerModel``."""

    @keyword_only
    def __init__(self, inputCol=None, outputCol=None, labels=None):
        super().__init__()
        self.labels = Param(self, 'labels', 'labels in index order')
        self._set(**{k: v for k, v in self._input_kwargs.items() if v is not None})

    def _transform(self, dataset):
        from .sql import Column
        labels = list(self.getOrDefault(self.labels))
        return dataset.withColumn(self.getOutputCol(), Column(self.getInputCol(), fn=lambda v: labels[int(v)]))


__all__ = ['VectorAssembler', 'MinMaxScaler', 'MinMaxScalerModel', 'OneHotEncoder', 'OneHotEncoderModel', 'Normalizer',
           'IndexToString', 'DoubleType', 'StringType']
