"""``pyspark.ml`` stand-in: ``Param`` / ``Params`` / ``keyword_only``, ``Estimator`` / ``Transformer`` /
``Model``, ``Pipeline`` / ``PipelineModel`` with save/load, the shared column mixins, and the two
feature stages the reference's Otto example uses (``StringIndexer``, ``StandardScaler``,
``examples/ml_pipeline_otto.py:49-50``).

Semantics that the Elephas stages rely on and that are reproduced here (``elephas/ml/params.py``,
``elephas/ml_model.py:37-42,72-76``): params are class-level descriptors copied per instance,
``_setDefault`` / ``_set`` / ``getOrDefault`` / ``isSet`` / ``isDefined``, ``_paramMap`` and
``_defaultParamMap`` keyed by ``Param``, ``@keyword_only`` storing ``_input_kwargs``.
"""
import copy
import functools
import importlib
import json
import os
import shutil
import uuid

import numpy as np

from .linalg import DenseVector, Vector
from .sql import Column, DoubleType, Row, StructField, StructType, VectorUDT


def keyword_only(func):
    """Force keyword arguments and stash them in ``self._input_kwargs`` (as ``pyspark.keyword_only``)."""

    @functools.wraps(func)
    def wrapper(self, *args, **kwargs):
        if args:
            raise TypeError("Method %s forces keyword arguments." % func.__name__)
        self._input_kwargs = kwargs
        return func(self, **kwargs)

    return wrapper


class Param:
    def __init__(self, parent, name, doc, typeConverter=None):
        self.parent = parent.uid if hasattr(parent, 'uid') else parent
        self.name = str(name)
        self.doc = str(doc)
        self.typeConverter = typeConverter or (lambda v: v)

    def __hash__(self):
        return hash((self.parent, self.name))

    def __eq__(self, other):
        return isinstance(other, Param) and self.parent == other.parent and self.name == other.name

    def __repr__(self):
        return "Param(parent=%r, name=%r, doc=%r)" % (self.parent, self.name, self.doc)


class Identifiable:
    def __init__(self):
        if not hasattr(self, 'uid'):
            self.uid = "%s_%s" % (type(self).__name__, uuid.uuid4().hex[-12:])


class Params(Identifiable):
    """Cooperative base: every mixin calls ``super().__init__()`` and then declares its ``Param``."""

    def __init__(self):
        super().__init__()
        if not hasattr(self, '_paramMap'):
            self._paramMap = {}
            self._defaultParamMap = {}

    @property
    def params(self):
        return sorted((v for v in self.__dict__.values() if isinstance(v, Param)), key=lambda p: p.name)

    def hasParam(self, name):
        return isinstance(self.__dict__.get(name), Param)

    def getParam(self, name):
        p = self.__dict__.get(name)
        if not isinstance(p, Param):
            raise ValueError("Cannot find param with name %s." % name)
        return p

    def _resolve(self, param):
        return self.getParam(param) if isinstance(param, str) else param

    def isSet(self, param):
        return self._resolve(param) in self._paramMap

    def hasDefault(self, param):
        return self._resolve(param) in self._defaultParamMap

    def isDefined(self, param):
        return self.isSet(param) or self.hasDefault(param)

    def getOrDefault(self, param):
        param = self._resolve(param)
        if param in self._paramMap:
            return self._paramMap[param]
        if param in self._defaultParamMap:
            return self._defaultParamMap[param]
        raise KeyError("Param %s is neither set nor has a default" % param.name)

    def getDefault(self, param):
        return self._defaultParamMap[self._resolve(param)]

    def set(self, param, value):
        self._paramMap[self._resolve(param)] = value
        return self

    def clear(self, param):
        self._paramMap.pop(self._resolve(param), None)

    def _set(self, **kwargs):
        for name, value in kwargs.items():
            if value is None:
                continue
            self._paramMap[self.getParam(name)] = value
        return self

    def _setDefault(self, **kwargs):
        for name, value in kwargs.items():
            self._defaultParamMap[self.getParam(name)] = value
        return self

    def extractParamMap(self, extra=None):
        merged = dict(self._defaultParamMap)
        merged.update(self._paramMap)
        if extra:
            merged.update(extra)
        return merged

    def explainParams(self):
        lines = []
        for p in self.params:
            state = []
            if self.hasDefault(p):
                state.append("default: %r" % (self.getDefault(p),))
            if self.isSet(p):
                state.append("current: %r" % (self._paramMap[p],))
            lines.append("%s: %s (%s)" % (p.name, p.doc, ", ".join(state) or "undefined"))
        return "\n".join(lines)

    def copy(self, extra=None):
        that = copy.copy(self)
        that._paramMap = dict(self._paramMap)
        that._defaultParamMap = dict(self._defaultParamMap)
        if extra:
            for p, v in extra.items():
                that._paramMap[that._resolve(p.name if isinstance(p, Param) else p)] = v
        return that


class HasFeaturesCol(Params):
    def __init__(self):
        super().__init__()
        self.featuresCol = Param(self, "featuresCol", "features column name.")
        self._setDefault(featuresCol='features')

    def getFeaturesCol(self):
        return self.getOrDefault(self.featuresCol)


class HasLabelCol(Params):
    def __init__(self):
        super().__init__()
        self.labelCol = Param(self, "labelCol", "label column name.")
        self._setDefault(labelCol='label')

    def getLabelCol(self):
        return self.getOrDefault(self.labelCol)


class HasOutputCol(Params):
    def __init__(self):
        super().__init__()
        self.outputCol = Param(self, "outputCol", "output column name.")
        self._setDefault(outputCol=self.uid + '__output')

    def getOutputCol(self):
        return self.getOrDefault(self.outputCol)


class HasInputCol(Params):
    def __init__(self):
        super().__init__()
        self.inputCol = Param(self, "inputCol", "input column name.")

    def getInputCol(self):
        return self.getOrDefault(self.inputCol)


# ------------------------------------------------------------- persistence
def _jsonable(value):
    try:
        json.dumps(value)
        return True
    except (TypeError, ValueError):
        return False


class MLWriter:
    def __init__(self, instance):
        self.instance = instance
        self._overwrite = False

    def overwrite(self):
        self._overwrite = True
        return self

    def save(self, path):
        if os.path.exists(path):
            if not self._overwrite:
                raise IOError("Path %s already exists. To overwrite it, use write().overwrite().save(path)" % path)
            shutil.rmtree(path)
        self.instance._save_impl(path)


class DefaultParamsWritable:
    """Persists class name, uid and every JSON-serialisable param under ``path/metadata``."""

    def write(self):
        return MLWriter(self)

    def save(self, path):
        self.write().save(path)

    def _metadata(self):
        params = {p.name: v for p, v in self._paramMap.items() if _jsonable(v)}
        defaults = {p.name: v for p, v in self._defaultParamMap.items() if _jsonable(v)}
        return {'class': type(self).__module__ + '.' + type(self).__name__, 'uid': self.uid, 'paramMap': params,
                'defaultParamMap': defaults}

    def _save_impl(self, path):
        os.makedirs(os.path.join(path, 'metadata'))
        with open(os.path.join(path, 'metadata', 'part-00000'), 'w') as fh:
            json.dump(self._metadata(), fh)


class DefaultParamsReadable:
    @classmethod
    def load(cls, path):
        with open(os.path.join(path, 'metadata', 'part-00000')) as fh:
            meta = json.load(fh)
        return _instantiate(meta, path)

    @classmethod
    def read(cls):
        return cls


def _instantiate(meta, path):
    module, _, name = meta['class'].rpartition('.')
    klass = getattr(importlib.import_module(module), name)
    if hasattr(klass, '_load_impl'):
        return klass._load_impl(meta, path)
    inst = klass()
    inst.uid = meta['uid']
    for k, v in meta.get('paramMap', {}).items():
        if inst.hasParam(k):
            inst._paramMap[inst.getParam(k)] = v
    return inst


# ------------------------------------------------------------ stage classes
class PipelineStage(Params):
    pass


class Transformer(PipelineStage):
    def transform(self, dataset, params=None):
        if params:
            return self.copy(params)._transform(dataset)
        return self._transform(dataset)

    def _transform(self, dataset):  # pragma: no cover - abstract
        raise NotImplementedError


class Estimator(PipelineStage):
    def fit(self, dataset, params=None):
        if params:
            return self.copy(params)._fit(dataset)
        return self._fit(dataset)

    def _fit(self, dataset):  # pragma: no cover - abstract
        raise NotImplementedError


class Model(Transformer):
    pass


class Pipeline(Estimator, DefaultParamsReadable, DefaultParamsWritable):
    @keyword_only
    def __init__(self, stages=None):
        super().__init__()
        self.stages = Param(self, "stages", "a list of pipeline stages")
        self._set(**self._input_kwargs)

    def setStages(self, value):
        return self._set(stages=value)

    def getStages(self):
        return self.getOrDefault(self.stages)

    def _fit(self, dataset):
        stages = self.getStages()
        last_estimator = max((i for i, s in enumerate(stages) if isinstance(s, Estimator)), default=-1)
        fitted = []
        for i, stage in enumerate(stages):
            if isinstance(stage, Estimator):
                model = stage.fit(dataset)
                fitted.append(model)
                if i < last_estimator:
                    dataset = model.transform(dataset)
            elif isinstance(stage, Transformer):
                fitted.append(stage)
                if i < last_estimator:
                    dataset = stage.transform(dataset)
            else:
                raise TypeError("Cannot recognize a pipeline stage of type %s." % type(stage))
        return PipelineModel(fitted)

    def _save_impl(self, path):
        os.makedirs(os.path.join(path, 'metadata'))
        stages = self.getStages() if self.isDefined(self.stages) else []
        meta = {'class': type(self).__module__ + '.' + type(self).__name__, 'uid': self.uid,
                'paramMap': {'stageUids': [s.uid for s in stages]}}
        with open(os.path.join(path, 'metadata', 'part-00000'), 'w') as fh:
            json.dump(meta, fh)
        for i, s in enumerate(stages):
            s.write().save(os.path.join(path, 'stages', '%d_%s' % (i, s.uid)))

    @classmethod
    def _load_impl(cls, meta, path):
        stage_dir = os.path.join(path, 'stages')
        stages = []
        if os.path.isdir(stage_dir):
            for d in sorted(os.listdir(stage_dir), key=lambda s: int(s.split('_')[0])):
                stages.append(DefaultParamsReadable.load(os.path.join(stage_dir, d)))
        inst = cls(stages=stages)
        inst.uid = meta['uid']
        return inst


class PipelineModel(Model, DefaultParamsReadable, DefaultParamsWritable):
    def __init__(self, stages=None):
        super().__init__()
        self.stages = list(stages or [])

    def _transform(self, dataset):
        for t in self.stages:
            dataset = t.transform(dataset)
        return dataset

    def _save_impl(self, path):
        os.makedirs(os.path.join(path, 'metadata'))
        meta = {'class': type(self).__module__ + '.' + type(self).__name__, 'uid': self.uid,
                'paramMap': {'stageUids': [s.uid for s in self.stages]}}
        with open(os.path.join(path, 'metadata', 'part-00000'), 'w') as fh:
            json.dump(meta, fh)
        for i, s in enumerate(self.stages):
            s.write().save(os.path.join(path, 'stages', '%d_%s' % (i, s.uid)))

    @classmethod
    def _load_impl(cls, meta, path):
        stage_dir = os.path.join(path, 'stages')
        stages = []
        if os.path.isdir(stage_dir):
            for d in sorted(os.listdir(stage_dir), key=lambda s: int(s.split('_')[0])):
                stages.append(DefaultParamsReadable.load(os.path.join(stage_dir, d)))
        inst = cls(stages)
        inst.uid = meta['uid']
        return inst


# ---------------------------------------------------------- feature stages
class _WithData:
    """Fitted stages keep their state in ``path/data/part-00000`` (JSON) next to Spark-style ``metadata``."""

    _state = ()

    def _save_impl(self, path):
        super()._save_impl(path)
        os.makedirs(os.path.join(path, 'data'))
        state = {k: (getattr(self, k).tolist() if isinstance(getattr(self, k), np.ndarray) else getattr(self, k))
                 for k in self._state}
        with open(os.path.join(path, 'data', 'part-00000'), 'w') as fh:
            json.dump(state, fh)

    @classmethod
    def _load_impl(cls, meta, path):
        inst = cls()
        inst.uid = meta['uid']
        for k, v in meta.get('paramMap', {}).items():
            if inst.hasParam(k):
                inst._paramMap[inst.getParam(k)] = v
        data = os.path.join(path, 'data', 'part-00000')
        if not os.path.exists(data):
            raise IOError("%s holds no fitted state (%s is missing): it was not written by this version" % (path, data))
        with open(data) as fh:
            state = json.load(fh)
        for k in cls._state:
            v = state.get(k)
            setattr(inst, k, np.asarray(v, dtype=np.float64) if isinstance(v, list) and k not in cls._lists else v)
        return inst

    _lists = ()


class StringIndexerModel(_WithData, Model, HasInputCol, HasOutputCol, DefaultParamsReadable, DefaultParamsWritable):
    _state = _lists = ('labels',)

    def __init__(self, labels=None, inputCol=None, outputCol=None):
        super().__init__()
        self.labels = list(labels or [])
        self._set(inputCol=inputCol, outputCol=outputCol)

    def _transform(self, dataset):
        index = {l: float(i) for i, l in enumerate(self.labels)}
        if hasattr(dataset, 'with_block_column') and self.getOutputCol() not in dataset.columns:
            # one pass over the label column, the indices appended as a block (rows are only rebuilt on demand)
            from .sql import DoubleType as _DoubleType
            factor = dataset.column_factor(self.getInputCol()) if hasattr(dataset, 'column_factor') else None
            if factor is not None:
                # the column factorised once per frame (cached there): one table lookup per DISTINCT label, then a gather
                uniques, inverse = factor
                table = np.asarray([index[str(u)] for u in uniques.tolist()], dtype=np.float64)
                return dataset.with_block_column(self.getOutputCol(), table[inverse], _DoubleType())
            i, get = dataset.columns.index(self.getInputCol()), tuple.__getitem__
            rows = dataset.rdd.collect()
            if rows and isinstance(rows[0], tuple):
                block = np.fromiter((index[str(get(r, i))] for r in rows), dtype=np.float64, count=len(rows))
                return dataset.with_block_column(self.getOutputCol(), block, _DoubleType())
        return dataset.withColumn(self.getOutputCol(), Column(self.getInputCol(), fn=lambda v: index[str(v)]))


class StringIndexer(Estimator, HasInputCol, HasOutputCol, DefaultParamsReadable, DefaultParamsWritable):
    """Labels ordered by descending frequency (ties alphabetical), as Spark's default ``frequencyDesc``."""

    @keyword_only
    def __init__(self, inputCol=None, outputCol=None):
        super().__init__()
        self._set(**self._input_kwargs)

    def _fit(self, dataset):
        i = dataset.columns.index(self.getInputCol())
        counts = {}
        factor = dataset.column_factor(self.getInputCol()) if hasattr(dataset, 'column_factor') else None
        if factor is not None:
            for u, c in zip(factor[0].tolist(), np.bincount(factor[1], minlength=len(factor[0])).tolist()):
                counts[str(u)] = counts.get(str(u), 0) + c
        else:
            for r in dataset.collect():
                counts[str(r[i])] = counts.get(str(r[i]), 0) + 1
        labels = [k for k, _ in sorted(counts.items(), key=lambda kv: (-kv[1], kv[0]))]
        return StringIndexerModel(labels, self.getInputCol(), self.getOutputCol())


class StandardScalerModel(_WithData, Model, HasInputCol, HasOutputCol, DefaultParamsReadable, DefaultParamsWritable):
    _state = ('mean', 'std', 'withMean', 'withStd')

    def __init__(self, mean=None, std=None, withMean=False, withStd=True, inputCol=None, outputCol=None):
        super().__init__()
        self.mean, self.std, self.withMean, self.withStd = mean, std, withMean, withStd
        self._set(inputCol=inputCol, outputCol=outputCol)

    def _transform(self, dataset):
        mean = self.mean if self.withMean else 0.0
        scale = np.where(self.std > 0, 1.0 / np.where(self.std > 0, self.std, 1.0), 0.0) if self.withStd else 1.0

        def fn(v):
            arr = v.toArray() if isinstance(v, Vector) else np.asarray(v, dtype=np.float64)
            return DenseVector((arr - mean) * scale)

        if hasattr(dataset, 'column_block') and self.getOutputCol() not in dataset.columns:
            try:
                block = dataset.column_block(self.getInputCol())
            except (TypeError, ValueError):
                block = None
            if block is not None and block.ndim == 2:
                # the whole column as one matrix in, one matrix out: later stages read the block, rows are built lazily
                from .sql import VectorUDT as _VectorUDT
                out = np.subtract(block, mean, dtype=np.float64)  # one temporary, scaled in place
                out *= scale
                return dataset.with_block_column(self.getOutputCol(), out, _VectorUDT())
        blocked = self._transform_blocks(dataset, mean, scale)
        if blocked is not None:
            return blocked
        return dataset.withColumn(self.getOutputCol(), Column(self.getInputCol(), fn=fn))

    def _transform_blocks(self, dataset, mean, scale):
        """Local DataFrames: scale a whole partition as one matrix (one NumPy expression per partition instead of three
        per row); the rows then share views of that matrix. None: not applicable, use the row path."""
        from .sql import DataFrame, _row_class
        from .context import RDD
        rdd = dataset.rdd
        out_col, in_col = self.getOutputCol(), self.getInputCol()
        if not (hasattr(rdd, 'partitions') and callable(rdd.partitions)) or out_col in dataset.columns:
            return None
        src = dataset.columns.index(in_col)
        names = list(dataset.columns) + [out_col]
        row_cls, new, get = _row_class(names), tuple.__new__, tuple.__getitem__
        parts = []
        for p in rdd.partitions():
            rows = p if isinstance(p, list) else list(p)
            if not rows:
                parts.append([])
                continue
            if not isinstance(rows[0], tuple):
                return None
            try:
                block = np.asarray([get(r, src).toArray() if isinstance(get(r, src), Vector) else
                                    np.asarray(get(r, src), dtype=np.float64).ravel() for r in rows], dtype=np.float64)
            except (TypeError, ValueError):
                return None
            if block.ndim != 2:
                return None
            block = (block - mean) * scale
            wrap = DenseVector._wrap
            parts.append([new(row_cls, tuple(r) + (wrap(block[i]),)) for i, r in enumerate(rows)])
        fields = list(dataset.schema.fields) + [StructField(out_col, VectorUDT())]
        return DataFrame(RDD(parts, rdd.context), StructType(fields), dataset.sparkSession)


class StandardScaler(Estimator, HasInputCol, HasOutputCol, DefaultParamsReadable, DefaultParamsWritable):
    @keyword_only
    def __init__(self, withMean=False, withStd=True, inputCol=None, outputCol=None):
        super().__init__()
        self.withMean, self.withStd = withMean, withStd
        self._set(inputCol=inputCol, outputCol=outputCol)

    def _fit(self, dataset):
        i = dataset.columns.index(self.getInputCol())
        rows = np.stack([r[i].toArray() if isinstance(r[i], Vector) else np.asarray(r[i], dtype=np.float64)
                         for r in dataset.collect()])
        std = rows.std(axis=0, ddof=1) if rows.shape[0] > 1 else np.zeros(rows.shape[1])
        return StandardScalerModel(rows.mean(axis=0), std, self.withMean, self.withStd, self.getInputCol(),
                                   self.getOutputCol())


__all__ = ['keyword_only', 'Param', 'Params', 'HasFeaturesCol', 'HasLabelCol', 'HasOutputCol', 'HasInputCol',
           'DefaultParamsReadable', 'DefaultParamsWritable', 'Estimator', 'Transformer', 'Model', 'Pipeline',
           'PipelineModel', 'StringIndexer', 'StandardScaler', 'Row', 'StructField', 'StructType', 'DoubleType',
           'VectorUDT']
