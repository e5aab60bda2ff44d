"""Minimal ``pyspark.sql`` stand-in: ``Row``, ``StructType`` & friends, ``DataFrame``, ``SparkSession``.

Only the surface Elephas' Spark ML stages and adapters use (``elephas/ml/adapter.py:11-47``,
``elephas/ml_model.py:191-256``): ``createDataFrame`` from an RDD of rows / LabeledPoints,
``df.rdd``, ``df.schema``, ``select`` / ``withColumnRenamed`` / ``withColumn``, temp views with the
one ``SELECT a AS x, b AS y FROM view`` statement shape Elephas issues, ``show`` / ``take`` /
``collect`` / ``count`` / ``first``.
"""
import copy
import re

import numpy as np

from .context import RDD, SparkContext
from .linalg import Vector
from .regression import LabeledPoint


# --------------------------------------------------------------------- types
class DataType:
    def __eq__(self, other):
        return type(self) is type(other) and self.__dict__ == other.__dict__

    def __hash__(self):
        return hash(type(self).__name__)

    def __repr__(self):
        return type(self).__name__ + "()"

    def simpleString(self):
        return type(self).__name__.replace('Type', '').lower()


class DoubleType(DataType):
    pass


class FloatType(DataType):
    pass


class IntegerType(DataType):
    pass


class LongType(DataType):
    pass


class StringType(DataType):
    pass


class BooleanType(DataType):
    pass


class VectorUDT(DataType):
    pass


class ArrayType(DataType):
    def __init__(self, elementType, containsNull=True):
        self.elementType = elementType
        self.containsNull = containsNull

    def __repr__(self):
        return "ArrayType(%r)" % (self.elementType,)


class StructField:
    def __init__(self, name, dataType, nullable=True, metadata=None):
        self.name, self.dataType, self.nullable, self.metadata = name, dataType, nullable, metadata or {}

    def __repr__(self):
        return "StructField(%r, %r, %r)" % (self.name, self.dataType, self.nullable)


class StructType(DataType):
    def __init__(self, fields=None):
        self.fields = list(fields or [])

    def add(self, field, data_type=None, nullable=True):
        if not isinstance(field, StructField):
            field = StructField(field, data_type, nullable)
        self.fields.append(field)
        return self

    @property
    def names(self):
        return [f.name for f in self.fields]

    def fieldNames(self):
        return self.names

    def __iter__(self):
        return iter(self.fields)

    def __len__(self):
        return len(self.fields)

    def __getitem__(self, key):
        if isinstance(key, str):
            return next(f for f in self.fields if f.name == key)
        return self.fields[key]

    def __repr__(self):
        return "StructType(%r)" % (self.fields,)


def _infer_type(value):
    if isinstance(value, bool):
        return BooleanType()
    if isinstance(value, (int, np.integer)):
        return LongType()
    if isinstance(value, (float, np.floating)):
        return DoubleType()
    if isinstance(value, str):
        return StringType()
    if isinstance(value, Vector):
        return VectorUDT()
    if isinstance(value, (list, tuple, np.ndarray)):
        return ArrayType(DoubleType())
    return StringType()


# ----------------------------------------------------------------------- Row
_ROW_CLASSES = {}


def _row_class(fields):
    """One ``Row`` subclass per schema (field names as a class attribute, like a namedtuple): building a row is then a
    bare ``tuple.__new__`` - no per-row field list, no instance dict. DataFrame transforms create one Row per record."""
    key = fields if isinstance(fields, tuple) else tuple(fields)
    cls = _ROW_CLASSES.get(key)
    if cls is None:
        cls = type('Row', (Row,), {'__fields__': list(key), '__slots__': ()})
        _ROW_CLASSES[key] = cls
    return cls


class Row(tuple):
    """Tuple with named fields: ``row.label``, ``row['label']``, ``row[0]``; ``row + tuple`` is a tuple."""

    def __new__(cls, *args, **kwargs):
        if args and kwargs:
            raise ValueError("Can not use both args and kwargs to create Row")
        if kwargs:
            row = tuple.__new__(cls, list(kwargs.values()))
            row.__fields__ = list(kwargs.keys())
            return row
        return tuple.__new__(cls, args)

    @classmethod
    def _make(cls, fields, values):
        return tuple.__new__(_row_class(fields), values)

    def __reduce__(self):
        fields = getattr(self, '__fields__', None)
        if fields is None:
            return (Row, tuple(self))
        return (Row._make, (list(fields), tuple(self)))

    def asDict(self):
        return dict(zip(self.__fields__, self))

    def __getitem__(self, item):
        if isinstance(item, str):
            try:
                return tuple.__getitem__(self, self.__fields__.index(item))
            except ValueError:
                raise KeyError(item)
        return tuple.__getitem__(self, item)

    def __getattr__(self, item):
        if item.startswith('__'):
            raise AttributeError(item)
        try:
            return tuple.__getitem__(self, self.__fields__.index(item))
        except (ValueError, AttributeError):
            raise AttributeError(item)

    def __repr__(self):
        if hasattr(self, '__fields__'):
            return "Row(%s)" % ", ".join("%s=%r" % kv for kv in zip(self.__fields__, self))
        return "<Row(%s)>" % ", ".join(repr(v) for v in self)


class Column:
    """A column reference with an optional alias, cast and row-wise function."""

    def __init__(self, name, fn=None, alias=None):
        self.name, self.fn, self._alias = name, fn, alias

    def alias(self, name):
        return Column(self.name, self.fn, name)

    def astype(self, data_type):
        caster = float if isinstance(data_type, (DoubleType, FloatType)) else (
            int if isinstance(data_type, (IntegerType, LongType)) else (lambda v: v))
        prev = self.fn
        return Column(self.name, (lambda v: caster(prev(v))) if prev else caster, self._alias)

    cast = astype

    @property
    def out_name(self):
        return self._alias or self.name

    # comparisons and arithmetic against literals (``df.filter(df.label > 0.5)``, ``col('x') * 2``)
    def _with(self, op):
        prev = self.fn
        return Column(self.name, (lambda v: op(prev(v))) if prev else op, self._alias)

    def __gt__(self, other):
        return self._with(lambda v: v > other)

    def __ge__(self, other):
        return self._with(lambda v: v >= other)

    def __lt__(self, other):
        return self._with(lambda v: v < other)

    def __le__(self, other):
        return self._with(lambda v: v <= other)

    def __eq__(self, other):  # noqa: D105 - builds a condition, like pyspark's Column
        return self._with(lambda v: v == other)

    def __ne__(self, other):
        return self._with(lambda v: v != other)

    __hash__ = object.__hash__

    def __add__(self, other):
        return self._with(lambda v: v + other)

    def __sub__(self, other):
        return self._with(lambda v: v - other)

    def __mul__(self, other):
        return self._with(lambda v: v * other)

    def __truediv__(self, other):
        return self._with(lambda v: v / other)

    def __invert__(self):
        return self._with(lambda v: not v)

    def isin(self, *values):
        pool = set(values[0]) if len(values) == 1 and isinstance(values[0], (list, tuple, set)) else set(values)
        return self._with(lambda v: v in pool)

    def isNull(self):
        return self._with(lambda v: v is None)

    def isNotNull(self):
        return self._with(lambda v: v is not None)


class functions:
    """Tiny ``pyspark.sql.functions``: ``col`` and ``array_argmax`` (the reference's tests spell the
    latter ``expr('array_position(c, array_max(c)) - 1')``, ``tests/test_ml_model.py:20-26``)."""

    @staticmethod
    def col(name):
        return Column(name)

    @staticmethod
    def array_argmax(name):
        return Column(name, fn=lambda v: int(np.argmax(np.asarray(v))))


# ----------------------------------------------------------------- DataFrame
class DataFrame:
    def __init__(self, rdd, schema, session, base=None, blocks=None):
        self._rdd = rdd  # None: rows are derived lazily from ``base``'s rows + the appended column blocks
        self.schema = schema
        self.sparkSession = session
        self.sql_ctx = session
        self._base, self._blocks = base, dict(blocks or {})
        self._block_cache = {}

    @property
    def rdd(self):
        if self._rdd is None:
            self._rdd = self._materialize_rows()
        return self._rdd

    # ----------------------------------------------------------------- column blocks
    # Pipeline stages that compute on whole columns (StandardScaler, ElephasTransformer) exchange dense matrices instead
    # of Python row objects: a stage appends its output as a *block* (one ndarray for all rows), the next stage reads
    # that block directly, and Row objects are only built if somebody asks for rows (collect / rdd / select). The
    # reference pays per row at every stage (np.asarray + Vectors.dense + Row per element, elephas/ml_model.py:222-250).
    def with_block_column(self, name, block, data_type):
        """New DataFrame = this one + column ``name`` whose values are the rows of ``block`` (``[rows]`` or
        ``[rows, width]``); ``data_type``: DoubleType (scalars), ArrayType(DoubleType()) (lists) or VectorUDT (vectors)."""
        block = np.asarray(block)
        if name in self.columns:
            raise ValueError("column %r already exists" % (name,))
        if block.shape[0] != self.count():
            raise ValueError("block has %d rows, the frame %d" % (block.shape[0], self.count()))
        schema = copy.deepcopy(self.schema)
        schema.add(StructField(name, data_type, True))
        return DataFrame(None, schema, self.sparkSession, base=self, blocks={name: block})

    def column_block(self, name):
        """The column as ONE float64 ndarray in row order (``[rows]`` for scalars, ``[rows, width]`` for vector / array
        columns). Free when the column was appended as a block; otherwise built from the rows once and cached."""
        if name in self._blocks:
            return self._blocks[name]
        if self._rdd is None and self._base is not None and name in self._base.columns:
            return self._base.column_block(name)
        cached = self._block_cache.get(name)
        if cached is None:
            i = self._col_index(name)
            get = tuple.__getitem__
            rows = self.rdd.collect()
            vals = [get(r, i) for r in rows] if rows and isinstance(rows[0], tuple) else [r[i] for r in rows]
            if vals and hasattr(vals[0], 'toArray'):
                width = len(vals[0].toArray())
                cached = np.empty((len(vals), width), dtype=np.float64)
                for k, v in enumerate(vals):
                    cached[k] = v.toArray()  # (row-wise copy into one matrix: no intermediate list of arrays)
            else:
                cached = np.asarray(vals, dtype=np.float64)
            self._block_cache[name] = cached
        return cached

    def column_factor(self, name):
        """``(uniques, inverse)`` of a scalar column (``np.unique(..., return_inverse=True)``), built from the rows once and
        cached: what a label indexer needs, without a Python-level lookup per row on later passes. None when the values do
        not sort (mixed types)."""
        key = ('factor', name)
        if key not in self._block_cache:
            if self._rdd is None and self._base is not None and name in self._base.columns and name not in self._blocks:
                return self._base.column_factor(name)
            i, get = self._col_index(name), tuple.__getitem__
            rows = self.rdd.collect()
            vals = [get(r, i) for r in rows] if rows and isinstance(rows[0], tuple) else [r[i] for r in rows]
            try:
                arr = np.asarray(vals)
                self._block_cache[key] = np.unique(arr, return_inverse=True) if arr.ndim == 1 and arr.dtype != object \
                    else None
            except TypeError:
                self._block_cache[key] = None
        return self._block_cache[key]

    def _materialize_rows(self):
        from .linalg import DenseVector
        base_rdd = self._base.rdd
        names = self.columns
        row_cls, new = _row_class(names), tuple.__new__
        cols = []
        for f in self.schema.fields[len(self._base.columns):]:
            block = self._blocks[f.name]
            if isinstance(f.dataType, VectorUDT):
                cols.append([DenseVector(v) for v in np.asarray(block, dtype=np.float64)])
            else:
                cols.append(np.asarray(block, dtype=np.float64).tolist())  # whole block -> Python floats at C speed
        parts, start = [], 0
        for p in base_rdd.partitions():
            rows = p if isinstance(p, list) else list(p)
            extra = [c[start:start + len(rows)] for c in cols]
            parts.append([new(row_cls, tuple(r) + tail) for r, tail in zip(rows, zip(*extra))] if rows else [])
            start += len(rows)
        return RDD(parts, base_rdd.context)

    @property
    def columns(self):
        return self.schema.names

    def count(self):
        if self._rdd is None and self._base is not None:
            return self._base.count()
        return self.rdd.count()

    def collect(self):
        return self.rdd.collect()

    def take(self, n):
        return self.rdd.take(n)

    def first(self):
        return self.rdd.first()

    def head(self, n=None):
        return self.first() if n is None else self.take(n)

    def _col_index(self, name):
        try:
            return self.columns.index(name)
        except ValueError:
            raise ValueError("cannot resolve column %r among %s" % (name, self.columns))

    def select(self, *cols):
        if len(cols) == 1 and isinstance(cols[0], (list, tuple)):
            cols = tuple(cols[0])
        cols = [Column(c) if isinstance(c, str) else c for c in cols]
        idx = [self._col_index(c.name) for c in cols]
        names = [c.out_name for c in cols]
        fns = [c.fn for c in cols]
        fields = []
        first = self.rdd.take(1)
        for c, i, fn in zip(cols, idx, fns):
            base = self.schema.fields[i]
            dtype = base.dataType if fn is None else _infer_type(fn(first[0][i])) if first else base.dataType
            fields.append(StructField(c.out_name, dtype, base.nullable))

        def project(row):
            return Row._make(names, [row[i] if fn is None else fn(row[i]) for i, fn in zip(idx, fns)])

        return DataFrame(self.rdd.map(project), StructType(fields), self.sparkSession)

    def withColumnRenamed(self, existing, new):
        names = [new if n == existing else n for n in self.columns]
        fields = [StructField(n, f.dataType, f.nullable) for n, f in zip(names, self.schema.fields)]
        return DataFrame(self.rdd.map(lambda r: Row._make(names, list(r))), StructType(fields), self.sparkSession)

    def withColumn(self, name, col):
        if not isinstance(col, Column):
            raise TypeError("withColumn expects a Column")
        src = self._col_index(col.name)
        fn = col.fn or (lambda v: v)
        names = list(self.columns)
        replace = names.index(name) if name in names else None
        if replace is None:
            names.append(name)
        first = self.rdd.take(1)
        dtype = _infer_type(fn(first[0][src])) if first else DoubleType()
        fields = list(self.schema.fields)
        if replace is None:
            fields.append(StructField(name, dtype))
        else:
            fields[replace] = StructField(name, dtype)

        row_cls, new, get = _row_class(names), tuple.__new__, tuple.__getitem__

        def extend(row):
            if replace is None:
                return new(row_cls, tuple(row) + (fn(get(row, src)),))
            vals = list(row)
            vals[replace] = fn(get(row, src))
            return new(row_cls, vals)

        return DataFrame(self.rdd.map(extend), StructType(fields), self.sparkSession)

    def drop(self, *cols):
        keep = [c for c in self.columns if c not in cols]
        return self.select(*keep)

    def __getitem__(self, name):
        self._col_index(name)
        return Column(name)

    def __getattr__(self, name):  # df.label, as pyspark allows
        if name.startswith('_') or name not in self.__dict__.get('schema', StructType([])).names:
            raise AttributeError(name)
        return Column(name)

    def filter(self, condition):
        """Rows for which ``condition`` holds: a Column compared with a literal (``df.x > 1``) or a function of the Row."""
        if isinstance(condition, Column):
            i, fn = self._col_index(condition.name), condition.fn or bool
            keep = lambda row: bool(fn(row[i]))  # noqa: E731
        elif callable(condition):
            keep = condition
        else:
            raise TypeError("filter expects a Column condition or a function of the row")
        return DataFrame(self.rdd.filter(keep), self.schema, self.sparkSession)

    where = filter

    def orderBy(self, *cols, ascending=True):
        cols = [c.name if isinstance(c, Column) else c for c in (cols[0] if len(cols) == 1 and isinstance(cols[0], (list, tuple)) else cols)]
        idx = [self._col_index(c) for c in cols]
        flags = list(ascending) if isinstance(ascending, (list, tuple)) else [ascending] * len(idx)
        rows = self.collect()
        for i, asc in reversed(list(zip(idx, flags))):   # stable sorts, last key first
            rows = sorted(rows, key=lambda r, i=i: r[i], reverse=not asc)
        return DataFrame(RDD([rows], self.rdd.context), self.schema, self.sparkSession)

    sort = orderBy

    def dropna(self, how='any', subset=None):
        idx = [self._col_index(c) for c in (subset or self.columns)]
        test = any if how == 'any' else all

        def missing(v):
            return v is None or (isinstance(v, float) and v != v)

        return self.filter(lambda row: not test(missing(row[i]) for i in idx))

    def groupBy(self, *cols):
        return GroupedData(self, [c.name if isinstance(c, Column) else c for c in cols])

    groupby = groupBy

    def distinct(self):
        seen, rows = set(), []
        for r in self.collect():
            key = tuple(repr(v) for v in r)
            if key not in seen:
                seen.add(key)
                rows.append(r)
        return DataFrame(RDD([rows], self.rdd.context), self.schema, self.sparkSession)

    def repartition(self, n):
        return DataFrame(self.rdd.repartition(n), self.schema, self.sparkSession)

    def limit(self, n):
        return DataFrame(RDD([self.take(int(n))], self.rdd.context), self.schema, self.sparkSession)

    def randomSplit(self, weights, seed=None):
        """Random split by normalised ``weights`` (every row lands in exactly one part)."""
        import random
        total = float(sum(weights))
        bounds, acc = [], 0.0
        for w in weights:
            acc += w / total
            bounds.append(acc)
        rng = random.Random(seed)
        parts = [[] for _ in weights]
        for row in self.collect():
            u = rng.random()
            for i, b in enumerate(bounds):
                if u < b or i == len(bounds) - 1:
                    parts[i].append(row)
                    break
        return [DataFrame(RDD([rows], self.rdd.context), self.schema, self.sparkSession) for rows in parts]

    def cache(self):
        return self

    def createOrReplaceTempView(self, name):
        self.sparkSession._views[name] = self

    registerTempTable = createOrReplaceTempView

    def show(self, n=20, truncate=True):
        rows = self.take(n)
        print(" | ".join(self.columns))
        for r in rows:
            cells = []
            for v in r:
                s = repr(v.toArray().tolist()) if isinstance(v, Vector) else repr(v)
                cells.append(s[:40] + '...' if truncate and len(s) > 43 else s)
            print(" | ".join(cells))
        if self.count() > n:
            print("only showing top %d rows" % n)

    def toPandas(self):
        import pandas as pd
        return pd.DataFrame([list(r) for r in self.collect()], columns=self.columns)

    def __repr__(self):
        return "DataFrame[%s]" % ", ".join("%s: %s" % (f.name, f.dataType.simpleString()) for f in self.schema.fields)


class DataFrameReader:
    """``spark.read``: CSV files (``header``, ``inferSchema``, ``sep``) and JSON-lines files into local DataFrames."""

    def __init__(self, session):
        self._session, self._format, self._options = session, 'csv', {}

    def format(self, source):  # noqa: A003 - pyspark's name
        if source not in ('csv', 'json', 'com.databricks.spark.csv'):
            raise NotImplementedError("the local reader knows 'csv' and 'json', not %r" % (source,))
        self._format = 'json' if source == 'json' else 'csv'
        return self

    def option(self, key, value):
        self._options[key] = value
        return self

    def options(self, **kwargs):
        self._options.update(kwargs)
        return self

    def load(self, path):
        return self.json(path) if self._format == 'json' else self.csv(path)

    @staticmethod
    def _flag(value):
        return str(value).lower() in ('true', '1', 'yes')

    def csv(self, path, header=None, inferSchema=None, sep=None, schema=None):
        import csv as _csv
        opts = self._options
        header = self._flag(opts.get('header', False) if header is None else header)
        infer = self._flag(opts.get('inferSchema', opts.get('inferschema', False)) if inferSchema is None else inferSchema)
        sep = sep or opts.get('sep') or opts.get('delimiter') or ','
        with open(path, newline='') as fh:
            records = [r for r in _csv.reader(fh, delimiter=sep) if r]
        names = records.pop(0) if header and records else ['_c%d' % i for i in range(len(records[0]) if records else 0)]
        if isinstance(schema, StructType):
            names, casts = schema.names, [_caster_of(f.dataType) for f in schema.fields]
        elif infer:
            casts = [_infer_text_caster([r[i] for r in records]) for i in range(len(names))]
        else:
            casts = [str] * len(names)
        rows = [tuple(None if v == '' and c is not str else c(v) for v, c in zip(r, casts)) for r in records]
        return self._session.createDataFrame(rows, schema if isinstance(schema, StructType) else names)

    def json(self, path):
        import json as _json
        with open(path) as fh:
            records = [_json.loads(line) for line in fh if line.strip()]
        names = sorted({k for r in records for k in r})
        return self._session.createDataFrame([tuple(r.get(k) for k in names) for r in records], names)


def _caster_of(data_type):
    if isinstance(data_type, (DoubleType, FloatType)):
        return float
    if isinstance(data_type, (IntegerType, LongType)):
        return int
    if isinstance(data_type, BooleanType):
        return lambda v: str(v).lower() == 'true'
    return str


def _infer_text_caster(values):
    values = [v for v in values if v != '']
    for cast in (int, float):
        try:
            for v in values:
                cast(v)
            return cast
        except ValueError:
            continue
    return str


class GroupedData:
    """``df.groupBy(cols)``: ``count()`` and the numeric aggregates ``avg / mean / sum / min / max`` of named columns."""

    def __init__(self, frame, keys):
        self.frame, self.keys = frame, keys
        self._idx = [frame._col_index(k) for k in keys]

    def _groups(self):
        groups = {}
        for row in self.frame.collect():
            groups.setdefault(tuple(row[i] for i in self._idx), []).append(row)
        return groups

    def _result(self, names, types, rows):
        fields = [self.frame.schema.fields[i] for i in self._idx] + [StructField(n, t) for n, t in zip(names, types)]
        all_names = self.keys + names
        data = [Row._make(all_names, list(r)) for r in rows]
        return DataFrame(RDD([data], self.frame.rdd.context), StructType(fields), self.frame.sparkSession)

    def count(self):
        return self._result(['count'], [LongType()], [key + (len(rows),) for key, rows in self._groups().items()])

    def _aggregate(self, label, fn, cols):
        idx = [self.frame._col_index(c) for c in cols]
        out = [key + tuple(float(fn([r[i] for r in rows])) for i in idx) for key, rows in self._groups().items()]
        return self._result(['%s(%s)' % (label, c) for c in cols], [DoubleType()] * len(cols), out)

    def avg(self, *cols):
        return self._aggregate('avg', lambda v: sum(v) / len(v), cols)

    mean = avg

    def sum(self, *cols):
        return self._aggregate('sum', sum, cols)

    def min(self, *cols):
        return self._aggregate('min', min, cols)

    def max(self, *cols):
        return self._aggregate('max', max, cols)


# -------------------------------------------------------------- SparkSession
_SELECT_RE = re.compile(r'^\s*select\s+(.*?)\s+from\s+(\w+)\s*$', re.IGNORECASE | re.DOTALL)


class SparkSession:
    _active = None

    class Builder:
        def __init__(self):
            self._master = None
            self._app = None

        def appName(self, name):
            self._app = name
            return self

        def master(self, master):
            self._master = master
            return self

        def config(self, *args, **kwargs):
            return self

        def getOrCreate(self):
            if SparkSession._active is None:
                if SparkContext._active is None:
                    SparkContext(master=self._master, appName=self._app)
                SparkSession._active = SparkSession(SparkContext._active)
            return SparkSession._active

    class _BuilderDescriptor:
        def __get__(self, obj, owner):
            return SparkSession.Builder()

    builder = _BuilderDescriptor()

    def __init__(self, sparkContext=None):
        self.sparkContext = sparkContext or SparkContext.getOrCreate()
        self._views = {}
        SparkSession._active = self

    @property
    def read(self):
        return DataFrameReader(self)

    def createDataFrame(self, data, schema=None):
        if type(data).__module__.startswith('pandas') and hasattr(data, 'itertuples'):   # a pandas.DataFrame
            schema = schema or [str(c) for c in data.columns]
            data = [tuple(v.item() if hasattr(v, 'item') else v for v in row) for row in data.itertuples(index=False, name=None)]
        rdd = data if isinstance(data, RDD) else self.sparkContext.parallelize(list(data))
        first = rdd.take(1)
        if first and isinstance(first[0], LabeledPoint):
            names = ['features', 'label']
            row_cls = _row_class(names)
            rdd = rdd.map(lambda lp: tuple.__new__(row_cls, (lp.features, lp.label)))
            schema = StructType([StructField('features', VectorUDT()), StructField('label', DoubleType())])
            return DataFrame(rdd, schema, self)
        if isinstance(schema, StructType):
            names = schema.names
            struct = copy.deepcopy(schema)
        else:
            if schema is None:
                if first and hasattr(first[0], '__fields__'):
                    names = list(first[0].__fields__)
                else:
                    names = ['_%d' % (i + 1) for i in range(len(first[0]))] if first else []
            else:
                names = list(schema)
            struct = StructType([StructField(n, _infer_type(first[0][i]) if first else StringType())
                                 for i, n in enumerate(names)])
        row_cls, new = _row_class(names), tuple.__new__
        rdd = rdd.map(lambda r: new(row_cls, r))
        return DataFrame(rdd, struct, self)

    def sql(self, query):
        m = _SELECT_RE.match(query)
        if not m:
            raise NotImplementedError("the local SparkSession only understands 'SELECT cols FROM view'")
        df = self._views[m.group(2)]
        cols = []
        for part in m.group(1).split(','):
            tokens = re.split(r'\s+as\s+', part.strip(), flags=re.IGNORECASE)
            col = Column(tokens[0].strip())
            cols.append(col.alias(tokens[1].strip()) if len(tokens) > 1 else col)
        return df.select(*cols)

    def stop(self):
        SparkSession._active = None
        self.sparkContext.stop()
