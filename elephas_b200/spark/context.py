"""Local stand-ins for ``SparkConf`` / ``SparkContext`` / ``RDD`` / ``Broadcast``.

Elephas uses Spark for three things only: shipping data partitions to workers
(``parallelize`` / ``repartition`` / ``mapPartitions`` / ``collect``), shipping weights
(``broadcast``), and bookkeeping transforms (``map``, ``zipWithIndex``, ``sortBy``, ``zip``,
``reduce``). On a single multi-GPU node none of that needs a JVM: an RDD here is an ordered list of
in-memory partitions, evaluated eagerly, and ``repartition`` is an order-preserving contiguous
split (so the ``zipWithIndex``/``sortBy`` dance of ``elephas/spark_model.py:257-266`` is not
needed to keep predictions aligned, although both methods exist).

Array fast path: ``to_simple_rdd`` (``elephas_b200/utils/rdd_utils.py``) stores each partition as a
:class:`PairBlock` of two dense arrays instead of a Python list of row tuples; the workers read the
arrays directly (no per-row Python work, ready for pinned-memory H2D copies) while generic
``map`` / ``collect`` callers still see ``(features, label)`` rows.
"""
import itertools
import os
from functools import reduce as _reduce

import numpy as np


class PairBlock:
    """A partition of ``(x_i, y_i)`` rows held as two arrays with equal leading dimension."""

    def __init__(self, x, y, x16=None):
        if len(x) != len(y):
            raise ValueError("features and labels differ in length: %d vs %d" % (len(x), len(y)))
        self.x, self.y = x, y
        # optional: the same features already cast to bf16 in pinned host memory ([rows, round_up(cols, 8)] torch tensor,
        # see rdd_utils.to_sharded_rdd(feature_dtype='bfloat16')): bf16 models stream these - half the PCIe bytes
        self.x16 = x16

    def __len__(self):
        return len(self.x)

    def __iter__(self):
        return zip(self.x, self.y)

    def __getitem__(self, item):
        if isinstance(item, slice):
            return PairBlock(self.x[item], self.y[item], self.x16[item] if self.x16 is not None else None)
        return (self.x[item], self.y[item])


class ArrayBlock:
    """A partition of rows held as one dense array."""

    def __init__(self, x):
        self.x = x

    def __len__(self):
        return len(self.x)

    def __iter__(self):
        return iter(self.x)

    def __getitem__(self, item):
        return ArrayBlock(self.x[item]) if isinstance(item, slice) else self.x[item]


def _split_bounds(n, parts):
    base, extra = divmod(n, parts)
    bounds, start = [], 0
    for i in range(parts):
        end = start + base + (1 if i < extra else 0)
        bounds.append((start, end))
        start = end
    return bounds


def _concat_blocks(partitions):
    """Merge partitions into one; keeps the array fast path when every partition is a block."""
    partitions = [p for p in partitions if len(p)]
    if len(partitions) == 1 and isinstance(partitions[0], (PairBlock, ArrayBlock)):
        return partitions[0]  # nothing to merge: the new partitions are views of the one block (no copy)
    if partitions and all(isinstance(p, PairBlock) for p in partitions):
        return PairBlock(np.concatenate([np.asarray(p.x) for p in partitions]),
                         np.concatenate([np.asarray(p.y) for p in partitions]))
    if partitions and all(isinstance(p, ArrayBlock) for p in partitions):
        return ArrayBlock(np.concatenate([np.asarray(p.x) for p in partitions]))
    return [row for p in partitions for row in p]


class Broadcast:
    def __init__(self, value):
        self._value = value

    @property
    def value(self):
        return self._value

    def unpersist(self, blocking=False):
        pass

    def destroy(self, blocking=False):
        self._value = None


class RDD:
    def __init__(self, partitions, context):
        self._partitions = list(partitions)
        self.context = context
        self.ctx = context

    # ----------------------------------------------------------- structure
    def getNumPartitions(self):
        return len(self._partitions)

    def partitions(self):
        """The raw partitions (lists or blocks) - the hook the elephas_b200 workers use."""
        return list(self._partitions)

    def glom(self):
        return RDD([[list(p)] for p in self._partitions], self.context)

    def repartition(self, numPartitions):
        numPartitions = max(1, int(numPartitions))
        if numPartitions == len(self._partitions) and getattr(self, '_pinned_partitioning', False):
            return self  # rank-sharded RDDs keep their one-partition-per-rank layout
        merged = _concat_blocks(self._partitions)
        return RDD([merged[a:b] for a, b in _split_bounds(len(merged), numPartitions)], self.context)

    coalesce = repartition

    def cache(self):
        return self

    persist = cache

    def unpersist(self, blocking=False):
        return self

    # ---------------------------------------------------------- transforms
    def mapPartitions(self, f, preservesPartitioning=False):
        out = []
        for p in self._partitions:
            res = f(iter(p))
            out.append(list(res) if res is not None else [])
        return RDD(out, self.context)

    def mapPartitionsWithIndex(self, f, preservesPartitioning=False):
        return RDD([list(f(i, iter(p)) or []) for i, p in enumerate(self._partitions)], self.context)

    def map(self, f):
        return RDD([[f(r) for r in p] for p in self._partitions], self.context)

    def flatMap(self, f):
        return RDD([[o for r in p for o in f(r)] for p in self._partitions], self.context)

    def filter(self, f):
        return RDD([[r for r in p if f(r)] for p in self._partitions], self.context)

    def zipWithIndex(self):
        out, start = [], 0
        for p in self._partitions:
            out.append([(r, start + i) for i, r in enumerate(p)])
            start += len(p)
        return RDD(out, self.context)

    def zip(self, other):
        if self.getNumPartitions() != other.getNumPartitions():
            raise ValueError("Can only zip RDDs with the same number of partitions")
        out = []
        for a, b in zip(self._partitions, other._partitions):
            a, b = list(a), list(b)
            if len(a) != len(b):
                raise ValueError("Can only zip RDDs with the same number of elements in each partition")
            out.append(list(zip(a, b)))
        return RDD(out, self.context)

    def sortBy(self, keyfunc, ascending=True, numPartitions=None):
        rows = sorted(self.collect(), key=keyfunc, reverse=not ascending)
        n = numPartitions or self.getNumPartitions()
        return RDD([rows[a:b] for a, b in _split_bounds(len(rows), n)], self.context)

    def union(self, other):
        return RDD(self._partitions + other._partitions, self.context)

    def distinct(self):
        seen, rows = set(), []
        for r in self.collect():
            if r not in seen:
                seen.add(r)
                rows.append(r)
        return RDD([rows], self.context)

    def keys(self):
        return self.map(lambda kv: kv[0])

    def values(self):
        return self.map(lambda kv: kv[1])

    def sample(self, withReplacement, fraction, seed=None):
        rng = np.random.default_rng(seed)
        return RDD([[r for r in p if rng.random() < fraction] for p in self._partitions], self.context)

    # ------------------------------------------------------------- actions
    def collect(self):
        return [r for p in self._partitions for r in p]

    def count(self):
        return sum(len(p) for p in self._partitions)

    def first(self):
        for p in self._partitions:
            for r in p:
                return r
        raise ValueError("RDD is empty")

    def take(self, n):
        return list(itertools.islice((r for p in self._partitions for r in p), n))

    def isEmpty(self):
        return self.count() == 0

    def reduce(self, f):
        rows = self.collect()
        if not rows:
            raise ValueError("Can not reduce() empty RDD")
        return _reduce(f, rows)

    def foreach(self, f):
        for r in self.collect():
            f(r)

    def max(self, key=None):
        return max(self.collect(), key=key) if key else max(self.collect())

    def min(self, key=None):
        return min(self.collect(), key=key) if key else min(self.collect())

    def sum(self):
        return sum(self.collect())

    def mean(self):
        rows = self.collect()
        return sum(rows) / len(rows)

    def toDF(self, schema=None):
        from .sql import SparkSession
        return SparkSession.builder.getOrCreate().createDataFrame(self, schema)


class SparkConf:
    def __init__(self, loadDefaults=True):
        self._conf = {}

    def setAppName(self, name):
        self._conf['spark.app.name'] = name
        return self

    def setMaster(self, master):
        self._conf['spark.master'] = master
        return self

    def set(self, key, value):
        self._conf[key] = value
        return self

    def get(self, key, default=None):
        return self._conf.get(key, default)


def _parallelism_from_master(master):
    if master and master.startswith('local['):
        inner = master[6:-1]
        if inner == '*':
            return os.cpu_count() or 1
        try:
            return max(1, int(inner.split(',')[0]))
        except ValueError:
            return 1
    return 1


class SparkContext:
    """One per process (like Spark); ``SparkContext.getOrCreate()`` returns the active one."""
    _active = None

    def __init__(self, master=None, appName=None, conf=None):
        conf = conf or SparkConf()
        if master:
            conf.setMaster(master)
        if appName:
            conf.setAppName(appName)
        self._conf = conf
        self.master = conf.get('spark.master', 'local[1]')
        self.appName = conf.get('spark.app.name', 'elephas_b200')
        self.defaultParallelism = _parallelism_from_master(self.master)
        SparkContext._active = self

    @classmethod
    def getOrCreate(cls, conf=None):
        if cls._active is None:
            cls(conf=conf)
        return cls._active

    def getConf(self):
        return self._conf

    def parallelize(self, c, numSlices=None):
        n_parts = max(1, int(numSlices or self.defaultParallelism))
        if isinstance(c, RDD):
            return c.repartition(n_parts)
        if isinstance(c, (PairBlock, ArrayBlock)):
            data = c
        elif isinstance(c, np.ndarray):
            data = ArrayBlock(c)
        else:
            data = list(c)
        return RDD([data[a:b] for a, b in _split_bounds(len(data), n_parts)], self)

    def broadcast(self, value):
        return Broadcast(value)

    def textFile(self, name, minPartitions=None):
        with open(name) as fh:
            lines = [ln.rstrip('\n') for ln in fh]
        return self.parallelize(lines, minPartitions)

    def stop(self):
        if SparkContext._active is self:
            SparkContext._active = None

    def setLogLevel(self, level):
        pass
