"""NumPy <-> DataFrame <-> RDD conversions for the Spark ML stages.

Reference: ``elephas/ml/adapter.py:11-47`` (``to_data_frame``, ``from_data_frame``,
``df_to_simple_rdd``). ``df_to_simple_rdd`` selects the two columns directly instead of going
through a temp view and SQL, so column names never have to be SQL-safe.
"""
from typing import Optional

import numpy as np

from ..spark import LabeledPoint, MLVector, SparkSession, Vector, Vectors
from ..utils.rdd_utils import from_labeled_point, lp_to_simple_rdd, to_labeled_point


def to_data_frame(sc, features: np.ndarray, labels: np.ndarray, categorical: bool = False):
    """Features and labels -> DataFrame with columns ``features`` (vector) and ``label`` (double)."""
    lp_rdd = to_labeled_point(sc, features, labels, categorical)
    return SparkSession.builder.getOrCreate().createDataFrame(lp_rdd)


def from_data_frame(df, categorical: bool = False, nb_classes: Optional[int] = None):
    """DataFrame -> ``(features, labels)`` arrays."""
    lp_rdd = df.rdd.map(lambda row: LabeledPoint(row.label, row.features))
    return from_labeled_point(lp_rdd, categorical, nb_classes)


def _as_mllib_vector(v):
    if isinstance(v, Vector):
        return v
    if MLVector is not Vector and isinstance(v, MLVector):
        return Vectors.fromML(v)
    return Vectors.dense(np.asarray(v, dtype=np.float64).ravel())


def df_to_simple_rdd(df, categorical: bool = False, nb_classes: Optional[int] = None,
                     features_col: str = 'features', label_col: str = 'label'):
    """DataFrame -> RDD of ``(features, label)`` pairs (labels one-hot when ``categorical``)."""
    rdd = df.rdd
    if hasattr(rdd, 'partitions') and callable(rdd.partitions):
        blocks = _columns_to_blocks(df, rdd, features_col, label_col, categorical, nb_classes)
        if blocks is not None:
            return blocks
    selected = df.select(features_col, label_col)
    lp_rdd = selected.rdd.map(lambda row: LabeledPoint(row[1], _as_mllib_vector(row[0])))
    return lp_to_simple_rdd(lp_rdd, categorical, nb_classes)


def _columns_to_blocks(df, rdd, features_col, label_col, categorical, nb_classes):
    """Local DataFrames: read the two columns of every partition straight into one feature matrix and one label
    array (a :class:`PairBlock`, the layout the workers consume) instead of building a ``LabeledPoint`` and a
    ``(features, one-hot)`` tuple per row. Returns None for anything it does not recognise (the row path handles it)."""
    from ..spark.context import RDD, PairBlock
    try:
        fi, li = df.columns.index(features_col), df.columns.index(label_col)
    except ValueError:
        return None
    get = tuple.__getitem__
    parts, label_parts = [], []
    for p in rdd.partitions():
        rows = p if isinstance(p, list) else list(p)
        if rows and not isinstance(rows[0], tuple):
            return None
        feats = []
        for r in rows:
            v = get(r, fi)
            feats.append(v.toArray() if isinstance(v, Vector) else np.asarray(v, dtype=np.float64).ravel())
        try:
            x = np.asarray(feats, dtype=np.float64).reshape(len(rows), -1) if rows else np.zeros((0, 0))
            y = np.asarray([get(r, li) for r in rows], dtype=np.float64).reshape(-1)
        except (TypeError, ValueError):
            return None
        parts.append(x)
        label_parts.append(y)
    if categorical:
        if not nb_classes:
            nb_classes = int(max((y.max() for y in label_parts if len(y)), default=0)) + 1
        eye = np.eye(int(nb_classes))
        label_parts = [eye[y.astype(np.int64)] if len(y) else np.zeros((0, int(nb_classes))) for y in label_parts]
    width = max((x.shape[1] for x in parts if len(x)), default=0)
    parts = [x if len(x) else np.zeros((0, width)) for x in parts]
    return RDD([PairBlock(x, y) for x, y in zip(parts, label_parts)], rdd.context)
