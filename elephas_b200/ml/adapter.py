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
    selected = df.select(features_col, label_col)
    lp_rdd = selected.rdd.map(lambda row: LabeledPoint(row[1], _as_mllib_vector(row[0])))
    return lp_to_simple_rdd(lp_rdd, categorical, nb_classes)
