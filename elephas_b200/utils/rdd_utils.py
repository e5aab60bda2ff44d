"""NumPy <-> RDD <-> LabeledPoint conversions.

Reference: ``elephas/utils/rdd_utils.py:10-85`` (``to_simple_rdd``, ``to_labeled_point``,
``from_labeled_point``, ``encode_label``, ``lp_to_simple_rdd``). Same signatures and results; with
the local Spark stand-in ``to_simple_rdd`` keeps each partition as two dense arrays
(:class:`elephas_b200.spark.context.PairBlock`) so that the GPU workers can stage a partition into
pinned memory without touching Python rows (generic RDD callers still see ``(x, y)`` pairs).
"""
from typing import Optional, Tuple

import numpy as np

from ..mllib.adapter import from_vector, to_vector
from ..spark import HAVE_PYSPARK, LabeledPoint, PairBlock


def to_simple_rdd(sc, features: np.ndarray, labels: np.ndarray):
    """Features and labels -> RDD of ``(feature_row, label_row)`` pairs."""
    if not HAVE_PYSPARK:
        return sc.parallelize(PairBlock(np.asarray(features), np.asarray(labels)))
    return sc.parallelize([(x, y) for x, y in zip(features, labels)])


def to_labeled_point(sc, features: np.ndarray, labels: np.ndarray, categorical: bool = False):
    """Features and labels -> RDD of ``LabeledPoint``; one-hot labels are argmax-decoded when ``categorical``."""
    points = [LabeledPoint(np.argmax(y) if categorical else y, to_vector(np.asarray(x).ravel()))
              for x, y in zip(features, labels)]
    return sc.parallelize(points)


def from_labeled_point(rdd, categorical: bool = False, nb_classes: Optional[int] = None) -> Tuple[np.ndarray, np.ndarray]:
    """LabeledPoint RDD -> ``(features, labels)`` arrays; labels one-hot encoded when ``categorical``."""
    pairs = rdd.map(lambda lp: (from_vector(lp.features), int(lp.label))).collect()
    features = np.array([p[0] for p in pairs])
    labels = np.array([p[1] for p in pairs])
    if categorical:
        if not nb_classes:
            nb_classes = int(np.max(labels)) + 1
        labels = np.stack([encode_label(l, nb_classes) for l in labels])
    return features, labels


def encode_label(label, nb_classes: int) -> np.ndarray:
    """One-hot vector for a single (integral) label."""
    encoded = np.zeros(nb_classes)
    encoded[int(label)] = 1.
    return encoded


def lp_to_simple_rdd(lp_rdd, categorical: bool = False, nb_classes: int = None):
    """LabeledPoint RDD -> RDD of ``(features, label)`` pairs (label one-hot when ``categorical``;
    ``nb_classes`` inferred as ``max + 1`` when not given, ``rdd_utils.py:80``)."""
    if categorical:
        if not nb_classes:
            nb_classes = int(lp_rdd.map(lambda lp: lp.label).map(int).max()) + 1
        return lp_rdd.map(lambda lp: (from_vector(lp.features), encode_label(lp.label, nb_classes)))
    return lp_rdd.map(lambda lp: (from_vector(lp.features), lp.label))


def to_sharded_rdd(sc, features: np.ndarray, labels: np.ndarray, rank: int = None, world_size: int = None,
                   feature_dtype: str = None):
    """Rank-sharded variant of :func:`to_simple_rdd` for SPMD (``torchrun``) programs (no counterpart in
    the reference, where Spark owns data placement): every rank passes only ITS OWN shard of the data and
    gets an RDD with one partition per rank in which partition ``rank`` holds that shard and the others
    are empty on this rank. ``SparkModel`` trains / predicts partition *i* on rank *i*, so the global
    dataset is never replicated in every process - the way to feed datasets sized for 180 GB GPUs."""
    from ..parallel.runtime import get_runtime
    from ..spark.context import RDD
    rt = get_runtime()
    rank = rt.rank if rank is None else rank
    world_size = rt.world_size if world_size is None else world_size
    features, labels = np.asarray(features), np.asarray(labels)
    x16 = None
    if feature_dtype in ('bfloat16', 'bf16'):
        # stage the shard ONCE as bf16 in pinned host memory (AVX-512 cast, csrc/host_cast.cpp): a model that computes in
        # bf16 then streams 2 bytes per feature and step instead of 4, with no per-step work on the CPU
        x16 = stage_features_bf16(features)
    elif feature_dtype not in (None, 'float32'):
        raise ValueError("feature_dtype must be None, 'float32' or 'bfloat16', got %r" % (feature_dtype,))
    empty = PairBlock(features[:0], labels[:0])
    rdd = RDD([PairBlock(features, labels, x16) if r == rank else empty for r in range(world_size)], sc)
    rdd._pinned_partitioning = True
    return rdd


def stage_features_bf16(features: np.ndarray):
    """``features`` (``[rows, ...]`` fp32) as a pinned ``torch.bfloat16`` tensor ``[rows, round_up(cols, 8)]`` (padding
    zeroed), or None when the native library is not available. Round to nearest even, bit-identical to the cast the
    device would do."""
    import os
    import torch
    from ..ops import native
    if not native.available():
        return None
    flat = np.ascontiguousarray(features, dtype=np.float32).reshape(len(features), -1)
    rows, cols = flat.shape
    ld = (cols + 7) // 8 * 8
    out = torch.empty(rows, ld, dtype=torch.bfloat16, pin_memory=torch.cuda.is_available())
    if rows:
        try:
            cores = len(os.sched_getaffinity(0))
        except (AttributeError, OSError):
            cores = os.cpu_count() or 1
        threads = max(1, min(16, cores // max(1, int(os.environ.get('LOCAL_WORLD_SIZE', '1')))))
        native.host_f32_to_bf16(torch.from_numpy(flat), out, rows, cols, threads=threads)
    return out
