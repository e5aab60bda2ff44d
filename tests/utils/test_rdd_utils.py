"""NumPy <-> RDD <-> LabeledPoint converters (reference: elephas/utils/rdd_utils.py:10-85). The checks are on VALUES and
on round trips, with features that differ from row to row, so a converter that dropped, reordered or mis-encoded rows
would be caught (the reference's own tests use all-ones features and mostly check shapes)."""
import numpy as np
import pytest

from elephas.utils import rdd_utils


def _data(rows=7, cols=5, classes=4, seed=0):
    rng = np.random.default_rng(seed)
    features = rng.normal(size=(rows, cols))
    classes_of = rng.integers(0, classes, rows)
    classes_of[0] = classes - 1  # the highest class is present, so nb_classes can be inferred
    return features, classes_of, np.eye(classes)[classes_of]


def test_to_simple_rdd_keeps_rows_and_order(spark_context):
    features, classes_of, _ = _data()
    rdd = rdd_utils.to_simple_rdd(spark_context, features, classes_of.astype(float))
    rows = rdd.collect()
    assert rdd.count() == len(rows) == 7
    assert np.array_equal(np.stack([r[0] for r in rows]), features)
    assert [float(r[1]) for r in rows] == [float(c) for c in classes_of]


@pytest.mark.parametrize('categorical', [True, False])
def test_labeled_points_round_trip(spark_context, categorical):
    features, classes_of, one_hot = _data(seed=1)
    labels = one_hot if categorical else classes_of.astype(float)
    lp_rdd = rdd_utils.to_labeled_point(spark_context, features, labels, categorical)
    points = lp_rdd.collect()
    assert [p.label for p in points] == [float(c) for c in classes_of]  # one-hot rows became class indices
    assert np.array_equal(np.stack([p.features.toArray() for p in points]), features)
    x, y = rdd_utils.from_labeled_point(lp_rdd, categorical, 4 if categorical else None)
    assert np.array_equal(x, features)
    assert np.array_equal(y, labels)


def test_encode_label_is_a_unit_vector():
    v = rdd_utils.encode_label(3, 10)
    assert v.shape == (10,) and v[3] == 1.0 and v.sum() == 1.0
    with pytest.raises(IndexError):
        rdd_utils.encode_label(10, 10)


def test_lp_to_simple_rdd_one_hot_regression_and_inferred_width(spark_context):
    features, classes_of, one_hot = _data(seed=2)
    lp_rdd = rdd_utils.to_labeled_point(spark_context, features, one_hot, True)
    rows = rdd_utils.lp_to_simple_rdd(lp_rdd, categorical=True, nb_classes=6).collect()
    assert all(r[1].shape == (6,) and r[1].sum() == 1.0 and int(np.argmax(r[1])) == c for r, c in zip(rows, classes_of))
    assert np.array_equal(np.stack([r[0] for r in rows]), features)
    # nb_classes omitted: max label + 1 (rdd_utils.py:80)
    inferred = rdd_utils.lp_to_simple_rdd(lp_rdd, categorical=True).collect()
    assert all(r[1].shape == (4,) for r in inferred)
    # regression: the label passes through as a scalar
    plain = rdd_utils.lp_to_simple_rdd(lp_rdd, categorical=False).collect()
    assert [float(r[1]) for r in plain] == [float(c) for c in classes_of]


def test_sharded_rdd_can_stage_features_as_bf16(spark_context):
    """``to_sharded_rdd(feature_dtype='bfloat16')`` keeps, next to the fp32 rows, a bf16 copy of the shard (round to
    nearest even, rows padded to a multiple of 8 columns with zeros) that bf16 models stream to the GPU as it is."""
    import torch
    from elephas_b200.ops import native
    if not native.available():
        pytest.skip("native library not built")
    rng = np.random.default_rng(3)
    x = rng.normal(size=(50, 93)).astype(np.float32)
    x[0, 0], x[1, 1] = np.inf, -0.0
    y = rng.normal(size=(50, 2)).astype(np.float32)
    rdd = rdd_utils.to_sharded_rdd(spark_context, x, y, rank=0, world_size=2, feature_dtype='bfloat16')
    mine, other = rdd.partitions()
    assert len(other) == 0 and len(mine) == 50 and np.array_equal(mine.x, x)
    assert mine.x16.dtype == torch.bfloat16 and tuple(mine.x16.shape) == (50, 96)
    assert torch.equal(mine.x16[:, :93], torch.from_numpy(x).to(torch.bfloat16)) and bool((mine.x16[:, 93:] == 0).all())
    part = mine[10:20]
    assert len(part) == 10 and torch.equal(part.x16, mine.x16[10:20])
    with pytest.raises(ValueError):
        rdd_utils.to_sharded_rdd(spark_context, x, y, rank=0, world_size=2, feature_dtype='float16')
