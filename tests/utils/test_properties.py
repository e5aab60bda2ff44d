"""Property tests (hypothesis) of the pure-Python building blocks: whatever the shapes, what goes in comes out."""
import socket
import threading

import numpy as np
import torch
from hypothesis import given, settings
from hypothesis import strategies as st
from hypothesis.extra import numpy as hnp

from elephas_b200.models import Dense, Sequential
from elephas_b200.parallel.flat import FlatLayout
from elephas_b200.utils import functional_utils as fu
from elephas_b200.utils.sockets import receive, send

widths = st.lists(st.integers(min_value=1, max_value=70), min_size=2, max_size=5)


def _chain(ws):
    return Sequential([Dense(ws[1], input_dim=ws[0])] + [Dense(w) for w in ws[2:]])


@settings(max_examples=30, deadline=None, derandomize=True)
@given(widths, st.integers(0, 2 ** 31 - 1))
def test_flat_layout_round_trips_keras_arrays_and_model_tensors(ws, seed):
    """pack_arrays / unpack_arrays (Keras layout, the list the reference moves around) and pack_model / unpack_to_model
    (the model's own tensors) are inverse to each other for any widths - padded leading dimensions included - and the
    padding stays zero."""
    model = _chain(ws)
    layout = FlatLayout(model)
    rng = np.random.default_rng(seed)
    arrays = [rng.normal(size=w.shape).astype(np.float32) for w in model.get_weights()]
    flat = layout.pack_arrays(arrays)
    assert flat.numel() == layout.numel and layout.numel % 64 == 0
    back = layout.unpack_arrays(flat)
    assert all(np.array_equal(a, b) for a, b in zip(arrays, back))
    assert float(flat.abs().sum()) == float(sum(np.abs(a).astype(np.float64).sum() for a in arrays)) or \
        np.isclose(float(flat.double().abs().sum()), sum(np.abs(a).astype(np.float64).sum() for a in arrays), rtol=1e-6)
    model.set_weights(arrays)
    again = layout.pack_model(model)
    assert torch.equal(again, flat)
    other = _chain(ws)
    layout.unpack_to_model(flat, other)
    assert all(np.array_equal(a, b) for a, b in zip(other.get_weights(), arrays))


@settings(max_examples=30, deadline=None, derandomize=True)
@given(st.lists(hnp.arrays(np.float32, hnp.array_shapes(max_dims=3, max_side=6), elements=st.floats(-100, 100, width=32)),
                min_size=1, max_size=5))
def test_parameter_list_arithmetic(arrays):
    """functional_utils on lists of arrays: (a + b) - b == a, a - a == neutral, divide_by scales every array."""
    doubled = fu.add_params(arrays, arrays)
    assert all(np.array_equal(d, 2 * a) for d, a in zip(doubled, arrays))
    assert all(np.array_equal(s, a) for s, a in zip(fu.subtract_params(doubled, arrays), arrays))
    zero = fu.get_neutral(arrays)
    assert all(z.shape == a.shape and not z.any() for z, a in zip(zero, arrays))
    assert all(np.allclose(h, a / 4.0) for h, a in zip(fu.divide_by(arrays, 4), arrays))


@settings(max_examples=20, deadline=None, derandomize=True)
@given(st.lists(hnp.arrays(np.float32, hnp.array_shapes(max_dims=2, max_side=40), elements=st.floats(-1, 1, width=32)),
                max_size=4), st.binary(max_size=3000))
def test_socket_frames_of_any_size(arrays, blob):
    """The 20-digit length header + pickle framing of the socket parameter server (elephas/utils/sockets.py:45-71) for
    payloads from empty to tens of kilobytes, sent while the other side reads."""
    a, b = socket.socketpair()
    payload = {'delta': arrays, 'blob': blob}
    t = threading.Thread(target=send, args=(a, payload))
    t.start()
    got = receive(b)
    t.join()
    a.close()
    b.close()
    assert got['blob'] == blob and len(got['delta']) == len(arrays)
    assert all(np.array_equal(x, y) and x.dtype == y.dtype for x, y in zip(got['delta'], arrays))


@settings(max_examples=25, deadline=None, derandomize=True)
@given(st.integers(1, 40), st.integers(1, 12), st.integers(2, 7), st.integers(0, 2 ** 31 - 1))
def test_labeled_point_and_data_frame_adapters_round_trip(rows, features, classes, seed):
    """features / labels -> LabeledPoint RDD / DataFrame -> back (``utils/rdd_utils.py``, ``ml/adapter.py``,
    ``mllib/adapter.py``), categorical and plain, for any sizes."""
    from elephas_b200.ml import adapter as ml_adapter
    from elephas_b200.mllib import adapter as mllib_adapter
    from elephas_b200.spark import SparkContext
    from elephas_b200.utils import rdd_utils
    sc = SparkContext.getOrCreate()
    rng = np.random.default_rng(seed)
    x = rng.normal(size=(rows, features))
    labels = rng.integers(0, classes, rows)
    labels[0] = classes - 1  # (lp_to_simple_rdd infers the class count from the largest label)
    onehot = np.eye(classes)[labels]
    lp = rdd_utils.to_labeled_point(sc, x, onehot, categorical=True)
    fx, fy = rdd_utils.from_labeled_point(lp, categorical=True, nb_classes=classes)
    assert np.allclose(fx, x) and np.array_equal(fy, onehot)
    pairs = rdd_utils.lp_to_simple_rdd(lp, categorical=True, nb_classes=classes).collect()
    assert len(pairs) == rows and all(np.allclose(p[0], x[i]) and np.array_equal(p[1], onehot[i]) for i, p in enumerate(pairs))
    # plain labels come back as integers: the reference reads them with int(lp.label) (elephas/utils/rdd_utils.py:47)
    fx2, fy2 = rdd_utils.from_labeled_point(rdd_utils.to_labeled_point(sc, x, labels.astype(float), categorical=False))
    assert np.allclose(fx2, x) and np.array_equal(np.asarray(fy2).reshape(-1), labels)
    df = ml_adapter.to_data_frame(sc, x, onehot, categorical=True)
    dx, dy = ml_adapter.from_data_frame(df, categorical=True, nb_classes=classes)
    assert np.allclose(dx, x) and np.array_equal(dy, onehot)
    simple = ml_adapter.df_to_simple_rdd(df, categorical=True, nb_classes=classes).collect()
    assert len(simple) == rows and all(np.allclose(s[0], x[i]) and np.array_equal(s[1], onehot[i]) for i, s in enumerate(simple))
    m = rng.normal(size=(rows, features))
    assert np.array_equal(mllib_adapter.from_matrix(mllib_adapter.to_matrix(m)), m)      # D6: values land where they were
    v = rng.normal(size=features)
    assert np.array_equal(mllib_adapter.from_vector(mllib_adapter.to_vector(v)), v)


@settings(max_examples=25, deadline=None, derandomize=True)
@given(st.lists(st.tuples(st.integers(1, 33), st.sampled_from([None, 'relu', 'tanh', 'sigmoid', 'softmax', 'elu']),
                          st.sampled_from([0.0, 0.0, 0.25, 0.5]), st.booleans()), min_size=1, max_size=4),
       st.integers(1, 17), st.integers(0, 2 ** 31 - 1))
def test_model_json_and_weights_round_trip(layers, in_dim, seed):
    """A random Dense / Activation / Dropout / BatchNormalization stack -> to_json + get_weights -> model_from_json +
    set_weights predicts the same values (the closure the reference ships to its workers, elephas/worker.py:27-34,
    elephas/utils/serialization.py)."""
    from elephas_b200.models import Activation, BatchNormalization, Dropout, model_from_json
    from elephas_b200.utils.serialization import dict_to_model, model_to_dict
    torch.manual_seed(seed % 1000)
    stack = []
    for i, (units, act, drop, norm) in enumerate(layers):
        kwargs = {'input_dim': in_dim} if i == 0 else {}
        if act in (None, 'elu'):
            stack.append(Dense(units, activation=act, **kwargs))
        else:
            stack += [Dense(units, **kwargs), Activation(act)]
        if norm:
            stack.append(BatchNormalization())
        if drop:
            stack.append(Dropout(drop))
    model = Sequential(stack)
    x = np.random.default_rng(seed).normal(size=(5, in_dim)).astype(np.float32)
    want = model.predict(x)
    twin = model_from_json(model.to_json())
    assert [type(l).__name__ for l in twin.layers] == [type(l).__name__ for l in model.layers]
    assert [w.shape for w in twin.get_weights()] == [w.shape for w in model.get_weights()]
    twin.set_weights(model.get_weights())
    assert np.allclose(twin.predict(x), want, atol=1e-6)
    third = dict_to_model(model_to_dict(model))
    assert np.allclose(third.predict(x), want, atol=1e-6)
    assert twin.to_json() == model_from_json(twin.to_json()).to_json()   # the JSON is a fixed point


@settings(max_examples=15, deadline=None, derandomize=True)
@given(st.integers(2, 60), st.integers(1, 9), st.integers(2, 5), st.integers(0, 2 ** 31 - 1))
def test_block_columns_equal_the_row_path(rows, features, classes, seed):
    """DataFrame stages that compute on whole columns (StringIndexer, StandardScaler) give the rows the row-wise
    definitions give: the lazily appended block columns materialise to the same values, in the same order."""
    from elephas_b200.spark import SparkSession, StandardScaler, StringIndexer, Vectors
    rng = np.random.default_rng(seed)
    x = rng.normal(size=(rows, features))
    cats = ['class_%d' % c for c in rng.integers(0, classes, rows)]
    spark = SparkSession.builder.getOrCreate()
    df = spark.createDataFrame([(Vectors.dense(f), c) for f, c in zip(x, cats)], ['features', 'category'])
    indexed = StringIndexer(inputCol='category', outputCol='idx').fit(df).transform(df)
    scaler = StandardScaler(inputCol='features', outputCol='scaled', withMean=True, withStd=True).fit(indexed)
    scaled = scaler.transform(indexed)
    out = scaled.collect()
    assert len(out) == rows and scaled.columns == ['features', 'category', 'idx', 'scaled']
    # label indices: descending frequency, ties alphabetical (Spark's default)
    counts = {}
    for c in cats:
        counts[c] = counts.get(c, 0) + 1
    order = [k for k, _ in sorted(counts.items(), key=lambda kv: (-kv[1], kv[0]))]
    assert [r['idx'] for r in out] == [float(order.index(c)) for c in cats]
    mean, std = x.mean(axis=0), x.std(axis=0, ddof=1) if rows > 1 else np.ones(features)
    want = (x - mean) / np.where(std > 0, std, 1.0) * (std > 0)
    got = np.stack([np.asarray(r['scaled'].toArray()) for r in out])
    assert np.allclose(got, want, atol=1e-9)
    assert [r['category'] for r in out] == cats and np.allclose(np.stack([r['features'].toArray() for r in out]), x)


@settings(max_examples=12, deadline=None, derandomize=True)
@given(widths, st.sampled_from(['sgd', 'momentum', 'adam', 'rmsprop', 'adagrad']), st.integers(0, 2 ** 31 - 1))
def test_checkpoints_resume_exactly(ws, opt_name, seed):
    """save -> load_model -> one more step == one more step without the detour: weights, optimizer slots (stored in Keras
    layout) and the iteration counter all survive the file (elephas/spark_model.py:92-134 relies on Keras for this)."""
    import os
    import tempfile
    from elephas_b200.models import SGD, Adagrad, Adam, RMSprop, load_model
    make = {'sgd': lambda: SGD(learning_rate=0.05), 'momentum': lambda: SGD(learning_rate=0.05, momentum=0.9, nesterov=True),
            'adam': lambda: Adam(learning_rate=0.01), 'rmsprop': lambda: RMSprop(learning_rate=0.01),
            'adagrad': lambda: Adagrad(learning_rate=0.05)}[opt_name]
    torch.manual_seed(seed % 997)
    model = _chain(ws)
    model.compile(make(), 'mse', ['mae'])
    rng = np.random.default_rng(seed)
    x = rng.normal(size=(24, ws[0])).astype(np.float32)
    y = rng.normal(size=(24, ws[-1])).astype(np.float32)
    for i in (0, 8):
        model.train_on_batch(x[i:i + 8], y[i:i + 8])
    with tempfile.TemporaryDirectory() as tmp:
        path = os.path.join(tmp, 'm.h5')
        model.save(path)
        resumed = load_model(path)
    assert resumed.optimizer.iterations == model.optimizer.iterations == 2
    a = model.train_on_batch(x[16:], y[16:])
    b = resumed.train_on_batch(x[16:], y[16:])
    assert np.allclose(a, b, atol=1e-6, rtol=1e-5)
    for u, v in zip(model.get_weights(), resumed.get_weights()):
        assert np.allclose(u, v, atol=1e-6, rtol=1e-5), np.abs(u - v).max()


@settings(max_examples=10, deadline=None, derandomize=True)
@given(st.integers(1, 97), st.integers(1, 6), st.integers(0, 2 ** 31 - 1))
def test_distributed_predict_and_evaluate_equal_the_local_ones(rows, workers, seed):
    """SparkModel.predict keeps the row order and SparkModel.evaluate the sample-weighted mean for ANY number of rows and
    partitions - empty and one-row partitions included (elephas/spark_model.py:235-308 does this with zipWithIndex /
    sortBy and a weighted reduce)."""
    from elephas.spark_model import SparkModel
    from elephas_b200.models import SGD, Activation
    from elephas_b200.spark import SparkContext
    torch.manual_seed(seed % 991)
    rng = np.random.default_rng(seed)
    x = rng.normal(size=(rows, 7)).astype(np.float32)
    y = np.eye(3, dtype=np.float32)[rng.integers(0, 3, rows)]
    model = Sequential([Dense(9, input_dim=7, activation='relu'), Dense(3), Activation('softmax')])
    model.compile(SGD(learning_rate=0.1), 'categorical_crossentropy', ['acc'])
    sm = SparkModel(model, mode='synchronous', num_workers=workers)
    sc = SparkContext.getOrCreate()
    local_pred = model.predict(x)
    assert np.allclose(sm.predict(x), local_pred, atol=1e-6)
    assert np.allclose(np.asarray(sm.predict(sc.parallelize(list(x), workers))), local_pred, atol=1e-6)
    local_eval = model.evaluate(x, y)
    assert np.allclose(sm.evaluate(x, y), local_eval, atol=1e-5)


@settings(max_examples=8, deadline=None, derandomize=True)
@given(st.integers(20, 90), st.integers(2, 4), st.sampled_from([4, 8, 16]), st.integers(0, 2 ** 31 - 1))
def test_synchronous_mode_is_the_mean_of_the_replicas(rows, workers, batch, seed):
    """The defining identity of the reference's synchronous mode (elephas/spark_model.py:217-231): after one round the
    master equals the mean of the replicas trained alone on their partitions from the same start - for any row count,
    worker count and batch size (partitions not larger than a batch contribute a zero delta, worker.py:41)."""
    import copy
    from elephas.spark_model import SparkModel
    from elephas_b200.models import SGD
    from elephas_b200.spark import SparkContext
    from elephas.utils.rdd_utils import to_simple_rdd
    torch.manual_seed(seed % 983)
    rng = np.random.default_rng(seed)
    x = rng.normal(size=(rows, 5)).astype(np.float32)
    y = rng.normal(size=(rows, 2)).astype(np.float32)
    model = Sequential([Dense(6, input_dim=5, activation='tanh'), Dense(2)])
    model.compile(SGD(learning_rate=0.05), 'mse')
    start = [w.copy() for w in model.get_weights()]
    sm = SparkModel(model, mode='synchronous', num_workers=workers, sync_rounds='fit')
    rdd = to_simple_rdd(SparkContext.getOrCreate(), x, y)
    parts = [p for p in rdd.repartition(workers).partitions()]
    sm.fit(rdd, epochs=2, batch_size=batch, verbose=0, shuffle=False)
    got = sm.master_network.get_weights()
    # replicas trained alone
    ends = []
    for part in parts:
        rows_p = list(part)
        replica = Sequential([Dense(6, input_dim=5, activation='tanh'), Dense(2)])
        replica.compile(SGD(learning_rate=0.05), 'mse')
        replica.set_weights(copy.deepcopy(start))
        if len(rows_p) > batch:
            px = np.stack([np.asarray(r[0]) for r in rows_p]).astype(np.float32)
            py = np.stack([np.asarray(r[1]) for r in rows_p]).astype(np.float32)
            replica.fit(px, py, batch_size=batch, epochs=2, shuffle=False, verbose=0)
        ends.append(replica.get_weights())
    for i, g in enumerate(got):
        want = np.mean([e[i] for e in ends], axis=0)
        assert np.allclose(g, want, atol=2e-6), np.abs(g - want).max()


@settings(max_examples=40, deadline=None, derandomize=True)
@given(st.lists(st.integers(-50, 50), max_size=60), st.integers(1, 7), st.integers(1, 7))
def test_rdd_stand_in_behaves_like_the_list_it_holds(items, slices, repart):
    """The local RDD (elephas_b200/spark/context.py) against plain Python on the same data: order-preserving
    transformations, partition bookkeeping, actions - for any contents and partition counts, empty partitions included."""
    from elephas_b200.spark import SparkContext
    rdd = SparkContext.getOrCreate().parallelize(items, slices)
    assert rdd.getNumPartitions() == slices and rdd.collect() == items and rdd.count() == len(items)
    assert sum(len(list(p)) for p in rdd.partitions()) == len(items)
    assert rdd.map(lambda v: v * 2).collect() == [v * 2 for v in items]
    assert rdd.filter(lambda v: v % 3 == 0).collect() == [v for v in items if v % 3 == 0]
    assert rdd.flatMap(lambda v: [v, -v]).collect() == [w for v in items for w in (v, -v)]
    assert rdd.zipWithIndex().collect() == [(v, i) for i, v in enumerate(items)]
    assert rdd.mapPartitions(lambda it: [sum(it)]).collect() == [sum(p) for p in rdd.glom().collect()]
    again = rdd.repartition(repart)
    assert again.getNumPartitions() == repart and again.collect() == items   # (the stand-in keeps the order)
    assert rdd.sortBy(lambda v: v).collect() == sorted(items)
    assert rdd.sortBy(lambda v: v, ascending=False).collect() == sorted(items, reverse=True)
    assert sorted(rdd.distinct().collect()) == sorted(set(items))
    assert rdd.zip(rdd.map(lambda v: v + 1)).collect() == [(v, v + 1) for v in items]
    assert rdd.union(rdd).collect() == items + items
    assert rdd.isEmpty() == (not items) and rdd.take(3) == items[:3]
    if items:
        assert rdd.reduce(lambda a, b: a + b) == sum(items) and rdd.sum() == sum(items)
        assert rdd.max() == max(items) and rdd.min() == min(items) and rdd.first() == items[0]
        assert np.isclose(rdd.mean(), np.mean(items))


@settings(max_examples=25, deadline=None, derandomize=True)
@given(st.integers(1, 30), st.integers(0, 35), st.integers(0, 2 ** 31 - 1))
def test_data_frame_operations_on_frames_with_block_columns(rows, limit, seed):
    """select / drop / limit / rename / withColumn / SQL on a frame whose newest columns exist only as blocks (appended by
    StringIndexer / StandardScaler): the lazily built rows equal what the eager definitions give."""
    from elephas_b200.spark import DoubleType, SparkSession, StandardScaler, StringIndexer, Vectors, functions
    rng = np.random.default_rng(seed)
    x = rng.normal(size=(rows, 3))
    cats = ['c%d' % c for c in rng.integers(0, 3, rows)]
    spark = SparkSession.builder.getOrCreate()
    df = spark.createDataFrame([(Vectors.dense(f), c, int(i)) for i, (f, c) in enumerate(zip(x, cats))],
                               ['features', 'category', 'row'])
    df = StringIndexer(inputCol='category', outputCol='idx').fit(df).transform(df)
    df = StandardScaler(inputCol='features', outputCol='scaled', withMean=True, withStd=False).fit(df).transform(df)
    idx = [r['idx'] for r in df.collect()]
    centred = x - x.mean(axis=0)
    sel = df.select('row', 'idx').collect()
    assert [(r['row'], r['idx']) for r in sel] == list(zip(range(rows), idx))
    assert df.drop('features', 'category').columns == ['row', 'idx', 'scaled']
    lim = df.limit(limit)
    assert lim.count() == min(limit, rows) and [r['row'] for r in lim.collect()] == list(range(min(limit, rows)))
    assert np.allclose(np.stack([r['scaled'].toArray() for r in lim.collect()]) if lim.count() else np.zeros((0, 3)),
                       centred[:limit])
    ren = df.withColumnRenamed('idx', 'label')
    assert ren.columns == ['features', 'category', 'row', 'label', 'scaled'] and [r['label'] for r in ren.collect()] == idx
    plus = df.withColumn('twice', functions.col('idx').astype(DoubleType()))
    assert [r['twice'] for r in plus.collect()] == idx
    df.createOrReplaceTempView('frame_under_test')
    out = spark.sql('SELECT scaled AS features, idx as label from frame_under_test').collect()
    assert [r['label'] for r in out] == idx and np.allclose(np.stack([r['features'].toArray() for r in out]), centred)


# ----------------------------------------------------------------------------- Keras-surface building blocks
small_floats = st.floats(min_value=-4.0, max_value=4.0, allow_nan=False, width=32)
matrices = hnp.arrays(np.float32, hnp.array_shapes(min_dims=2, max_dims=2, min_side=1, max_side=6), elements=small_floats)


@settings(max_examples=60, deadline=None, derandomize=True)
@given(matrices, st.sampled_from(['MaxNorm', 'UnitNorm', 'NonNeg', 'MinMaxNorm']), st.integers(0, 1))
def test_constraints_are_projections(w, name, axis):
    """Applying a constraint twice equals applying it once, and the constrained weight satisfies the bound."""
    from elephas_b200.models import constraints
    con = {'MaxNorm': constraints.MaxNorm(1.5, axis=axis), 'UnitNorm': constraints.UnitNorm(axis=axis),
           'NonNeg': constraints.NonNeg(), 'MinMaxNorm': constraints.MinMaxNorm(0.5, 1.5, axis=axis)}[name]
    w = np.where(np.abs(w) < 1e-2, 0.0, w).astype(np.float32)   # (norms of the order of Keras' epsilon 1e-7 take two steps)
    once = con(torch.tensor(w))
    twice = con(once)
    np.testing.assert_allclose(twice.numpy(), once.numpy(), rtol=1e-4, atol=1e-5)
    norms = once.norm(dim=axis)
    if name == 'MaxNorm':
        assert bool((norms <= 1.5 + 1e-4).all())
    if name == 'NonNeg':
        assert bool((once >= 0).all())
    if name == 'UnitNorm':
        nonzero = torch.tensor(w).norm(dim=axis) > 1e-3
        np.testing.assert_allclose(norms[nonzero].numpy(), np.ones(int(nonzero.sum())), rtol=1e-3)


@settings(max_examples=40, deadline=None, derandomize=True)
@given(matrices, st.floats(0.001, 0.5), st.floats(0.0, 0.5))
def test_regularizer_gradients_match_their_closed_form(w, l1, l2):
    from elephas_b200.models import regularizers
    t = torch.tensor(w, requires_grad=True)
    regularizers.L1L2(l1, l2)(t).backward()
    np.testing.assert_allclose(t.grad.numpy(), l1 * np.sign(w) + 2.0 * l2 * w, rtol=1e-5, atol=1e-6)


@settings(max_examples=60, deadline=None, derandomize=True)
@given(st.lists(st.lists(st.integers(1, 50), max_size=9), min_size=1, max_size=8), st.integers(1, 10),
       st.sampled_from(['pre', 'post']), st.sampled_from(['pre', 'post']))
def test_pad_sequences_keeps_the_right_end_of_every_sequence(seqs, maxlen, padding, truncating):
    from elephas_b200.models.utils import pad_sequences
    out = pad_sequences(seqs, maxlen=maxlen, padding=padding, truncating=truncating)
    assert out.shape == (len(seqs), maxlen)
    for row, seq in zip(out, seqs):
        kept = seq[-maxlen:] if truncating == 'pre' else seq[:maxlen]
        body = row[maxlen - len(kept):] if padding == 'pre' else row[:len(kept)]
        rest = row[:maxlen - len(kept)] if padding == 'pre' else row[len(kept):]
        assert body.tolist() == kept and not rest.any()


@settings(max_examples=40, deadline=None, derandomize=True)
@given(st.floats(1e-4, 1.0), st.integers(1, 500), st.floats(0.05, 0.99), st.integers(0, 2000), st.integers(0, 2000))
def test_decaying_schedules_never_increase_and_never_go_negative(lr, steps, rate, a, b):
    from elephas_b200.models import schedules
    lo, hi = min(a, b), max(a, b)
    for sched in (schedules.ExponentialDecay(lr, steps, rate), schedules.ExponentialDecay(lr, steps, rate, staircase=True),
                  schedules.InverseTimeDecay(lr, steps, rate), schedules.CosineDecay(lr, steps, alpha=0.1),
                  schedules.PolynomialDecay(lr, steps, end_learning_rate=lr * 0.01)):
        assert 0.0 <= sched(hi) <= sched(lo) * (1 + 1e-9) <= lr * (1 + 1e-9)   # (0.5 ** 1074 underflows to zero)


@settings(max_examples=25, deadline=None, derandomize=True)
@given(st.integers(1, 6), st.integers(1, 5), st.integers(2, 7), st.integers(0, 2 ** 31 - 1))
def test_masked_recurrent_rows_do_not_depend_on_their_padding(batch, features, steps, seed):
    """Whatever follows the last unmasked step of a row - other rows' lengths, the padded values - the row's result is the
    result of the bare sequence."""
    from elephas_b200.models import GRU, Masking, Sequential as Seq
    rng = np.random.default_rng(seed)
    torch.manual_seed(seed % 1000)
    m = Seq([Masking(0.0, input_shape=(steps, features)), GRU(3)])
    lengths = rng.integers(1, steps + 1, batch)
    x = np.zeros((batch, steps, features), dtype=np.float32)
    for r, n in enumerate(lengths):
        x[r, :n] = rng.uniform(0.5, 1.5, (n, features))
    out = m.predict(x)
    for r, n in enumerate(lengths):
        np.testing.assert_allclose(out[r], m.predict(x[r:r + 1, :n])[0], rtol=1e-5, atol=1e-6)
