"""The smaller corners of the ``keras.Model`` surface: row weights, ``*_on_batch``, ``get_layer``, ``pop``, weight files."""
import numpy as np
import pytest
import torch

from elephas_b200.models import Dense, Sequential, load_model
from elephas_b200.models.optimizers import SGD


def _xy(n=96, seed=0):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((n, 6)).astype(np.float32)
    y = np.eye(3, dtype=np.float32)[rng.integers(0, 3, n)]
    return x, y


def _model(lr=0.1, names=('a', 'b')):
    m = Sequential([Dense(8, activation='relu', input_shape=(6,), name=names[0]), Dense(3, activation='softmax', name=names[1])])
    m.compile(SGD(learning_rate=lr), 'categorical_crossentropy', ['acc'])
    return m


def _oracle_step(weights, x, y, w, lr):
    ws = [torch.tensor(v, requires_grad=True) for v in weights]
    h = torch.relu(torch.tensor(x) @ ws[0] + ws[1])
    p = torch.softmax(h @ ws[2] + ws[3], dim=1)
    per_row = -(torch.tensor(y) * torch.log(p)).sum(1)
    loss = (per_row * torch.tensor(w)).sum() / len(x)
    loss.backward()
    return float(loss.detach()), [(v - lr * v.grad).detach().numpy() for v in ws]


def test_sample_and_class_weights_scale_the_rows():
    x, y = _xy()
    rng = np.random.default_rng(1)
    sw = rng.uniform(0.0, 2.0, len(x)).astype(np.float32)
    cw = {0: 1.0, 1: 3.0, 2: 0.5}
    m = _model()
    w0 = m.get_weights()
    rows = sw * np.array([cw[int(k)] for k in y.argmax(1)], dtype=np.float32)
    loss, after = _oracle_step(w0, x, y, rows, 0.1)
    got = m.train_on_batch(x, y, sample_weight=sw, class_weight=cw, return_dict=True)
    assert got['loss'] == pytest.approx(loss, rel=1e-5) and 0.0 <= got['acc'] <= 1.0
    for a, b in zip(m.get_weights(), after):
        np.testing.assert_allclose(a, b, rtol=1e-4, atol=1e-6)
    # through fit: one batch covering every row, no shuffling, is the same step
    m2 = _model()
    m2.set_weights(w0)
    h = m2.fit(x, y, batch_size=len(x), epochs=1, shuffle=False, sample_weight=sw, class_weight=cw)
    assert h.history['loss'][0] == pytest.approx(loss, rel=1e-5)
    for a, b in zip(m2.get_weights(), after):
        np.testing.assert_allclose(a, b, rtol=1e-4, atol=1e-6)
    # zero weights: nothing moves; integer labels are read as class indices
    m3 = _model()
    m3.set_weights(w0)
    m3.fit(x, y, batch_size=32, epochs=1, sample_weight=np.zeros(len(x)))
    for a, b in zip(m3.get_weights(), w0):
        np.testing.assert_array_equal(a, b)
    with pytest.raises(ValueError):
        m3.fit(x, y, sample_weight=np.ones(3))


def test_class_weight_with_integer_labels_and_validation_split():
    from elephas_b200.models.model import _row_weights
    w = _row_weights(np.array([0, 2, 1, 2]), None, {2: 4.0})
    np.testing.assert_array_equal(w, [1.0, 4.0, 1.0, 4.0])
    assert _row_weights(np.zeros(3), None, None) is None
    x, y = _xy()
    h = _model().fit(x, y, batch_size=16, epochs=2, validation_split=0.25, class_weight={0: 2.0})
    assert len(h.history['val_loss']) == 2


def test_on_batch_calls_and_return_dict():
    x, y = _xy(32)
    m = _model()
    before = m.get_weights()
    t = m.test_on_batch(x, y, return_dict=True)
    assert set(t) == {'loss', 'acc'}
    for a, b in zip(m.get_weights(), before):
        np.testing.assert_array_equal(a, b)
    e = m.evaluate(x, y, batch_size=8, return_dict=True)
    assert e['loss'] == pytest.approx(t['loss'], rel=1e-5)
    np.testing.assert_allclose(m.predict_on_batch(x), m.predict(x), rtol=1e-6)


def test_get_layer_and_pop():
    m = _model()
    assert m.get_layer('a').units == 8 and m.get_layer(index=-1).name == 'b'
    with pytest.raises(ValueError):
        m.get_layer('c')
    with pytest.raises(ValueError):
        m.get_layer(index=9)
    x, _ = _xy(8)
    hidden = m.get_layer('a')
    last = m.pop()
    assert last.name == 'b' and m.output_shape == (None, 8)
    np.testing.assert_allclose(m.predict(x), np.maximum(x @ m.get_weights()[0] + m.get_weights()[1], 0), rtol=1e-5, atol=1e-6)
    m.add(Dense(2, activation='softmax', name='new_head'))
    m.compile(SGD(learning_rate=0.1), 'categorical_crossentropy')
    assert m.output_shape == (None, 2) and m.get_layer('a') is hidden
    assert m.predict(x).shape == (8, 2)
    m.pop(), m.pop()
    with pytest.raises(TypeError):
        m.pop()


def test_weight_files_by_position_and_by_name(tmp_path):
    x, _ = _xy(8)
    m = _model()
    path = str(tmp_path / 'w.h5')
    m.save_weights(path)
    from elephas_b200.io import hdf5
    root = hdf5.load_tree(path)
    assert 'model_weights' not in root and 'a' in root     # the layout of keras.Model.save_weights
    other = _model(names=('p', 'q'))
    other.load_weights(path)                               # by position: the names differ
    np.testing.assert_allclose(other.predict(x), m.predict(x), rtol=1e-6)
    # by name: only 'a' exists in both
    third = Sequential([Dense(8, activation='relu', input_shape=(6,), name='a'), Dense(5, name='head')])
    head = [w.copy() for w in third.get_weights()[2:]]
    third.load_weights(path, by_name=True)
    np.testing.assert_array_equal(third.get_weights()[0], m.get_weights()[0])
    np.testing.assert_array_equal(third.get_weights()[2], head[0])
    # a full model file works as a weight file too
    full = str(tmp_path / 'full.h5')
    m.save(full)
    fourth = _model(names=('r', 's'))
    fourth.load_weights(full)
    np.testing.assert_allclose(fourth.predict(x), load_model(full).predict(x), rtol=1e-6)
    # shape mismatches are errors unless asked to skip them by name
    clash = Sequential([Dense(4, input_shape=(6,), name='a')])
    with pytest.raises(ValueError):
        clash.load_weights(path, by_name=True)
    clash.load_weights(path, by_name=True, skip_mismatch=True)
    with pytest.raises(ValueError):
        clash.load_weights(path)
    with pytest.raises(IOError):
        m.save_weights(path, overwrite=False)


def test_summary_lists_shapes_and_parameter_totals():
    from elephas_b200.models import BatchNormalization
    m = Sequential([Dense(8, input_shape=(4,), name='d0'), BatchNormalization(name='bn'), Dense(2, name='d1')])
    lines = []
    m.summary(print_fn=lines.append)
    text = '\n'.join(lines)
    assert 'd0 (Dense)' in text and '(None, 8)' in text and 'bn (BatchNormalization)' in text
    assert 'Total params: 90' in text and 'Trainable params: 74' in text and 'Non-trainable params: 16' in text
    assert all(len(line) <= 65 for line in lines)


def test_pad_sequences_and_normalize():
    from elephas_b200.models.utils import normalize, pad_sequences, to_categorical
    seqs = [[1, 2, 3], [4], [], [5, 6, 7, 8, 9]]
    np.testing.assert_array_equal(pad_sequences(seqs, maxlen=4), [[0, 1, 2, 3], [0, 0, 0, 4], [0, 0, 0, 0], [6, 7, 8, 9]])
    np.testing.assert_array_equal(pad_sequences(seqs, maxlen=4, padding='post', truncating='post', value=-1),
                                  [[1, 2, 3, -1], [4, -1, -1, -1], [-1, -1, -1, -1], [5, 6, 7, 8]])
    assert pad_sequences(seqs).shape == (4, 5) and pad_sequences(seqs).dtype == np.int32
    vec = pad_sequences([np.ones((2, 3)), np.ones((1, 3))], dtype='float32')
    assert vec.shape == (2, 2, 3) and vec[1, 0].sum() == 0 and vec[1, 1].sum() == 3
    with pytest.raises(ValueError):
        pad_sequences(seqs, padding='left')
    n = normalize(np.array([[3.0, 4.0], [0.0, 0.0]]))
    np.testing.assert_allclose(n, [[0.6, 0.8], [0.0, 0.0]])
    assert to_categorical([0, 2]).tolist() == [[1, 0, 0], [0, 0, 1]]


def test_steps_per_epoch_validation_freq_and_unknown_keywords():
    x, y = _xy(96)
    m = _model()
    h = m.fit(x, y, batch_size=16, epochs=4, steps_per_epoch=2, validation_data=(x[:32], y[:32]), validation_freq=2)
    assert m.optimizer.iterations == 8                       # 2 steps in each of 4 epochs
    assert len(h.history['loss']) == 4 and len(h.history['val_loss']) == 2
    h = _model().fit(x, y, batch_size=16, epochs=3, validation_split=0.25, validation_freq=[1, 3], workers=4)
    assert len(h.history['val_loss']) == 2
    with pytest.raises(TypeError):
        m.fit(x, y, batchsize=16)


def test_symbolic_ends_build_feature_extractors_and_new_heads_on_shared_layers():
    from elephas_b200.models import Model
    x, y = _xy(64)
    base = _model(names=('h', 'o'))
    base.fit(x, y, batch_size=16, epochs=1)
    features = Model(inputs=base.input, outputs=base.get_layer('h').output)
    w = base.get_weights()
    np.testing.assert_allclose(features.predict(x), np.maximum(x @ w[0] + w[1], 0.0), rtol=1e-5, atol=1e-6)
    assert base.inputs[0] is base.input and base.outputs[0] is base.output and base.get_layer('o').input is base.get_layer('h').output
    # a new head on the frozen trunk: only the head moves, and the base model sees nothing of it
    base.get_layer('h').trainable = False
    head = Dense(2, activation='softmax', name='new_head')(base.get_layer('h').output)
    tuned = Model(inputs=base.input, outputs=head)
    tuned.compile(SGD(learning_rate=0.5), 'categorical_crossentropy')
    assert len(tuned.trainable_weights) == 2
    y2 = np.eye(2, dtype=np.float32)[(x[:, 0] > 0).astype(int)]
    tuned.fit(x, y2, batch_size=16, epochs=2)
    after = base.get_weights()
    np.testing.assert_array_equal(after[0], w[0])
    np.testing.assert_array_equal(after[2], w[2])
    assert tuned.predict(x).shape == (64, 2) and base.predict(x).shape == (64, 3)


def test_a_frozen_batch_normalization_uses_its_moving_statistics_while_training():
    from elephas_b200.models import BatchNormalization
    rng = np.random.default_rng(0)
    x = (3.0 + 2.0 * rng.standard_normal((64, 5))).astype(np.float32)
    y = rng.standard_normal((64, 2)).astype(np.float32)
    m = Sequential([BatchNormalization(input_shape=(5,), name='bn'), Dense(2, name='d')])
    m.get_layer('bn').trainable = False
    m.compile(SGD(learning_rate=0.1), 'mse')
    before = m.get_weights()
    m.fit(x, y, batch_size=16, epochs=2)
    after = m.get_weights()
    for a, b in zip(after[:4], before[:4]):       # gamma, beta AND the moving statistics stay as they were
        np.testing.assert_array_equal(a, b)
    assert not np.array_equal(after[4], before[4])


def test_layer_level_weights_and_freezing_a_whole_model():
    m = _model(names=('h', 'o'))
    h = m.get_layer('h')
    k, b = h.get_weights()
    assert k.shape == (6, 8) and b.shape == (8,) and h.count_params() == 56
    assert h.input_shape == (None, 6) and h.output_shape == (None, 8)
    h.set_weights([np.ones((6, 8), np.float32), np.zeros(8, np.float32)])
    np.testing.assert_array_equal(m.get_weights()[0], np.ones((6, 8), np.float32))
    x = np.ones((2, 6), np.float32)
    np.testing.assert_allclose(np.maximum(x @ h.get_weights()[0], 0), 6.0)
    assert h._live_owners() == [m]
    with pytest.raises(ValueError):
        h.set_weights([np.ones((8, 6), np.float32), np.zeros(8, np.float32)])
    with pytest.raises(ValueError):
        h.set_weights([np.ones((6, 8), np.float32)])
    assert m.trainable and len(m.non_trainable_weights) == 0
    m.trainable = False
    assert not m.trainable and m.trainable_weights == [] and len(m.non_trainable_weights) == 4
    assert all(not w.requires_grad for w in m.weights)
    m.trainable = True
    assert len(m.trainable_weights) == 4 and all(w.requires_grad for w in m.weights)


def test_models_stay_picklable_and_copies_own_their_layers():
    import copy
    import pickle
    m = _model()
    twin, back = copy.deepcopy(m), pickle.loads(pickle.dumps(m))
    x = np.ones((2, 6), np.float32)
    for other in (twin, back):
        assert other.layers[1]._live_owners() == [other]
        np.testing.assert_allclose(other.predict(x), m.predict(x), rtol=1e-6)
    assert m.layers[1]._live_owners() == [m]


def test_evaluate_takes_sample_weights():
    x, y = _xy(40)
    m = _model()
    rng = np.random.default_rng(4)
    w = rng.uniform(0, 2, 40).astype(np.float32)
    p = np.clip(m.predict(x), 1e-7, 1.0)
    per_row = -(y * np.log(p)).sum(1)
    loss, acc = m.evaluate(x, y, batch_size=16, sample_weight=w)
    assert loss == pytest.approx(float((per_row * w).mean()), rel=1e-5)
    assert acc == pytest.approx(float((p.argmax(1) == y.argmax(1)).mean()))
    assert m.evaluate(x, y, sample_weight=np.ones(40))[0] == pytest.approx(m.evaluate(x, y)[0], rel=1e-6)
