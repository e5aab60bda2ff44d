"""The wider Keras layer vocabulary (``elephas_b200/models/layers_extra.py``) against NumPy restatements of the Keras
formulas, through JSON and HDF5, and in training."""
import numpy as np
import pytest
import torch

from elephas_b200.models import (ELU, GRU, LSTM, Activation, Adam, Average, AveragePooling1D, Concatenate, Conv1D, Dense,
                                 Embedding, Flatten, GlobalAveragePooling1D, GlobalMaxPooling1D, GlobalMaxPooling2D, Input,
                                 LayerNormalization, LeakyReLU, Maximum, MaxPooling1D, Minimum, Model, Multiply, Sequential,
                                 SimpleRNN, Subtract, load_model, model_from_json)


def _sigmoid(v):
    return 1.0 / (1.0 + np.exp(-v))


def _run(layer, x):
    return layer(torch.from_numpy(np.asarray(x))).detach().numpy()


def test_recurrent_layers_follow_the_keras_formulas():
    rng = np.random.default_rng(0)
    x = rng.normal(size=(3, 5, 4)).astype(np.float32)
    u = 6
    # LSTM: gates i, f, c, o
    lstm = LSTM(u, return_sequences=True)
    got = _run(lstm, x)
    k, r, b = [w.detach().numpy() for w in lstm.weights]
    assert k.shape == (4, 4 * u) and r.shape == (u, 4 * u) and np.array_equal(b[u:2 * u], np.ones(u)) and not b[:u].any()
    h, c = np.zeros((3, u), np.float32), np.zeros((3, u), np.float32)
    want = []
    for t in range(5):
        z = x[:, t] @ k + h @ r + b
        i, f, g, o = z[:, :u], z[:, u:2 * u], z[:, 2 * u:3 * u], z[:, 3 * u:]
        c = _sigmoid(f) * c + _sigmoid(i) * np.tanh(g)
        h = _sigmoid(o) * np.tanh(c)
        want.append(h)
    assert np.allclose(got, np.stack(want, axis=1), atol=1e-5)
    assert np.allclose(_run(LSTM(u), x).shape, (3, u))
    # GRU, reset_after (TF 2 default): gates z, r, h; bias [2, 3u]
    gru = GRU(u)
    got = _run(gru, x)
    k, r, b = [w.detach().numpy() for w in gru.weights]
    assert b.shape == (2, 3 * u)
    b[:] = rng.normal(size=b.shape) * 0.1          # (a view of the layer's tensor: the layer sees the new bias)
    got = _run(gru, x)
    h = np.zeros((3, u), np.float32)
    for t in range(5):
        xz, hz = x[:, t] @ k + b[0], h @ r + b[1]
        z = _sigmoid(xz[:, :u] + hz[:, :u])
        rr = _sigmoid(xz[:, u:2 * u] + hz[:, u:2 * u])
        cand = np.tanh(xz[:, 2 * u:] + rr * hz[:, 2 * u:])
        h = z * h + (1 - z) * cand
    assert np.allclose(got, h, atol=1e-5)
    # classic GRU (reset_after=False)
    gru1 = GRU(u, reset_after=False)
    got = _run(gru1, x)
    k, r, b = [w.detach().numpy() for w in gru1.weights]
    h = np.zeros((3, u), np.float32)
    for t in range(5):
        xz = x[:, t] @ k + b
        z = _sigmoid(xz[:, :u] + h @ r[:, :u])
        rr = _sigmoid(xz[:, u:2 * u] + h @ r[:, u:2 * u])
        cand = np.tanh(xz[:, 2 * u:] + (rr * h) @ r[:, 2 * u:])
        h = z * h + (1 - z) * cand
    assert np.allclose(got, h, atol=1e-5)
    # SimpleRNN, backwards
    rnn = SimpleRNN(u, go_backwards=True)
    got = _run(rnn, x)
    k, r, b = [w.detach().numpy() for w in rnn.weights]
    h = np.zeros((3, u), np.float32)
    for t in range(4, -1, -1):
        h = np.tanh(x[:, t] @ k + h @ r + b)
    assert np.allclose(got, h, atol=1e-5)
    assert [n.split('/', 1)[1] for n in lstm.weight_names()] == ['lstm_cell/kernel:0', 'lstm_cell/recurrent_kernel:0',
                                                                  'lstm_cell/bias:0']


def test_conv1d_pooling_normalisation_embedding_and_merges():
    rng = np.random.default_rng(1)
    x = rng.normal(size=(2, 9, 3)).astype(np.float32)
    for padding, stride, dilation in (('valid', 1, 1), ('same', 2, 1), ('causal', 1, 2), ('same', 1, 1)):
        conv = Conv1D(4, 3, strides=stride, padding=padding, dilation_rate=dilation)
        got = _run(conv, x)
        kernel = conv.to_keras_layout(0, conv.kernel.detach()).numpy()          # (width, channels, filters)
        bias = conv.bias.detach().numpy()
        eff = 2 * dilation + 1
        if padding == 'causal':
            xp = np.pad(x, ((0, 0), (eff - 1, 0), (0, 0)))
        elif padding == 'same':
            total = max((-(-9 // stride) - 1) * stride + eff - 9, 0)
            xp = np.pad(x, ((0, 0), (total // 2, total - total // 2), (0, 0)))
        else:
            xp = x
        steps = (xp.shape[1] - eff) // stride + 1
        want = np.stack([sum(xp[:, t * stride + j * dilation] @ kernel[j] for j in range(3)) + bias for t in range(steps)], 1)
        assert got.shape == want.shape == tuple(conv.compute_output_shape((2, 9, 3))) and np.allclose(got, want, atol=1e-5)
    assert np.allclose(_run(MaxPooling1D(2), x), np.maximum(x[:, 0:8:2], x[:, 1:9:2]))
    avg_same = _run(AveragePooling1D(2, padding='same'), x)
    assert avg_same.shape == (2, 5, 3) and np.allclose(avg_same[:, -1], x[:, -1]) and \
        np.allclose(avg_same[:, 0], x[:, :2].mean(1), atol=1e-6)
    assert np.allclose(_run(GlobalAveragePooling1D(), x), x.mean(1), atol=1e-6)
    assert np.allclose(_run(GlobalMaxPooling1D(), x), x.max(1))
    img = rng.normal(size=(2, 4, 5, 3)).astype(np.float32)
    assert np.allclose(_run(GlobalMaxPooling2D(), img), img.max(axis=(1, 2)))
    ln = LayerNormalization()
    got = _run(ln, x)
    want = (x - x.mean(-1, keepdims=True)) / np.sqrt(x.var(-1, keepdims=True) + 1e-3)
    assert np.allclose(got, want, atol=1e-5)
    emb = Embedding(11, 4, input_length=6)
    idx = rng.integers(0, 11, (3, 6))
    assert np.array_equal(_run(emb, idx), emb.embeddings.detach().numpy()[idx])
    assert np.allclose(_run(LeakyReLU(0.2), x), np.where(x > 0, x, 0.2 * x))
    assert np.allclose(_run(ELU(), x), np.where(x > 0, x, np.exp(x) - 1), atol=1e-6)
    a, b = torch.from_numpy(x), torch.from_numpy(x[::-1].copy())
    for layer, fn in ((Multiply(), np.multiply), (Maximum(), np.maximum), (Minimum(), np.minimum),
                      (Subtract(), np.subtract), (Average(), lambda p, q: (p + q) / 2)):
        assert np.allclose(layer([a, b]).numpy(), fn(a.numpy(), b.numpy()), atol=1e-6)
    with pytest.raises(ValueError):
        Subtract()([a, b, a])


def test_sequence_models_train_and_survive_json_and_hdf5(tmp_path):
    """A model mixing the new layers (Embedding -> Conv1D -> LSTM | GRU branches -> merge) learns a toy sequence task,
    and comes back from JSON and from a Keras-layout HDF5 file with the same predictions."""
    torch.manual_seed(3)
    rng = np.random.default_rng(3)
    tokens = rng.integers(1, 20, (256, 12))
    labels = (np.sum(tokens == 7, axis=1) > 0).astype(int)          # does the sequence contain token 7?
    y = np.eye(2, dtype=np.float32)[labels]
    inp = Input(shape=(12,))
    e = Embedding(20, 8, name='emb')(inp)
    c = LeakyReLU(0.1)(Conv1D(12, 3, padding='same', name='conv')(e))
    a = LSTM(10, name='lstm')(c)
    b = GRU(10, name='gru')(LayerNormalization(name='ln')(c))
    p = GlobalMaxPooling1D()(c)
    merged = Concatenate()([Maximum()([a, b]), p])
    out = Activation('softmax')(Dense(2, name='head')(merged))
    model = Model(inputs=inp, outputs=out)
    model.compile(Adam(learning_rate=0.01), 'categorical_crossentropy', ['acc'])
    before = model.evaluate(tokens, y)
    model.fit(tokens, y, batch_size=32, epochs=12, verbose=0)
    after = model.evaluate(tokens, y)
    assert after[0] < before[0] and after[1] > 0.9
    want = model.predict(tokens[:16])
    twin = model_from_json(model.to_json())
    twin.set_weights(model.get_weights())
    assert np.allclose(twin.predict(tokens[:16]), want, atol=1e-6)
    path = str(tmp_path / 'seq.h5')
    model.save(path)
    loaded = load_model(path)
    assert np.allclose(loaded.predict(tokens[:16]), want, atol=1e-6)
    assert [type(l).__name__ for l in loaded.layers] == [type(l).__name__ for l in model.layers]
    # SimpleRNN in a Sequential stack with return_sequences feeding a second recurrent layer
    seq = Sequential([SimpleRNN(6, return_sequences=True, input_shape=(5, 3)), GRU(4, reset_after=False), Flatten(),
                      Dense(1)])
    seq.compile(Adam(learning_rate=0.01), 'mse')
    xs = rng.normal(size=(64, 5, 3)).astype(np.float32)
    ys = xs.sum(axis=(1, 2), keepdims=False).reshape(-1, 1).astype(np.float32)
    first = seq.evaluate(xs, ys)
    seq.fit(xs, ys, batch_size=16, epochs=15, verbose=0)
    assert seq.evaluate(xs, ys) < first


def test_more_keras_losses_and_metrics_match_their_definitions():
    from elephas_b200.models import losses, metrics
    rng = np.random.default_rng(4)
    p = rng.random((6, 4)).astype(np.float32)
    p /= p.sum(1, keepdims=True)
    y = np.eye(4, dtype=np.float32)[rng.integers(0, 4, 6)]
    t = lambda a: torch.from_numpy(np.asarray(a, dtype=np.float32))  # noqa: E731
    signed = 2 * y - 1
    assert np.allclose(losses.hinge(t(y), t(p)).numpy(), np.maximum(1 - signed * p, 0).mean(1), atol=1e-6)
    assert np.allclose(losses.hinge(t(signed), t(p)).numpy(), np.maximum(1 - signed * p, 0).mean(1), atol=1e-6)
    assert np.allclose(losses.squared_hinge(t(y), t(p)).numpy(), (np.maximum(1 - signed * p, 0) ** 2).mean(1), atol=1e-6)
    pos, neg = (y * p).sum(1), ((1 - y) * p).max(1)
    assert np.allclose(losses.categorical_hinge(t(y), t(p)).numpy(), np.maximum(neg - pos + 1, 0), atol=1e-6)
    q = rng.random((6, 4)).astype(np.float32)
    q /= q.sum(1, keepdims=True)
    assert np.allclose(losses.kl_divergence(t(q), t(p)).numpy(), (q * np.log(q / p)).sum(1), atol=1e-5)
    counts = rng.poisson(3.0, (6, 4)).astype(np.float32)
    rate = rng.random((6, 4)).astype(np.float32) * 4 + 0.1
    assert np.allclose(losses.poisson(t(counts), t(rate)).numpy(), (rate - counts * np.log(rate + 1e-7)).mean(1), atol=1e-5)
    err = rng.normal(size=(6, 4)).astype(np.float32) * 2
    want = np.where(np.abs(err) <= 1, 0.5 * err ** 2, np.abs(err) - 0.5).mean(1)
    assert np.allclose(losses.huber(t(np.zeros_like(err)), t(err)).numpy(), want, atol=1e-6)
    assert losses.get('kld') is losses.kl_divergence and losses.canonical_name('huber_loss') == 'huber'
    top2 = metrics.top_k_categorical_accuracy(t(y), t(p), k=2).numpy()
    assert np.array_equal(top2, [float(y[i].argmax() in np.argsort(-p[i])[:2]) for i in range(6)])
    sparse = metrics.sparse_top_k_categorical_accuracy(t(y.argmax(1)), t(p), k=2).numpy()
    assert np.array_equal(sparse, top2)
    # they compile and train like any built-in loss
    model = Sequential([Dense(8, input_dim=5, activation='relu'), Dense(1)])
    model.compile(Adam(learning_rate=0.02), 'huber', ['mae'])
    xs = rng.normal(size=(64, 5)).astype(np.float32)
    ys = xs[:, :1] * 2
    first = model.evaluate(xs, ys)[0]
    model.fit(xs, ys, batch_size=16, epochs=20, verbose=0)
    assert model.evaluate(xs, ys)[0] < first
