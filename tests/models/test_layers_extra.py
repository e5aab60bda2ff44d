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


def test_second_batch_of_layers_and_device_moves(tmp_path):
    from elephas_b200.models import BatchNormalization, Conv2D
    from elephas_b200.models.layers_extra import (Bidirectional, Conv2DTranspose, DepthwiseConv2D, Dot, GaussianNoise,
                                                  Permute, RepeatVector, SeparableConv2D, SpatialDropout1D,
                                                  SpatialDropout2D, TimeDistributed, UpSampling2D)
    rng = np.random.default_rng(6)
    img = rng.normal(size=(2, 7, 6, 3)).astype(np.float32)
    # depthwise: per-channel correlation, Keras kernel (kh, kw, channels, multiplier), output channel c * m + j
    dw = DepthwiseConv2D(3, depth_multiplier=2, padding='same', strides=2)
    got = _run(dw, img)
    k = dw.to_keras_layout(0, dw.depthwise_kernel.detach()).numpy()
    b = dw.bias.detach().numpy()
    assert k.shape == (3, 3, 3, 2) and got.shape == (2, 4, 3, 6) == tuple(dw.compute_output_shape((2, 7, 6, 3)))
    xp = np.pad(img, ((0, 0), (1, 1), (0, 1), (0, 0)))          # TF 'same' for 7 -> 4 (stride 2): 1 + 1; for 6 -> 3: 0 + 1
    want = np.zeros((2, 4, 3, 6), np.float32)
    for p in range(4):
        for q in range(3):
            patch = xp[:, 2 * p:2 * p + 3, 2 * q:2 * q + 3, :]   # [n, 3, 3, c]
            want[:, p, q, :] = np.einsum('nhwc,hwcm->ncm', patch, k).reshape(2, 6) + b
    assert np.allclose(got, want, atol=1e-5)
    assert np.array_equal(dw.from_keras_layout(0, k), dw.depthwise_kernel.detach().numpy())
    sep = SeparableConv2D(5, 3, padding='valid')
    got = _run(sep, img)
    kd = sep.to_keras_layout(0, sep.depthwise_kernel.detach()).numpy()
    kp = sep.to_keras_layout(1, sep.pointwise_kernel.detach()).numpy()
    assert kp.shape == (1, 1, 3, 5) and got.shape == (2, 5, 4, 5)
    mid = np.zeros((2, 5, 4, 3), np.float32)
    for p in range(5):
        for q in range(4):
            mid[:, p, q, :] = np.einsum('nhwc,hwcm->ncm', img[:, p:p + 3, q:q + 3, :], kd).reshape(2, 3)
    assert np.allclose(got, mid @ kp[0, 0] + sep.bias.detach().numpy(), atol=1e-5)
    assert np.array_equal(sep.from_keras_layout(1, kp), sep.pointwise_kernel.detach().numpy())
    # transposed convolution: the adjoint of the forward convolution with the same kernel (TF output sizes)
    for padding, k_, s_ in (('valid', 3, 2), ('same', 3, 2), ('same', 4, 2), ('valid', 2, 3), ('same', 3, 1)):
        up = Conv2DTranspose(4, k_, strides=s_, padding=padding, use_bias=False)
        small = rng.normal(size=(2, 4, 5, 3)).astype(np.float32)
        y = _run(up, small)
        out_shape = tuple(up.compute_output_shape((2, 4, 5, 3)))
        assert y.shape == out_shape
        if k_ >= s_:
            # <conv_T(x), z> == <x, conv(z)> with conv = Conv2D over the transposed-kernel roles
            z = rng.normal(size=y.shape).astype(np.float32)
            kern = up.to_keras_layout(0, up.kernel.detach()).numpy()                 # (kh, kw, filters, cin)
            fwd = Conv2D(3, k_, strides=s_, padding=padding, use_bias=False)
            fwd.build(z.shape)
            fwd.kernel.data.copy_(torch.from_numpy(np.transpose(kern, (3, 0, 1, 2))))  # OHWI with O = cin, I = filters
            back = _run(fwd, z)
            assert back.shape == small.shape
            assert np.isclose((y * z).sum(), (small * back).sum(), rtol=1e-4)
    assert np.array_equal(_run(UpSampling2D((2, 3)), img), img.repeat(2, axis=1).repeat(3, axis=2))
    assert _run(UpSampling2D(2, interpolation='bilinear'), img).shape == (2, 14, 12, 3)
    seq = rng.normal(size=(3, 5, 4)).astype(np.float32)
    assert np.array_equal(_run(Permute((2, 1)), seq), seq.transpose(0, 2, 1))
    assert np.array_equal(_run(RepeatVector(3), seq[:, 0]), np.repeat(seq[:, :1], 3, axis=1))
    a, b2 = torch.from_numpy(seq[:, 0]), torch.from_numpy(seq[:, 1])
    assert np.allclose(Dot()([a, b2]).numpy()[:, 0], (seq[:, 0] * seq[:, 1]).sum(-1), atol=1e-6)
    cos = Dot(normalize=True)([a, b2]).numpy()[:, 0]
    assert np.allclose(cos, (seq[:, 0] * seq[:, 1]).sum(-1) / np.linalg.norm(seq[:, 0], axis=1) / np.linalg.norm(seq[:, 1], axis=1),
                       atol=1e-6)
    torch.manual_seed(0)
    dropped = SpatialDropout1D(0.5)(torch.ones(4, 6, 8), training=True).numpy()
    assert set(np.unique(dropped)) <= {0.0, 2.0} and all(len(np.unique(dropped[n, :, c])) == 1 for n in range(4) for c in range(8))
    assert np.array_equal(SpatialDropout2D(0.5)(torch.ones(2, 3, 3, 4), training=False).numpy(), np.ones((2, 3, 3, 4)))
    assert np.array_equal(GaussianNoise(0.5)(torch.ones(2, 3), training=False).numpy(), np.ones((2, 3)))
    assert not np.array_equal(GaussianNoise(0.5)(torch.ones(2, 3), training=True).numpy(), np.ones((2, 3)))
    # wrappers: TimeDistributed(Dense) == Dense on every step; Bidirectional = forward | backward halves
    td = TimeDistributed(Dense(3))
    got = _run(td, seq)
    kern, bias = [w.detach().numpy() for w in td.weights]
    assert np.allclose(got, seq @ kern.T + bias, atol=1e-6)
    bi = Bidirectional(LSTM(4, return_sequences=True))
    got = _run(bi, seq)
    assert got.shape == (3, 5, 8) and len(bi.weights) == 6
    assert np.allclose(got[..., :4], _run(bi.forward, seq), atol=1e-6)
    assert np.allclose(got[..., 4:], _run(bi.backward, seq)[:, ::-1], atol=1e-6)
    assert bi.backward.go_backwards and not bi.forward.go_backwards
    assert bi.weight_names()[0].endswith('forward_' + bi.forward.name[len('forward_'):] + '/lstm_cell/kernel:0')
    # a model with wrappers: JSON / HDF5 round trip, training, and a device move that re-binds every weight reference
    torch.manual_seed(2)
    model = Sequential([Bidirectional(GRU(5, return_sequences=True), merge_mode='sum', input_shape=(5, 4)),
                        TimeDistributed(Dense(3, activation='relu')), GlobalMaxPooling1D(), Dense(2), Activation('softmax')])
    model.compile(Adam(learning_rate=0.02), 'categorical_crossentropy', ['acc'])
    xs = rng.normal(size=(96, 5, 4)).astype(np.float32)
    ys = np.eye(2, dtype=np.float32)[(xs[:, :, 0].sum(1) > 0).astype(int)]
    first = model.evaluate(xs, ys)[0]
    model.fit(xs, ys, batch_size=16, epochs=10, verbose=0)
    assert model.evaluate(xs, ys)[0] < first
    want = model.predict(xs[:8])
    twin = model_from_json(model.to_json())
    twin.set_weights(model.get_weights())
    assert np.allclose(twin.predict(xs[:8]), want, atol=1e-6)
    path = str(tmp_path / 'wrapped.h5')
    model.save(path)
    assert np.allclose(load_model(path).predict(xs[:8]), want, atol=1e-6)
    cnn = Sequential([SeparableConv2D(4, 3, padding='same', input_shape=(7, 6, 3)), BatchNormalization(),
                      Conv2DTranspose(2, 2, strides=2), DepthwiseConv2D(3), GlobalMaxPooling2D(), Dense(2)])
    assert cnn.predict(img).shape == (2, 2)
    for m in (model, cnn):
        m.to('meta')    # no data, but every reference must follow: the weight lists AND the attributes the layers use
        stack = list(m.layers)
        while stack:
            layer = stack.pop()
            stack.extend(v for v in vars(layer).values() if hasattr(v, '_weights') and v is not layer)
            assert all(w.device.type == 'meta' for w in layer.weights)
            assert all(v.device.type == 'meta' for v in vars(layer).values() if isinstance(v, torch.Tensor))
