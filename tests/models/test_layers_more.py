"""Third batch of layers (``models/layers_more.py``): values against torch oracles, shapes, JSON round trips."""
import json

import numpy as np
import pytest
import torch

from elephas_b200.models import (AlphaDropout, Attention, Cropping1D, Cropping2D, Dense, Flatten, GaussianDropout,
                                 GlobalAveragePooling1D, Input, Lambda, LayerNormalization, Model, MultiHeadAttention,
                                 PReLU, Rescaling, Sequential, ThresholdedReLU, UpSampling1D, ZeroPadding1D, Add,
                                 model_from_json)
from elephas_b200.models.optimizers import Adam


def _double_plus(x, shift=0.0):
    return 2.0 * x + shift


def test_lambda_travels_by_name_through_custom_objects():
    m = Sequential([Dense(3, input_shape=(4,)), Lambda(_double_plus, arguments={'shift': 1.0}), Dense(2)])
    x = np.random.default_rng(0).standard_normal((5, 4)).astype(np.float32)
    cfg = [l for l in json.loads(m.to_json())['config']['layers'] if l['class_name'] == 'Lambda'][0]['config']
    assert cfg['function'] == '_double_plus' and cfg['arguments'] == {'shift': 1.0}
    with pytest.raises(ValueError):
        model_from_json(m.to_json())
    back = model_from_json(m.to_json(), custom_objects={'_double_plus': _double_plus})
    back.set_weights(m.get_weights())
    np.testing.assert_allclose(back.predict(x), m.predict(x), rtol=1e-6)
    w = m.get_weights()
    np.testing.assert_allclose(m.predict(x), (2.0 * (x @ w[0] + w[1]) + 1.0) @ w[2] + w[3], rtol=1e-5, atol=1e-6)
    # a function that changes the shape: the output shape is found by probing
    s = Sequential([Lambda(lambda t: t.sum(dim=1, keepdim=True), input_shape=(4,)), Dense(1)])
    assert s.output_shape == (None, 1) and s.layers[-1].kernel.shape == (1, 1)


def test_activations_and_noise_layers():
    x = torch.randn(64, 6)
    p = PReLU(alpha_initializer={'class_name': 'Constant', 'config': {'value': 0.25}})
    np.testing.assert_allclose(p(x).detach().numpy(), torch.nn.functional.prelu(x, torch.tensor(0.25)).numpy(), rtol=1e-6)
    assert p.alpha.shape == (6,)
    shared = PReLU(shared_axes=[1, 2])
    shared(torch.randn(2, 5, 5, 3))
    assert shared.alpha.shape == (1, 1, 3)
    t = ThresholdedReLU(theta=0.5)(x)
    np.testing.assert_array_equal(t.numpy(), np.where(x.numpy() > 0.5, x.numpy(), 0.0))
    big = torch.nn.functional.selu(torch.randn(200000))
    torch.manual_seed(0)
    dropped = AlphaDropout(0.3).call(big, training=True)
    assert float(dropped.mean()) == pytest.approx(float(big.mean()), abs=0.02)
    assert float(dropped.std()) == pytest.approx(float(big.std()), rel=0.03)
    assert AlphaDropout(0.3).call(big, training=False) is big
    noisy = GaussianDropout(0.2).call(torch.ones(200000), training=True)
    assert float(noisy.mean()) == pytest.approx(1.0, abs=0.01) and float(noisy.var()) == pytest.approx(0.25, rel=0.05)


def test_shape_layers():
    x = torch.arange(2 * 5 * 3, dtype=torch.float32).reshape(2, 5, 3)
    assert ZeroPadding1D((1, 2))(x).shape == (2, 8, 3) and float(ZeroPadding1D(1)(x)[0, 0].abs().sum()) == 0.0
    np.testing.assert_array_equal(Cropping1D((1, 2))(x).numpy(), x[:, 1:3].numpy())
    np.testing.assert_array_equal(UpSampling1D(2)(x)[:, :4].numpy(), x[:, [0, 0, 1, 1]].numpy())
    img = torch.randn(2, 8, 9, 3)
    np.testing.assert_array_equal(Cropping2D(((1, 2), (3, 0)))(img).numpy(), img[:, 1:6, 3:].numpy())
    assert Cropping2D(2).compute_output_shape((None, 8, 9, 3)) == (None, 4, 5, 3)
    np.testing.assert_allclose(Rescaling(1 / 255.0, offset=-0.5)(img).numpy(), img.numpy() / 255.0 - 0.5, rtol=1e-6)


def test_attention_layers_match_torch():
    torch.manual_seed(1)
    q, v, k = torch.randn(3, 4, 8), torch.randn(3, 6, 8), torch.randn(3, 6, 8)
    out = Attention()([q, v, k])
    ref = torch.softmax(q @ k.transpose(1, 2), dim=-1) @ v
    np.testing.assert_allclose(out.numpy(), ref.numpy(), rtol=1e-5, atol=1e-6)
    causal = Attention(causal=True, use_scale=True)([q[:, :4], v[:, :4]])
    assert causal.shape == (3, 4, 8)
    np.testing.assert_allclose(causal[:, 0].detach().numpy(), v[:, 0].numpy(), rtol=1e-5, atol=1e-6)   # step 0 sees only itself

    heads, key_dim, dim = 2, 4, 8
    mha = MultiHeadAttention(heads, key_dim)
    got = mha(q, v)                                # query from q, keys and values from v
    assert [tuple(w.shape) for w in mha.weights] == [(8, 2, 4), (2, 4), (8, 2, 4), (2, 4), (8, 2, 4), (2, 4), (2, 4, 8), (8,)]
    ref = torch.nn.MultiheadAttention(dim, heads, batch_first=True)
    with torch.no_grad():                          # the same weights in torch's packed layout
        packed = torch.cat([w.reshape(dim, heads * key_dim).t() for w in (mha.wq, mha.wk, mha.wv)], dim=0)
        ref.in_proj_weight.copy_(packed)
        ref.in_proj_bias.copy_(torch.cat([b.reshape(-1) for b in (mha.bq, mha.bk, mha.bv)]))
        ref.out_proj.weight.copy_(mha.wo.reshape(heads * key_dim, dim).t())
        ref.out_proj.bias.copy_(mha.bo)
    want, _ = ref(q, v, v)
    np.testing.assert_allclose(got.detach().numpy(), want.detach().numpy(), rtol=1e-4, atol=1e-5)


def test_a_small_transformer_block_trains_and_round_trips():
    rng = np.random.default_rng(0)
    x = rng.standard_normal((256, 6, 8)).astype(np.float32)
    y = np.eye(2, dtype=np.float32)[(x[:, :, 0].max(axis=1) > 1.0).astype(int)]
    torch.manual_seed(0)
    inp = Input(shape=(6, 8))
    att = MultiHeadAttention(2, 4, name='mha')([inp, inp])
    h = LayerNormalization()(Add()([inp, att]))
    out = Dense(2, activation='softmax')(GlobalAveragePooling1D()(Dense(16, activation='relu')(h)))
    m = Model(inputs=inp, outputs=out)
    m.compile(Adam(learning_rate=0.01), 'categorical_crossentropy', ['acc'])
    hist = m.fit(x, y, batch_size=32, epochs=15)
    assert hist.history['loss'][-1] < 0.8 * hist.history['loss'][0]
    back = model_from_json(m.to_json())
    back.set_weights(m.get_weights())
    np.testing.assert_allclose(back.predict(x[:16]), m.predict(x[:16]), rtol=1e-5, atol=1e-6)
    names = [n for n in back.get_layer('mha').weight_names()]
    assert names[0] == 'mha/query/kernel:0' and names[-1] == 'mha/attention_output/bias:0'


def test_zoo_sequence_models_train_through_spark_model():
    from elephas_b200.models import zoo
    from elephas_b200.spark import SparkContext
    from elephas_b200.spark_model import SparkModel
    from elephas_b200.utils.rdd_utils import to_simple_rdd
    rng = np.random.default_rng(0)
    tokens = rng.integers(1, 30, (192, 10)).astype(np.float32)
    tokens[:, 7:] *= (rng.random((192, 3)) < 0.5)            # some trailing padding
    y = np.eye(2, dtype=np.float32)[(np.sum(tokens < 10, axis=1) >= 3).astype(int)]
    sc = SparkContext.getOrCreate()
    for build in (lambda: zoo.lstm_classifier(vocab=30, maxlen=10, embed=8, units=12, bidirectional=True, dropout=0.0),
                  lambda: zoo.transformer_classifier(vocab=30, maxlen=10, embed=16, heads=2, ff=32, blocks=1, dropout=0.0)):
        torch.manual_seed(0)
        model = build()
        model.compile(Adam(learning_rate=0.02), 'categorical_crossentropy', ['acc'])
        before = model.evaluate(tokens, y, batch_size=64)[0]
        sm = SparkModel(model, mode='synchronous', num_workers=2)
        sm.fit(to_simple_rdd(sc, tokens, y), epochs=6, batch_size=16, verbose=0)
        after = sm.evaluate(tokens, y)
        assert after[0] < before and len(sm.predict(tokens[:5])) == 5
        again = model_from_json(model.to_json())
        assert again.count_params() == model.count_params()


def test_normalization_layer_adapts_and_round_trips(tmp_path):
    from elephas_b200.models import Normalization, load_model
    rng = np.random.default_rng(0)
    x = (5.0 + 3.0 * rng.standard_normal((500, 4)) * [1, 2, 3, 4]).astype(np.float32)
    norm = Normalization(input_shape=(4,))
    m = Sequential([norm, Dense(1)])
    norm.adapt(x)
    z = (x - norm.get_weights()[0]) / np.sqrt(norm.get_weights()[1])
    np.testing.assert_allclose(z.mean(0), 0.0, atol=1e-4), np.testing.assert_allclose(z.std(0), 1.0, rtol=1e-4)
    assert [tuple(w.shape) for w in norm.get_weights()] == [(4,), (4,), ()] and norm.trainable_weights == []
    w = m.get_weights()
    np.testing.assert_allclose(m.predict(x[:8]), z[:8] @ w[3] + w[4], rtol=1e-4, atol=1e-5)
    m.compile('sgd', 'mse')
    before = [a.copy() for a in m.get_weights()[:3]]
    m.fit(x, x[:, :1], batch_size=50, epochs=1)
    for a, b in zip(m.get_weights()[:3], before):      # the statistics are not trained
        np.testing.assert_array_equal(a, b)
    path = str(tmp_path / 'n.h5')
    m.save(path)
    np.testing.assert_allclose(load_model(path).predict(x[:8]), m.predict(x[:8]), rtol=1e-6)
    fixed = Sequential([Normalization(mean=[1.0, 2.0], variance=[4.0, 9.0], input_shape=(2,))])
    np.testing.assert_allclose(fixed.predict(np.array([[3.0, 8.0]], np.float32)), [[1.0, 2.0]], rtol=1e-6)
    assert model_from_json(fixed.to_json()).predict(np.array([[3.0, 8.0]], np.float32)).tolist() == [[1.0, 2.0]]
    with pytest.raises(ValueError):
        Normalization(mean=[0.0])
