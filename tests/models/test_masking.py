"""Masks from ``Embedding(mask_zero=True)`` / ``Masking`` reach the recurrent layers (``Model._forward``): a padded batch
computes, row by row, what the unpadded sequences compute on their own."""
import numpy as np
import pytest
import torch

from elephas_b200.models import (GRU, LSTM, Bidirectional, Dense, Dropout, Embedding, GlobalAveragePooling1D, Masking,
                                 Sequential, SimpleRNN, model_from_json)


def _padded_tokens(rng, rows=6, steps=7, vocab=11, left=False):
    lengths = rng.integers(1, steps + 1, rows)
    lengths[0] = steps
    x = np.zeros((rows, steps), dtype=np.float32)
    for r, n in enumerate(lengths):
        seq = rng.integers(1, vocab, n)
        if left:
            x[r, steps - n:] = seq
        else:
            x[r, :n] = seq
    return x, lengths


@pytest.mark.parametrize('cell', [SimpleRNN, LSTM, GRU])
@pytest.mark.parametrize('left', [False, True])
def test_padded_rows_equal_their_unpadded_sequences(cell, left):
    rng = np.random.default_rng(0)
    x, lengths = _padded_tokens(rng, left=left)
    m = Sequential([Embedding(11, 5, mask_zero=True, input_length=7), Dropout(0.0), cell(4)])
    out = m.predict(x)
    for r, n in enumerate(lengths):
        seq = x[r, 7 - n:] if left else x[r, :n]
        alone = m.predict(seq[None, :])          # the same weights on the bare sequence (any length works)
        np.testing.assert_allclose(out[r], alone[0], rtol=1e-5, atol=1e-6)
    # without the mask the padding steps do change the state
    plain = Sequential([Embedding(11, 5, input_length=7), cell(4)])
    plain.set_weights(m.get_weights())
    short = int(np.argmin(lengths))
    assert not np.allclose(plain.predict(x)[short], out[short], atol=1e-6)


def test_masked_steps_repeat_the_previous_output_and_bidirectional_uses_the_mask():
    rng = np.random.default_rng(1)
    x, lengths = _padded_tokens(rng)
    seq = Sequential([Embedding(11, 5, mask_zero=True, input_length=7), LSTM(3, return_sequences=True)])
    out = seq.predict(x)
    for r, n in enumerate(lengths):
        for t in range(n, 7):
            np.testing.assert_array_equal(out[r, t], out[r, n - 1])
    bi = Sequential([Embedding(11, 5, mask_zero=True, input_length=7), Bidirectional(GRU(3)), Dense(2)])
    got = bi.predict(x)
    for r, n in enumerate(lengths):
        np.testing.assert_allclose(got[r], bi.predict(x[r:r + 1, :n])[0], rtol=1e-5, atol=1e-6)


def test_masking_layer_average_pooling_and_json():
    rng = np.random.default_rng(2)
    x = rng.standard_normal((5, 6, 3)).astype(np.float32)
    x[1, 4:] = 0.0
    x[3, 2:] = 0.0
    m = Sequential([Masking(mask_value=0.0, input_shape=(6, 3)), LSTM(4, return_sequences=True), GlobalAveragePooling1D()])
    out = m.predict(x)
    np.testing.assert_allclose(out[1], m.predict(x[1:2, :4])[0], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(out[3], m.predict(x[3:4, :2])[0], rtol=1e-5, atol=1e-6)
    back = model_from_json(m.to_json())
    back.set_weights(m.get_weights())
    np.testing.assert_allclose(back.predict(x), out, rtol=1e-6)
    assert [type(l).__name__ for l in back.layers][-3:] == ['Masking', 'LSTM', 'GlobalAveragePooling1D']


def test_a_masked_model_trains():
    rng = np.random.default_rng(3)
    x, lengths = _padded_tokens(rng, rows=256, steps=8, vocab=12)
    y = np.eye(2, dtype=np.float32)[(np.array([row[:n].mean() for row, n in zip(x, lengths)]) > 6).astype(int)]
    torch.manual_seed(0)
    m = Sequential([Embedding(12, 8, mask_zero=True, input_length=8), GRU(12), Dense(2, activation='softmax')])
    m.compile('adam', 'categorical_crossentropy', ['acc'])
    h = m.fit(x, y, batch_size=32, epochs=12)
    assert h.history['loss'][-1] < h.history['loss'][0] and h.history['acc'][-1] > 0.75
