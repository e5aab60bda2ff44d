"""The helpers of ``keras.utils`` / ``keras.preprocessing`` that scripts around the reference's examples use to shape their
arrays before ``to_simple_rdd`` (``examples/mnist_mlp_spark.py:30-31`` one-hot encodes with ``to_categorical``)."""
import numpy as np

from .model import to_categorical  # noqa: F401


def pad_sequences(sequences, maxlen=None, dtype='int32', padding='pre', truncating='pre', value=0.0):
    """``keras.preprocessing.sequence.pad_sequences``: a list of sequences of unequal lengths -> one ``[n, maxlen]`` array
    (or ``[n, maxlen, ...]`` for sequences of vectors), padded / truncated at the front (``'pre'``) or the back."""
    if padding not in ('pre', 'post') or truncating not in ('pre', 'post'):
        raise ValueError("padding and truncating must be 'pre' or 'post'")
    lengths = [len(s) for s in sequences]
    if maxlen is None:
        maxlen = max(lengths) if lengths else 0
    tail = ()
    for s in sequences:
        if len(s):
            tail = np.asarray(s).shape[1:]
            break
    out = np.full((len(sequences), int(maxlen)) + tuple(tail), value, dtype=dtype)
    for i, s in enumerate(sequences):
        if not len(s) or not maxlen:
            continue
        s = np.asarray(s[-maxlen:] if truncating == 'pre' else s[:maxlen], dtype=dtype)
        if s.shape[1:] != tuple(tail):
            raise ValueError("sequence %d has element shape %s, expected %s" % (i, s.shape[1:], tuple(tail)))
        if padding == 'post':
            out[i, :len(s)] = s
        else:
            out[i, maxlen - len(s):] = s
    return out


def normalize(x, axis=-1, order=2):
    """``keras.utils.normalize``: scale to unit ``order``-norm along ``axis`` (zero vectors stay zero)."""
    x = np.asarray(x, dtype=np.float64 if np.asarray(x).dtype == np.float64 else np.float32)
    norms = np.atleast_1d(np.linalg.norm(x, order, axis))
    norms[norms == 0] = 1
    return x / np.expand_dims(norms, axis)


def set_random_seed(seed):
    """``keras.utils.set_random_seed``: seeds Python's, NumPy's and torch's generators (initial weights, shuffling and
    dropout masks of this package draw from the torch one)."""
    import random

    import torch
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
