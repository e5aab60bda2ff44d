"""Keras-named activations on torch.

The reference never defines activations itself; it forwards whatever the user's Keras model uses
(``tests/conftest.py:12-18,25-27`` use relu / softmax / linear, ``tests/integration/
test_custom_models.py:16-21`` a custom callable). ``get`` resolves a name, a callable or a
``custom_objects`` entry exactly like ``keras.activations.get`` + ``custom_objects`` would.
"""
import torch
import torch.nn.functional as F

# ids shared with the CUDA epilogues (csrc/math.cuh, include/elephas_b200.h)
NATIVE_ACT_IDS = {
    'linear': 0, 'relu': 1, 'sigmoid': 2, 'tanh': 3, 'softplus': 4, 'softsign': 5, 'elu': 6, 'selu': 7,
    'exponential': 8, 'hard_sigmoid': 9,
}


def linear(x):
    return x


def relu(x):
    return torch.relu(x)


def sigmoid(x):
    return torch.sigmoid(x)


def tanh(x):
    return torch.tanh(x)


def softmax(x, axis=-1):
    return torch.softmax(x, dim=axis)


def softplus(x):
    return F.softplus(x)


def softsign(x):
    return F.softsign(x)


def elu(x, alpha=1.0):
    return F.elu(x, alpha=alpha)


def selu(x):
    return F.selu(x)


def exponential(x):
    return torch.exp(x)


def hard_sigmoid(x):
    return torch.clamp(0.2 * x + 0.5, 0.0, 1.0)


def swish(x):
    return x * torch.sigmoid(x)


def gelu(x):
    return F.gelu(x)


def relu6(x):
    return torch.clamp(x, 0.0, 6.0)


def leaky_relu(x, negative_slope=0.2):   # (the Keras 3 activation of that name; the LeakyReLU layer defaults to 0.3)
    return F.leaky_relu(x, negative_slope)


def mish(x):
    return x * torch.tanh(F.softplus(x))


def hard_silu(x):
    return x * torch.clamp(x / 6.0 + 0.5, 0.0, 1.0)


def log_softmax(x, axis=-1):
    return torch.log_softmax(x, dim=axis)


def softmax_v2(x):
    return softmax(x)


_BUILTIN = {
    'relu6': relu6, 'leaky_relu': leaky_relu, 'mish': mish, 'hard_silu': hard_silu, 'hard_swish': hard_silu,
    'log_softmax': log_softmax, 'softmax_v2': softmax_v2,
    'linear': linear, 'relu': relu, 'sigmoid': sigmoid, 'tanh': tanh, 'softmax': softmax, 'softplus': softplus,
    'softsign': softsign, 'elu': elu, 'selu': selu, 'exponential': exponential, 'hard_sigmoid': hard_sigmoid,
    'swish': swish, 'silu': swish, 'gelu': gelu,
}


def get(identifier, custom_objects=None):
    """Resolve ``identifier`` (None, name, callable) to ``(name, fn)``."""
    if identifier is None:
        return 'linear', linear
    if callable(identifier):
        name = getattr(identifier, '__name__', 'custom_activation')
        if name in _BUILTIN and _BUILTIN[name] is identifier:
            return name, identifier
        return name, identifier
    if isinstance(identifier, dict):  # {"class_name": ..., "config": ...} form
        identifier = identifier.get('config', identifier.get('class_name'))
        if isinstance(identifier, dict):
            identifier = identifier.get('name')
    if custom_objects and identifier in custom_objects:
        return identifier, custom_objects[identifier]
    if identifier in _BUILTIN:
        return identifier, _BUILTIN[identifier]
    raise ValueError("Unknown activation function: %r (pass it through custom_objects)" % (identifier,))


def serialize(name, fn):
    """Keras serialises an activation as its name (custom callables by ``__name__``)."""
    return name


def is_builtin(name, fn) -> bool:
    return name in _BUILTIN and _BUILTIN[name] is fn
