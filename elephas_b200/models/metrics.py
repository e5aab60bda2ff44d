"""Keras-named metrics on torch. ``'acc'`` / ``'accuracy'`` resolve against the loss and the
output shape the way ``keras.Model.compile`` does (binary / categorical / sparse categorical).

Reference use: ``['acc']`` (``tests/integration/test_end_to_end.py:40``), ``['mae',
'mean_absolute_percentage_error']`` (``:92``), ``['mae']`` (``tests/test_ml_model.py:151``).
Each metric returns a per-sample vector; callers average.
"""

from . import losses as _losses


def binary_accuracy(y_true, y_pred, threshold=0.5):
    y_true = _losses._match(y_true, y_pred)
    hit = ((y_pred > threshold).to(y_pred.dtype) == y_true).to(y_pred.dtype)
    return hit.flatten(1).mean(dim=1) if hit.dim() > 1 else hit


def categorical_accuracy(y_true, y_pred):
    return (y_true.argmax(dim=-1) == y_pred.argmax(dim=-1)).to(y_pred.dtype)


def sparse_categorical_accuracy(y_true, y_pred):
    idx = y_true.reshape(y_pred.shape[0], -1)[:, 0].long()
    return (idx == y_pred.argmax(dim=-1)).to(y_pred.dtype)


def top_k_categorical_accuracy(y_true, y_pred, k=5):
    k = min(int(k), y_pred.shape[-1])
    top = y_pred.topk(k, dim=-1).indices
    return (top == y_true.argmax(dim=-1, keepdim=True)).any(dim=-1).to(y_pred.dtype)


def sparse_top_k_categorical_accuracy(y_true, y_pred, k=5):
    k = min(int(k), y_pred.shape[-1])
    idx = y_true.reshape(y_pred.shape[0], -1)[:, :1].long()
    return (y_pred.topk(k, dim=-1).indices == idx).any(dim=-1).to(y_pred.dtype)


_BUILTIN = {
    'binary_accuracy': binary_accuracy, 'categorical_accuracy': categorical_accuracy,
    'sparse_categorical_accuracy': sparse_categorical_accuracy,
    'top_k_categorical_accuracy': top_k_categorical_accuracy,
    'sparse_top_k_categorical_accuracy': sparse_top_k_categorical_accuracy,
}


def resolve(identifier, loss_name, output_dim, custom_objects=None):
    """Return ``(display_name, fn)`` for one entry of ``compile(metrics=[...])``."""
    if callable(identifier):
        return getattr(identifier, '__name__', 'metric'), identifier
    if identifier in ('acc', 'accuracy'):
        if loss_name == 'sparse_categorical_crossentropy':
            fn = sparse_categorical_accuracy
        elif loss_name == 'binary_crossentropy' or output_dim == 1:
            fn = binary_accuracy
        else:
            fn = categorical_accuracy
        return identifier, fn
    if custom_objects and identifier in custom_objects:
        return identifier, custom_objects[identifier]
    if identifier in _BUILTIN:
        return identifier, _BUILTIN[identifier]
    if identifier in _losses._BUILTIN:
        return identifier, _losses._BUILTIN[identifier]
    raise ValueError("Unknown metric: %r" % (identifier,))
