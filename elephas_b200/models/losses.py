"""Keras-named losses on torch (``sum_over_batch_size`` reduction, epsilon 1e-7 as in Keras).

The reference passes ``model.loss`` strings straight to ``keras.Model.compile``
(``elephas/worker.py:32``, ``elephas/spark_model.py:193``); the names exercised by its tests are
``categorical_crossentropy``, ``binary_crossentropy``, ``mse``, ``mae``
(``tests/integration/test_end_to_end.py:40,92``, ``test_custom_models.py:24``,
``tests/test_ml_model.py:150``). Each loss returns the per-sample loss vector; callers average.
"""
import torch

EPS = 1e-7


def _match(y_true, y_pred):
    y_true = y_true.to(y_pred.dtype)
    if y_true.dim() == y_pred.dim() - 1:
        y_true = y_true.unsqueeze(-1)
    return y_true


def mean_squared_error(y_true, y_pred):
    y_true = _match(y_true, y_pred)
    return ((y_pred - y_true) ** 2).flatten(1).mean(dim=1) if y_pred.dim() > 1 else (y_pred - y_true) ** 2


def mean_absolute_error(y_true, y_pred):
    y_true = _match(y_true, y_pred)
    return (y_pred - y_true).abs().flatten(1).mean(dim=1) if y_pred.dim() > 1 else (y_pred - y_true).abs()


def mean_absolute_percentage_error(y_true, y_pred):
    y_true = _match(y_true, y_pred)
    diff = ((y_true - y_pred) / torch.clamp(y_true.abs(), min=EPS)).abs()
    return 100.0 * (diff.flatten(1).mean(dim=1) if diff.dim() > 1 else diff)


def mean_squared_logarithmic_error(y_true, y_pred):
    y_true = _match(y_true, y_pred)
    a = torch.log(torch.clamp(y_pred, min=EPS) + 1.0)
    b = torch.log(torch.clamp(y_true, min=EPS) + 1.0)
    d = (a - b) ** 2
    return d.flatten(1).mean(dim=1) if d.dim() > 1 else d


def categorical_crossentropy(y_true, y_pred):
    """``y_pred`` are probabilities; Keras renormalises, clips to [eps, 1-eps] and takes -sum(y log p)."""
    y_true = y_true.to(y_pred.dtype)
    p = y_pred / y_pred.sum(dim=-1, keepdim=True)
    p = torch.clamp(p, EPS, 1.0 - EPS)
    return -(y_true * torch.log(p)).sum(dim=-1)


def sparse_categorical_crossentropy(y_true, y_pred):
    idx = y_true.reshape(y_pred.shape[0], -1)[:, 0].long()
    p = torch.clamp(y_pred, EPS, 1.0 - EPS)
    return -torch.log(p.gather(1, idx.unsqueeze(1))).squeeze(1)


def binary_crossentropy(y_true, y_pred):
    y_true = _match(y_true, y_pred)
    p = torch.clamp(y_pred, EPS, 1.0 - EPS)
    bce = -(y_true * torch.log(p) + (1.0 - y_true) * torch.log(1.0 - p))
    return bce.flatten(1).mean(dim=1) if bce.dim() > 1 else bce


def logcosh(y_true, y_pred):
    y_true = _match(y_true, y_pred)
    d = y_pred - y_true
    v = d + torch.nn.functional.softplus(-2.0 * d) - 0.6931471805599453
    return v.flatten(1).mean(dim=1) if v.dim() > 1 else v


def cosine_proximity(y_true, y_pred):
    y_true = _match(y_true, y_pred)
    a = torch.nn.functional.normalize(y_true, dim=-1)
    b = torch.nn.functional.normalize(y_pred, dim=-1)
    return -(a * b).sum(dim=-1)


def _mean_last(v):
    return v.flatten(1).mean(dim=1) if v.dim() > 1 else v


def hinge(y_true, y_pred):
    """Labels in {-1, 1} (0 / 1 labels are converted, as Keras does): mean(max(1 - y * p, 0))."""
    y_true = _match(y_true, y_pred)
    if bool(((y_true == 0) | (y_true == 1)).all()):
        y_true = 2.0 * y_true - 1.0
    return _mean_last(torch.clamp(1.0 - y_true * y_pred, min=0.0))


def squared_hinge(y_true, y_pred):
    y_true = _match(y_true, y_pred)
    if bool(((y_true == 0) | (y_true == 1)).all()):
        y_true = 2.0 * y_true - 1.0
    return _mean_last(torch.clamp(1.0 - y_true * y_pred, min=0.0) ** 2)


def categorical_hinge(y_true, y_pred):
    y_true = y_true.to(y_pred.dtype)
    pos = (y_true * y_pred).sum(dim=-1)
    neg = ((1.0 - y_true) * y_pred).max(dim=-1).values
    return torch.clamp(neg - pos + 1.0, min=0.0)


def kl_divergence(y_true, y_pred):
    y_true = torch.clamp(y_true.to(y_pred.dtype), EPS, 1.0)
    y_pred = torch.clamp(y_pred, EPS, 1.0)
    return (y_true * torch.log(y_true / y_pred)).sum(dim=-1)


def poisson(y_true, y_pred):
    y_true = _match(y_true, y_pred)
    return _mean_last(y_pred - y_true * torch.log(y_pred + EPS))


def huber(y_true, y_pred, delta=1.0):
    y_true = _match(y_true, y_pred)
    err = (y_pred - y_true).abs()
    quad = torch.clamp(err, max=delta)
    return _mean_last(0.5 * quad ** 2 + delta * (err - quad))


_BUILTIN = {
    'hinge': hinge, 'squared_hinge': squared_hinge, 'categorical_hinge': categorical_hinge,
    'kl_divergence': kl_divergence, 'kld': kl_divergence, 'kullback_leibler_divergence': kl_divergence,
    'poisson': poisson, 'huber': huber, 'huber_loss': huber,
    'mean_squared_error': mean_squared_error, 'mse': mean_squared_error, 'MSE': mean_squared_error,
    'mean_absolute_error': mean_absolute_error, 'mae': mean_absolute_error, 'MAE': mean_absolute_error,
    'mean_absolute_percentage_error': mean_absolute_percentage_error, 'mape': mean_absolute_percentage_error,
    'mean_squared_logarithmic_error': mean_squared_logarithmic_error, 'msle': mean_squared_logarithmic_error,
    'categorical_crossentropy': categorical_crossentropy,
    'sparse_categorical_crossentropy': sparse_categorical_crossentropy,
    'binary_crossentropy': binary_crossentropy,
    'logcosh': logcosh, 'log_cosh': logcosh,
    'cosine_proximity': cosine_proximity, 'cosine_similarity': cosine_proximity,
}

CANONICAL = {
    'mse': 'mean_squared_error', 'MSE': 'mean_squared_error', 'mae': 'mean_absolute_error',
    'MAE': 'mean_absolute_error', 'mape': 'mean_absolute_percentage_error',
    'msle': 'mean_squared_logarithmic_error', 'log_cosh': 'logcosh', 'cosine_similarity': 'cosine_proximity',
    'kld': 'kl_divergence', 'kullback_leibler_divergence': 'kl_divergence', 'huber_loss': 'huber',
}


def canonical_name(identifier):
    if isinstance(identifier, dict):
        identifier = deserialize(identifier)
    if isinstance(identifier, Loss):
        return identifier.canonical_name()
    if callable(identifier):
        return getattr(identifier, '__name__', 'custom_loss')
    return CANONICAL.get(identifier, identifier)


def family_name(identifier):
    """The named function a loss belongs to, whatever its options (``SparseCategoricalCrossentropy(from_logits=True)`` ->
    ``'sparse_categorical_crossentropy'``): what ``'accuracy'`` is resolved against."""
    if isinstance(identifier, dict) and 'class_name' in identifier:
        identifier = deserialize(identifier)
    if isinstance(identifier, Loss):
        return identifier.function
    return canonical_name(identifier)


def get(identifier, custom_objects=None):
    if isinstance(identifier, dict):
        identifier = deserialize(identifier, custom_objects)
    if isinstance(identifier, Loss) and identifier.is_default:
        return _BUILTIN[identifier.function]
    if callable(identifier):
        return identifier
    if custom_objects and identifier in custom_objects:
        return custom_objects[identifier]
    if identifier in _BUILTIN:
        return _BUILTIN[identifier]
    raise ValueError("Unknown loss function: %r" % (identifier,))


# ------------------------------------------------------------------------------------------------------ loss objects
class Loss:
    """``keras.losses.Loss`` objects: ``compile(loss=CategoricalCrossentropy(from_logits=True))``. Calling one returns the
    per-sample vector like the functions above. An object with default arguments is the named function (the fused engines
    recognise it by ``canonical_name``); any other configuration runs on the autograd path."""

    function = None          # canonical name of the function a default-configured object equals
    _defaults = {}

    def __init__(self, reduction='auto', name=None, **params):
        unknown = set(params) - set(self._defaults)
        if unknown:
            raise TypeError("%s got unexpected argument(s) %s" % (type(self).__name__, sorted(unknown)))
        if reduction not in ('auto', 'sum_over_batch_size'):
            raise ValueError("only the default reduction (sum over batch size) is supported, got %r" % (reduction,))
        self.reduction, self.name = reduction, name or self.function
        self.params = dict(self._defaults)
        self.params.update(params)

    @property
    def is_default(self):
        return self.params == self._defaults

    @property
    def __name__(self):  # what History keys and saved training configs call it
        return self.name

    def canonical_name(self):
        if self.is_default:
            return self.function
        return "%s(%s)" % (type(self).__name__, ", ".join("%s=%r" % kv for kv in sorted(self.params.items())))

    def __call__(self, y_true, y_pred):
        return _BUILTIN[self.function](y_true, y_pred)

    def get_config(self):
        cfg = {'reduction': self.reduction, 'name': self.name}
        cfg.update(self.params)
        return cfg

    @classmethod
    def from_config(cls, config):
        return cls(**config)


def _smooth(y_true, amount, classes):
    return y_true * (1.0 - amount) + amount / classes if amount else y_true


class CategoricalCrossentropy(Loss):
    function = 'categorical_crossentropy'
    _defaults = {'from_logits': False, 'label_smoothing': 0.0, 'axis': -1}

    def __call__(self, y_true, y_pred):
        if self.params['axis'] not in (-1, y_pred.dim() - 1):
            raise ValueError("CategoricalCrossentropy: only the last axis is supported")
        y_true = _smooth(y_true.to(y_pred.dtype), float(self.params['label_smoothing']), y_pred.shape[-1])
        if self.params['from_logits']:
            return -(y_true * torch.log_softmax(y_pred, dim=-1)).sum(dim=-1)
        return categorical_crossentropy(y_true, y_pred)


class SparseCategoricalCrossentropy(Loss):
    function = 'sparse_categorical_crossentropy'
    _defaults = {'from_logits': False}

    def __call__(self, y_true, y_pred):
        if self.params['from_logits']:
            idx = y_true.reshape(y_pred.shape[0], -1)[:, 0].long()
            return torch.nn.functional.cross_entropy(y_pred, idx, reduction='none')
        return sparse_categorical_crossentropy(y_true, y_pred)


class BinaryCrossentropy(Loss):
    function = 'binary_crossentropy'
    _defaults = {'from_logits': False, 'label_smoothing': 0.0, 'axis': -1}

    def __call__(self, y_true, y_pred):
        y_true = _match(y_true, y_pred)
        y_true = _smooth(y_true, float(self.params['label_smoothing']), 2.0)
        if self.params['from_logits']:
            bce = torch.clamp(y_pred, min=0.0) - y_pred * y_true + torch.log1p(torch.exp(-y_pred.abs()))
            return bce.flatten(1).mean(dim=1) if bce.dim() > 1 else bce
        return binary_crossentropy(y_true, y_pred)


class Huber(Loss):
    function = 'huber'
    _defaults = {'delta': 1.0}

    def __call__(self, y_true, y_pred):
        return huber(y_true, y_pred, delta=float(self.params['delta']))


def _plain(class_name, function):
    return type(class_name, (Loss,), {'function': function, '__doc__': "``keras.losses.%s``." % class_name})


MeanSquaredError = _plain('MeanSquaredError', 'mean_squared_error')
MeanAbsoluteError = _plain('MeanAbsoluteError', 'mean_absolute_error')
MeanAbsolutePercentageError = _plain('MeanAbsolutePercentageError', 'mean_absolute_percentage_error')
MeanSquaredLogarithmicError = _plain('MeanSquaredLogarithmicError', 'mean_squared_logarithmic_error')
Hinge = _plain('Hinge', 'hinge')
SquaredHinge = _plain('SquaredHinge', 'squared_hinge')
CategoricalHinge = _plain('CategoricalHinge', 'categorical_hinge')
KLDivergence = _plain('KLDivergence', 'kl_divergence')
Poisson = _plain('Poisson', 'poisson')
LogCosh = _plain('LogCosh', 'logcosh')


class CosineSimilarity(Loss):
    function = 'cosine_proximity'
    _defaults = {'axis': -1}


_LOSS_CLASSES = {c.__name__: c for c in (
    CategoricalCrossentropy, SparseCategoricalCrossentropy, BinaryCrossentropy, Huber, MeanSquaredError, MeanAbsoluteError,
    MeanAbsolutePercentageError, MeanSquaredLogarithmicError, Hinge, SquaredHinge, CategoricalHinge, KLDivergence, Poisson,
    LogCosh, CosineSimilarity)}


def serialize(loss):
    """What ``training_config['loss']`` of a saved model holds: the name, or ``{"class_name", "config"}`` for an object."""
    if isinstance(loss, Loss):
        return {'class_name': type(loss).__name__, 'config': loss.get_config()}
    return loss if isinstance(loss, str) else getattr(loss, '__name__', None)


def deserialize(config, custom_objects=None):
    if not isinstance(config, dict):
        return config
    name = config.get('class_name')
    cls = (custom_objects or {}).get(name) or _LOSS_CLASSES.get(name)
    if cls is None:
        raise ValueError("Unknown loss class: %r" % (name,))
    return cls.from_config(dict(config.get('config') or {}))
