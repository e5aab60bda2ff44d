"""Weight constraints with the Keras config schema; applied to a weight right after every optimizer step.

``axis`` refers to the KERAS layout of the weight (a Dense kernel is ``(in, out)``: ``axis=0`` bounds the norm of the
weights coming into each unit); ``Model`` hands a Keras-layout view of its own storage to the constraint.
"""
import torch

_EPS = 1e-7


def _norms(w, axis):
    axis = tuple(axis) if isinstance(axis, (list, tuple)) else (int(axis),)
    return torch.sqrt((w * w).sum(dim=axis, keepdim=True))


class Constraint:
    def __call__(self, w):
        return w

    def get_config(self):
        return {}

    @classmethod
    def from_config(cls, config):
        return cls(**config)


class MaxNorm(Constraint):
    def __init__(self, max_value=2, axis=0):
        self.max_value, self.axis = float(max_value), axis

    def __call__(self, w):
        norms = _norms(w, self.axis)
        return w * (norms.clamp(0.0, self.max_value) / (_EPS + norms))

    def get_config(self):
        return {'max_value': self.max_value, 'axis': self.axis}


class NonNeg(Constraint):
    def __call__(self, w):
        return w * (w >= 0).to(w.dtype)


class UnitNorm(Constraint):
    def __init__(self, axis=0):
        self.axis = axis

    def __call__(self, w):
        return w / (_EPS + _norms(w, self.axis))

    def get_config(self):
        return {'axis': self.axis}


class MinMaxNorm(Constraint):
    def __init__(self, min_value=0.0, max_value=1.0, rate=1.0, axis=0):
        self.min_value, self.max_value, self.rate, self.axis = float(min_value), float(max_value), float(rate), axis

    def __call__(self, w):
        norms = _norms(w, self.axis)
        desired = self.rate * norms.clamp(self.min_value, self.max_value) + (1.0 - self.rate) * norms
        return w * (desired / (_EPS + norms))

    def get_config(self):
        return {'min_value': self.min_value, 'max_value': self.max_value, 'rate': self.rate, 'axis': self.axis}


max_norm, non_neg, unit_norm, min_max_norm = MaxNorm, NonNeg, UnitNorm, MinMaxNorm

_CLASSES = {'MaxNorm': MaxNorm, 'NonNeg': NonNeg, 'UnitNorm': UnitNorm, 'MinMaxNorm': MinMaxNorm}
_ALIASES = {'max_norm': MaxNorm, 'maxnorm': MaxNorm, 'non_neg': NonNeg, 'nonneg': NonNeg, 'unit_norm': UnitNorm,
            'unitnorm': UnitNorm, 'min_max_norm': MinMaxNorm, 'minmaxnorm': MinMaxNorm}


def serialize(constraint):
    if constraint is None:
        return None
    return {'class_name': type(constraint).__name__, 'config': constraint.get_config()}


def deserialize(config, custom_objects=None):
    name = config['class_name']
    cls = (custom_objects or {}).get(name) or _CLASSES.get(name)
    if cls is None:
        raise ValueError("Unknown constraint: %r" % (name,))
    return cls.from_config(dict(config.get('config') or {}))


def get(identifier, custom_objects=None):
    if identifier is None:
        return None
    if isinstance(identifier, dict):
        return deserialize(identifier, custom_objects)
    if isinstance(identifier, str):
        cls = (custom_objects or {}).get(identifier) or _CLASSES.get(identifier) or _ALIASES.get(identifier.lower())
        if cls is None:
            raise ValueError("Unknown constraint: %r" % (identifier,))
        return cls()
    if callable(identifier):
        return identifier
    raise ValueError("Could not interpret constraint identifier: %r" % (identifier,))
