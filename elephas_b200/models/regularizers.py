"""Weight / activity penalties with the Keras config schema (``{"class_name": "L2", "config": {"l2": 0.01}}``).

A model that reaches the workers as JSON (``elephas/spark_model.py:193`` -> ``elephas/worker.py:29``) carries the
regularizers of its layers in the layer configs; ``Model`` adds the penalties to the loss it minimises and reports, as
``keras.Model.fit`` does. Penalties of activations are divided by the batch size (Keras' ``activity_regularizer`` rule).
"""
import torch


class Regularizer:
    def __call__(self, w):
        raise NotImplementedError

    def get_config(self):
        return {}

    @classmethod
    def from_config(cls, config):
        return cls(**config)


class L1L2(Regularizer):
    def __init__(self, l1=0.0, l2=0.0):
        self.l1, self.l2 = float(l1 or 0.0), float(l2 or 0.0)

    def __call__(self, w):
        w = w.float()
        total = w.new_zeros(())
        if self.l1:
            total = total + self.l1 * w.abs().sum()
        if self.l2:
            total = total + self.l2 * (w * w).sum()
        return total

    def get_config(self):
        return {'l1': self.l1, 'l2': self.l2}


class L1(L1L2):
    def __init__(self, l1=0.01, **kwargs):
        super().__init__(l1=kwargs.pop('l', l1))

    def get_config(self):
        return {'l1': self.l1}


class L2(L1L2):
    def __init__(self, l2=0.01, **kwargs):
        super().__init__(l2=kwargs.pop('l', l2))

    def get_config(self):
        return {'l2': self.l2}


class OrthogonalRegularizer(Regularizer):
    """Penalises the off-diagonal of the Gram matrix of the rows (or columns) of a 2-D weight, after L2-normalising."""

    def __init__(self, factor=0.01, mode='rows'):
        if mode not in ('rows', 'columns'):
            raise ValueError("mode must be 'rows' or 'columns'")
        self.factor, self.mode = float(factor), mode

    def __call__(self, w):
        if w.dim() != 2:
            raise ValueError("OrthogonalRegularizer needs a rank-2 weight")
        w = w.float()
        w = torch.nn.functional.normalize(w, dim=1 if self.mode == 'rows' else 0)
        gram = w @ w.t() if self.mode == 'rows' else w.t() @ w
        size = gram.shape[0]
        off = gram * (1.0 - torch.eye(size, dtype=gram.dtype, device=gram.device))
        return self.factor * 0.5 * off.abs().sum() / (size * (size - 1) / 2.0)

    def get_config(self):
        return {'factor': self.factor, 'mode': self.mode}


def l1(l1=0.01):  # noqa: A001 - the Keras spelling
    return L1(l1)


def l2(l2=0.01):
    return L2(l2)


def l1_l2(l1=0.01, l2=0.01):
    return L1L2(l1, l2)


_CLASSES = {'L1': L1, 'L2': L2, 'L1L2': L1L2, 'OrthogonalRegularizer': OrthogonalRegularizer}
_ALIASES = {'l1': L1, 'l2': L2, 'l1_l2': L1L2, 'l1l2': L1L2, 'orthogonal_regularizer': OrthogonalRegularizer}


def serialize(reg):
    if reg is None:
        return None
    return {'class_name': type(reg).__name__, 'config': reg.get_config()}


def deserialize(config, custom_objects=None):
    name = config['class_name']
    cls = (custom_objects or {}).get(name) or _CLASSES.get(name)
    if cls is None:
        raise ValueError("Unknown regularizer: %r" % (name,))
    return cls.from_config(dict(config.get('config') or {}))


def get(identifier, custom_objects=None):
    if identifier is None:
        return None
    if isinstance(identifier, dict):
        return deserialize(identifier, custom_objects)
    if isinstance(identifier, str):
        cls = (custom_objects or {}).get(identifier) or _CLASSES.get(identifier) or _ALIASES.get(identifier.lower())
        if cls is None:
            raise ValueError("Unknown regularizer: %r" % (identifier,))
        return cls() if cls is not L1L2 else L1L2(0.01, 0.01)
    if callable(identifier):
        return identifier
    raise ValueError("Could not interpret regularizer identifier: %r" % (identifier,))
