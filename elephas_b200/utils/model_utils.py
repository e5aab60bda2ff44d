"""Model-type bookkeeping for the Spark ML transformer.

Reference: ``elephas/utils/model_utils.py:9-70`` (``ModelType``, ``LossModelTypeMapper`` singleton
with ``register_loss``, ``ModelTypeEncoder`` / ``as_enum`` JSON round trip).
"""
import json
import threading
from enum import Enum


class ModelType(Enum):
    CLASSIFICATION = 1
    REGRESSION = 2


class Singleton:
    """Base class whose subclasses have exactly one instance (constructed on first call)."""
    _instances = {}
    _guard = threading.Lock()

    def __new__(cls, *args, **kwargs):
        with Singleton._guard:
            inst = Singleton._instances.get(cls)
            if inst is None:
                inst = super().__new__(cls)
                inst._singleton_ready = False
                Singleton._instances[cls] = inst
        return inst


_REGRESSION_LOSSES = (
    'mean_squared_error', 'mse', 'mean_absolute_error', 'mae', 'mean_absolute_percentage_error',
    'mean_squared_logarithmic_error', 'cosine_proximity', 'logcosh')
_CLASSIFICATION_LOSSES = (
    'binary_crossentropy', 'categorical_crossentropy', 'sparse_categorical_crossentropy')


class LossModelTypeMapper(Singleton):
    """Maps a Keras loss name (or a registered callable) to CLASSIFICATION / REGRESSION."""

    def __init__(self):
        if self._singleton_ready:
            return
        self._mapping = {name: ModelType.REGRESSION for name in _REGRESSION_LOSSES}
        self._mapping.update({name: ModelType.CLASSIFICATION for name in _CLASSIFICATION_LOSSES})
        self._singleton_ready = True

    @staticmethod
    def _key(loss):
        return loss.__name__ if callable(loss) else loss

    def get_model_type(self, loss):
        return self._mapping.get(self._key(loss))

    def register_loss(self, loss, model_type: ModelType):
        self._mapping[self._key(loss)] = model_type


class ModelTypeEncoder(json.JSONEncoder):
    """Encodes ``ModelType`` members as ``{"__enum__": "ModelType.X"}``."""

    def default(self, obj):
        if isinstance(obj, ModelType):
            return {"__enum__": str(obj)}
        return super().default(obj)


def as_enum(d):
    """``object_hook`` undoing :class:`ModelTypeEncoder`."""
    if "__enum__" in d:
        _, member = d["__enum__"].split(".")
        return getattr(ModelType, member)
    return d
