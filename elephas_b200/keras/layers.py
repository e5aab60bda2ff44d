"""``keras.layers``: every layer class of ``elephas_b200.models`` plus ``Input``."""
from ..models.layers import LAYER_CLASSES as _CLASSES
from ..models.layers import Input, Layer  # noqa: F401
from ..models import layers_extra as _extra, layers_more as _more  # noqa: F401  (register their classes)

globals().update(_CLASSES)
__all__ = sorted(_CLASSES) + ['Input', 'Layer']
