"""``keras.regularizers``."""
from ..models.regularizers import (L1, L1L2, L2, OrthogonalRegularizer, Regularizer, deserialize, get, l1, l1_l2, l2,  # noqa: F401
                                   serialize)
