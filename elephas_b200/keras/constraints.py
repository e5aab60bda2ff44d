"""``keras.constraints``."""
from ..models.constraints import (Constraint, MaxNorm, MinMaxNorm, NonNeg, UnitNorm, deserialize, get, max_norm,  # noqa: F401
                                  min_max_norm, non_neg, serialize, unit_norm)
