"""``keras.losses``: the functions and the loss classes."""
from ..models.losses import *  # noqa: F401,F403
from ..models.losses import (BinaryCrossentropy, CategoricalCrossentropy, CategoricalHinge, CosineSimilarity, Hinge,  # noqa: F401
                             Huber, KLDivergence, LogCosh, Loss, MeanAbsoluteError, MeanAbsolutePercentageError,
                             MeanSquaredError, MeanSquaredLogarithmicError, Poisson, SparseCategoricalCrossentropy,
                             SquaredHinge, deserialize, get, serialize)
