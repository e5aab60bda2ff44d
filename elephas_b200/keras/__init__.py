"""``tensorflow.keras``-shaped import paths over ``elephas_b200.models``: a script written against Keras changes the prefix
of its imports and nothing else.

    from tensorflow.keras.models import Sequential        ->  from elephas_b200.keras.models import Sequential
    from tensorflow.keras.layers import Dense, Dropout    ->  from elephas_b200.keras.layers import Dense, Dropout
    from tensorflow.keras import optimizers, callbacks    ->  from elephas_b200.keras import optimizers, callbacks

The reference has no model code of its own: every script it runs builds a ``tensorflow.keras`` model first
(``examples/mnist_mlp_spark.py:1-10``). These modules are thin re-exports; the implementation lives in ``models/``.
"""
from . import (backend, callbacks, constraints, initializers, layers, losses, metrics, models, optimizers, preprocessing,  # noqa: F401
               regularizers, utils)
from ..models import mixed_precision  # noqa: F401
from .layers import Input  # noqa: F401
from .models import Model, Sequential  # noqa: F401
