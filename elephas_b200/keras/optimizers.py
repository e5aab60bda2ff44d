"""``keras.optimizers`` (+ ``keras.optimizers.schedules``)."""
from ..models.optimizers import (SGD, Adadelta, Adagrad, Adam, Adamax, AdamW, Nadam, Optimizer, RMSprop, deserialize,  # noqa: F401
                                 get, serialize)
from ..models import schedules  # noqa: F401
