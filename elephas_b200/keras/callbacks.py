"""``keras.callbacks``."""
from ..models.callbacks import (Callback, CSVLogger, EarlyStopping, LambdaCallback, LearningRateScheduler,  # noqa: F401
                                ModelCheckpoint, ReduceLROnPlateau, TerminateOnNaN)
from ..models.model import History  # noqa: F401
