"""``keras.utils``."""
from ..models.utils import normalize, pad_sequences, to_categorical  # noqa: F401
