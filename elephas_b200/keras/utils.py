"""``keras.utils``."""
from ..models.utils import normalize, pad_sequences, set_random_seed, to_categorical  # noqa: F401
