"""``keras.preprocessing``."""
from . import sequence  # noqa: F401
