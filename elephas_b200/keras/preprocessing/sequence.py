"""``keras.preprocessing.sequence``."""
from ...models.utils import pad_sequences  # noqa: F401
