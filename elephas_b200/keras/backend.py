"""The handful of ``keras.backend`` calls scripts make around model building."""
from ..models.layers import reset_name_counters
from ..models.losses import EPS as _EPS


def clear_session():
    """Start layer auto-naming over (``dense``, ``dense_1``, ... count from zero again). Models hold their own device
    memory; dropping the last reference to a model releases it."""
    reset_name_counters()


def epsilon():
    return _EPS


def floatx():
    return 'float32'


def image_data_format():
    return 'channels_last'
