"""``keras.models``."""
from ..models.model import Model, Sequential, load_model, model_from_config, model_from_json  # noqa: F401


def clone_model(model, custom_objects=None):
    """A new model with the same architecture and freshly initialised weights (``keras.models.clone_model``)."""
    import json
    return model_from_json(json.dumps(json.loads(model.to_json())), custom_objects or getattr(model, 'custom_objects', None))
