"""Model <-> dictionary helpers (reference: ``elephas/utils/serialization.py:6-25``)."""
from typing import Any, Dict, Optional


def model_to_dict(model) -> Dict[str, Any]:
    """``{'model': <architecture JSON>, 'weights': <list of arrays>}`` - key order matters to callers."""
    return dict(model=model.to_json(), weights=model.get_weights())


def dict_to_model(_dict: Dict[str, Any], custom_objects: Optional[Dict[str, Any]] = None):
    """Rebuild the model described by :func:`model_to_dict`."""
    from ..models import model_from_json
    model = model_from_json(_dict['model'], custom_objects)
    model.set_weights(_dict['weights'])
    return model
