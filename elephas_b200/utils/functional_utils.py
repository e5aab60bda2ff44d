"""Element-wise arithmetic over *lists of arrays* (one entry per model weight).

Reference: ``elephas/utils/functional_utils.py:6-43`` (``add_params``, ``subtract_params``,
``get_neutral``, ``divide_by``). The reference only handles NumPy arrays on the host; here every
helper also accepts torch tensors (CPU or CUDA) and stays on the device it was given, so the
delta arithmetic of the workers never has to leave the GPU.
"""
from typing import List, Sequence

import numpy as np

try:  # torch is optional for the pure-NumPy users of this module
    import torch
except Exception:  # pragma: no cover
    torch = None


def _is_tensor(x) -> bool:
    return torch is not None and isinstance(x, torch.Tensor)


def _pairwise(left: Sequence, right: Sequence, op):
    if len(left) != len(right):
        raise ValueError("parameter lists differ in length: %d vs %d" % (len(left), len(right)))
    return [op(a, b) for a, b in zip(left, right)]


def add_params(param_list_left: Sequence, param_list_right: Sequence) -> List:
    """Add two parameter lists entry by entry."""
    return _pairwise(param_list_left, param_list_right, lambda a, b: a + b)


def subtract_params(param_list_left: Sequence, param_list_right: Sequence) -> List:
    """Subtract the right parameter list from the left one entry by entry."""
    return _pairwise(param_list_left, param_list_right, lambda a, b: a - b)


def get_neutral(array_list: Sequence) -> List:
    """Zero-valued arrays with the shapes (and dtypes / devices) of ``array_list``."""
    return [torch.zeros_like(x) if _is_tensor(x) else np.zeros_like(x) for x in array_list]


def divide_by(array_list: Sequence, num_workers: int) -> List:
    """Divide every entry by ``num_workers``."""
    return [x / num_workers for x in array_list]
