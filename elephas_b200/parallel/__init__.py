"""One-rank-per-GPU runtime, fused collectives and the device-resident parameter server."""
from .runtime import Runtime, default_device, get_runtime, reset_runtime  # noqa: F401
