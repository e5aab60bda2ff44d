"""``keras.mixed_precision``: the global dtype policy, mapped onto this package's compute dtype.

``set_global_policy('mixed_bfloat16')`` makes every model built afterwards compute in bf16 on CUDA devices with fp32
master weights and an fp32 loss (``Model.set_compute_dtype('bfloat16')``: tcgen05 GEMMs / implicit-GEMM convolutions,
``ops/cnn.py``); ``'float32'`` is the default. ``'mixed_float16'`` is not offered: B200 needs no loss scaling with bf16,
and the native kernels have no fp16 form. The reference inherits whatever policy its Keras has (no code of its own).
"""

_POLICIES = {'float32': 'float32', 'mixed_bfloat16': 'bfloat16', 'bfloat16': 'bfloat16'}
_current = 'float32'


class Policy:
    def __init__(self, name):
        if name not in _POLICIES:
            raise ValueError("dtype policy must be one of %s, got %r" % (sorted(_POLICIES), name))
        self.name = name

    @property
    def compute_dtype(self):
        return _POLICIES[self.name]

    @property
    def variable_dtype(self):
        return 'float32'   # master weights always stay fp32

    def get_config(self):
        return {'name': self.name}

    def __repr__(self):
        return '<Policy "%s">' % self.name


def set_global_policy(policy):
    global _current
    name = policy.name if isinstance(policy, Policy) else policy
    Policy(name)  # validates
    _current = name


def global_policy():
    return Policy(_current)


def compute_dtype():
    return _POLICIES[_current]
