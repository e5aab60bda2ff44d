"""Keras-compatible layers on torch.

The reference has no layer code: users hand it Keras models (``tests/conftest.py:8-40``,
``examples/*.py``) and it round-trips them through ``to_json`` / ``model_from_json`` /
``get_weights`` / ``set_weights`` (``elephas/worker.py:30-34``, ``elephas/utils/serialization.py``).
This module provides the layer vocabulary those models use (Dense / Activation / Dropout /
InputLayer, plus the convolutional set needed for the ResNet-50-shaped workload of BASELINE.json)
with Keras' config schema, weight order and weight *layouts* at the API boundary.

Internally weights live in the layout the B200 kernels want: a Dense kernel is stored
``[out, in]`` (K-major for the tcgen05 GEMM ``Y = X W^T``) and exposed as Keras' ``(in, out)``
through ``get_weights`` / ``set_weights``; Conv2D is stored OHWI (channels-last, the implicit-GEMM layout) and exposed HWIO.
"""
import math
import re
import threading
from collections import defaultdict

import numpy as np
import torch
import torch.nn.functional as F

from . import activations as _act

_name_lock = threading.Lock()
_name_counters = defaultdict(int)


def _to_snake(name: str) -> str:
    s = re.sub('(.)([A-Z][a-z0-9]+)', r'\1_\2', name)
    return re.sub('([a-z])([A-Z])', r'\1_\2', s).lower()


def unique_name(prefix: str) -> str:
    """Keras-style automatic names: ``dense``, ``dense_1``, ``dense_2`` ..."""
    with _name_lock:
        n = _name_counters[prefix]
        _name_counters[prefix] += 1
    return prefix if n == 0 else "%s_%d" % (prefix, n)


def reset_name_counters():
    with _name_lock:
        _name_counters.clear()


def _init_config(identifier, default):
    if identifier is None:
        identifier = default
    if isinstance(identifier, dict):
        return identifier
    table = {
        'glorot_uniform': 'GlorotUniform', 'glorot_normal': 'GlorotNormal', 'he_normal': 'HeNormal',
        'he_uniform': 'HeUniform', 'zeros': 'Zeros', 'ones': 'Ones', 'lecun_normal': 'LecunNormal',
        'lecun_uniform': 'LecunUniform', 'random_normal': 'RandomNormal', 'random_uniform': 'RandomUniform',
        'normal': 'RandomNormal', 'uniform': 'RandomUniform', 'truncated_normal': 'TruncatedNormal',
        'variance_scaling': 'VarianceScaling', 'orthogonal': 'Orthogonal', 'identity': 'Identity',
        'constant': 'Constant', 'zero': 'Zeros', 'one': 'Ones',
    }
    cls = table.get(identifier, identifier)
    defaults = {
        'Zeros': {}, 'Ones': {}, 'Constant': {'value': 0},
        'RandomNormal': {'mean': 0.0, 'stddev': 0.05, 'seed': None},
        'TruncatedNormal': {'mean': 0.0, 'stddev': 0.05, 'seed': None},
        'RandomUniform': {'minval': -0.05, 'maxval': 0.05, 'seed': None},
        'VarianceScaling': {'scale': 1.0, 'mode': 'fan_in', 'distribution': 'truncated_normal', 'seed': None},
        'Orthogonal': {'gain': 1.0, 'seed': None}, 'Identity': {'gain': 1.0},
    }
    return {'class_name': cls, 'config': dict(defaults.get(cls, {'seed': None}))}


def _initialize(tensor, init_cfg, fan_in, fan_out):
    cls = init_cfg['class_name']
    cfg = init_cfg.get('config', {})
    gen = None
    if cfg.get('seed') is not None:
        gen = torch.Generator().manual_seed(int(cfg['seed']))
    with torch.no_grad():
        if cls == 'Zeros':
            tensor.zero_()
        elif cls == 'Ones':
            tensor.fill_(1.0)
        elif cls in ('GlorotUniform', 'HeUniform', 'LecunUniform'):
            scale = {'GlorotUniform': 6.0 / (fan_in + fan_out), 'HeUniform': 6.0 / fan_in,
                     'LecunUniform': 3.0 / fan_in}[cls]
            tensor.uniform_(-math.sqrt(scale), math.sqrt(scale), generator=gen)
        elif cls in ('GlorotNormal', 'HeNormal', 'LecunNormal'):
            var = {'GlorotNormal': 2.0 / (fan_in + fan_out), 'HeNormal': 2.0 / fan_in,
                   'LecunNormal': 1.0 / fan_in}[cls]
            std = math.sqrt(var) / 0.87962566103423978  # Keras draws from a truncated normal
            torch.nn.init.trunc_normal_(tensor, mean=0.0, std=std, a=-2 * std, b=2 * std, generator=gen)
        elif cls == 'RandomNormal':
            tensor.normal_(cfg.get('mean', 0.0), cfg.get('stddev', 0.05), generator=gen)
        elif cls == 'RandomUniform':
            tensor.uniform_(cfg.get('minval', -0.05), cfg.get('maxval', 0.05), generator=gen)
        elif cls == 'TruncatedNormal':
            mean, std = float(cfg.get('mean', 0.0)), float(cfg.get('stddev', 0.05))
            torch.nn.init.trunc_normal_(tensor, mean=mean, std=std, a=mean - 2 * std, b=mean + 2 * std, generator=gen)
        elif cls == 'Constant':
            value = cfg.get('value', 0)
            if isinstance(value, (int, float)):
                tensor.fill_(float(value))
            else:  # a Keras-layout array: the caller's layouts differ, only shapes that match as they are can be taken
                value = torch.as_tensor(value, dtype=tensor.dtype)
                if tuple(value.shape) != tuple(tensor.shape):
                    raise ValueError("Constant initializer: value of shape %s for a weight stored as %s"
                                     % (tuple(value.shape), tuple(tensor.shape)))
                tensor.copy_(value)
        elif cls == 'VarianceScaling':
            fan = {'fan_in': fan_in, 'fan_out': fan_out, 'fan_avg': (fan_in + fan_out) / 2.0}[cfg.get('mode', 'fan_in')]
            var = float(cfg.get('scale', 1.0)) / max(1.0, fan)
            dist = cfg.get('distribution', 'truncated_normal')
            if dist in ('truncated_normal', 'normal'):  # ('normal' is the Keras 2.2 spelling of the truncated form)
                std = math.sqrt(var) / 0.87962566103423978
                torch.nn.init.trunc_normal_(tensor, mean=0.0, std=std, a=-2 * std, b=2 * std, generator=gen)
            elif dist == 'untruncated_normal':
                tensor.normal_(0.0, math.sqrt(var), generator=gen)
            elif dist == 'uniform':
                limit = math.sqrt(3.0 * var)
                tensor.uniform_(-limit, limit, generator=gen)
            else:
                raise ValueError("VarianceScaling: unknown distribution %r" % (dist,))
        elif cls == 'Orthogonal':
            if tensor.dim() < 2:
                raise ValueError("Orthogonal initializer needs a weight of rank >= 2")
            flat = torch.empty(tensor.shape[0], tensor.numel() // tensor.shape[0]).normal_(0.0, 1.0, generator=gen)
            q, r = torch.linalg.qr(flat if flat.shape[0] >= flat.shape[1] else flat.t())
            q = q * torch.sign(torch.diagonal(r))
            tensor.copy_((q if flat.shape[0] >= flat.shape[1] else q.t()).reshape(tensor.shape) * float(cfg.get('gain', 1.0)))
        elif cls == 'Identity':
            if tensor.dim() != 2:
                raise ValueError("Identity initializer needs a rank-2 weight")
            tensor.copy_(torch.eye(tensor.shape[0], tensor.shape[1]) * float(cfg.get('gain', 1.0)))
        else:
            raise ValueError("Unsupported initializer %r" % (cls,))
    return tensor


_OPTION_SUFFIXES = ('_regularizer', '_constraint')


class KerasTensor:
    """Symbolic tensor of the functional API: produced by ``Input`` or by calling a layer."""

    def __init__(self, shape, layer, node_index=0, tensor_index=0):
        self.shape = tuple(shape)
        self.layer = layer
        self.node_index = node_index
        self.tensor_index = tensor_index

    def __repr__(self):
        return "KerasTensor(shape=%s, from=%s)" % (self.shape, self.layer.name)


class Layer:
    """Base class. Subclasses define ``build``, ``call``, ``compute_output_shape`` and ``_extra_config``."""

    def __init__(self, name=None, trainable=True, dtype='float32', input_shape=None, input_dim=None,
                 batch_input_shape=None, **kwargs):
        # `<weight>_regularizer`, `<weight>_constraint`, `activity_regularizer`: every layer takes them the same way
        self._weight_options = {}
        self._take_weight_options(**{k: kwargs.pop(k) for k in list(kwargs) if k.endswith(_OPTION_SUFFIXES)})
        kwargs.pop('seed', None)  # (Keras 3 writes a seed into the config of every layer that draws random numbers)
        if kwargs:
            raise TypeError("%s got unexpected keyword argument(s) %s" % (type(self).__name__, sorted(kwargs)))
        self.name = name or unique_name(_to_snake(type(self).__name__))
        self.trainable = trainable
        self.dtype = dtype
        if batch_input_shape is None and input_shape is not None:
            batch_input_shape = (None,) + tuple(input_shape)
        if batch_input_shape is None and input_dim is not None:
            batch_input_shape = (None, int(input_dim))
        self.batch_input_shape = tuple(batch_input_shape) if batch_input_shape is not None else None
        self.built = False
        self._weights = []  # list of (keras_name, tensor, trainable) in Keras order
        self.inbound = []  # functional API: list (per call) of lists of KerasTensor
        self.outputs = []  # functional API: KerasTensor per call

    # -- regularizers / constraints -------------------------------------------
    def _take_weight_options(self, custom_objects=None, **options):
        from . import constraints, regularizers
        for key, value in options.items():
            module = regularizers if key.endswith('_regularizer') else constraints
            obj = module.get(value, custom_objects)
            if obj is not None:
                self._weight_options[key] = obj
            else:
                self._weight_options.pop(key, None)

    @property
    def activity_regularizer(self):
        return self._weight_options.get('activity_regularizer')

    def weight_options(self):
        """``[(kind, option, index, tensor)]``: the regularizers ('regularizer') and constraints ('constraint') of this
        layer's own weights, resolved to positions in ``_weights`` (``recurrent_regularizer`` -> ``recurrent_kernel``)."""
        out = []
        for key, obj in self._weight_options.items():
            target, kind = key.rsplit('_', 1)
            if target == 'activity':
                continue
            names = {target, target + '_kernel'}
            for idx, (name, tensor, _) in enumerate(self._weights):
                if name.rsplit('/', 1)[-1] in names:
                    out.append((kind, obj, idx, tensor))
        return out

    # -- freezing ------------------------------------------------------------
    @property
    def trainable(self):
        return self._trainable

    @trainable.setter
    def trainable(self, value):
        """``layer.trainable = False`` freezes the layer: its weights leave ``trainable_weights`` and stop asking autograd
        for gradients (recompile the model afterwards, as in Keras)."""
        self._trainable = bool(value)
        for _, w, t in self.__dict__.get('_weights', ()):
            if w.is_leaf:
                w.requires_grad_(bool(t and value))

    # -- symbolic ends (functional API: Model(inputs=base.input, outputs=base.get_layer('x').output)) --------------
    @property
    def output(self):
        if not self.outputs:
            raise AttributeError("layer %r has not been called on a symbolic input yet" % self.name)
        return self.outputs[0]

    @property
    def input(self):  # noqa: A003 - the Keras name
        if not self.inbound:
            raise AttributeError("layer %r has not been called on a symbolic input yet" % self.name)
        ins = self.inbound[0]
        return ins[0] if len(ins) == 1 else list(ins)

    # -- weights -------------------------------------------------------------
    def add_weight(self, name, tensor, trainable=True):
        tensor.requires_grad_(bool(trainable and self.trainable))
        self._weights.append((name, tensor, trainable))
        return tensor

    @property
    def weights(self):
        return [w for _, w, _ in self._weights]

    @property
    def trainable_weights(self):
        return [w for _, w, t in self._weights if t and self.trainable]

    @property
    def non_trainable_weights(self):
        return [w for _, w, t in self._weights if not (t and self.trainable)]

    def count_params(self):
        return int(sum(w.numel() for w in self.weights))

    def get_weights(self):
        """This layer's weights as NumPy arrays in the Keras layouts (a Dense kernel is ``(in, out)``)."""
        for model in self._live_owners():
            model._sync_from_engine()   # a native engine may hold newer values than the layer's tensors
        return [self.to_keras_layout(i, w.detach()).cpu().numpy().copy() for i, (_, w, _) in enumerate(self._weights)]

    def __getstate__(self):
        state = dict(self.__dict__)
        state.pop('_owners', None)   # weak references: the model that owns a copy registers itself again
        return state

    def _live_owners(self):
        """The models this layer is part of (weak references kept by ``Model._init_graph``)."""
        return [m for m in (ref() for ref in self.__dict__.get('_owners', ())) if m is not None]

    def set_weights(self, weights):
        if len(weights) != len(self._weights):
            raise ValueError("layer %r expects %d weight arrays, got %d" % (self.name, len(self._weights), len(weights)))
        import numpy as _np
        owners = self._live_owners()
        for model in owners:
            model._sync_from_engine()
        with torch.no_grad():
            for i, ((_, w, _), value) in enumerate(zip(self._weights, weights)):
                value = _np.asarray(self.from_keras_layout(i, _np.asarray(value, dtype=_np.float32)), order='C')
                if tuple(value.shape) != tuple(w.shape):
                    raise ValueError("layer %r: weight %d expects shape %s, got %s"
                                     % (self.name, i, tuple(self.to_keras_layout(i, w).shape), tuple(weights[i].shape)))
                w.copy_(torch.from_numpy(value))
        for model in owners:
            model._push_to_engine()

    @property
    def output_shape(self):
        return self.output.shape

    @property
    def input_shape(self):
        ins = self.input
        return [t.shape for t in ins] if isinstance(ins, list) else ins.shape

    def weight_names(self):
        return ["%s/%s:0" % (self.name, n) for n, _, _ in self._weights]

    def to_keras_layout(self, index, tensor):
        return tensor

    def from_keras_layout(self, index, array):
        return array

    # -- shapes / config -----------------------------------------------------
    def build(self, input_shape):
        self.built = True

    def compute_output_shape(self, input_shape):
        return input_shape

    def call(self, x, training=False):
        return x

    def _extra_config(self):
        return {}

    def get_config(self):
        cfg = {'name': self.name, 'trainable': self.trainable}
        if self.batch_input_shape is not None:
            cfg['batch_input_shape'] = list(self.batch_input_shape)
        cfg['dtype'] = self.dtype
        cfg.update(self._extra_config())
        from . import constraints, regularizers
        for key, obj in self._weight_options.items():
            module = regularizers if key.endswith('_regularizer') else constraints
            cfg[key] = module.serialize(obj) if hasattr(obj, 'get_config') else obj
        return cfg

    @classmethod
    def from_config(cls, config, custom_objects=None):
        cfg = dict(config)
        if 'batch_input_shape' in cfg and cfg['batch_input_shape'] is not None:
            cfg['batch_input_shape'] = tuple(cfg['batch_input_shape'])
        return cls(**cfg)

    # -- functional API ------------------------------------------------------
    def __call__(self, inputs, training=False):
        if isinstance(inputs, KerasTensor) or (isinstance(inputs, (list, tuple)) and inputs
                                                and isinstance(inputs[0], KerasTensor)):
            ins = list(inputs) if isinstance(inputs, (list, tuple)) else [inputs]
            shapes = [t.shape for t in ins]
            in_shape = shapes if len(shapes) > 1 else shapes[0]
            if not self.built:
                self.build(in_shape)
                self.built = True
            out = KerasTensor(self.compute_output_shape(in_shape), self, node_index=len(self.inbound))
            self.inbound.append(ins)
            self.outputs.append(out)
            return out
        x = inputs
        if not self.built:
            shape = [tuple(t.shape) for t in x] if isinstance(x, (list, tuple)) else tuple(x.shape)
            self.build(shape)
            self.built = True
        return self.call(x, training=training)


class InputLayer(Layer):
    def __init__(self, input_shape=None, batch_input_shape=None, sparse=False, ragged=False, **kwargs):
        kwargs.setdefault('name', None)
        name = kwargs.pop('name') or unique_name('input')
        super().__init__(name=name, input_shape=input_shape, batch_input_shape=batch_input_shape, **kwargs)
        self.sparse = sparse
        self.ragged = ragged
        self.built = True

    def _extra_config(self):
        return {'sparse': self.sparse, 'ragged': self.ragged}

    def get_config(self):
        return {'batch_input_shape': list(self.batch_input_shape), 'dtype': self.dtype, 'sparse': self.sparse,
                'ragged': self.ragged, 'name': self.name}

    @classmethod
    def from_config(cls, config, custom_objects=None):
        cfg = dict(config)
        cfg.pop('trainable', None)
        return cls(batch_input_shape=tuple(cfg.pop('batch_input_shape')), **cfg)


def Input(shape=None, batch_shape=None, name=None, dtype='float32'):
    """Functional-API entry point: returns the symbolic output of a fresh ``InputLayer``."""
    if batch_shape is None:
        batch_shape = (None,) + tuple(shape)
    layer = InputLayer(batch_input_shape=tuple(batch_shape), name=name, dtype=dtype)
    out = KerasTensor(tuple(batch_shape), layer)
    layer.inbound.append([])
    layer.outputs.append(out)
    return out


class Dense(Layer):
    def __init__(self, units, activation=None, use_bias=True, kernel_initializer='glorot_uniform',
                 bias_initializer='zeros', kernel_regularizer=None, bias_regularizer=None,
                 activity_regularizer=None, kernel_constraint=None, bias_constraint=None, custom_objects=None,
                 **kwargs):
        super().__init__(**kwargs)
        self._take_weight_options(custom_objects, kernel_regularizer=kernel_regularizer, bias_regularizer=bias_regularizer,
                                  activity_regularizer=activity_regularizer, kernel_constraint=kernel_constraint,
                                  bias_constraint=bias_constraint)
        self.units = int(units)
        self.activation_name, self.activation = _act.get(activation, custom_objects)
        self.use_bias = use_bias
        self.kernel_initializer = _init_config(kernel_initializer, 'glorot_uniform')
        self.bias_initializer = _init_config(bias_initializer, 'zeros')
        self.kernel = None
        self.bias = None

    def build(self, input_shape):
        in_dim = int(input_shape[-1])
        self.kernel = self.add_weight('kernel', _initialize(torch.empty(self.units, in_dim), self.kernel_initializer,
                                                            in_dim, self.units))
        if self.use_bias:
            self.bias = self.add_weight('bias', _initialize(torch.empty(self.units), self.bias_initializer, in_dim,
                                                            self.units))
        self.built = True

    def to_keras_layout(self, index, tensor):
        return tensor.t() if index == 0 else tensor

    def from_keras_layout(self, index, array):
        return array.T if index == 0 else array

    def compute_output_shape(self, input_shape):
        return tuple(input_shape[:-1]) + (self.units,)

    def call(self, x, training=False):
        if x.is_cuda and x.dim() == 2 and self.use_bias:
            from ..ops import cnn as _cnn
            if _cnn.usable(x) and _cnn.linear_supported(self.kernel):
                return self.activation(_cnn.linear(x, self.kernel, self.bias))  # bf16 compute: tcgen05 GEMM (K1)
        return self.activation(F.linear(x, self.kernel, self.bias))

    def _extra_config(self):
        return {'units': self.units, 'activation': self.activation_name, 'use_bias': self.use_bias,
                'kernel_initializer': self.kernel_initializer, 'bias_initializer': self.bias_initializer,
                'kernel_regularizer': None, 'bias_regularizer': None, 'activity_regularizer': None,
                'kernel_constraint': None, 'bias_constraint': None}

    @classmethod
    def from_config(cls, config, custom_objects=None):
        cfg = dict(config)
        if cfg.get('batch_input_shape') is not None:
            cfg['batch_input_shape'] = tuple(cfg['batch_input_shape'])
        return cls(custom_objects=custom_objects, **cfg)


class Activation(Layer):
    def __init__(self, activation, custom_objects=None, **kwargs):
        super().__init__(**kwargs)
        self.activation_name, self.activation = _act.get(activation, custom_objects)

    def call(self, x, training=False):
        return self.activation(x)

    def _extra_config(self):
        return {'activation': self.activation_name}

    @classmethod
    def from_config(cls, config, custom_objects=None):
        cfg = dict(config)
        if cfg.get('batch_input_shape') is not None:
            cfg['batch_input_shape'] = tuple(cfg['batch_input_shape'])
        return cls(custom_objects=custom_objects, **cfg)


class ReLU(Layer):
    def __init__(self, max_value=None, negative_slope=0.0, threshold=0.0, **kwargs):
        super().__init__(**kwargs)
        self.max_value, self.negative_slope, self.threshold = max_value, negative_slope, threshold

    def call(self, x, training=False):
        y = torch.where(x >= self.threshold, x, self.negative_slope * (x - self.threshold))
        return torch.clamp(y, max=self.max_value) if self.max_value is not None else y

    def _extra_config(self):
        return {'max_value': self.max_value, 'negative_slope': self.negative_slope, 'threshold': self.threshold}


class Softmax(Layer):
    def __init__(self, axis=-1, **kwargs):
        super().__init__(**kwargs)
        self.axis = axis

    def call(self, x, training=False):
        return torch.softmax(x, dim=self.axis)

    def _extra_config(self):
        return {'axis': self.axis}


class Dropout(Layer):
    """Inverted dropout: kept units are scaled by ``1 / (1 - rate)`` at training time."""

    def __init__(self, rate, noise_shape=None, seed=None, **kwargs):
        super().__init__(**kwargs)
        self.rate = float(rate)
        self.noise_shape = noise_shape
        self.seed = seed

    def call(self, x, training=False):
        return F.dropout(x, p=self.rate, training=True) if (training and self.rate > 0.0) else x

    def _extra_config(self):
        return {'rate': self.rate, 'noise_shape': self.noise_shape, 'seed': self.seed}


class Flatten(Layer):
    def __init__(self, data_format=None, **kwargs):
        super().__init__(**kwargs)
        self.data_format = data_format or 'channels_last'

    def compute_output_shape(self, input_shape):
        return (input_shape[0], int(np.prod([d for d in input_shape[1:]])))

    def call(self, x, training=False):
        return x.reshape(x.shape[0], -1)

    def _extra_config(self):
        return {'data_format': self.data_format}


class Reshape(Layer):
    def __init__(self, target_shape, **kwargs):
        super().__init__(**kwargs)
        self.target_shape = tuple(target_shape)

    def compute_output_shape(self, input_shape):
        return (input_shape[0],) + self.target_shape

    def call(self, x, training=False):
        return x.reshape((x.shape[0],) + self.target_shape)

    def _extra_config(self):
        return {'target_shape': list(self.target_shape)}


def _pair(v):
    return (int(v), int(v)) if isinstance(v, int) else tuple(int(i) for i in v)


def _conv_out(size, k, stride, padding, dilation=1):
    if size is None:
        return None
    if padding == 'same':
        return -(-size // stride)
    eff = (k - 1) * dilation + 1
    return (size - eff) // stride + 1


def _same_pad(x, k, stride, dilation=(1, 1)):
    """TensorFlow 'same' padding on an NHWC tensor (extra pixel goes bottom/right)."""
    h, w = x.shape[1], x.shape[2]
    pads = []
    for size, kk, s, d in ((w, k[1], stride[1], dilation[1]), (h, k[0], stride[0], dilation[0])):
        eff = (kk - 1) * d + 1
        total = max((-(-size // s) - 1) * s + eff - size, 0)
        pads += [total // 2, total - total // 2]
    return pads  # (w_left, w_right, h_top, h_bottom)


class Conv2D(Layer):
    """channels_last convolution. Kernel exposed as Keras HWIO, stored OHWI ``[filters, kh, kw, cin]``: the K-major weight
    operand of the implicit-GEMM kernels (``ops/cnn.py``) as it is, and a channels-last OIHW view for the torch ops."""

    def __init__(self, filters, kernel_size, strides=(1, 1), padding='valid', data_format=None,
                 dilation_rate=(1, 1), groups=1, activation=None, use_bias=True,
                 kernel_initializer='glorot_uniform', bias_initializer='zeros', kernel_regularizer=None,
                 bias_regularizer=None, activity_regularizer=None, kernel_constraint=None, bias_constraint=None,
                 custom_objects=None, **kwargs):
        super().__init__(**kwargs)
        self._take_weight_options(custom_objects, kernel_regularizer=kernel_regularizer, bias_regularizer=bias_regularizer,
                                  activity_regularizer=activity_regularizer, kernel_constraint=kernel_constraint,
                                  bias_constraint=bias_constraint)
        if (data_format or 'channels_last') != 'channels_last':
            raise ValueError("only data_format='channels_last' is supported")
        self.filters = int(filters)
        self.kernel_size = _pair(kernel_size)
        self.strides = _pair(strides)
        self.padding = padding
        self.dilation_rate = _pair(dilation_rate)
        self.groups = groups
        self.activation_name, self.activation = _act.get(activation, custom_objects)
        self.use_bias = use_bias
        self.kernel_initializer = _init_config(kernel_initializer, 'glorot_uniform')
        self.bias_initializer = _init_config(bias_initializer, 'zeros')
        self.kernel = None
        self.bias = None

    def build(self, input_shape):
        cin = int(input_shape[-1])
        kh, kw = self.kernel_size
        fan_in, fan_out = kh * kw * cin // self.groups, kh * kw * self.filters
        w = torch.empty(self.filters, kh, kw, cin // self.groups)
        self.kernel = self.add_weight('kernel', _initialize(w, self.kernel_initializer, fan_in, fan_out))
        if self.use_bias:
            self.bias = self.add_weight('bias', _initialize(torch.empty(self.filters), self.bias_initializer, fan_in,
                                                            fan_out))
        self.built = True

    def to_keras_layout(self, index, tensor):
        return tensor.permute(1, 2, 3, 0) if index == 0 else tensor

    def from_keras_layout(self, index, array):
        return np.transpose(array, (3, 0, 1, 2)) if index == 0 else array

    def compute_output_shape(self, input_shape):
        n, h, w, _ = input_shape
        return (n, _conv_out(h, self.kernel_size[0], self.strides[0], self.padding, self.dilation_rate[0]),
                _conv_out(w, self.kernel_size[1], self.strides[1], self.padding, self.dilation_rate[1]), self.filters)

    def call(self, x, training=False):
        pads = (0, 0, 0, 0)
        if self.padding == 'same':
            pads = tuple(_same_pad(x, self.kernel_size, self.strides, self.dilation_rate))  # (left, right, top, bottom)
        from ..ops import cnn as _cnn
        if _cnn.usable(x) and _cnn.conv2d_supported(x, self.kernel, self.strides, self.dilation_rate, self.groups, pads):
            # bf16 compute on a B200: implicit GEMM on tcgen05; 'same' padding is the TMA's out-of-bounds zero fill
            y = _cnn.conv2d(x, self.kernel, self.bias, self.strides[0], (pads[2], pads[3], pads[0], pads[1]))
            return self.activation(y)
        if any(pads):
            x = F.pad(x, (0, 0) + pads)
        y = F.conv2d(x.permute(0, 3, 1, 2), self.kernel.permute(0, 3, 1, 2), self.bias, stride=self.strides,
                     dilation=self.dilation_rate, groups=self.groups)
        return self.activation(y.permute(0, 2, 3, 1))

    def _extra_config(self):
        return {'filters': self.filters, 'kernel_size': list(self.kernel_size), 'strides': list(self.strides),
                'padding': self.padding, 'data_format': 'channels_last', 'dilation_rate': list(self.dilation_rate),
                'groups': self.groups, 'activation': self.activation_name, 'use_bias': self.use_bias,
                'kernel_initializer': self.kernel_initializer, 'bias_initializer': self.bias_initializer,
                'kernel_regularizer': None, 'bias_regularizer': None, 'activity_regularizer': None,
                'kernel_constraint': None, 'bias_constraint': None}

    @classmethod
    def from_config(cls, config, custom_objects=None):
        cfg = dict(config)
        if cfg.get('batch_input_shape') is not None:
            cfg['batch_input_shape'] = tuple(cfg['batch_input_shape'])
        return cls(custom_objects=custom_objects, **cfg)


class BatchNormalization(Layer):
    """Keras weight order: gamma, beta, moving_mean, moving_variance (the last two non-trainable)."""

    def __init__(self, axis=-1, momentum=0.99, epsilon=1e-3, center=True, scale=True,
                 beta_initializer='zeros', gamma_initializer='ones', moving_mean_initializer='zeros',
                 moving_variance_initializer='ones', beta_regularizer=None, gamma_regularizer=None,
                 beta_constraint=None, gamma_constraint=None, **kwargs):
        super().__init__(**kwargs)
        self._take_weight_options(beta_regularizer=beta_regularizer, gamma_regularizer=gamma_regularizer,
                                  beta_constraint=beta_constraint, gamma_constraint=gamma_constraint)
        self.axis = axis[0] if isinstance(axis, (list, tuple)) else axis
        self.momentum, self.epsilon, self.center, self.scale = momentum, epsilon, center, scale
        self.gamma = self.beta = self.moving_mean = self.moving_variance = None

    def build(self, input_shape):
        c = int(input_shape[self.axis])
        if self.scale:
            self.gamma = self.add_weight('gamma', torch.ones(c))
        if self.center:
            self.beta = self.add_weight('beta', torch.zeros(c))
        self.moving_mean = self.add_weight('moving_mean', torch.zeros(c), trainable=False)
        self.moving_variance = self.add_weight('moving_variance', torch.ones(c), trainable=False)
        self.built = True

    def call(self, x, training=False, residual=None, fuse_relu=False):
        """``residual`` / ``fuse_relu``: the model's executor hands a following ``Add`` / ``Activation('relu')`` to this
        layer when nothing else reads the intermediate values (one kernel instead of three passes over the tensor)."""
        nd = x.dim()
        axis = self.axis % nd
        training = training and self.trainable   # a frozen BatchNormalization runs in inference mode (TF 2 semantics)
        if axis == nd - 1 and x.is_cuda and x.shape[-1] % 8 == 0 and (training or not torch.is_grad_enabled()):
            from ..ops import cnn as _cnn
            if _cnn.usable(x):
                return _cnn.batch_norm(x, self.gamma, self.beta, self.moving_mean, self.moving_variance, self.epsilon,
                                       self.momentum, training, residual=residual, relu=fuse_relu)
        xc = x.movedim(axis, 1) if axis != 1 else x
        y = F.batch_norm(xc, self.moving_mean, self.moving_variance, self.gamma, self.beta, training=training,
                         momentum=1.0 - self.momentum, eps=self.epsilon)
        y = y.movedim(1, axis) if axis != 1 else y
        if residual is not None:
            y = y + residual
        return torch.relu(y) if fuse_relu else y

    def _extra_config(self):
        return {'axis': [self.axis], 'momentum': self.momentum, 'epsilon': self.epsilon, 'center': self.center,
                'scale': self.scale, 'beta_initializer': _init_config('zeros', 'zeros'),
                'gamma_initializer': _init_config('ones', 'ones'),
                'moving_mean_initializer': _init_config('zeros', 'zeros'),
                'moving_variance_initializer': _init_config('ones', 'ones'), 'beta_regularizer': None,
                'gamma_regularizer': None, 'beta_constraint': None, 'gamma_constraint': None}


class _Pool2D(Layer):
    def __init__(self, pool_size=(2, 2), strides=None, padding='valid', data_format=None, **kwargs):
        super().__init__(**kwargs)
        self.pool_size = _pair(pool_size)
        self.strides = _pair(strides) if strides is not None else self.pool_size
        self.padding = padding

    def compute_output_shape(self, input_shape):
        n, h, w, c = input_shape
        return (n, _conv_out(h, self.pool_size[0], self.strides[0], self.padding),
                _conv_out(w, self.pool_size[1], self.strides[1], self.padding), c)

    def _extra_config(self):
        return {'pool_size': list(self.pool_size), 'padding': self.padding, 'strides': list(self.strides),
                'data_format': 'channels_last'}


class MaxPooling2D(_Pool2D):
    def call(self, x, training=False, zero_pads=None):
        """``zero_pads`` = ((top, bottom), (left, right)) of a ZeroPadding2D folded into this layer (``Model._fusion_plan``)."""
        from ..ops import cnn as _cnn
        if x.dim() == 4 and _cnn.usable(x) and (zero_pads is None or self.padding == 'valid'):
            # native kernel (csrc/pool_kernels.cu): padding handled inside the window walk, no padded copy
            if zero_pads is not None:
                pads = (zero_pads[0][0], zero_pads[0][1], zero_pads[1][0], zero_pads[1][1])
            elif self.padding == 'same':
                l, r, t, b = _same_pad(x, self.pool_size, self.strides)
                pads = (t, b, l, r)
            else:
                pads = (0, 0, 0, 0)
            if _cnn.max_pool_supported(x, self.pool_size, self.strides, pads):
                return _cnn.max_pool2d(x, self.pool_size, self.strides[0], pads, pad_is_zero=zero_pads is not None)
        if zero_pads is not None:
            (t, b), (l, r) = zero_pads
            x = F.pad(x, (0, 0, l, r, t, b))
        if self.padding == 'same':
            pads = _same_pad(x, self.pool_size, self.strides)
            if any(pads):
                x = F.pad(x, (0, 0) + tuple(pads), value=float('-inf'))
        return F.max_pool2d(x.permute(0, 3, 1, 2), self.pool_size, self.strides).permute(0, 2, 3, 1)


class AveragePooling2D(_Pool2D):
    def call(self, x, training=False):
        return F.avg_pool2d(x.permute(0, 3, 1, 2), self.pool_size, self.strides).permute(0, 2, 3, 1)


class GlobalAveragePooling2D(Layer):
    def __init__(self, data_format=None, keepdims=False, **kwargs):
        super().__init__(**kwargs)
        self.keepdims = keepdims

    def compute_output_shape(self, input_shape):
        return (input_shape[0], input_shape[-1])

    def call(self, x, training=False):
        from ..ops import cnn as _cnn
        if x.dim() == 4 and x.shape[-1] % 8 == 0 and _cnn.usable(x):
            return _cnn.global_avg_pool(x)  # native kernel, fp32 accumulation (csrc/pool_kernels.cu)
        return x.mean(dim=(1, 2))

    def _extra_config(self):
        return {'data_format': 'channels_last', 'keepdims': self.keepdims}


class ZeroPadding2D(Layer):
    def __init__(self, padding=(1, 1), data_format=None, **kwargs):
        super().__init__(**kwargs)
        if isinstance(padding, int):
            padding = ((padding, padding), (padding, padding))
        elif isinstance(padding[0], int):
            padding = ((padding[0], padding[0]), (padding[1], padding[1]))
        self.padding = tuple(tuple(int(v) for v in p) for p in padding)

    def compute_output_shape(self, input_shape):
        n, h, w, c = input_shape
        (t, b), (l, r) = self.padding
        return (n, None if h is None else h + t + b, None if w is None else w + l + r, c)

    def call(self, x, training=False):
        (t, b), (l, r) = self.padding
        return F.pad(x, (0, 0, l, r, t, b))

    def _extra_config(self):
        return {'padding': [list(p) for p in self.padding], 'data_format': 'channels_last'}


class Add(Layer):
    def compute_output_shape(self, input_shape):
        return input_shape[0]

    def call(self, xs, training=False):
        out = xs[0]
        for t in xs[1:]:
            out = out + t
        return out


class Concatenate(Layer):
    def __init__(self, axis=-1, **kwargs):
        super().__init__(**kwargs)
        self.axis = axis

    def compute_output_shape(self, input_shape):
        shape = list(input_shape[0])
        shape[self.axis] = sum(s[self.axis] for s in input_shape)
        return tuple(shape)

    def call(self, xs, training=False):
        return torch.cat(list(xs), dim=self.axis)

    def _extra_config(self):
        return {'axis': self.axis}


LAYER_CLASSES = {c.__name__: c for c in (
    InputLayer, Dense, Activation, ReLU, Softmax, Dropout, Flatten, Reshape, Conv2D, BatchNormalization,
    MaxPooling2D, AveragePooling2D, GlobalAveragePooling2D, ZeroPadding2D, Add, Concatenate)}


def deserialize_layer(spec, custom_objects=None):
    cls_name = spec['class_name']
    if custom_objects and cls_name in custom_objects:
        cls = custom_objects[cls_name]
    elif cls_name in LAYER_CLASSES:
        cls = LAYER_CLASSES[cls_name]
    else:
        raise ValueError("Unknown layer class %r (pass it through custom_objects)" % (cls_name,))
    return cls.from_config(spec['config'], custom_objects=custom_objects)
