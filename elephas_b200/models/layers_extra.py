"""More of the Keras layer vocabulary: advanced activations, merge layers, 1-D convolution / pooling, normalisation over
features, embeddings and recurrent layers.

The reference hands an arbitrary compiled Keras model to its workers (``elephas/worker.py:27-42``: ``model_from_json`` +
``model.fit``), so the layer set decides which user models can be trained at all. These layers follow the Keras 2.10
conventions for configuration keys, weight names, weight layouts and formulas (so that JSON and HDF5 files mean the same
thing on both sides); they compute with torch operators and take part in the captured training step of
``ops/torch_step.py`` like any other layer outside the native Dense / convolution engines.
"""
import numpy as np
import torch
import torch.nn.functional as F

from . import activations as _act
from .layers import LAYER_CLASSES, Layer, _conv_out, _init_config, _initialize


# ----------------------------------------------------------------------------------------------- advanced activations
class LeakyReLU(Layer):
    def __init__(self, alpha=0.3, **kwargs):
        super().__init__(**kwargs)
        self.alpha = float(alpha)

    def call(self, x, training=False):
        return F.leaky_relu(x, self.alpha)

    def _extra_config(self):
        return {'alpha': self.alpha}


class ELU(Layer):
    def __init__(self, alpha=1.0, **kwargs):
        super().__init__(**kwargs)
        self.alpha = float(alpha)

    def call(self, x, training=False):
        return F.elu(x, self.alpha)

    def _extra_config(self):
        return {'alpha': self.alpha}


# ------------------------------------------------------------------------------------------------------ merge layers
class _Merge(Layer):
    def compute_output_shape(self, input_shape):
        return tuple(input_shape[0])

    def call(self, xs, training=False):
        out = xs[0]
        for t in xs[1:]:
            out = self._combine(out, t)
        return self._finish(out, len(xs))

    def _finish(self, out, n):
        return out


class Multiply(_Merge):
    _combine = staticmethod(torch.mul)


class Subtract(_Merge):
    def call(self, xs, training=False):
        if len(xs) != 2:
            raise ValueError("A `Subtract` layer should be called on exactly 2 inputs")
        return xs[0] - xs[1]


class Maximum(_Merge):
    _combine = staticmethod(torch.maximum)


class Minimum(_Merge):
    _combine = staticmethod(torch.minimum)


class Average(_Merge):
    _combine = staticmethod(torch.add)

    def _finish(self, out, n):
        return out / float(n)


# ----------------------------------------------------------------------------------------------------------- pooling
class GlobalMaxPooling2D(Layer):
    def __init__(self, data_format=None, keepdims=False, **kwargs):
        super().__init__(**kwargs)
        self.keepdims = keepdims

    def compute_output_shape(self, input_shape):
        return (input_shape[0], input_shape[-1])

    def call(self, x, training=False):
        return x.amax(dim=(1, 2))

    def _extra_config(self):
        return {'data_format': 'channels_last', 'keepdims': self.keepdims}


class GlobalAveragePooling1D(Layer):
    def __init__(self, data_format=None, keepdims=False, **kwargs):
        super().__init__(**kwargs)
        self.keepdims = keepdims

    def compute_output_shape(self, input_shape):
        return (input_shape[0], input_shape[-1])

    def call(self, x, training=False):
        return x.mean(dim=1)

    def _extra_config(self):
        return {'data_format': 'channels_last', 'keepdims': self.keepdims}


class GlobalMaxPooling1D(GlobalAveragePooling1D):
    def call(self, x, training=False):
        return x.amax(dim=1)


class _Pool1D(Layer):
    def __init__(self, pool_size=2, strides=None, padding='valid', data_format=None, **kwargs):
        super().__init__(**kwargs)
        self.pool_size = int(pool_size[0] if isinstance(pool_size, (list, tuple)) else pool_size)
        strides = self.pool_size if strides is None else strides
        self.strides = int(strides[0] if isinstance(strides, (list, tuple)) else strides)
        self.padding = padding

    def compute_output_shape(self, input_shape):
        n, t, c = input_shape
        return (n, _conv_out(t, self.pool_size, self.strides, self.padding), c)

    def _padded(self, x, value):
        if self.padding != 'same':
            return x
        t = x.shape[1]
        total = max((-(-t // self.strides) - 1) * self.strides + self.pool_size - t, 0)
        return F.pad(x, (0, 0, total // 2, total - total // 2), value=value) if total else x

    def _extra_config(self):
        return {'pool_size': [self.pool_size], 'padding': self.padding, 'strides': [self.strides],
                'data_format': 'channels_last'}


class MaxPooling1D(_Pool1D):
    def call(self, x, training=False):
        x = self._padded(x, float('-inf'))
        return F.max_pool1d(x.transpose(1, 2), self.pool_size, self.strides).transpose(1, 2)


class AveragePooling1D(_Pool1D):
    def call(self, x, training=False):
        if self.padding == 'same':  # Keras averages over the valid elements of a border window only
            ones = self._padded(torch.ones_like(x[:1, :, :1]), 0.0)
            total = F.avg_pool1d(self._padded(x, 0.0).transpose(1, 2), self.pool_size, self.strides)
            count = F.avg_pool1d(ones.transpose(1, 2), self.pool_size, self.strides)
            return (total / count).transpose(1, 2)
        return F.avg_pool1d(x.transpose(1, 2), self.pool_size, self.strides).transpose(1, 2)


# ----------------------------------------------------------------------------------------------------- normalisation
class LayerNormalization(Layer):
    """Normalises over the last axis (the only form the Keras default uses), then ``gamma * x + beta``."""

    def __init__(self, axis=-1, epsilon=1e-3, center=True, scale=True, beta_initializer='zeros',
                 gamma_initializer='ones', beta_regularizer=None, gamma_regularizer=None, beta_constraint=None,
                 gamma_constraint=None, **kwargs):
        super().__init__(**kwargs)
        axis = axis[0] if isinstance(axis, (list, tuple)) else axis
        self.axis, self.epsilon, self.center, self.scale = int(axis), float(epsilon), bool(center), bool(scale)
        self.gamma = self.beta = None

    def build(self, input_shape):
        if self.axis not in (-1, len(input_shape) - 1):
            raise ValueError("LayerNormalization is implemented for the last axis only (got axis=%r)" % (self.axis,))
        c = int(input_shape[-1])
        if self.scale:
            self.gamma = self.add_weight('gamma', torch.ones(c))
        if self.center:
            self.beta = self.add_weight('beta', torch.zeros(c))
        self.built = True

    def call(self, x, training=False):
        return F.layer_norm(x, (x.shape[-1],), self.gamma, self.beta, self.epsilon)

    def _extra_config(self):
        return {'axis': [self.axis], 'epsilon': self.epsilon, 'center': self.center, 'scale': self.scale,
                'beta_initializer': _init_config('zeros', 'zeros'), 'gamma_initializer': _init_config('ones', 'ones'),
                'beta_regularizer': None, 'gamma_regularizer': None, 'beta_constraint': None, 'gamma_constraint': None}


# --------------------------------------------------------------------------------------------------------- embedding
class Embedding(Layer):
    """Integer indices ``[batch, time]`` -> vectors ``[batch, time, output_dim]``; weight ``embeddings`` as in Keras."""

    def __init__(self, input_dim, output_dim, embeddings_initializer='uniform', embeddings_regularizer=None,
                 activity_regularizer=None, embeddings_constraint=None, mask_zero=False, input_length=None, **kwargs):
        if input_length is not None and 'input_shape' not in kwargs and 'batch_input_shape' not in kwargs:
            kwargs['input_shape'] = (int(input_length),)
        super().__init__(**kwargs)
        self.input_dim, self.output_dim = int(input_dim), int(output_dim)
        self.mask_zero, self.input_length = bool(mask_zero), input_length
        self.embeddings = None

    def build(self, input_shape):
        table = torch.empty(self.input_dim, self.output_dim).uniform_(-0.05, 0.05)  # Keras 'uniform'
        self.embeddings = self.add_weight('embeddings', table)
        self.built = True

    def compute_output_shape(self, input_shape):
        return tuple(input_shape) + (self.output_dim,)

    def call(self, x, training=False):
        return F.embedding(x.long(), self.embeddings)

    def _extra_config(self):
        return {'input_dim': self.input_dim, 'output_dim': self.output_dim,
                'embeddings_initializer': {'class_name': 'RandomUniform', 'config': {'minval': -0.05, 'maxval': 0.05,
                                                                                     'seed': None}},
                'embeddings_regularizer': None, 'activity_regularizer': None, 'embeddings_constraint': None,
                'mask_zero': self.mask_zero, 'input_length': self.input_length}


# ------------------------------------------------------------------------------------------------------------ Conv1D
class Conv1D(Layer):
    """channels_last 1-D convolution over ``[batch, time, channels]``. The kernel is kept in the torch layout
    ``[filters, channels, width]`` and exposed in the Keras layout ``(width, channels, filters)``."""

    def __init__(self, filters, kernel_size, strides=1, padding='valid', data_format=None, dilation_rate=1, groups=1,
                 activation=None, use_bias=True, kernel_initializer='glorot_uniform', bias_initializer='zeros',
                 kernel_regularizer=None, bias_regularizer=None, activity_regularizer=None, kernel_constraint=None,
                 bias_constraint=None, custom_objects=None, **kwargs):
        super().__init__(**kwargs)
        first = lambda v: int(v[0] if isinstance(v, (list, tuple)) else v)  # noqa: E731
        self.filters, self.kernel_size, self.strides = int(filters), first(kernel_size), first(strides)
        self.dilation_rate, self.groups, self.padding = first(dilation_rate), int(groups), padding
        self.activation_name, self.activation = _act.get(activation, custom_objects)
        self.use_bias = use_bias
        self.kernel_initializer = _init_config(kernel_initializer, 'glorot_uniform')
        self.bias_initializer = _init_config(bias_initializer, 'zeros')
        self.kernel = self.bias = None

    def build(self, input_shape):
        cin = int(input_shape[-1]) // self.groups
        fan_in, fan_out = cin * self.kernel_size, self.filters * self.kernel_size
        self.kernel = self.add_weight('kernel', _initialize(torch.empty(self.filters, cin, self.kernel_size),
                                                            self.kernel_initializer, fan_in, fan_out))
        if self.use_bias:
            self.bias = self.add_weight('bias', _initialize(torch.empty(self.filters), self.bias_initializer, fan_in,
                                                            fan_out))
        self.built = True

    def to_keras_layout(self, index, tensor):
        return tensor.permute(2, 1, 0) if index == 0 else tensor

    def from_keras_layout(self, index, array):
        return np.transpose(array, (2, 1, 0)) if index == 0 else array

    def compute_output_shape(self, input_shape):
        n, t, _ = input_shape
        if self.padding == 'causal':
            return (n, None if t is None else -(-t // self.strides), self.filters)
        return (n, _conv_out(t, self.kernel_size, self.strides, self.padding, self.dilation_rate), self.filters)

    def call(self, x, training=False):
        eff = (self.kernel_size - 1) * self.dilation_rate + 1
        t = x.shape[1]
        if self.padding == 'causal':
            x = F.pad(x, (0, 0, eff - 1, 0))
        elif self.padding == 'same':
            total = max((-(-t // self.strides) - 1) * self.strides + eff - t, 0)
            if total:
                x = F.pad(x, (0, 0, total // 2, total - total // 2))
        y = F.conv1d(x.transpose(1, 2), self.kernel, self.bias, stride=self.strides, dilation=self.dilation_rate,
                     groups=self.groups)
        return self.activation(y.transpose(1, 2))

    def _extra_config(self):
        return {'filters': self.filters, 'kernel_size': [self.kernel_size], 'strides': [self.strides],
                'padding': self.padding, 'data_format': 'channels_last', 'dilation_rate': [self.dilation_rate],
                'groups': self.groups, 'activation': self.activation_name, 'use_bias': self.use_bias,
                'kernel_initializer': self.kernel_initializer, 'bias_initializer': self.bias_initializer,
                'kernel_regularizer': None, 'bias_regularizer': None, 'activity_regularizer': None,
                'kernel_constraint': None, 'bias_constraint': None}


# --------------------------------------------------------------------------------------------------------- recurrent
class _Recurrent(Layer):
    """Shared part of SimpleRNN / LSTM / GRU: weights in the Keras layouts (``kernel [in, g*units]``,
    ``recurrent_kernel [units, g*units]``, ``bias``), named below the cell like Keras 2.10 does
    (``lstm/lstm_cell/kernel:0``), a time loop over ``[batch, time, features]`` inputs."""

    gates = 1
    cell_name = 'cell'

    def __init__(self, units, activation='tanh', recurrent_activation='sigmoid', use_bias=True,
                 kernel_initializer='glorot_uniform', recurrent_initializer='orthogonal', bias_initializer='zeros',
                 return_sequences=False, return_state=False, go_backwards=False, stateful=False, unroll=False,
                 dropout=0.0, recurrent_dropout=0.0, custom_objects=None, **kwargs):
        for key in ('kernel_regularizer', 'recurrent_regularizer', 'bias_regularizer', 'activity_regularizer',
                    'kernel_constraint', 'recurrent_constraint', 'bias_constraint', 'time_major', 'implementation'):
            kwargs.pop(key, None)
        super().__init__(**kwargs)
        if return_state or stateful:
            raise ValueError("%s: return_state / stateful are not supported" % type(self).__name__)
        self.units = int(units)
        self.activation_name, self.activation = _act.get(activation, custom_objects)
        self.recurrent_activation_name, self.recurrent_activation = _act.get(recurrent_activation, custom_objects)
        self.use_bias, self.return_sequences, self.go_backwards = bool(use_bias), bool(return_sequences), bool(go_backwards)
        self.dropout, self.recurrent_dropout = float(dropout), float(recurrent_dropout)
        self.kernel_initializer = _init_config(kernel_initializer, 'glorot_uniform')
        self.kernel = self.recurrent_kernel = self.bias = None

    def _bias_shape(self):
        return (self.gates * self.units,)

    def _initial_bias(self):
        return torch.zeros(self._bias_shape())

    def build(self, input_shape):
        in_dim, u, g = int(input_shape[-1]), self.units, self.gates
        self.kernel = self.add_weight(self.cell_name + '/kernel', _initialize(torch.empty(in_dim, g * u),
                                                                             self.kernel_initializer, in_dim, g * u))
        rec = torch.empty(u, g * u)
        for k in range(g):  # Keras: one orthogonal matrix per gate
            torch.nn.init.orthogonal_(rec[:, k * u:(k + 1) * u])
        self.recurrent_kernel = self.add_weight(self.cell_name + '/recurrent_kernel', rec)
        if self.use_bias:
            self.bias = self.add_weight(self.cell_name + '/bias', self._initial_bias())
        self.built = True

    def compute_output_shape(self, input_shape):
        n, t, _ = input_shape
        return (n, t, self.units) if self.return_sequences else (n, self.units)

    def _masks(self, x, training):
        """Input / recurrent dropout masks, one per sequence and reused at every time step (as Keras does)."""
        mask_in = mask_rec = None
        if training and self.dropout > 0:
            mask_in = F.dropout(torch.ones(x.shape[0], x.shape[2], device=x.device, dtype=x.dtype), self.dropout, True)
        if training and self.recurrent_dropout > 0:
            mask_rec = F.dropout(torch.ones(x.shape[0], self.units, device=x.device, dtype=x.dtype),
                                 self.recurrent_dropout, True)
        return mask_in, mask_rec

    def call(self, x, training=False):
        steps = range(x.shape[1] - 1, -1, -1) if self.go_backwards else range(x.shape[1])
        mask_in, mask_rec = self._masks(x, training)
        state = self._initial_state(x)
        outputs = []
        for t in steps:
            xt = x[:, t] if mask_in is None else x[:, t] * mask_in
            h, state = self._step(xt, state, mask_rec)
            if self.return_sequences:
                outputs.append(h)
        return torch.stack(outputs, dim=1) if self.return_sequences else h

    def _extra_config(self):
        return {'units': self.units, 'activation': self.activation_name, 'use_bias': self.use_bias,
                'kernel_initializer': self.kernel_initializer,
                'recurrent_initializer': {'class_name': 'Orthogonal', 'config': {'gain': 1.0, 'seed': None}},
                'bias_initializer': _init_config('zeros', 'zeros'), 'return_sequences': self.return_sequences,
                'return_state': False, 'go_backwards': self.go_backwards, 'stateful': False, 'unroll': False,
                'dropout': self.dropout, 'recurrent_dropout': self.recurrent_dropout}


class SimpleRNN(_Recurrent):
    gates, cell_name = 1, 'simple_rnn_cell'

    def __init__(self, units, activation='tanh', **kwargs):
        kwargs.pop('recurrent_activation', None)
        super().__init__(units, activation=activation, **kwargs)

    def _initial_state(self, x):
        return x.new_zeros(x.shape[0], self.units)

    def _step(self, xt, h, mask_rec):
        z = xt @ self.kernel + (h if mask_rec is None else h * mask_rec) @ self.recurrent_kernel
        if self.bias is not None:
            z = z + self.bias
        h = self.activation(z)
        return h, h


class LSTM(_Recurrent):
    """Gate order i, f, c, o; ``unit_forget_bias`` starts the forget gate's bias at one."""

    gates, cell_name = 4, 'lstm_cell'

    def __init__(self, units, unit_forget_bias=True, **kwargs):
        self.unit_forget_bias = bool(unit_forget_bias)
        super().__init__(units, **kwargs)

    def _initial_bias(self):
        b = torch.zeros(4 * self.units)
        if self.unit_forget_bias:
            b[self.units:2 * self.units] = 1.0
        return b

    def _initial_state(self, x):
        return (x.new_zeros(x.shape[0], self.units), x.new_zeros(x.shape[0], self.units))

    def _step(self, xt, state, mask_rec):
        h, c = state
        z = xt @ self.kernel + (h if mask_rec is None else h * mask_rec) @ self.recurrent_kernel
        if self.bias is not None:
            z = z + self.bias
        i, f, g, o = z.chunk(4, dim=-1)
        i, f, o = self.recurrent_activation(i), self.recurrent_activation(f), self.recurrent_activation(o)
        c = f * c + i * self.activation(g)
        h = o * self.activation(c)
        return h, (h, c)

    def _extra_config(self):
        cfg = super()._extra_config()
        cfg.update({'recurrent_activation': self.recurrent_activation_name, 'unit_forget_bias': self.unit_forget_bias,
                    'implementation': 2})
        return cfg


class GRU(_Recurrent):
    """Gate order z, r, h. ``reset_after=True`` (the TF 2 default: the reset gate is applied after the recurrent matrix
    product, bias of shape ``[2, 3 * units]``) and the classic ``reset_after=False`` form."""

    gates, cell_name = 3, 'gru_cell'

    def __init__(self, units, reset_after=True, **kwargs):
        self.reset_after = bool(reset_after)
        super().__init__(units, **kwargs)

    def _bias_shape(self):
        return (2, 3 * self.units) if self.reset_after else (3 * self.units,)

    def _initial_state(self, x):
        return x.new_zeros(x.shape[0], self.units)

    def _step(self, xt, h, mask_rec):
        u = self.units
        hm = h if mask_rec is None else h * mask_rec
        xz = xt @ self.kernel
        if self.reset_after:
            hz = hm @ self.recurrent_kernel
            if self.bias is not None:
                xz, hz = xz + self.bias[0], hz + self.bias[1]
            z = self.recurrent_activation(xz[:, :u] + hz[:, :u])
            r = self.recurrent_activation(xz[:, u:2 * u] + hz[:, u:2 * u])
            cand = self.activation(xz[:, 2 * u:] + r * hz[:, 2 * u:])
        else:
            if self.bias is not None:
                xz = xz + self.bias
            z = self.recurrent_activation(xz[:, :u] + hm @ self.recurrent_kernel[:, :u])
            r = self.recurrent_activation(xz[:, u:2 * u] + hm @ self.recurrent_kernel[:, u:2 * u])
            cand = self.activation(xz[:, 2 * u:] + (r * hm) @ self.recurrent_kernel[:, 2 * u:])
        h = z * h + (1.0 - z) * cand
        return h, h

    def _extra_config(self):
        cfg = super()._extra_config()
        cfg.update({'recurrent_activation': self.recurrent_activation_name, 'reset_after': self.reset_after,
                    'implementation': 2})
        return cfg


EXTRA_LAYERS = (LeakyReLU, ELU, Multiply, Subtract, Maximum, Minimum, Average, GlobalMaxPooling2D, GlobalAveragePooling1D,
                GlobalMaxPooling1D, MaxPooling1D, AveragePooling1D, LayerNormalization, Embedding, Conv1D, SimpleRNN, LSTM,
                GRU)
LAYER_CLASSES.update({c.__name__: c for c in EXTRA_LAYERS})
