"""More of the Keras layer vocabulary: advanced activations, merge layers, 1-D convolution / pooling, normalisation over
features, embeddings, recurrent layers and their wrappers (Bidirectional, TimeDistributed), depthwise / separable /
transposed convolutions, up-sampling, shape and noise layers.

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
        return (input_shape[0], 1, input_shape[-1]) if self.keepdims else (input_shape[0], input_shape[-1])

    accepts_mask = True

    def compute_mask(self, x, mask):
        return None

    def call(self, x, training=False, mask=None):
        if mask is None:
            return x.mean(dim=1, keepdim=bool(self.keepdims))
        m = mask.to(x.dtype).unsqueeze(-1)   # mean over the steps that are not masked
        return (x * m).sum(dim=1, keepdim=bool(self.keepdims)) / m.sum(dim=1, keepdim=bool(self.keepdims)).clamp(min=1.0)

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
        self._take_weight_options(beta_regularizer=beta_regularizer, gamma_regularizer=gamma_regularizer,
                                  beta_constraint=beta_constraint, gamma_constraint=gamma_constraint)
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
        self._take_weight_options(embeddings_regularizer=embeddings_regularizer, activity_regularizer=activity_regularizer,
                                  embeddings_constraint=embeddings_constraint)
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

    def compute_mask(self, x, mask):
        return (x != 0) if self.mask_zero else None

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
        self._take_weight_options(custom_objects, kernel_regularizer=kernel_regularizer, bias_regularizer=bias_regularizer,
                                  activity_regularizer=activity_regularizer, kernel_constraint=kernel_constraint,
                                  bias_constraint=bias_constraint)
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
def _from_config_with_custom_objects(cls, config, custom_objects=None):
    cfg = dict(config)
    if cfg.get('batch_input_shape') is not None:
        cfg['batch_input_shape'] = tuple(cfg['batch_input_shape'])
    return cls(custom_objects=custom_objects, **cfg)


Conv1D.from_config = classmethod(_from_config_with_custom_objects)


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
        for key in ('time_major', 'implementation', 'seed', 'zero_output_for_mask', 'use_cudnn'):
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

    accepts_mask = True

    def compute_mask(self, x, mask):
        return mask if self.return_sequences else None

    def _input_bias(self):
        return self.bias

    def call(self, x, training=False, mask=None):
        """``mask`` (``[batch, time]``, from ``Embedding(mask_zero=True)`` / ``Masking``): a masked step leaves the state
        untouched and repeats the previous output, as Keras does."""
        steps = range(x.shape[1] - 1, -1, -1) if self.go_backwards else range(x.shape[1])
        mask_in, mask_rec = self._masks(x, training)
        # the input half of every step in ONE product over [batch * time, features]; the loop keeps the recurrent half
        xz = (x if mask_in is None else x * mask_in.unsqueeze(1)) @ self.kernel
        bias = self._input_bias()
        if bias is not None:
            xz = xz + bias
        state = self._initial_state(x)
        h = x.new_zeros(x.shape[0], self.units)
        outputs = []
        for t in steps:
            h_new, state_new = self._step(xz[:, t], state, mask_rec)
            if mask is None:
                h, state = h_new, state_new
            else:
                keep = mask[:, t].reshape(-1, 1).to(torch.bool)
                h = torch.where(keep, h_new, h)
                state = tuple(torch.where(keep, n, o) for n, o in zip(state_new, state)) \
                    if isinstance(state, tuple) else torch.where(keep, state_new, state)
            if self.return_sequences:
                outputs.append(h)
        return torch.stack(outputs, dim=1) if self.return_sequences else h

    from_config = classmethod(_from_config_with_custom_objects)

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

    def _step(self, xz, h, mask_rec):
        h = self.activation(xz + (h if mask_rec is None else h * mask_rec) @ self.recurrent_kernel)
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

    def _step(self, xz, state, mask_rec):
        h, c = state
        z = xz + (h if mask_rec is None else h * mask_rec) @ self.recurrent_kernel
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

    def _input_bias(self):
        if self.bias is None:
            return None
        return self.bias[0] if self.reset_after else self.bias

    def _step(self, xz, h, mask_rec):
        u = self.units
        hm = h if mask_rec is None else h * mask_rec
        if self.reset_after:
            hz = hm @ self.recurrent_kernel
            if self.bias is not None:
                hz = hz + self.bias[1]
            z = self.recurrent_activation(xz[:, :u] + hz[:, :u])
            r = self.recurrent_activation(xz[:, u:2 * u] + hz[:, u:2 * u])
            cand = self.activation(xz[:, 2 * u:] + r * hz[:, 2 * u:])
        else:
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


# ------------------------------------------------------------------------------- more convolutions and resampling
def _tf_same_pads(size, k, stride, dilation=1):
    eff = (k - 1) * dilation + 1
    total = max((-(-size // stride) - 1) * stride + eff - size, 0)
    return total // 2, total - total // 2


class DepthwiseConv2D(Layer):
    """One filter stack per input channel. Weight ``depthwise_kernel`` in Keras is ``(kh, kw, channels, multiplier)``; it is
    kept as the grouped-convolution weight ``[channels * multiplier, 1, kh, kw]`` (same output channel order)."""

    def __init__(self, kernel_size, strides=(1, 1), padding='valid', depth_multiplier=1, data_format=None,
                 dilation_rate=(1, 1), activation=None, use_bias=True, depthwise_initializer='glorot_uniform',
                 bias_initializer='zeros', custom_objects=None, **kwargs):
        kwargs.pop('groups', None)
        super().__init__(**kwargs)
        pair = lambda v: (int(v), int(v)) if isinstance(v, int) else tuple(int(i) for i in v)  # noqa: E731
        self.kernel_size, self.strides, self.dilation_rate = pair(kernel_size), pair(strides), pair(dilation_rate)
        self.padding, self.depth_multiplier, self.use_bias = padding, int(depth_multiplier), bool(use_bias)
        self.activation_name, self.activation = _act.get(activation, custom_objects)
        self.depthwise_initializer = _init_config(depthwise_initializer, 'glorot_uniform')
        self.depthwise_kernel = self.bias = None

    def build(self, input_shape):
        self.channels = cin = int(input_shape[-1])
        kh, kw = self.kernel_size
        w = torch.empty(cin * self.depth_multiplier, 1, kh, kw)
        self.depthwise_kernel = self.add_weight('depthwise_kernel', _initialize(w, self.depthwise_initializer, kh * kw,
                                                                                kh * kw * self.depth_multiplier))
        if self.use_bias:
            self.bias = self.add_weight('bias', torch.zeros(cin * self.depth_multiplier))
        self.built = True

    def to_keras_layout(self, index, tensor):
        if index != 0:
            return tensor
        kh, kw = self.kernel_size
        return tensor.reshape(self.channels, self.depth_multiplier, kh, kw).permute(2, 3, 0, 1)

    def from_keras_layout(self, index, array):
        if index != 0:
            return array
        kh, kw, c, m = array.shape
        return np.transpose(array, (2, 3, 0, 1)).reshape(c * m, 1, kh, kw)

    def compute_output_shape(self, input_shape):
        n, h, w, c = input_shape
        return (n, _conv_out(h, self.kernel_size[0], self.strides[0], self.padding, self.dilation_rate[0]),
                _conv_out(w, self.kernel_size[1], self.strides[1], self.padding, self.dilation_rate[1]),
                c * self.depth_multiplier)

    def _depthwise(self, x):
        if self.padding == 'same':
            t, b = _tf_same_pads(x.shape[1], self.kernel_size[0], self.strides[0], self.dilation_rate[0])
            l, r = _tf_same_pads(x.shape[2], self.kernel_size[1], self.strides[1], self.dilation_rate[1])
            if t or b or l or r:
                x = F.pad(x, (0, 0, l, r, t, b))
        return F.conv2d(x.permute(0, 3, 1, 2), self.depthwise_kernel, None, stride=self.strides,
                        dilation=self.dilation_rate, groups=x.shape[-1])

    def call(self, x, training=False):
        y = self._depthwise(x)
        if self.bias is not None:
            y = y + self.bias.view(1, -1, 1, 1)
        return self.activation(y.permute(0, 2, 3, 1))

    def _extra_config(self):
        return {'kernel_size': list(self.kernel_size), 'strides': list(self.strides), 'padding': self.padding,
                'data_format': 'channels_last', 'dilation_rate': list(self.dilation_rate), 'groups': 1,
                'activation': self.activation_name, 'use_bias': self.use_bias, 'depth_multiplier': self.depth_multiplier,
                'depthwise_initializer': self.depthwise_initializer, 'bias_initializer': _init_config('zeros', 'zeros'),
                'depthwise_regularizer': None, 'bias_regularizer': None, 'activity_regularizer': None,
                'depthwise_constraint': None, 'bias_constraint': None}

    @classmethod
    def from_config(cls, config, custom_objects=None):
        cfg = dict(config)
        if cfg.get('batch_input_shape') is not None:
            cfg['batch_input_shape'] = tuple(cfg['batch_input_shape'])
        return cls(custom_objects=custom_objects, **cfg)


class SeparableConv2D(DepthwiseConv2D):
    """Depthwise convolution followed by a 1 x 1 convolution (``pointwise_kernel`` ``(1, 1, channels * multiplier,
    filters)`` in Keras, kept as ``[filters, channels * multiplier]``)."""

    def __init__(self, filters, kernel_size, pointwise_initializer='glorot_uniform', **kwargs):
        super().__init__(kernel_size, **kwargs)
        self.filters = int(filters)
        self.pointwise_initializer = _init_config(pointwise_initializer, 'glorot_uniform')
        self.pointwise_kernel = None

    def build(self, input_shape):
        self.channels = cin = int(input_shape[-1])
        kh, kw = self.kernel_size
        mid = cin * self.depth_multiplier
        self.depthwise_kernel = self.add_weight('depthwise_kernel', _initialize(
            torch.empty(mid, 1, kh, kw), self.depthwise_initializer, kh * kw, kh * kw * self.depth_multiplier))
        self.pointwise_kernel = self.add_weight('pointwise_kernel', _initialize(
            torch.empty(self.filters, mid), self.pointwise_initializer, mid, self.filters))
        if self.use_bias:
            self.bias = self.add_weight('bias', torch.zeros(self.filters))
        self.built = True

    def to_keras_layout(self, index, tensor):
        if index == 1:
            return tensor.t().reshape(1, 1, tensor.shape[1], tensor.shape[0])
        return super().to_keras_layout(index, tensor)

    def from_keras_layout(self, index, array):
        if index == 1:
            return np.ascontiguousarray(array.reshape(array.shape[2], array.shape[3]).T)
        return super().from_keras_layout(index, array)

    def compute_output_shape(self, input_shape):
        return tuple(super().compute_output_shape(input_shape)[:3]) + (self.filters,)

    def call(self, x, training=False):
        y = self._depthwise(x).permute(0, 2, 3, 1)
        return self.activation(F.linear(y, self.pointwise_kernel, self.bias))

    def _extra_config(self):
        cfg = super()._extra_config()
        cfg.update({'filters': self.filters, 'pointwise_initializer': self.pointwise_initializer,
                    'pointwise_regularizer': None, 'pointwise_constraint': None})
        return cfg


class Conv2DTranspose(Layer):
    """Transposed convolution with TensorFlow's output geometry ('valid': ``in * s + max(k - s, 0)``, 'same':
    ``in * s``). Keras kernel ``(kh, kw, filters, channels)``, kept as torch's ``[channels, filters, kh, kw]``."""

    def __init__(self, filters, kernel_size, strides=(1, 1), padding='valid', output_padding=None, data_format=None,
                 dilation_rate=(1, 1), activation=None, use_bias=True, kernel_initializer='glorot_uniform',
                 bias_initializer='zeros', custom_objects=None, **kwargs):
        kwargs.pop('groups', None)
        super().__init__(**kwargs)
        if output_padding is not None or tuple(np.atleast_1d(dilation_rate)) not in ((1,), (1, 1)):
            raise ValueError("Conv2DTranspose: output_padding / dilation are not supported")
        pair = lambda v: (int(v), int(v)) if isinstance(v, int) else tuple(int(i) for i in v)  # noqa: E731
        self.filters, self.kernel_size, self.strides, self.padding = int(filters), pair(kernel_size), pair(strides), padding
        self.activation_name, self.activation = _act.get(activation, custom_objects)
        self.use_bias = bool(use_bias)
        self.kernel_initializer = _init_config(kernel_initializer, 'glorot_uniform')
        self.kernel = self.bias = None

    def build(self, input_shape):
        cin = int(input_shape[-1])
        kh, kw = self.kernel_size
        self.kernel = self.add_weight('kernel', _initialize(torch.empty(cin, self.filters, kh, kw), self.kernel_initializer,
                                                            kh * kw * cin, kh * kw * self.filters))
        if self.use_bias:
            self.bias = self.add_weight('bias', torch.zeros(self.filters))
        self.built = True

    def to_keras_layout(self, index, tensor):
        return tensor.permute(2, 3, 1, 0) if index == 0 else tensor

    def from_keras_layout(self, index, array):
        return np.transpose(array, (3, 2, 0, 1)) if index == 0 else array

    def _out(self, size, k, s):
        if size is None:
            return None
        return size * s if self.padding == 'same' else size * s + max(k - s, 0)

    def compute_output_shape(self, input_shape):
        n, h, w, _ = input_shape
        return (n, self._out(h, self.kernel_size[0], self.strides[0]), self._out(w, self.kernel_size[1], self.strides[1]),
                self.filters)

    def call(self, x, training=False):
        full = F.conv_transpose2d(x.permute(0, 3, 1, 2), self.kernel, self.bias, stride=self.strides)
        out = []
        for dim, (size, k, s) in enumerate(zip(x.shape[1:3], self.kernel_size, self.strides)):
            want = self._out(size, k, s)
            start = max(k - s, 0) // 2 if self.padding == 'same' else 0
            out.append((start, want))
        (h0, hh), (w0, ww) = out
        full = full[:, :, h0:h0 + hh, w0:w0 + ww]
        pad_h, pad_w = hh - full.shape[2], ww - full.shape[3]
        if pad_h > 0 or pad_w > 0:  # stride larger than the kernel: the tail of the output sees no input
            full = F.pad(full, (0, max(pad_w, 0), 0, max(pad_h, 0)))
            if self.bias is not None:  # ... but still the bias
                if pad_h > 0:
                    full[:, :, -pad_h:, :] = self.bias.view(1, -1, 1, 1)
                if pad_w > 0:
                    full[:, :, :, -pad_w:] = self.bias.view(1, -1, 1, 1)
        return self.activation(full.permute(0, 2, 3, 1))

    def _extra_config(self):
        return {'filters': self.filters, 'kernel_size': list(self.kernel_size), 'strides': list(self.strides),
                'padding': self.padding, 'data_format': 'channels_last', 'dilation_rate': [1, 1], 'groups': 1,
                'activation': self.activation_name, 'use_bias': self.use_bias, 'kernel_initializer': self.kernel_initializer,
                'bias_initializer': _init_config('zeros', 'zeros'), 'kernel_regularizer': None, 'bias_regularizer': None,
                'activity_regularizer': None, 'kernel_constraint': None, 'bias_constraint': None, 'output_padding': None}

    from_config = classmethod(DepthwiseConv2D.from_config.__func__)


class UpSampling2D(Layer):
    def __init__(self, size=(2, 2), data_format=None, interpolation='nearest', **kwargs):
        super().__init__(**kwargs)
        self.size = (int(size), int(size)) if isinstance(size, int) else tuple(int(v) for v in size)
        if interpolation not in ('nearest', 'bilinear'):
            raise ValueError("UpSampling2D: interpolation must be 'nearest' or 'bilinear'")
        self.interpolation = interpolation

    def compute_output_shape(self, input_shape):
        n, h, w, c = input_shape
        return (n, None if h is None else h * self.size[0], None if w is None else w * self.size[1], c)

    def call(self, x, training=False):
        if self.interpolation == 'nearest':
            return x.repeat_interleave(self.size[0], dim=1).repeat_interleave(self.size[1], dim=2)
        y = F.interpolate(x.permute(0, 3, 1, 2), scale_factor=tuple(float(v) for v in self.size), mode='bilinear',
                          align_corners=False)
        return y.permute(0, 2, 3, 1)

    def _extra_config(self):
        return {'size': list(self.size), 'data_format': 'channels_last', 'interpolation': self.interpolation}


# ------------------------------------------------------------------------------------------ shape and noise layers
class Permute(Layer):
    def __init__(self, dims, **kwargs):
        super().__init__(**kwargs)
        self.dims = tuple(int(d) for d in dims)

    def compute_output_shape(self, input_shape):
        return (input_shape[0],) + tuple(input_shape[d] for d in self.dims)

    def call(self, x, training=False):
        return x.permute((0,) + self.dims)

    def _extra_config(self):
        return {'dims': list(self.dims)}


class RepeatVector(Layer):
    def __init__(self, n, **kwargs):
        super().__init__(**kwargs)
        self.n = int(n)

    def compute_output_shape(self, input_shape):
        return (input_shape[0], self.n, input_shape[1])

    def call(self, x, training=False):
        return x.unsqueeze(1).expand(-1, self.n, -1)

    def _extra_config(self):
        return {'n': self.n}


class _ChannelDropout(Layer):
    """Drops whole feature maps (one decision per sample and channel), scaled like Dropout."""

    def __init__(self, rate, data_format=None, noise_shape=None, seed=None, **kwargs):
        super().__init__(**kwargs)
        self.rate = float(rate)

    def call(self, x, training=False):
        if not training or self.rate <= 0.0:
            return x
        shape = (x.shape[0],) + (1,) * (x.dim() - 2) + (x.shape[-1],)
        return x * F.dropout(torch.ones(shape, dtype=x.dtype, device=x.device), self.rate, True)

    def _extra_config(self):
        return {'rate': self.rate, 'noise_shape': None, 'seed': None}


class SpatialDropout1D(_ChannelDropout):
    pass


class SpatialDropout2D(_ChannelDropout):
    def _extra_config(self):
        return dict(super()._extra_config(), data_format='channels_last')


class GaussianNoise(Layer):
    def __init__(self, stddev, seed=None, **kwargs):
        super().__init__(**kwargs)
        self.stddev = float(stddev)

    def call(self, x, training=False):
        return x + torch.randn_like(x) * self.stddev if training else x

    def _extra_config(self):
        return {'stddev': self.stddev, 'seed': None}


class Dot(Layer):
    """Batch-wise dot product of two ``[batch, features]`` inputs (``axes=1`` / ``-1``), optionally L2-normalised."""

    def __init__(self, axes=-1, normalize=False, **kwargs):
        super().__init__(**kwargs)
        self.axes, self.normalize = axes, bool(normalize)

    def compute_output_shape(self, input_shape):
        return (input_shape[0][0], 1)

    def call(self, xs, training=False):
        a, b = xs
        if a.dim() != 2 or b.dim() != 2:
            raise ValueError("Dot is implemented for 2-D inputs")
        if self.normalize:
            a, b = F.normalize(a, dim=-1), F.normalize(b, dim=-1)
        return (a * b).sum(dim=-1, keepdim=True)

    def _extra_config(self):
        return {'axes': self.axes, 'normalize': self.normalize}


# ------------------------------------------------------------------------------------------------------- wrappers
class _Wrapper(Layer):
    """A layer that owns other layers: their weights are listed under this layer's name, as Keras does."""

    def _adopt(self, inner, prefix=''):
        for name, tensor, trainable in inner._weights:
            self._weights.append((prefix + name, tensor, trainable))

    @staticmethod
    def _inner_from(spec, custom_objects=None):
        from .layers import deserialize_layer
        return spec if isinstance(spec, Layer) else deserialize_layer(spec, custom_objects)

    @staticmethod
    def _spec(inner):
        return {'class_name': type(inner).__name__, 'config': inner.get_config()}


class TimeDistributed(_Wrapper):
    """Applies ``layer`` to every time step of ``[batch, time, ...]`` (one call on the flattened batch)."""

    def __init__(self, layer, custom_objects=None, **kwargs):
        super().__init__(**kwargs)
        self.layer = self._inner_from(layer, custom_objects)

    def build(self, input_shape):
        inner_shape = (input_shape[0],) + tuple(input_shape[2:])
        if not self.layer.built:
            self.layer.build(inner_shape)
            self.layer.built = True
        self._adopt(self.layer)
        self.built = True

    def to_keras_layout(self, index, tensor):
        return self.layer.to_keras_layout(index, tensor)

    def from_keras_layout(self, index, array):
        return self.layer.from_keras_layout(index, array)

    def compute_output_shape(self, input_shape):
        inner = self.layer.compute_output_shape((input_shape[0],) + tuple(input_shape[2:]))
        return (input_shape[0], input_shape[1]) + tuple(inner[1:])

    def call(self, x, training=False):
        b, t = x.shape[0], x.shape[1]
        y = self.layer.call(x.reshape((b * t,) + tuple(x.shape[2:])), training=training)
        return y.reshape((b, t) + tuple(y.shape[1:]))

    def _extra_config(self):
        return {'layer': self._spec(self.layer)}

    @classmethod
    def from_config(cls, config, custom_objects=None):
        cfg = dict(config)
        if cfg.get('batch_input_shape') is not None:
            cfg['batch_input_shape'] = tuple(cfg['batch_input_shape'])
        return cls(custom_objects=custom_objects, **cfg)


class Bidirectional(_Wrapper):
    """Runs a recurrent layer forwards and a copy of it backwards and merges the two (``concat`` / ``sum`` / ``mul`` /
    ``ave``); with ``return_sequences`` the backward outputs are put back into forward time order first."""

    def __init__(self, layer, merge_mode='concat', weights=None, backward_layer=None, custom_objects=None, **kwargs):
        super().__init__(**kwargs)
        if merge_mode not in ('concat', 'sum', 'mul', 'ave'):
            raise ValueError("Bidirectional: merge_mode must be one of concat / sum / mul / ave")
        self.merge_mode = merge_mode
        self.forward = self._inner_from(layer, custom_objects)
        if not isinstance(self.forward, _Recurrent):
            raise ValueError("Bidirectional wraps SimpleRNN / LSTM / GRU layers")
        if backward_layer is not None:
            self.backward = self._inner_from(backward_layer, custom_objects)
        else:
            cfg = self.forward.get_config()
            cfg['go_backwards'] = not self.forward.go_backwards
            cfg['name'] = 'backward_' + self.forward.name
            self.backward = type(self.forward).from_config(cfg, custom_objects=custom_objects)
        if not self.forward.name.startswith('forward_'):
            self.forward.name = 'forward_' + self.forward.name

    def build(self, input_shape):
        for inner in (self.forward, self.backward):
            if not inner.built:
                inner.build(input_shape)
                inner.built = True
            self._adopt(inner, prefix=inner.name + '/')
        self.built = True

    def compute_output_shape(self, input_shape):
        shape = list(self.forward.compute_output_shape(input_shape))
        if self.merge_mode == 'concat':
            shape[-1] *= 2
        return tuple(shape)

    accepts_mask = True

    def compute_mask(self, x, mask):
        return mask if self.forward.return_sequences else None

    def call(self, x, training=False, mask=None):
        a = self.forward.call(x, training=training, mask=mask)
        b = self.backward.call(x, training=training, mask=mask)
        if self.forward.return_sequences:
            b = b.flip(1)
        if self.merge_mode == 'concat':
            return torch.cat([a, b], dim=-1)
        if self.merge_mode == 'sum':
            return a + b
        if self.merge_mode == 'mul':
            return a * b
        return (a + b) / 2.0

    def _extra_config(self):
        fwd = self._spec(self.forward)
        fwd['config']['name'] = self.forward.name[len('forward_'):] if self.forward.name.startswith('forward_') \
            else self.forward.name
        return {'layer': fwd, 'merge_mode': self.merge_mode}

    from_config = classmethod(TimeDistributed.from_config.__func__)


class Masking(Layer):
    """Marks the time steps whose features all equal ``mask_value``: they are zeroed here and skipped by the recurrent
    layers downstream (``Model._forward`` carries the mask along)."""

    def __init__(self, mask_value=0.0, **kwargs):
        super().__init__(**kwargs)
        self.mask_value = float(mask_value)

    def compute_mask(self, x, mask):
        return (x != self.mask_value).any(dim=-1)

    def call(self, x, training=False):
        return x * self.compute_mask(x, None).unsqueeze(-1).to(x.dtype)

    def _extra_config(self):
        return {'mask_value': self.mask_value}


EXTRA_LAYERS = (Masking, LeakyReLU, ELU, Multiply, Subtract, Maximum, Minimum, Average, GlobalMaxPooling2D, GlobalAveragePooling1D,
                GlobalMaxPooling1D, MaxPooling1D, AveragePooling1D, LayerNormalization, Embedding, Conv1D, SimpleRNN, LSTM,
                GRU, DepthwiseConv2D, SeparableConv2D, Conv2DTranspose, UpSampling2D, Permute, RepeatVector,
                SpatialDropout1D, SpatialDropout2D, GaussianNoise, Dot, TimeDistributed, Bidirectional)
LAYER_CLASSES.update({c.__name__: c for c in EXTRA_LAYERS})
