"""Third part of the Keras layer vocabulary: ``Lambda``, learnable / thresholded activations, the dropout variants that keep
SELU statistics or multiply by noise, 1-D padding / cropping / up-sampling, 2-D cropping, input scaling, and attention
(``Attention``, ``MultiHeadAttention`` with the Keras weight layouts).

Same conventions as ``layers_extra.py`` (Keras 2.10 config keys, weight names and layouts); see there for why the layer set
matters to the reference's workers (``elephas/worker.py:27-42``).
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

from .layers import LAYER_CLASSES, Layer, _init_config, _initialize


# ------------------------------------------------------------------------------------------------------------ Lambda
class Lambda(Layer):
    """A user function as a layer. Keras pickles the function's byte code into the model JSON; here the JSON carries the
    function's NAME and the function itself travels in ``custom_objects`` (``SparkModel(..., custom_objects={...})``, as
    the reference does for custom activations: ``tests/integration/test_custom_models.py``)."""

    def __init__(self, function, output_shape=None, arguments=None, custom_objects=None, **kwargs):
        kwargs.pop('mask', None), kwargs.pop('function_type', None), kwargs.pop('output_shape_type', None)
        kwargs.pop('module', None), kwargs.pop('output_shape_module', None)
        super().__init__(**kwargs)
        if isinstance(function, str):
            if not custom_objects or function not in custom_objects:
                raise ValueError("Lambda: function %r must be passed through custom_objects" % (function,))
            self.function_name, self.function = function, custom_objects[function]
        elif callable(function):
            self.function_name, self.function = getattr(function, '__name__', 'lambda'), function
        else:
            raise ValueError("Lambda: function must be a callable or the name of one in custom_objects")
        self.output_shape_arg = tuple(output_shape) if output_shape is not None else None
        self.arguments = dict(arguments or {})

    def compute_output_shape(self, input_shape):
        if self.output_shape_arg is not None:
            return (input_shape[0],) + self.output_shape_arg
        dims = lambda shape: tuple(1 if d is None else int(d) for d in shape[1:])  # noqa: E731 (free axes probed with length 1)
        if isinstance(input_shape, list):   # run the function on one zero row of every input to learn the shape
            probe = [torch.zeros((1,) + dims(s)) for s in input_shape]
        else:
            probe = torch.zeros((1,) + dims(input_shape))
        out = self.function(probe, **self.arguments)
        return (None,) + tuple(out.shape[1:])

    def call(self, x, training=False):
        return self.function(x, **self.arguments)

    def _extra_config(self):
        return {'function': self.function_name, 'output_shape': list(self.output_shape_arg) if self.output_shape_arg else None,
                'arguments': self.arguments}

    @classmethod
    def from_config(cls, config, custom_objects=None):
        cfg = dict(config)
        if cfg.get('batch_input_shape') is not None:
            cfg['batch_input_shape'] = tuple(cfg['batch_input_shape'])
        if isinstance(cfg.get('function'), (list, tuple)):
            raise ValueError("Lambda: this file carries pickled byte code; re-create the layer with a named function "
                             "and pass it through custom_objects")
        return cls(custom_objects=custom_objects, **cfg)


# ------------------------------------------------------------------------------------------------------- activations
class PReLU(Layer):
    """``max(0, x) + alpha * min(0, x)`` with a learned ``alpha`` per feature (shared over ``shared_axes``)."""

    def __init__(self, alpha_initializer='zeros', alpha_regularizer=None, alpha_constraint=None, shared_axes=None, **kwargs):
        super().__init__(**kwargs)
        self._take_weight_options(alpha_regularizer=alpha_regularizer, alpha_constraint=alpha_constraint)
        self.alpha_initializer = _init_config(alpha_initializer, 'zeros')
        self.shared_axes = None if shared_axes is None else [int(a) for a in np.atleast_1d(shared_axes)]
        self.alpha = None

    def build(self, input_shape):
        shape = [int(d) for d in input_shape[1:]]
        for axis in self.shared_axes or []:
            shape[axis - 1] = 1
        self.alpha = self.add_weight('alpha', _initialize(torch.empty(shape), self.alpha_initializer, 1, 1))
        self.built = True

    def call(self, x, training=False):
        return torch.relu(x) - self.alpha * torch.relu(-x)

    def _extra_config(self):
        return {'alpha_initializer': self.alpha_initializer, 'alpha_regularizer': None, 'alpha_constraint': None,
                'shared_axes': self.shared_axes}


class ThresholdedReLU(Layer):
    def __init__(self, theta=1.0, **kwargs):
        super().__init__(**kwargs)
        self.theta = float(theta)

    def call(self, x, training=False):
        return x * (x > self.theta).to(x.dtype)

    def _extra_config(self):
        return {'theta': self.theta}


# ----------------------------------------------------------------------------------------------------------- dropout
class AlphaDropout(Layer):
    """Dropout that keeps mean and variance of SELU activations (dropped units go to ``-alpha'``, then an affine fix)."""

    def __init__(self, rate, noise_shape=None, seed=None, **kwargs):
        super().__init__(**kwargs)
        self.rate = float(rate)

    def call(self, x, training=False):
        if not training or self.rate <= 0.0:
            return x
        alpha_p = -1.7580993408473766
        keep = (torch.rand_like(x) >= self.rate).to(x.dtype)
        a = ((1.0 - self.rate) * (1.0 + self.rate * alpha_p ** 2)) ** -0.5
        b = -a * alpha_p * self.rate
        return a * (x * keep + alpha_p * (1.0 - keep)) + b

    def _extra_config(self):
        return {'rate': self.rate, 'noise_shape': None, 'seed': None}


class GaussianDropout(Layer):
    """Multiplies by ``N(1, rate / (1 - rate))`` while training."""

    def __init__(self, rate, seed=None, **kwargs):
        super().__init__(**kwargs)
        self.rate = float(rate)

    def call(self, x, training=False):
        if not training or self.rate <= 0.0:
            return x
        return x * (1.0 + torch.randn_like(x) * math.sqrt(self.rate / (1.0 - self.rate)))

    def _extra_config(self):
        return {'rate': self.rate, 'seed': None}


# ---------------------------------------------------------------------------------------- padding / cropping / scale
def _pair(v):
    return (int(v), int(v)) if isinstance(v, int) else tuple(int(i) for i in v)


class ZeroPadding1D(Layer):
    def __init__(self, padding=1, **kwargs):
        super().__init__(**kwargs)
        self.padding = _pair(padding)

    def compute_output_shape(self, input_shape):
        n, t, c = input_shape
        return (n, None if t is None else t + sum(self.padding), c)

    def call(self, x, training=False):
        return F.pad(x, (0, 0, self.padding[0], self.padding[1]))

    def _extra_config(self):
        return {'padding': list(self.padding)}


class Cropping1D(Layer):
    def __init__(self, cropping=(1, 1), **kwargs):
        super().__init__(**kwargs)
        self.cropping = _pair(cropping)

    def compute_output_shape(self, input_shape):
        n, t, c = input_shape
        return (n, None if t is None else t - sum(self.cropping), c)

    def call(self, x, training=False):
        return x[:, self.cropping[0]:x.shape[1] - self.cropping[1]]

    def _extra_config(self):
        return {'cropping': list(self.cropping)}


class Cropping2D(Layer):
    def __init__(self, cropping=((0, 0), (0, 0)), data_format=None, **kwargs):
        super().__init__(**kwargs)
        if isinstance(cropping, int):
            cropping = ((cropping, cropping), (cropping, cropping))
        elif isinstance(cropping[0], int):
            cropping = ((cropping[0], cropping[0]), (cropping[1], cropping[1]))
        self.cropping = tuple(tuple(int(v) for v in c) for c in cropping)

    def compute_output_shape(self, input_shape):
        n, h, w, c = input_shape
        (t, b), (l, r) = self.cropping
        return (n, None if h is None else h - t - b, None if w is None else w - l - r, c)

    def call(self, x, training=False):
        (t, b), (l, r) = self.cropping
        return x[:, t:x.shape[1] - b, l:x.shape[2] - r]

    def _extra_config(self):
        return {'cropping': [list(c) for c in self.cropping], 'data_format': 'channels_last'}


class UpSampling1D(Layer):
    def __init__(self, size=2, **kwargs):
        super().__init__(**kwargs)
        self.size = int(size)

    def compute_output_shape(self, input_shape):
        n, t, c = input_shape
        return (n, None if t is None else t * self.size, c)

    def call(self, x, training=False):
        return x.repeat_interleave(self.size, dim=1)

    def _extra_config(self):
        return {'size': self.size}


class Rescaling(Layer):
    """``x * scale + offset`` (the usual first layer of an image model: ``Rescaling(1 / 255)``)."""

    def __init__(self, scale, offset=0.0, **kwargs):
        super().__init__(**kwargs)
        self.scale, self.offset = float(scale), float(offset)

    def call(self, x, training=False):
        return x * self.scale + self.offset

    def _extra_config(self):
        return {'scale': self.scale, 'offset': self.offset}


class Normalization(Layer):
    """``keras.layers.Normalization``: ``(x - mean) / sqrt(variance)`` over the feature axis, with the statistics either
    given (``mean=`` / ``variance=``) or learned from data by ``adapt(data)`` before training. Weights (non-trainable, the
    Keras order): ``mean``, ``variance``, ``count``."""

    def __init__(self, axis=-1, mean=None, variance=None, invert=False, **kwargs):
        super().__init__(**kwargs)
        axis = axis[0] if isinstance(axis, (list, tuple)) else axis
        if axis not in (-1, None):
            raise ValueError("Normalization: only axis=-1 (per feature) or None (one scalar) is supported")
        if (mean is None) != (variance is None):
            raise ValueError("Normalization: pass both mean and variance, or neither (and call adapt)")
        self.axis, self.invert = axis, bool(invert)
        self._given = None if mean is None else (np.asarray(mean, dtype=np.float32), np.asarray(variance, dtype=np.float32))
        self.mean = self.variance = self.count = None

    def build(self, input_shape):
        width = 1 if self.axis is None else int(input_shape[-1])
        mean, var = (np.zeros(width, np.float32), np.ones(width, np.float32)) if self._given is None else \
            (np.broadcast_to(self._given[0].reshape(-1), (width,)).copy(), np.broadcast_to(self._given[1].reshape(-1), (width,)).copy())
        self.mean = self.add_weight('mean', torch.from_numpy(mean), trainable=False)
        self.variance = self.add_weight('variance', torch.from_numpy(var), trainable=False)
        self.count = self.add_weight('count', torch.zeros(()), trainable=False)
        self.built = True

    def adapt(self, data, batch_size=None, steps=None):
        data = np.asarray(data, dtype=np.float64)
        if not self.built:
            self.build((None,) + tuple(data.shape[1:]))
        flat = data.reshape(-1, 1) if self.axis is None else data.reshape(-1, data.shape[-1])
        with torch.no_grad():
            self.mean.copy_(torch.from_numpy(flat.mean(axis=0).astype(np.float32)))
            self.variance.copy_(torch.from_numpy(flat.var(axis=0).astype(np.float32)))
            self.count.fill_(float(flat.shape[0]))
        for model in self._live_owners():
            model._push_to_engine()

    def call(self, x, training=False):
        scale = torch.clamp(torch.sqrt(self.variance), min=1e-7)
        return x * scale + self.mean if self.invert else (x - self.mean) / scale

    def _extra_config(self):
        given = self._given
        return {'axis': None if self.axis is None else [self.axis], 'invert': self.invert,
                'mean': None if given is None else given[0].tolist(), 'variance': None if given is None else given[1].tolist()}


# --------------------------------------------------------------------------------------------------------- attention
class Attention(Layer):
    """Dot-product (Luong) attention over ``[query, value]`` or ``[query, value, key]``; ``use_scale`` learns one scalar."""

    def __init__(self, use_scale=False, score_mode='dot', dropout=0.0, causal=False, **kwargs):
        super().__init__(**kwargs)
        if score_mode != 'dot':
            raise ValueError("Attention: only score_mode='dot' is supported")
        self.use_scale, self.dropout, self.causal = bool(use_scale), float(dropout), bool(causal)
        self.scale = None

    def build(self, input_shape):
        if self.use_scale:
            self.scale = self.add_weight('scale', torch.ones(()))
        self.built = True

    def compute_output_shape(self, input_shape):
        q, v = input_shape[0], input_shape[1]
        return (q[0], q[1], v[-1])

    def call(self, xs, training=False):
        q, v = xs[0], xs[1]
        k = xs[2] if len(xs) > 2 else v
        scores = q @ k.transpose(1, 2)
        if self.scale is not None:
            scores = scores * self.scale
        if self.causal:
            tq, tk = scores.shape[1], scores.shape[2]
            scores = scores.masked_fill(~torch.ones(tq, tk, dtype=torch.bool, device=scores.device).tril(), float('-inf'))
        w = torch.softmax(scores, dim=-1)
        if training and self.dropout > 0:
            w = F.dropout(w, self.dropout, True)
        return w @ v

    def _extra_config(self):
        return {'use_scale': self.use_scale, 'score_mode': 'dot', 'dropout': self.dropout, 'causal': self.causal}


class MultiHeadAttention(Layer):
    """``keras.layers.MultiHeadAttention``: called on ``[query, value]`` or ``[query, value, key]`` (``layer(q, v)`` works
    too). Weights in the Keras layouts and order: ``query / key / value`` kernels ``[dim, heads, key_dim]`` with biases
    ``[heads, key_dim]``, ``attention_output`` kernel ``[heads, value_dim, out_dim]`` with bias ``[out_dim]``. The scaled
    dot-product core is ``torch``'s fused attention."""

    def __init__(self, num_heads, key_dim, value_dim=None, dropout=0.0, use_bias=True, output_shape=None,
                 attention_axes=None, kernel_initializer='glorot_uniform', bias_initializer='zeros', **kwargs):
        super().__init__(**kwargs)
        if attention_axes is not None or output_shape is not None:
            raise ValueError("MultiHeadAttention: attention_axes / output_shape are not supported")
        self.num_heads, self.key_dim = int(num_heads), int(key_dim)
        self.value_dim = int(value_dim) if value_dim else self.key_dim
        self.dropout, self.use_bias = float(dropout), bool(use_bias)
        self.kernel_initializer = _init_config(kernel_initializer, 'glorot_uniform')
        self.bias_initializer = _init_config(bias_initializer, 'zeros')

    def __call__(self, inputs, value=None, key=None, training=False):
        if value is not None:
            inputs = [inputs, value] + ([key] if key is not None else [])
        return super().__call__(inputs, training=training)

    def build(self, input_shape):
        q_dim, v_dim = int(input_shape[0][-1]), int(input_shape[1][-1])
        k_dim = int(input_shape[2][-1]) if len(input_shape) > 2 else v_dim
        h = self.num_heads

        def dense(name, in_dim, per_head):
            w = _initialize(torch.empty(in_dim, h * per_head), self.kernel_initializer, in_dim, h * per_head)
            kernel = self.add_weight(name + '/kernel', w.reshape(in_dim, h, per_head).clone())
            bias = self.add_weight(name + '/bias', torch.zeros(h, per_head)) if self.use_bias else None
            return kernel, bias

        self.wq, self.bq = dense('query', q_dim, self.key_dim)
        self.wk, self.bk = dense('key', k_dim, self.key_dim)
        self.wv, self.bv = dense('value', v_dim, self.value_dim)
        wo = _initialize(torch.empty(h * self.value_dim, q_dim), self.kernel_initializer, h * self.value_dim, q_dim)
        self.wo = self.add_weight('attention_output/kernel', wo.reshape(h, self.value_dim, q_dim).clone())
        self.bo = self.add_weight('attention_output/bias', torch.zeros(q_dim)) if self.use_bias else None
        self.built = True

    def compute_output_shape(self, input_shape):
        return tuple(input_shape[0])

    def call(self, xs, training=False):
        q_in, v_in = xs[0], xs[1]
        k_in = xs[2] if len(xs) > 2 else v_in

        def project(x, w, b):   # [b, t, d] x [d, h, e] -> [b, h, t, e]
            y = torch.einsum('btd,dhe->bhte', x, w)
            return y if b is None else y + b.unsqueeze(0).unsqueeze(2)

        q, k, v = project(q_in, self.wq, self.bq), project(k_in, self.wk, self.bk), project(v_in, self.wv, self.bv)
        ctx = F.scaled_dot_product_attention(q, k, v, dropout_p=self.dropout if training else 0.0)
        out = torch.einsum('bhte,heo->bto', ctx, self.wo)
        return out if self.bo is None else out + self.bo

    def _extra_config(self):
        return {'num_heads': self.num_heads, 'key_dim': self.key_dim, 'value_dim': self.value_dim, 'dropout': self.dropout,
                'use_bias': self.use_bias, 'output_shape': None, 'attention_axes': None,
                'kernel_initializer': self.kernel_initializer, 'bias_initializer': self.bias_initializer,
                'kernel_regularizer': None, 'bias_regularizer': None, 'activity_regularizer': None,
                'kernel_constraint': None, 'bias_constraint': None}


MORE_LAYERS = (Normalization, Lambda, PReLU, ThresholdedReLU, AlphaDropout, GaussianDropout, ZeroPadding1D, Cropping1D, Cropping2D,
               UpSampling1D, Rescaling, Attention, MultiHeadAttention)
LAYER_CLASSES.update({c.__name__: c for c in MORE_LAYERS})
