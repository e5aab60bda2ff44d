"""Keras-compatible optimizers (config schema, defaults and update rules of Keras 2.10) on torch.

The reference serialises the compiled model's optimizer on the driver and re-creates it on every
worker (``elephas/spark_model.py:54,193,200``, ``elephas/worker.py:29-33,88-89``) through
``keras.optimizers.{get,serialize,deserialize}``; the same three functions exist here and speak the
same ``{"class_name": ..., "config": {...}}`` dictionaries, so an optimizer config written by
Keras can be fed to ``ElephasEstimator.set_optimizer_config`` unchanged
(``tests/test_ml_model.py:69-70``).

Update rules follow ``tf.keras.optimizers`` (OptimizerV2): time-based decay
``lr_t = lr / (1 + decay * iterations)``, epsilon 1e-7, RMSprop/Adam with epsilon *outside* the
square root. The CUDA multi-tensor kernels (``csrc/optim_kernels.cu``) implement the same formulas;
this module is both the CPU path and their numerical oracle.
"""
import copy
import math

import torch


class Optimizer:
    _class_name = 'Optimizer'
    _defaults = {}
    _default_weight_decay = None

    def __init__(self, learning_rate=None, **kwargs):
        if learning_rate is not None:  # Keras takes the rate as the first positional argument: Adam(1e-3)
            kwargs['learning_rate'] = learning_rate
        if 'lr' in kwargs:  # legacy Keras spelling, used by the reference tests (SGD(lr=0.1))
            kwargs['learning_rate'] = kwargs.pop('lr')
        self.clipnorm = kwargs.pop('clipnorm', None)
        self.clipvalue = kwargs.pop('clipvalue', None)
        self.global_clipnorm = kwargs.pop('global_clipnorm', None)
        # keys of the Keras 2.11+ / Keras 3 optimizer configs: decoupled weight decay is applied, the switches that have
        # no meaning here are accepted when they are off
        self.weight_decay = kwargs.pop('weight_decay', self._default_weight_decay) or None
        self.gradient_accumulation_steps = int(kwargs.pop('gradient_accumulation_steps', None) or 0) or None
        if self.gradient_accumulation_steps is not None and self.gradient_accumulation_steps < 2:
            raise ValueError("gradient_accumulation_steps must be >= 2 (or None), got %r" % self.gradient_accumulation_steps)
        for key in ('use_ema', 'loss_scale_factor'):
            if kwargs.pop(key, None):
                raise ValueError("%s: %s is not supported" % (self._class_name, key))
        for key in ('ema_momentum', 'ema_overwrite_frequency', 'jit_compile', 'is_legacy_optimizer'):
            kwargs.pop(key, None)
        cfg = dict(self._defaults)
        unknown = set(kwargs) - set(cfg) - {'name'}
        if unknown:
            raise TypeError("Unexpected keyword argument(s) for %s: %s" % (self._class_name, sorted(unknown)))
        cfg.update(kwargs)
        cfg.setdefault('name', self._class_name)
        if isinstance(cfg['learning_rate'], dict):  # a serialised schedule ({"class_name": "ExponentialDecay", ...})
            from . import schedules
            cfg['learning_rate'] = schedules.deserialize(cfg['learning_rate'])
        self._config = cfg
        self.iterations = 0
        self._slots = {}
        self._hyper_version = 0  # bumped when a hyper-parameter is assigned: captured CUDA graphs bake them in

    # ------------------------------------------------------------------ config
    def __getattr__(self, item):
        cfg = self.__dict__.get('_config')
        if cfg is not None and item in cfg:
            return cfg[item]
        raise AttributeError(item)

    @property
    def learning_rate(self):
        return self._config['learning_rate']

    @learning_rate.setter
    def learning_rate(self, value):
        value = value if callable(value) else float(value)
        if value != self._config['learning_rate']:
            self._config['learning_rate'] = value
            self._hyper_version = self.__dict__.get('_hyper_version', 0) + 1

    lr = learning_rate

    def get_config(self):
        cfg = {'name': self._config['name']}
        cfg.update({k: v for k, v in self._config.items() if k != 'name'})
        if self.schedule is not None:
            from . import schedules
            cfg['learning_rate'] = schedules.serialize(self.schedule)
        for k in ('clipnorm', 'clipvalue', 'global_clipnorm', 'weight_decay', 'gradient_accumulation_steps'):
            if getattr(self, k) is not None:
                cfg[k] = getattr(self, k)
        return cfg

    @classmethod
    def from_config(cls, config):
        return cls(**copy.deepcopy(config))

    # ------------------------------------------------------------------- state
    @property
    def schedule(self):
        """The learning-rate schedule (``models/schedules.py``), or None for a plain number."""
        lr = self._config['learning_rate']
        return lr if callable(lr) else None

    def _base_lr(self):
        lr = self._config['learning_rate']
        return float(lr(self.iterations)) if callable(lr) else float(lr)

    def _lr_t(self):
        decay = self._config.get('decay', 0.0) or 0.0
        lr = self._base_lr()
        return lr / (1.0 + decay * self.iterations) if decay else lr

    def _slot(self, idx, name, like, init=0.0):
        key = (idx, name)
        s = self._slots.get(key)
        if s is None or s.shape != like.shape or s.device != like.device:
            s = torch.full_like(like, init)
            self._slots[key] = s
        return s

    def reset_state(self):
        self.iterations = 0
        self._slots = {}
        self._held = None

    def _clip(self, grads):
        if self.clipvalue is not None:
            grads = [g.clamp(-self.clipvalue, self.clipvalue) for g in grads]
        if self.clipnorm is not None:
            grads = [g * min(1.0, self.clipnorm / (float(g.norm()) + 1e-12)) for g in grads]
        if self.global_clipnorm is not None:
            total = math.sqrt(sum(float(g.norm()) ** 2 for g in grads))
            scale = min(1.0, self.global_clipnorm / (total + 1e-12))
            grads = [g * scale for g in grads]
        return grads

    @torch.no_grad()
    def apply_gradients(self, params, grads):
        """``params[i] -= step(grads[i])`` in place; advances ``iterations``."""
        grads, params = list(grads), list(params)
        if self.gradient_accumulation_steps:
            # the gradients of k calls are averaged and applied on the k-th (an update every k batches, as Keras 3 does)
            held = self.__dict__.setdefault('_held', None)
            if held is None:
                self._held, self._held_calls = [g.clone() for g in grads], 1
            else:
                for h, g in zip(held, grads):
                    h.add_(g)
                self._held_calls += 1
            if self._held_calls < self.gradient_accumulation_steps:
                return
            grads = [h.div_(self._held_calls) for h in self._held]
            self._held = None
        grads = self._clip(grads)
        if self.weight_decay:  # decoupled (AdamW-style): w <- w - lr * wd * w, before the gradient step
            shrink = 1.0 - self._lr_t() * float(self.weight_decay)
            for p in params:
                p.mul_(shrink)
        self._apply(params, grads)
        self.iterations += 1

    def _apply(self, params, grads):  # pragma: no cover - abstract
        raise NotImplementedError


class SGD(Optimizer):
    _class_name = 'SGD'
    _defaults = {'learning_rate': 0.01, 'decay': 0.0, 'momentum': 0.0, 'nesterov': False}

    def _apply(self, params, grads):
        lr = self._lr_t()
        mom = float(self._config['momentum'])
        for i, (p, g) in enumerate(zip(params, grads)):
            if mom == 0.0:
                p.add_(g, alpha=-lr)
                continue
            v = self._slot(i, 'momentum', p)
            v.mul_(mom).add_(g, alpha=-lr)
            if self._config['nesterov']:
                p.add_(v, alpha=mom).add_(g, alpha=-lr)
            else:
                p.add_(v)


class RMSprop(Optimizer):
    _class_name = 'RMSprop'
    _defaults = {'learning_rate': 0.001, 'decay': 0.0, 'rho': 0.9, 'momentum': 0.0, 'epsilon': 1e-07,
                 'centered': False}

    def _apply(self, params, grads):
        lr = self._lr_t()
        rho, eps, mom = float(self._config['rho']), float(self._config['epsilon']), float(self._config['momentum'])
        for i, (p, g) in enumerate(zip(params, grads)):
            rms = self._slot(i, 'rms', p)
            rms.mul_(rho).addcmul_(g, g, value=1.0 - rho)
            denom = rms
            if self._config['centered']:
                mg = self._slot(i, 'mg', p)
                mg.mul_(rho).add_(g, alpha=1.0 - rho)
                denom = rms - mg * mg
            step = g / (denom.sqrt() + eps)
            if mom > 0.0:
                m = self._slot(i, 'momentum', p)
                m.mul_(mom).add_(step, alpha=lr)
                p.sub_(m)
            else:
                p.add_(step, alpha=-lr)


class Adam(Optimizer):
    _class_name = 'Adam'
    _defaults = {'learning_rate': 0.001, 'decay': 0.0, 'beta_1': 0.9, 'beta_2': 0.999, 'epsilon': 1e-07,
                 'amsgrad': False}

    def _apply(self, params, grads):
        t = self.iterations + 1
        b1, b2, eps = float(self._config['beta_1']), float(self._config['beta_2']), float(self._config['epsilon'])
        lr = self._lr_t() * math.sqrt(1.0 - b2 ** t) / (1.0 - b1 ** t)
        for i, (p, g) in enumerate(zip(params, grads)):
            m = self._slot(i, 'm', p)
            v = self._slot(i, 'v', p)
            m.mul_(b1).add_(g, alpha=1.0 - b1)
            v.mul_(b2).addcmul_(g, g, value=1.0 - b2)
            if self._config['amsgrad']:
                vhat = self._slot(i, 'vhat', p)
                torch.maximum(vhat, v, out=vhat)
                p.addcdiv_(m, vhat.sqrt() + eps, value=-lr)
            else:
                p.addcdiv_(m, v.sqrt() + eps, value=-lr)


class AdamW(Adam):
    """Adam with decoupled weight decay (``keras.optimizers.AdamW``, default ``weight_decay=0.004``)."""
    _class_name = 'AdamW'
    _default_weight_decay = 0.004


class Adagrad(Optimizer):
    _class_name = 'Adagrad'
    _defaults = {'learning_rate': 0.001, 'decay': 0.0, 'initial_accumulator_value': 0.1, 'epsilon': 1e-07}

    def _apply(self, params, grads):
        lr, eps = self._lr_t(), float(self._config['epsilon'])
        for i, (p, g) in enumerate(zip(params, grads)):
            acc = self._slot(i, 'accumulator', p, init=float(self._config['initial_accumulator_value']))
            acc.addcmul_(g, g)
            p.addcdiv_(g, acc.sqrt() + eps, value=-lr)


class Adadelta(Optimizer):
    _class_name = 'Adadelta'
    _defaults = {'learning_rate': 0.001, 'decay': 0.0, 'rho': 0.95, 'epsilon': 1e-07}

    def _apply(self, params, grads):
        lr, rho, eps = self._lr_t(), float(self._config['rho']), float(self._config['epsilon'])
        for i, (p, g) in enumerate(zip(params, grads)):
            acc_g = self._slot(i, 'accum_grad', p)
            acc_d = self._slot(i, 'accum_var', p)
            acc_g.mul_(rho).addcmul_(g, g, value=1.0 - rho)
            update = g * (acc_d + eps).sqrt() / (acc_g + eps).sqrt()
            acc_d.mul_(rho).addcmul_(update, update, value=1.0 - rho)
            p.add_(update, alpha=-lr)


class Adamax(Optimizer):
    _class_name = 'Adamax'
    _defaults = {'learning_rate': 0.001, 'decay': 0.0, 'beta_1': 0.9, 'beta_2': 0.999, 'epsilon': 1e-07}

    def _apply(self, params, grads):
        t = self.iterations + 1
        b1, b2, eps = float(self._config['beta_1']), float(self._config['beta_2']), float(self._config['epsilon'])
        lr = self._lr_t() / (1.0 - b1 ** t)
        for i, (p, g) in enumerate(zip(params, grads)):
            m = self._slot(i, 'm', p)
            u = self._slot(i, 'v', p)
            m.mul_(b1).add_(g, alpha=1.0 - b1)
            torch.maximum(u * b2, g.abs(), out=u)
            p.addcdiv_(m, u + eps, value=-lr)


class Nadam(Optimizer):
    _class_name = 'Nadam'
    _defaults = {'learning_rate': 0.001, 'decay': 0.004, 'beta_1': 0.9, 'beta_2': 0.999, 'epsilon': 1e-07}

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self._m_schedule = 1.0

    def reset_state(self):
        super().reset_state()
        self._m_schedule = 1.0

    def _lr_t(self):  # Nadam's `decay` is the momentum-schedule decay, not an lr decay
        return self._base_lr()

    def _apply(self, params, grads):
        t = self.iterations + 1
        b1, b2, eps = float(self._config['beta_1']), float(self._config['beta_2']), float(self._config['epsilon'])
        decay = float(self._config['decay'])
        mu_t = b1 * (1.0 - 0.5 * (0.96 ** (decay * t)))
        mu_t1 = b1 * (1.0 - 0.5 * (0.96 ** (decay * (t + 1))))
        sched = self._m_schedule * mu_t
        sched_next = sched * mu_t1
        self._m_schedule = sched
        lr = self._lr_t()
        for i, (p, g) in enumerate(zip(params, grads)):
            m = self._slot(i, 'm', p)
            v = self._slot(i, 'v', p)
            m.mul_(b1).add_(g, alpha=1.0 - b1)
            v.mul_(b2).addcmul_(g, g, value=1.0 - b2)
            g_prime = g / (1.0 - sched)
            m_prime = m / (1.0 - sched_next)
            v_prime = v / (1.0 - b2 ** t)
            m_bar = (1.0 - mu_t) * g_prime + mu_t1 * m_prime
            p.addcdiv_(m_bar, v_prime.sqrt() + eps, value=-lr)


_CLASSES = {c._class_name.lower(): c for c in (SGD, RMSprop, Adam, AdamW, Adagrad, Adadelta, Adamax, Nadam)}


def serialize(optimizer: Optimizer) -> dict:
    return {'class_name': optimizer._class_name, 'config': optimizer.get_config()}


def deserialize(config: dict, custom_objects=None) -> Optimizer:
    name = config['class_name']
    if custom_objects and name in custom_objects:
        return custom_objects[name].from_config(config.get('config', {}))
    cls = _CLASSES.get(name.lower())
    if cls is None:
        raise ValueError("Unknown optimizer: %r" % (name,))
    return cls.from_config(config.get('config', {}))


def get(identifier) -> Optimizer:
    """``keras.optimizers.get``: instance, config dict or (case-insensitive) name."""
    if isinstance(identifier, Optimizer):
        return identifier
    if isinstance(identifier, dict):
        return deserialize(identifier)
    if isinstance(identifier, str):
        cls = _CLASSES.get(identifier.lower())
        if cls is None:
            raise ValueError("Unknown optimizer: %r" % (identifier,))
        return cls()
    raise ValueError("Could not interpret optimizer identifier: %r" % (identifier,))
