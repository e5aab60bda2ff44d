"""Learning-rate schedules with the config schema of ``tf.keras.optimizers.schedules``.

An optimizer serialised by Keras carries its schedule as ``"learning_rate": {"class_name": "ExponentialDecay", "config":
{...}}``; the reference ships that dictionary to every worker (``elephas/spark_model.py:54``, ``elephas/worker.py:29-33``).
``schedule(step)`` is evaluated on the host before every update; a step that bakes the rate into a captured graph cannot
follow it, such optimizers run on the eager path (``ops/engine.py::_optimizer_spec``).
"""
import bisect
import math


class LearningRateSchedule:
    def __call__(self, step):
        raise NotImplementedError

    def get_config(self):
        raise NotImplementedError

    @classmethod
    def from_config(cls, config):
        return cls(**config)


class ExponentialDecay(LearningRateSchedule):
    def __init__(self, initial_learning_rate, decay_steps, decay_rate, staircase=False, name=None):
        self.initial_learning_rate, self.decay_steps = float(initial_learning_rate), float(decay_steps)
        self.decay_rate, self.staircase, self.name = float(decay_rate), bool(staircase), name

    def __call__(self, step):
        p = step / self.decay_steps
        return self.initial_learning_rate * self.decay_rate ** (math.floor(p) if self.staircase else p)

    def get_config(self):
        return {'initial_learning_rate': self.initial_learning_rate, 'decay_steps': self.decay_steps,
                'decay_rate': self.decay_rate, 'staircase': self.staircase, 'name': self.name}


class InverseTimeDecay(ExponentialDecay):
    def __call__(self, step):
        p = step / self.decay_steps
        return self.initial_learning_rate / (1.0 + self.decay_rate * (math.floor(p) if self.staircase else p))


class PiecewiseConstantDecay(LearningRateSchedule):
    def __init__(self, boundaries, values, name=None):
        if len(boundaries) != len(values) - 1:
            raise ValueError("The length of boundaries should be 1 less than the length of values")
        self.boundaries, self.values, self.name = list(boundaries), [float(v) for v in values], name

    def __call__(self, step):
        return self.values[bisect.bisect_left(self.boundaries, step)]   # values[i] while step <= boundaries[i]

    def get_config(self):
        return {'boundaries': self.boundaries, 'values': self.values, 'name': self.name}


class PolynomialDecay(LearningRateSchedule):
    def __init__(self, initial_learning_rate, decay_steps, end_learning_rate=0.0001, power=1.0, cycle=False, name=None):
        self.initial_learning_rate, self.decay_steps = float(initial_learning_rate), float(decay_steps)
        self.end_learning_rate, self.power, self.cycle, self.name = float(end_learning_rate), float(power), bool(cycle), name

    def __call__(self, step):
        steps = self.decay_steps
        if self.cycle:
            steps = steps * max(1.0, math.ceil(step / steps))
        else:
            step = min(step, steps)
        return (self.initial_learning_rate - self.end_learning_rate) * (1.0 - step / steps) ** self.power \
            + self.end_learning_rate

    def get_config(self):
        return {'initial_learning_rate': self.initial_learning_rate, 'decay_steps': self.decay_steps,
                'end_learning_rate': self.end_learning_rate, 'power': self.power, 'cycle': self.cycle, 'name': self.name}


class CosineDecay(LearningRateSchedule):
    def __init__(self, initial_learning_rate, decay_steps, alpha=0.0, name=None):
        self.initial_learning_rate, self.decay_steps = float(initial_learning_rate), float(decay_steps)
        self.alpha, self.name = float(alpha), name

    def __call__(self, step):
        frac = min(step, self.decay_steps) / self.decay_steps
        cosine = 0.5 * (1.0 + math.cos(math.pi * frac))
        return self.initial_learning_rate * ((1.0 - self.alpha) * cosine + self.alpha)

    def get_config(self):
        return {'initial_learning_rate': self.initial_learning_rate, 'decay_steps': self.decay_steps,
                'alpha': self.alpha, 'name': self.name}


class CosineDecayRestarts(LearningRateSchedule):
    def __init__(self, initial_learning_rate, first_decay_steps, t_mul=2.0, m_mul=1.0, alpha=0.0, name=None):
        self.initial_learning_rate, self.first_decay_steps = float(initial_learning_rate), float(first_decay_steps)
        self.t_mul, self.m_mul, self.alpha, self.name = float(t_mul), float(m_mul), float(alpha), name

    def __call__(self, step):
        done = step / self.first_decay_steps
        if self.t_mul == 1.0:
            restarts = math.floor(done)
            frac = done - restarts
        else:
            restarts = math.floor(math.log(1.0 - done * (1.0 - self.t_mul)) / math.log(self.t_mul))
            before = (1.0 - self.t_mul ** restarts) / (1.0 - self.t_mul)
            frac = (done - before) / self.t_mul ** restarts
        cosine = 0.5 * self.m_mul ** restarts * (1.0 + math.cos(math.pi * frac))
        return self.initial_learning_rate * ((1.0 - self.alpha) * cosine + self.alpha)

    def get_config(self):
        return {'initial_learning_rate': self.initial_learning_rate, 'first_decay_steps': self.first_decay_steps,
                't_mul': self.t_mul, 'm_mul': self.m_mul, 'alpha': self.alpha, 'name': self.name}


_CLASSES = {c.__name__: c for c in (ExponentialDecay, InverseTimeDecay, PiecewiseConstantDecay, PolynomialDecay,
                                    CosineDecay, CosineDecayRestarts)}


def serialize(schedule):
    return {'class_name': type(schedule).__name__, 'config': schedule.get_config()}


def deserialize(config, custom_objects=None):
    name = config['class_name']
    cls = (custom_objects or {}).get(name) or _CLASSES.get(name)
    if cls is None:
        raise ValueError("Unknown learning-rate schedule: %r" % (name,))
    return cls.from_config(dict(config.get('config') or {}))
