"""Training step of Dense / Activation / Dropout chains on the CPU without the autograd tape.

Every model the reference ships is such a chain (``examples/*.py``, ``tests/conftest.py:8-28``), and on machines without
a GPU they are small: a step of the MNIST MLP is a dozen matrix products of a few MFLOP, so what a step costs is not
arithmetic but per-operator overhead - graph construction, the backward engine's dispatch, temporaries. This module
writes the forward and backward pass out by hand with in-place torch operators (the CPU counterpart of what
``ops/engine.py`` does with the sm_100a kernels): about 30 operator calls per step instead of the tape's ~90, no graph.
The optimizer is the model's own (``Optimizer.apply_gradients``), the loss value and the metrics come from the same
functions as on the tape path, so both paths report and apply the same numbers (``tests/models/test_cpu_mlp.py``).

Covered: hidden activations relu / linear / sigmoid / tanh, Dropout after a hidden layer, outputs softmax +
categorical cross-entropy, sigmoid + binary cross-entropy, or linear / sigmoid / tanh / relu + mean squared / absolute
error. Anything else stays on the tape.
``ELEPHAS_B200_CPU_ENGINE=0`` switches the module off.

Reference call sites: ``model.fit`` / ``train_on_batch`` inside the workers (``elephas/worker.py:42,108,123``).
"""
import os

import numpy as np
import torch

_HIDDEN = ('relu', 'linear', 'sigmoid', 'tanh')
_EPS = 1e-7  # Keras' clipping constant of the cross-entropies (models/losses.py)


def enabled():
    return os.environ.get('ELEPHAS_B200_CPU_ENGINE', '1') != '0'


def _activate_(z, name):
    """In-place activation of the pre-activations ``z``; returns the activations."""
    if name == 'relu':
        return z.relu_()
    if name == 'sigmoid':
        return z.sigmoid_()
    if name == 'tanh':
        return z.tanh_()
    return z


def _times_derivative_(grad, a, name):
    """``grad *= act'(z)`` expressed through the activations ``a = act(z)`` (in place on ``grad``)."""
    if name == 'relu':
        return grad.mul_(a > 0)
    if name == 'sigmoid':
        return grad.mul_(a * (1.0 - a))
    if name == 'tanh':
        return grad.mul_(1.0 - a * a)
    return grad


def _dropout_mask(rng, like, keep):
    """Inverted-dropout mask (``1 / keep`` with probability ``keep``, else 0) in the shape of ``like``. Drawn with NumPy:
    torch's CPU generator is the cost of a small step (a ``bernoulli_`` of a 64 x 128 tensor takes as long as the layer's
    matrix product, NumPy's float32 uniforms a third of that). ``rng`` belongs to the step object - one per model
    replica, so worker threads do not interleave their draws - and is seeded from torch's generator when the step is
    built, which keeps ``torch.manual_seed`` in charge of reproducibility."""
    u = rng.random(tuple(like.shape), dtype=np.float32)
    return torch.from_numpy((u < keep).astype(np.float32) * np.float32(1.0 / keep))


class CpuMlpStep:
    """``step(x, y) -> [loss, metrics...]``: one optimizer step of ``model`` on a CPU batch."""

    def __init__(self, model, ops, out_act, loss_name):
        self.model, self.ops, self.out_act, self.loss_name = model, ops, out_act, loss_name
        self.rng = np.random.default_rng(int(torch.randint(0, 2 ** 31 - 1, (1,))))  # dropout masks (see _dropout_mask)
        penalties, constraints, _ = model._weight_options()
        self.regularized, self.constrained = bool(penalties), bool(constraints)

    @torch.no_grad()
    def __call__(self, xb, yb):
        model, ops = self.model, self.ops
        yb = yb.to(torch.float32)
        if yb.dim() == 1:
            yb = yb.unsqueeze(-1)
        saved = []  # (layer input, activations, dropout mask or None) per layer
        h = xb
        for i, op in enumerate(ops):
            layer = op.layer
            a = torch.addmm(layer.bias, h, layer.kernel.t())
            last = i == len(ops) - 1
            if last and self.out_act == 'softmax':
                a = torch.softmax(a, dim=-1)
            else:
                a = _activate_(a, self.out_act if last else op.act)
            mask = None
            nxt = a
            if not last and op.drop_p > 0.0:
                keep = 1.0 - op.drop_p
                mask = _dropout_mask(self.rng, a, keep)  # inverted dropout, as Keras scales it
                nxt = a * mask
            saved.append((h, a, mask))
            h = nxt
        out = h
        rows = xb.shape[0]
        # dL/dz of the output layer for the mean-over-batch loss
        if self.loss_name == 'categorical_crossentropy':
            dz = (out - yb).div_(rows)
        elif self.loss_name == 'binary_crossentropy':  # sigmoid outputs; Keras clips p to [eps, 1 - eps]: no gradient there
            dz = (out - yb).div_(rows * out.shape[-1]).mul_((out > _EPS) & (out < 1.0 - _EPS))
        else:  # mean squared / absolute error: mean over the outputs, then over the batch
            if self.loss_name == 'mean_squared_error':
                dz = (out - yb).mul_(2.0 / (rows * out.shape[-1]))
            else:
                dz = torch.sign(out - yb).div_(rows * out.shape[-1])
            _times_derivative_(dz, out, self.out_act)
        grads = [None] * (2 * len(ops))
        for i in range(len(ops) - 1, -1, -1):
            layer = ops[i].layer
            h_in, a, mask = saved[i]
            if i < len(ops) - 1:
                if mask is not None:
                    dz.mul_(mask)
                _times_derivative_(dz, a, ops[i].act)
            grads[2 * i] = dz.t().mm(h_in)      # [units, in]: the layout the kernel is stored in
            grads[2 * i + 1] = dz.sum(dim=0)
            if i > 0:
                dz = dz.mm(layer.kernel)
        params = []
        for op in ops:
            params += [op.layer.kernel, op.layer.bias]
        penalty = 0.0
        if self.regularized:  # L1 / L2 penalties of the layer configs: closed-form gradient, value added to the loss
            where = {id(p): k for k, p in enumerate(params)}
            for w, reg in model._weight_options()[0]:
                g = grads[where[id(w)]]
                if reg.l2:
                    g.add_(w, alpha=2.0 * reg.l2)
                if reg.l1:
                    g.add_(torch.sign(w), alpha=reg.l1)
                penalty += float(reg(w))
        model.optimizer.apply_gradients(params, grads)
        if self.constrained:
            model._apply_constraints()
        stats = [float(model._loss_fn(yb, out).mean()) + penalty]
        for _, fn in model._metric_fns:
            stats.append(float(fn(yb, out).mean()))
        return stats


def maybe_create(model):
    """A :class:`CpuMlpStep` for a compiled Dense chain on the CPU, or None (autograd path)."""
    if not enabled() or model.device.type != 'cpu' or not getattr(model, '_compiled', False):
        return None
    if any(getattr(fn, 'stateful', False) for _, fn in model._metric_fns):
        return None  # count-based metrics live on the autograd path
    from ..models.regularizers import L1L2
    penalties, _, activity = model._weight_options()
    if activity or any(not isinstance(reg, L1L2) for _, reg in penalties):
        return None  # activity penalties and penalties without a closed-form gradient here: autograd
    from . import engine as _engine  # (only its pure-Python plan builder is used: no native library needed)
    try:
        plan = _engine._plan(model)
    except Exception:  # pragma: no cover - a model the plan builder does not understand stays on the tape
        return None
    if plan is None:
        return None
    ops, out_act = plan
    loss_name = getattr(model, '_loss_name', None)
    if any(op.act not in _HIDDEN for op in ops[:-1]):
        return None
    if loss_name == 'categorical_crossentropy':
        if out_act != 'softmax':
            return None
    elif loss_name == 'binary_crossentropy':
        if out_act != 'sigmoid':
            return None
    elif loss_name in ('mean_squared_error', 'mean_absolute_error'):
        if out_act not in _HIDDEN:
            return None
    else:
        return None
    opt = model.optimizer
    if opt is None or opt.clipnorm is not None or opt.clipvalue is not None or opt.global_clipnorm is not None:
        return None
    # the chain must hold ALL trainable weights of the model (a kernel and a bias per Dense layer, in order)
    expected = []
    for op in ops:
        expected += [op.layer.kernel, op.layer.bias]
    trainable = model.trainable_weights
    if len(trainable) != len(expected) or any(a is not b for a, b in zip(trainable, expected)):
        return None
    return CpuMlpStep(model, ops, out_act, loss_name)
