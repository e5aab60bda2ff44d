"""Graph-captured training step for models the native Dense engine does not cover (CNNs, custom activations,
functional graphs): torch autograd as the tape - the convolution / BatchNorm / pooling / Dense ops on it are the native
sm_100a kernels of ``ops/cnn.py`` when the model computes in bf16 - everything around it B200-style.

* all trainable weights live in ONE flat fp32 buffer (the model's tensors become views of it), gradients accumulate
  into a second flat buffer through pre-assigned ``.grad`` views, and the optimizer is ONE launch of the native
  multi-tensor kernel (K10, ``csrc/optim_kernels.cu``) over the flat buffers - Keras formulas, device-side step
  counter - instead of four small kernels per weight tensor;
* forward + loss + backward + optimizer + loss/metric accumulation are recorded once per batch shape into a CUDA graph
  and replayed: one host launch per step, no per-step host synchronisation (running sums stay on the device and are
  read once per epoch);
* ``model.set_compute_dtype('bfloat16')`` runs the forward/backward under bf16 autocast, master weights and loss fp32.

The reference trains these models through ``keras.Model.fit`` / ``train_on_batch`` (``elephas/worker.py:42,123``).
"""
import contextlib
import gc
import os

import torch

from . import cnn as _cnn
from . import native
from .engine import _CAPTURE_LOCK, _optimizer_spec, native_enabled

_ALIGN = 64  # elements: every tensor starts on a 256-byte boundary of the flat buffers


def graphs_enabled():
    return os.environ.get('ELEPHAS_B200_TORCH_GRAPHS', '1') != '0'


def maybe_create(model):
    """A :class:`TorchStepEngine` for a compiled model on a CUDA device, or None (plain eager path)."""
    if not native_enabled() or not graphs_enabled() or model.device.type != 'cuda' or not model._compiled:
        return None
    kind = _optimizer_spec(model.optimizer)
    params = model.trainable_weights
    if kind is None or not params or any(p.dtype != torch.float32 for p in params) or len(model._metric_fns) > 4:
        return None
    if any(getattr(fn, 'stateful', False) for _, fn in model._metric_fns):
        return None  # count-based metrics (precision, AUC, ...) keep their state on the host
    from ..utils import tracing
    with tracing.span('torch_step/create'):
        return TorchStepEngine(model, kind)


class TorchStepEngine:
    def __init__(self, model, kind):
        self.model, self.kind, self.device = model, kind, model.device
        self.params = list(model.trainable_weights)
        offs, total = [], 0
        for p in self.params:
            offs.append(total)
            total += (p.numel() + _ALIGN - 1) // _ALIGN * _ALIGN
        self.offsets, self.numel = offs, total
        dev = self.device
        self.w32 = torch.zeros(total, dtype=torch.float32, device=dev)
        self.g32 = torch.zeros(total, dtype=torch.float32, device=dev)
        with torch.no_grad():
            for p, off in zip(self.params, offs):
                view = self.w32[off:off + p.numel()].view(p.shape)
                view.copy_(p)
                p.data = view                                   # the model now trains inside the flat buffer
                p.grad = self.g32[off:off + p.numel()].view(p.shape)   # autograd accumulates in place
                p._eb_flat_grad = True  # native ops may add their parameter gradients straight into this (zeroed) view
        # bf16 copy of all weights for the native conv / linear ops: refreshed by ONE cast kernel at the start of a step
        # (instead of one cast per layer), the layers read views of it (ops/cnn.py: weights_bf16)
        self.w16 = torch.zeros(total, dtype=torch.bfloat16, device=dev)
        for p, off in zip(self.params, offs):
            p._eb_w16 = self.w16[off:off + p.numel()].view(p.shape)
        self.step = torch.zeros(1, dtype=torch.int64, device=dev)
        self.stats = torch.zeros(8, dtype=torch.float64, device=dev)   # [0] loss sum, [1..4] metric sums, [7] rows
        self.slots = [None, None, None]
        self._in_warmup = False
        self._alloc_slots()
        self.import_optimizer_state()
        self.graphs = {}
        self.failed = False
        # synchronous data parallel, frequency='batch' (parallel/sync_torch.py): an object with ``delta(flip)`` (bf16 flat
        # buffer in peer-mapped memory the optimizer writes its step into) and ``reduce(flip)`` (fused all-reduce + update
        # of w32); both double buffered, ``flip`` alternates every step and is part of the graph key
        self.sync = None
        self.flip = 0

    # ------------------------------------------------------------- optimizer state
    def _alloc_slots(self):
        cfg = self.model.optimizer._config
        name = self.model.optimizer._class_name
        need = {'SGD': 1 if cfg.get('momentum') else 0, 'RMSprop': 3, 'Adam': 3 if cfg.get('amsgrad') else 2,
                'Adagrad': 1, 'Adadelta': 2, 'Adamax': 2}[name]
        for i in range(need):
            self.slots[i] = torch.zeros(self.numel, dtype=torch.float32, device=self.device)
        if name == 'Adagrad':
            self.slots[0].fill_(float(cfg.get('initial_accumulator_value', 0.1)))

    def reset_optimizer(self):
        """Fresh optimizer state (slots, step counter): what a newly built worker replica starts from."""
        for s in self.slots:
            if s is not None:
                s.zero_()
        if self.model.optimizer._class_name == 'Adagrad':
            self.slots[0].fill_(float(self.model.optimizer._config.get('initial_accumulator_value', 0.1)))
        self.step.zero_()

    _SLOT_NAMES = {'SGD': ('momentum',), 'RMSprop': ('rms', 'momentum', 'mg'), 'Adam': ('m', 'v', 'vhat'),
                   'Adagrad': ('accumulator',), 'Adadelta': ('accum_grad', 'accum_var'), 'Adamax': ('m', 'v')}

    def import_optimizer_state(self):
        """Python optimizer -> flat slots (resume from a checkpoint, or a model that trained eagerly before)."""
        opt = self.model.optimizer
        self.step.fill_(int(opt.iterations))
        names = self._SLOT_NAMES[opt._class_name]
        for i, (p, off) in enumerate(zip(self.params, self.offsets)):
            for j, nm in enumerate(names):
                s = opt._slots.get((i, nm))
                if s is not None and self.slots[j] is not None and s.numel() == p.numel():
                    self.slots[j][off:off + p.numel()].copy_(s.reshape(-1))

    def export_optimizer_state(self):
        """Flat slots -> Python optimizer (what ``save`` / ``get_config`` / an eager continuation read)."""
        opt = self.model.optimizer
        opt.iterations = int(self.step.item())
        names = self._SLOT_NAMES[opt._class_name]
        for i, (p, off) in enumerate(zip(self.params, self.offsets)):
            for j, nm in enumerate(names):
                if self.slots[j] is not None:
                    opt._slots[(i, nm)] = self.slots[j][off:off + p.numel()].view(p.shape).clone()

    def _opt_kwargs(self):
        cfg = self.model.optimizer._config
        return dict(lr=cfg['learning_rate'], decay=cfg.get('decay', 0.0) or 0.0,
                    momentum=cfg.get('momentum', 0.0) or 0.0, rho=cfg.get('rho', 0.9), eps=cfg.get('epsilon', 1e-7),
                    beta1=cfg.get('beta_1', 0.9), beta2=cfg.get('beta_2', 0.999), nesterov=cfg.get('nesterov', False),
                    centered=cfg.get('centered', False), amsgrad=cfg.get('amsgrad', False))

    def still_valid(self):
        """False once somebody replaced the model's weight tensors (``model.to``) or detached the gradient views."""
        cur = self.model.trainable_weights
        if len(cur) != len(self.params):
            return False
        base = self.w32.data_ptr()
        for p, q, off in zip(cur, self.params, self.offsets):
            if p is not q or p.data_ptr() != base + 4 * off:
                return False
        return True

    # ------------------------------------------------------------------- one step
    def _body(self, x, y):
        model = self.model
        self.g32.zero_()
        use16 = self._bf16_compute()
        if use16:
            native.cast_f32_to_bf16(self.w32, self.w16)
        with model._autocast(x), (_cnn.fresh_weight_copy() if use16 else contextlib.nullcontext()):
            out = model(x, training=True)
        out = out.float()
        per_sample = model._loss_fn(y, out)
        penalty = model._regularization_loss()  # None unless a layer carries a regularizer
        if penalty is None:
            per_sample.mean().backward()
        else:
            (per_sample.mean() + penalty).backward()
        if self.sync is None:
            native.optimizer_step(self.kind, self.numel, self.g32, self.step, w32=self.w32, slots=self.slots,
                                  apply=True, advance=True, **self._opt_kwargs())
        else:
            # the step of this replica goes out as a bf16 delta; ONE fused kernel averages the deltas of all ranks over
            # NVLink and applies the mean to the fp32 weights (the reference's collect + average, per batch)
            native.optimizer_step(self.kind, self.numel, self.g32, self.step, slots=self.slots, apply=False,
                                  delta_bf16=self.sync.delta(self.flip), advance=True, **self._opt_kwargs())
            if not self._in_warmup:
                self.sync.reduce(self.flip)
        model._apply_constraints()
        with torch.no_grad():
            self.stats[0] += per_sample.detach().double().sum()
            if penalty is not None:
                self.stats[0] += penalty.detach().double() * float(x.shape[0])
            for i, (_, fn) in enumerate(model._metric_fns[:4]):
                self.stats[1 + i] += fn(y, out.detach()).double().sum()
            self.stats[7] += float(x.shape[0])

    def _bf16_compute(self):
        return getattr(self.model, '_compute_dtype', 'float32') == 'bfloat16' and _cnn.enabled()

    def _record(self, key, x, y):
        from ..utils import tracing
        with tracing.span('torch_step/capture'):
            return self._record_inner(key, x, y)

    def _record_inner(self, key, x, y):
        sx, sy = torch.empty_like(x), torch.empty_like(y)
        sx.copy_(x)
        sy.copy_(y)
        stream = torch.cuda.Stream(self.device)
        stream.wait_stream(torch.cuda.current_stream(self.device))
        # warm-up outside of capture (cuDNN algorithm selection, lazy module loads) on saved state
        saved = (self.w32.clone(), [s.clone() if s is not None else None for s in self.slots], self.step.clone(),
                 self.stats.clone(), self._buffers_snapshot())
        self._in_warmup = True  # (collectives run exactly once per step on every rank: never in the warm-up passes)
        try:
            with torch.cuda.stream(stream):
                for _ in range(2):
                    self._body(sx, sy)
        finally:
            self._in_warmup = False
        torch.cuda.current_stream(self.device).wait_stream(stream)
        stream.synchronize()
        self._restore(saved)
        graph = torch.cuda.CUDAGraph()
        launched = native.launch_count()
        gc.collect()
        was_enabled = gc.isenabled()
        gc.disable()
        try:
            with torch.cuda.graph(graph, stream=stream, capture_error_mode='thread_local'):
                self._body(sx, sy)
        finally:
            if was_enabled:
                gc.enable()
        entry = (graph, sx, sy, native.launch_count() - launched)
        self.graphs[key] = entry
        self._restore(saved)
        return entry

    def _buffers_snapshot(self):
        """Non-trainable state the forward pass mutates (BatchNormalization moving statistics)."""
        return [(w, w.clone()) for w in self.model.weights if not w.requires_grad]

    def _restore(self, saved):
        w32, slots, step, stats, buffers = saved
        with torch.no_grad():
            self.w32.copy_(w32)
            for dst, src in zip(self.slots, slots):
                if dst is not None:
                    dst.copy_(src)
            self.step.copy_(step)
            self.stats.copy_(stats)
            for w, value in buffers:
                w.copy_(value)

    def train_step(self, x, y):
        """One captured step on device tensors ``x`` / ``y`` (running sums go to ``self.stats``)."""
        key = (tuple(x.shape), tuple(y.shape), x.dtype, y.dtype, self.flip if self.sync is not None else -1)
        entry = self.graphs.get(key)
        if entry is None:
            with _CAPTURE_LOCK:
                entry = self._record(key, x, y)
        graph, sx, sy, launches = entry
        sx.copy_(x)
        sy.copy_(y)
        graph.replay()
        native.count_replay(launches)
        if self.sync is not None:
            self.flip ^= 1

    def read_stats(self, reset=True):
        """``([loss, metrics...], rows)`` means of the running sums since the last reset (ONE device->host read)."""
        host = self.stats.cpu()
        rows = float(host[7])
        n = 1 + len(self.model._metric_fns[:4])
        vals = [float(host[i]) / rows if rows else 0.0 for i in range(n)]
        if reset:
            self.stats.zero_()
        return vals, rows

    def invalidate_graphs(self):
        with _CAPTURE_LOCK:
            self.graphs = {}
