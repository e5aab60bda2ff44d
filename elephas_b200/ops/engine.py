"""Native sm_100a training / inference engine for MLP-shaped models.

Every workload the reference ships is a chain of Dense / Activation / Dropout layers
(``tests/conftest.py:8-40``, ``examples/*.py``). For such models on a CUDA device
``Model.fit / train_on_batch / predict / evaluate`` run here instead of torch autograd:

* forward:   one tcgen05 GEMM per Dense with bias + activation + dropout fused in the epilogue
             (``csrc/gemm_sm100.cu``); the last Dense writes fp32 pre-activations
* loss:      one kernel: output activation + Keras loss + metrics + dL/dz (``csrc/loss_kernels.cu``)
* backward:  per Dense one wgrad GEMM (``dW = dZ^T X``, both operands MN-major, fp32 output straight
             into the flat gradient buffer), one column-sum for ``db``, and one dgrad GEMM whose
             epilogue multiplies by the previous layer's activation derivative and regenerated
             dropout mask (``dZ_prev = (dZ W) * act'(A_prev) * mask``)
* update:    one fused optimizer kernel over the flat parameter buffer (``csrc/optim_kernels.cu``)
             which also refreshes the bf16 compute copy - or, in the synchronous data-parallel path,
             emits the step as a bf16 delta that the fused P2P all-reduce + update kernel consumes.

The whole step is recorded once per batch size into a CUDA graph and replayed (iteration counter,
learning-rate schedule and dropout streams live in device memory), so the small MNIST-sized models
are not launch-bound and Python is off the hot path.

Weights live in one flat fp32 master buffer plus an index-identical bf16 compute copy
(``parallel/flat.py``); ``export_weights`` / ``import_weights`` sync them with the Keras-facing
``Model`` tensors lazily.
"""
import gc
import os
import threading

import numpy as np
import torch

from ..models import layers as L
from ..parallel.flat import FlatLayout
from . import native

_round8 = lambda v: (v + 7) // 8 * 8  # noqa: E731
_CAPTURE_LOCK = threading.Lock()  # one CUDA-graph capture at a time per process
_SIDE_STREAM = os.environ.get('ELEPHAS_B200_SIDE_STREAM', '1') != '0'  # bias-gradient sums next to the GEMMs
# Kernels for layers with <= 8 output units (csrc/narrow_head.cu). Opt-in: on the wide-MLP step they measure the same as
# the tensor-core path (2.446 vs 2.449 ms/step, A/B on one B200) and their weight gradient uses atomics, which costs
# run-to-run bit reproducibility.
_NARROW_HEAD = os.environ.get('ELEPHAS_B200_NARROW_HEAD', '0') != '0'
_FUSED_MLP = os.environ.get('ELEPHAS_B200_FUSED_MLP', '1') != '0'  # K8: whole-MLP forward kernel for predict / evaluate

_METRIC_FOR_FN = {'categorical_accuracy': 'categorical_accuracy', 'binary_accuracy': 'binary_accuracy',
                  'mean_absolute_error': 'mean_absolute_error', 'mean_squared_error': 'mean_squared_error',
                  'mean_absolute_percentage_error': 'mean_absolute_percentage_error'}


class _DenseOp:
    __slots__ = ('layer', 'in_dim', 'out_dim', 'ld_in', 'ld_out', 'act', 'drop_p', 'w_seg', 'b_seg', 'index')


def _plan(model):
    """Translate the layer chain into fused Dense ops; returns None when the model does not qualify."""
    from ..models import activations as A
    chain = [l for l in model.layers]
    if not chain or not isinstance(chain[0], L.InputLayer):
        return None
    if len(getattr(model, '_inputs', [])) != 1 or len(getattr(model, '_outputs', [])) != 1:
        return None
    in_shape = chain[0].batch_input_shape
    if in_shape is None or len(in_shape) != 2 or in_shape[1] is None:
        return None
    # every layer must consume exactly the previous layer's single output
    prev = chain[0]
    for layer in chain[1:]:
        if len(layer.inbound) != 1 or len(layer.inbound[0]) != 1 or layer.inbound[0][0].layer is not prev:
            return None
        prev = layer
    ops = []
    out_act = None
    i = 1
    while i < len(chain):
        layer = chain[i]
        if not isinstance(layer, L.Dense) or not layer.use_bias:
            return None
        if not A.is_builtin(layer.activation_name, layer.activation):
            return None
        act = layer.activation_name
        drop = 0.0
        j = i + 1
        if j < len(chain) and isinstance(chain[j], L.Activation) and act == 'linear':
            if not A.is_builtin(chain[j].activation_name, chain[j].activation):
                return None
            act = chain[j].activation_name
            j += 1
        if j < len(chain) and isinstance(chain[j], L.Dropout):
            drop = chain[j].rate
            j += 1
        op = _DenseOp()
        op.layer, op.in_dim, op.out_dim = layer, int(layer.kernel.shape[1]), int(layer.units)
        op.act, op.drop_p, op.index = act, float(drop), len(ops)
        ops.append(op)
        i = j
    if not ops:
        return None
    last = ops[-1]
    if last.drop_p > 0.0:
        return None
    out_act = last.act
    if out_act not in native.ACT_IDS:
        return None
    for op in ops[:-1]:
        if op.act not in native.ACT_IDS or op.act == 'softmax':
            return None
        if op.drop_p > 0.0 and op.out_dim % 4:
            return None
    return ops, out_act


def _training_spec(model, out_act, out_dim):
    """(loss_id, metric_ids) for the fused loss kernel, or None when loss/metrics are not native."""
    loss_name = getattr(model, '_loss_name', None)
    if loss_name not in native.LOSS_IDS:
        return None
    if loss_name == 'categorical_crossentropy' and out_act != 'softmax':
        return None
    if out_act == 'softmax' and loss_name != 'categorical_crossentropy':
        return None
    metric_ids = []
    from ..models import losses as _losses
    from ..models import metrics as _metrics
    for name, fn in model._metric_fns:
        key = None
        if fn is _metrics.categorical_accuracy:
            key = 'categorical_accuracy'
        elif fn is _metrics.binary_accuracy:
            key = 'binary_accuracy'
        elif fn is _losses.mean_absolute_error:
            key = 'mean_absolute_error'
        elif fn is _losses.mean_squared_error:
            key = 'mean_squared_error'
        elif fn is _losses.mean_absolute_percentage_error:
            key = 'mean_absolute_percentage_error'
        if key is None:
            return None
        metric_ids.append(native.METRIC_IDS[key])
    if len(metric_ids) > 4:
        return None
    return native.LOSS_IDS[loss_name], metric_ids


def _optimizer_spec(opt):
    if opt is None or opt._class_name not in native.OPT_IDS:
        return None
    if opt.clipnorm is not None or opt.clipvalue is not None or opt.global_clipnorm is not None:
        return None
    if getattr(opt, 'weight_decay', None) or getattr(opt, 'gradient_accumulation_steps', None):
        return None  # decoupled weight decay / accumulation over several batches are not in the optimizer kernels
    if getattr(opt, 'schedule', None) is not None:
        return None  # the kernels take the rate as a launch argument that a captured graph freezes
    return native.OPT_IDS[opt._class_name]


def native_enabled():
    return os.environ.get('ELEPHAS_B200_NATIVE', '1') != '0'


def maybe_create(model):
    """An :class:`MLPEngine` for a compiled model on a CUDA device, or None (torch path)."""
    if not native_enabled() or model.device.type != 'cuda':
        return None
    if model.has_weight_options():
        return None  # weight / activity penalties and constraints: the captured torch step applies them
    plan = _plan(model)
    if plan is None:
        return None
    ops, out_act = plan
    if any(not op.layer.trainable for op in ops):
        return None  # frozen layers: the fused step updates the whole flat buffer; the captured torch step skips them
    spec = _training_spec(model, out_act, ops[-1].out_dim)
    kind = _optimizer_spec(model.optimizer)
    if spec is None or kind is None:
        return None
    return MLPEngine(model, ops, out_act, spec, kind)


def maybe_create_inference(model):
    """Inference-only engine (uncompiled models: ``ElephasTransformer``, ``SparkModel.predict``)."""
    if not native_enabled() or model.device.type != 'cuda':
        return None
    cached = getattr(model, '_infer_engine', None)
    if cached is not None:
        return cached
    plan = _plan(model)
    if plan is None:
        return None
    ops, out_act = plan
    eng = MLPEngine(model, ops, out_act, None, None)
    model._infer_engine = eng
    return eng


class _BatchBuffers:
    """Activations / gradients / graph for one batch size."""

    def __init__(self, eng, B):
        dev = eng.device
        self.B = B
        self.x32 = torch.zeros(B, eng.ops[0].in_dim, dtype=torch.float32, device=dev)
        self.y32 = torch.zeros(B, eng.y_cols, dtype=torch.float32, device=dev)
        self.x16 = torch.zeros(B, eng.ops[0].ld_in, dtype=torch.bfloat16, device=dev)
        self.acts = [torch.zeros(B, op.ld_out, dtype=torch.bfloat16, device=dev) for op in eng.ops[:-1]]
        n_out = eng.ops[-1].out_dim
        self.z = torch.zeros(B, max(4, (n_out + 3) // 4 * 4), dtype=torch.float32, device=dev)
        self.out = torch.zeros(B, n_out, dtype=torch.float32, device=dev)
        self.dz = [torch.zeros(B, op.ld_out, dtype=torch.bfloat16, device=dev) for op in eng.ops]
        self.graphs = {}


class MLPEngine:
    def __init__(self, model, ops, out_act, spec, opt_kind):
        native.require()
        self.model = model
        self.device = model.device
        self.ops = ops
        self.out_act = out_act
        self.layout = FlatLayout(model)
        seg_of = {}
        for seg in self.layout.segments:
            seg_of[(id(seg.layer), seg.index)] = seg
        for op in ops:
            op.w_seg = seg_of[(id(op.layer), 0)]
            op.b_seg = seg_of[(id(op.layer), 1)]
            op.ld_in = op.w_seg.ld
            op.ld_out = _round8(op.out_dim)
        for a, b in zip(ops[:-1], ops[1:]):
            assert a.ld_out == b.ld_in, "activation / weight leading dimensions must agree"
        n = self.layout.numel
        dev = self.device
        with torch.cuda.device(dev):
            self.w32 = self.layout.pack_model(model, device=dev)
            self.w16 = torch.zeros(n, dtype=torch.bfloat16, device=dev)
            native.cast_f32_to_bf16(self.w32, self.w16)
        self.trainable = spec is not None
        self.y_cols = ops[-1].out_dim
        self.use_graphs = os.environ.get('ELEPHAS_B200_GRAPHS', '1') != '0'
        self._buffers = {}
        self._graph_launches = {}
        self._weights_dirty_in_model = False  # True: engine holds newer weights than the Model tensors
        self.seed = int(torch.randint(0, 2 ** 62, (1,)).item())
        self.stats = torch.zeros(8, dtype=torch.float64, device=dev)
        self._pinned_stats = torch.zeros(8, dtype=torch.float64).pin_memory()
        self.step = torch.zeros(1, dtype=torch.int64, device=dev)
        self.delta_target = None  # (bf16 tensor) when the optimizer must emit deltas instead of applying
        self.comm_hook = None     # callable(lo, n): a delta range is ready (synchronous data parallel)
        self._aux_stream = None
        self.input_bf16 = False   # training batches are delivered straight into buf.x16 (already bf16)
        self.pull_source = None   # per Dense op: device address of the bf16 mirror of its kernel in a parameter server's
        #                           stripe (usually peer memory) that the training forward GEMM pulls from (K2); None = off
        self.grad_sink = None     # callable(warmup): consumes the complete flat gradient g32 INSTEAD of the local optimizer
        #                           (parameter server: optimizer step fused with the push into peer memory)
        self.pre_step = None      # callables run (and captured) at the start / end of every training step
        self.post_step = None
        self.mid_step = None      # ... and between the forward pass and the loss
        self.comm_join = None     # callable(): make the compute stream wait for all communication of the step
        self.last_graph_kernels = 0  # native kernels in the most recently recorded graph
        self._hook_generation = 0  # bumped by set_step_hooks(): part of the graph key (never an object address)
        self._in_warmup = False
        if self.trainable:
            self.loss_id, self.metric_ids = spec
            self.opt_kind = opt_kind
            self.g32 = torch.zeros(n, dtype=torch.float32, device=dev)
            self._init_optimizer_state()

    # ------------------------------------------------------------- optimizer state
    def _init_optimizer_state(self):
        opt = self.model.optimizer
        n, dev = self.layout.numel, self.device
        cfg = opt._config
        name = opt._class_name
        need = {'SGD': (cfg.get('momentum', 0.0) != 0.0, False, False),
                'RMSprop': (True, cfg.get('momentum', 0.0) > 0.0, bool(cfg.get('centered'))),
                'Adam': (True, True, bool(cfg.get('amsgrad'))), 'Adagrad': (True, False, False),
                'Adadelta': (True, True, False), 'Adamax': (True, True, False)}[name]
        init0 = float(cfg.get('initial_accumulator_value', 0.0)) if name == 'Adagrad' else 0.0
        self.slots = [torch.full((n,), init0 if i == 0 else 0.0, dtype=torch.float32, device=dev) if flag else None
                      for i, flag in enumerate(need)]
        self.step.fill_(int(opt.iterations))
        self._import_slots()

    def reset_optimizer(self, optimizer=None):
        """Fresh optimizer state (zero slots, iteration 0), as every reference worker starts a fit with
        a newly built optimizer (``elephas/worker.py:29,88``). CUDA graphs stay valid: pointers do not change."""
        if not self.trainable:
            return
        if optimizer is not None:
            self.model.optimizer = optimizer
        opt = self.model.optimizer
        init0 = float(opt._config.get('initial_accumulator_value', 0.0)) if opt._class_name == 'Adagrad' else 0.0
        for i, sl in enumerate(self.slots):
            if sl is not None:
                sl.fill_(init0 if i == 0 else 0.0)
        self.step.zero_()
        opt.iterations = 0
        opt._slots = {}

    _SLOT_NAMES = {'SGD': ('momentum',), 'RMSprop': ('rms', 'momentum', 'mg'), 'Adam': ('m', 'v', 'vhat'),
                   'Adagrad': ('accumulator',), 'Adadelta': ('accum_grad', 'accum_var'), 'Adamax': ('m', 'v')}

    def _trainable_segments(self):
        params = self.model.trainable_weights
        out = []
        for seg, (_, _, w) in zip(self.layout.segments, self.model._weight_entries()):
            for i, p in enumerate(params):
                if p is w:
                    out.append((i, seg))
        return out

    def _import_slots(self):
        """Bring optimizer slots restored from a checkpoint (torch tensors) into the flat buffers."""
        opt = self.model.optimizer
        names = self._SLOT_NAMES[opt._class_name]
        for i, seg in self._trainable_segments():
            for k, sname in enumerate(names):
                t = opt._slots.get((i, sname))
                if t is not None and k < len(self.slots) and self.slots[k] is not None:
                    self.layout.view(self.slots[k], seg).copy_(t.to(self.device).reshape(seg.rows, seg.cols))

    def export_optimizer_state(self):
        """Mirror the flat slots / step counter back into the Keras-facing optimizer object."""
        if not self.trainable:
            return
        opt = self.model.optimizer
        opt.iterations = int(self.step.item())
        names = self._SLOT_NAMES[opt._class_name]
        for i, seg in self._trainable_segments():
            for k, sname in enumerate(names):
                if k < len(self.slots) and self.slots[k] is not None:
                    opt._slots[(i, sname)] = self.layout.view(self.slots[k], seg).reshape(seg.shape).clone()

    # ------------------------------------------------------------------ weights
    def export_weights(self):
        """Engine -> Model tensors (called lazily by ``Model.get_weights`` / ``save``)."""
        if self._weights_dirty_in_model:
            self.layout.unpack_to_model(self.w32, self.model)
            self._weights_dirty_in_model = False
            self.export_optimizer_state()

    def import_weights(self):
        """Model tensors -> engine (called by ``Model.set_weights``)."""
        with torch.cuda.device(self.device):
            self.layout.pack_model(self.model, out=self.w32)
            native.cast_f32_to_bf16(self.w32, self.w16)
        self._weights_dirty_in_model = False

    def refresh_bf16(self):
        native.cast_f32_to_bf16(self.w32, self.w16)

    # ------------------------------------------------------------------ kernels
    def _buf(self, B):
        buf = self._buffers.get(B)
        if buf is None:
            with torch.cuda.device(self.device):
                buf = _BatchBuffers(self, B)
            self._buffers[B] = buf
        return buf

    def _w16(self, op):
        return self.w16[op.w_seg.offset:op.w_seg.offset + op.w_seg.span].view(op.out_dim, op.ld_in)

    def _forward(self, buf, training):
        B = buf.B
        if not (self.input_bf16 and training):  # (streamed training inputs arrive already cast to bf16 by the host: half the PCIe bytes)
            native.stage_f32_to_bf16(buf.x32, buf.x16, B, self.ops[0].in_dim)
        a = buf.x16
        if not training and self._fused_mlp_ok():
            # K8: the whole chain in one persistent kernel, hidden activations never leave the SM
            native.mlp_forward(buf.x16, B, self.ops[0].ld_in,
                               [(self._w16(op), op.out_dim, self.w32[op.b_seg.offset:], native.ACT_IDS[op.act])
                                for op in self.ops], buf.z, ldx=self.ops[0].ld_in)
            return
        pull = self.pull_source if training else None

        def weights(op):
            # K2: with a pull source (address of a parameter server's bf16 mirror, usually peer memory) the GEMM reads
            # the weight tiles from there through TMA and side-writes them into the local copy the backward pass uses
            if pull is None or pull[op.index] is None:
                return self._w16(op), None
            return int(pull[op.index]), self._w16(op)

        for op in self.ops[:-1]:
            out = buf.acts[op.index]
            w, w_copy = weights(op)
            native.gemm(a, w, B, op.out_dim, op.in_dim, lda=op.ld_in, ldb=op.ld_in,
                        out_bf16=out, ldo=op.ld_out, bias=self.w32[op.b_seg.offset:],
                        act=native.ACT_IDS[op.act], drop_p=op.drop_p if training else 0.0, seed=self.seed,
                        drop_stream=op.index, step_ptr=self.step if training else None, b_copy=w_copy)
            a = out
        op = self.ops[-1]
        w, w_copy = weights(op)
        if self._narrow(op) and w_copy is None:
            # a handful of output units: one pass over the activations at HBM speed (csrc/narrow_head.cu)
            native.head_fwd(a, w, B, op.in_dim, op.out_dim, buf.z, bias=self.w32[op.b_seg.offset:], lda=op.ld_in,
                            ldw=op.ld_in)
        else:
            native.gemm(a, w, B, op.out_dim, op.in_dim, lda=op.ld_in, ldb=op.ld_in,
                        out_f32=buf.z, ldo32=buf.z.stride(0), bias=self.w32[op.b_seg.offset:], act=0, b_copy=w_copy)

    def _fused_mlp_ok(self):
        """Inference forward through ``eb_mlp_forward`` (K8)? Small Dense chains only (see its limits)."""
        ok = getattr(self, '_fused_mlp', None)
        if ok is None:
            units = [op.out_dim for op in self.ops]
            ok = _FUSED_MLP and native.mlp_forward_supported(self.ops[0].ld_in, units)
            if ok and max(units) > 128:
                # streamed-weights variant (layers up to 512 wide): the better choice for small row counts; on big blocks the
                # chain of K1 GEMMs overlaps MMA and epilogue of a layer, which one CTA per row tile cannot
                ok = os.environ.get('ELEPHAS_B200_FUSED_MLP_WIDE', '0') == '1'
            self._fused_mlp = ok
        return ok

    def _narrow(self, op):
        """Last layer with at most 8 output units and an 8-aligned input width: served by the narrow-head kernels."""
        return (_NARROW_HEAD and op is self.ops[-1] and op.out_dim <= native.HEAD_MAX_OUT and op.in_dim % 8 == 0
                and op.in_dim * op.out_dim * 2 <= 96 * 1024)

    def _loss(self, buf, with_grad, rows=None, grad_rows=None):
        op = self.ops[-1]
        rows = buf.B if rows is None else rows
        native.loss_fwd_bwd(buf.z, buf.y32, rows, op.out_dim, self.loss_id, native.ACT_IDS[self.out_act],
                            out=buf.out, dz=buf.dz[op.index] if with_grad else None, stats=self.stats,
                            metrics=self.metric_ids, grad_scale=1.0 / float(grad_rows or rows),
                            ldy=buf.y32.stride(0))

    def _fused_sgd(self):
        """Plain SGD (no momentum, no decay): the weight update rides in the wgrad GEMM epilogue."""
        if not self.trainable or self.opt_kind != native.OPT_IDS['SGD']:
            return False
        cfg = self.model.optimizer._config
        return not cfg.get('momentum') and not cfg.get('decay')

    def _opt_kwargs(self):
        cfg = self.model.optimizer._config
        return dict(lr=cfg['learning_rate'], decay=cfg.get('decay', 0.0) or 0.0,
                    momentum=cfg.get('momentum', 0.0) or 0.0, rho=cfg.get('rho', 0.9), eps=cfg.get('epsilon', 1e-7),
                    beta1=cfg.get('beta_1', 0.9), beta2=cfg.get('beta_2', 0.999), nesterov=cfg.get('nesterov', False),
                    centered=cfg.get('centered', False), amsgrad=cfg.get('amsgrad', False))

    def _optimizer_range(self, lo, n, advance):
        """Optimizer kernel over flat elements [lo, lo+n): applies the step, or - in delta mode - writes it as
        a bf16 delta into the communication buffer."""
        delta = self.delta_target
        native.optimizer_step(self.opt_kind, n, self.g32[lo:], self.step, w32=self.w32[lo:], w16=self.w16[lo:],
                              slots=[sl[lo:] if sl is not None else None for sl in self.slots],
                              delta_bf16=delta[lo:] if delta is not None else None, apply=delta is None,
                              advance=advance, **self._opt_kwargs())

    def _emit(self, lo, n):
        """Delta range [lo, lo+n) is complete: hand it to the communication hook (per-bucket fused all-reduce)."""
        # (skipped during the pre-capture warm-up run: collectives must execute exactly once per step on
        # every rank, and a rank without a batch at this step does not run a warm-up)
        if self.comm_hook is not None and self.delta_target is not None and not self._in_warmup:
            self.comm_hook(lo, n)

    def bucket_ranges(self):
        """The order in which delta ranges are emitted during a step: one bucket per Dense kernel, last layer
        first, then one bucket with all vectors (ranks without a batch replay exactly this sequence)."""
        out = []
        for op in reversed(self.ops):
            off = op.w_seg.offset
            out.append((off, (op.w_seg.span + 63) // 64 * 64))
        out.append((self.layout.matrix_end, self.layout.numel - self.layout.matrix_end))
        return out

    def _backward(self, buf):
        B = buf.B
        fused = self._fused_sgd() and self.grad_sink is None  # a gradient sink wants the plain gradient in g32
        delta_mode = self.delta_target is not None and self.grad_sink is None
        lr = float(self.model.optimizer._config['learning_rate']) if fused else 0.0
        # bias gradients (column sums of dZ, bandwidth-bound) run on a side stream next to the GEMMs: the
        # clustered GEMM grid leaves 16 of the 148 SMs idle, and nothing needs the sums before the optimizer
        side = self._side_stream()
        cur = torch.cuda.current_stream(self.device)
        for op in reversed(self.ops):
            dz = buf.dz[op.index]
            if side is not None:
                ready = torch.cuda.Event()
                ready.record(cur)
                side.wait_event(ready)
                with torch.cuda.stream(side):
                    native.colsum_bf16(dz, B, op.out_dim, self.g32[op.b_seg.offset:], ld=op.ld_out)
            a_prev = buf.x16 if op.index == 0 else buf.acts[op.index - 1]
            off = op.w_seg.offset
            span = (op.w_seg.span + 63) // 64 * 64
            narrow = self._narrow(op)
            if op.index > 0:
                prev = self.ops[op.index - 1]
                # dZ_prev[B, in] = (dZ[B, out] x W[out, in]) * act'(A_prev) * dropout mask
                # (before the wgrad of this layer: the fused-SGD epilogue overwrites W)
                if narrow and prev.drop_p == 0.0:
                    native.head_dgrad(dz, self._w16(op), a_prev, buf.dz[prev.index], B, op.in_dim, op.out_dim,
                                      act=native.ACT_IDS[prev.act], lddz=op.ld_out, ldw=op.ld_in, ldaux=prev.ld_out,
                                      lddzp=prev.ld_out)
                else:
                    native.gemm(dz, self._w16(op), B, op.in_dim, op.out_dim, b_mn=True, lda=op.ld_out, ldb=op.ld_in,
                                out_bf16=buf.dz[prev.index], ldo=prev.ld_out, epi=native.EPI_ACTGRAD,
                                act=native.ACT_IDS[prev.act], aux=a_prev, ldaux=prev.ld_out, drop_p=prev.drop_p,
                                seed=self.seed, drop_stream=prev.index, step_ptr=self.step)
            # dW[out, in] = dZ^T[out, B] x A_prev[B, in]: both operands MN-major
            if narrow:
                # dZ-weighted column sums into the gradient buffer, then the (tiny) optimizer range of this kernel
                native.head_wgrad(dz, a_prev, B, op.in_dim, op.out_dim, self.g32[off:], lddz=op.ld_out, lda=op.ld_in,
                                  lddw=op.ld_in)
                if fused or delta_mode:
                    self._optimizer_range(off, span, advance=False)
            elif not fused:
                native.gemm(dz, a_prev, op.out_dim, op.in_dim, B, a_mn=True, b_mn=True, lda=op.ld_out,
                            ldb=op.ld_in, out_f32=self.g32[off:], ldo32=op.ld_in)
                if delta_mode:
                    self._optimizer_range(off, span, advance=False)
            elif not delta_mode:
                # W32 <- W32 - lr * dW and W16 <- bf16(W32) in the epilogue: no gradient buffer, no extra pass
                native.gemm(dz, a_prev, op.out_dim, op.in_dim, B, a_mn=True, b_mn=True, lda=op.ld_out,
                            ldb=op.ld_in, out_f32=self.w32[off:], ldo32=op.ld_in, beta=1.0, alpha=-lr,
                            out_bf16=self.w16[off:], ldo=op.ld_in, bf16_from_f32=True)
            else:
                # synchronous data parallel: the step lr * dW goes straight into the peer-mapped delta buffer
                native.gemm(dz, a_prev, op.out_dim, op.in_dim, B, a_mn=True, b_mn=True, lda=op.ld_out,
                            ldb=op.ld_in, alpha=lr, out_bf16=self.delta_target[off:], ldo=op.ld_in)
            if delta_mode:
                self._emit(off, span)  # this layer's delta can travel while the earlier layers are computed
            if side is None:
                native.colsum_bf16(dz, B, op.out_dim, self.g32[op.b_seg.offset:], ld=op.ld_out)
        if side is not None:
            done = torch.cuda.Event()
            done.record(side)
            cur.wait_event(done)

    def _side_stream(self):
        if not _SIDE_STREAM:
            return None
        if self._aux_stream is None:
            self._aux_stream = torch.cuda.Stream(self.device)
        return self._aux_stream

    def _pull_key(self):
        return tuple(int(a or 0) for a in self.pull_source) if self.pull_source is not None else 0

    def _optimizer(self):
        if self.grad_sink is not None:
            self.grad_sink(self._in_warmup)
            return
        if self.delta_target is not None or self._fused_sgd():
            # matrices were handled layer by layer (fused epilogue or per-layer delta): only vectors are left
            lo = self.layout.matrix_end
        else:
            lo = 0
        n = self.layout.numel - lo
        self._optimizer_range(lo, n, advance=True)
        if self.delta_target is not None:
            self._emit(lo, n)
        if self.comm_join is not None and self.delta_target is not None and not self._in_warmup:
            self.comm_join()

    def _train_body(self, buf):
        if self.pre_step is not None:
            self.pre_step()   # e.g. pull of the bias tail from a parameter server (captured with the step)
        self._forward(buf, training=True)
        if self.mid_step is not None:
            self.mid_step()   # e.g. release of a parameter-server lock held while the forward GEMMs pulled the weights
        self._loss(buf, with_grad=True)
        self._backward(buf)
        self._optimizer()
        if self.post_step is not None and not self._in_warmup:
            self.post_step()  # e.g. push of the step's delta into a parameter server

    def set_step_hooks(self, pre=None, mid=None, post=None):
        """Install the callables captured at the start / between forward and loss / at the end of every training step
        and return the previous triple. Every call starts a new hook generation, which is part of the CUDA-graph key:
        a graph recorded with one set of hooks is never replayed for another, whatever addresses the callables have."""
        previous = (self.pre_step, self.mid_step, self.post_step)
        self.pre_step, self.mid_step, self.post_step = pre, mid, post
        self._hook_generation += 1
        return previous

    # -------------------------------------------------------------- graph capture
    def _run(self, buf, which, body):
        """Run ``body`` through a CUDA graph recorded on first use. One graph per batch size, kind
        (train / eval / infer) and - for training - delta target, because pointers are baked in."""
        if not self.use_graphs:
            body()
            return
        is_train = which == 'train'
        key = self._graph_key(which)
        graph = buf.graphs.get(key)
        if graph is None:
            with _CAPTURE_LOCK:
                graph = self._record(buf, key, body, is_train)
        graph.replay()
        native.count_replay(self._graph_launches.get(id(graph), 0))

    def _graph_key(self, which):
        is_train = which == 'train'
        return (which, self.delta_target.data_ptr() if (is_train and self.delta_target is not None) else 0,
                self._pull_key() if is_train else 0, self._hook_generation if is_train else 0,
                bool(self.input_bf16) if is_train else False, (self.grad_sink is not None) if is_train else False)

    def _record(self, buf, key, body, is_train):
        # warm-up outside capture (module loading, function attributes), then record
        stream = torch.cuda.Stream(self.device)
        stream.wait_stream(torch.cuda.current_stream(self.device))
        saved = self._snapshot_state() if is_train else None
        self._in_warmup = True
        try:
            with torch.cuda.stream(stream):
                body()
        finally:
            self._in_warmup = False
        torch.cuda.current_stream(self.device).wait_stream(stream)
        if saved is not None:
            self._restore_state(saved)
        graph = torch.cuda.CUDAGraph()
        launched = native.launch_count()
        # No garbage collection while capturing: a collected CUDAGraph / stream / event of an earlier fit would be
        # destroyed from inside the capture, which CUDA forbids and which invalidates the capture.
        gc.collect()
        gc_was_enabled = gc.isenabled()
        gc.disable()
        try:
            with torch.cuda.graph(graph, stream=stream, capture_error_mode='thread_local'):
                body()
        finally:
            if gc_was_enabled:
                gc.enable()
        # kernels recorded into the graph: every replay launches this many native kernels
        self._graph_launches[id(graph)] = self.last_graph_kernels = native.launch_count() - launched
        buf.graphs[key] = graph
        if saved is not None:
            self._restore_state(saved)
        return graph

    def _snapshot_state(self):
        # stream-level syncs only: another worker thread may be capturing a graph on its own stream
        torch.cuda.current_stream(self.device).synchronize()
        return (self.w32.clone(), self.w16.clone(), [s.clone() if s is not None else None for s in self.slots],
                self.step.clone(), self.stats.clone())

    def _restore_state(self, saved):
        torch.cuda.current_stream(self.device).synchronize()
        w32, w16, slots, step, stats = saved
        self.w32.copy_(w32)
        self.w16.copy_(w16)
        for dst, src in zip(self.slots, slots):
            if dst is not None:
                dst.copy_(src)
        self.step.copy_(step)
        self.stats.copy_(stats)

    def invalidate_graphs(self):
        with _CAPTURE_LOCK:  # never destroy graph objects while another worker thread is capturing
            for buf in self._buffers.values():
                buf.graphs = {}

    # ------------------------------------------------------------------ public
    def _to_device_xy(self, x, y, buf, rows):
        xt = x if isinstance(x, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32))
        xt = xt.reshape(rows, -1)
        buf.x32[:rows].copy_(xt, non_blocking=True)
        if y is not None:
            yt = y if isinstance(y, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(y, dtype=np.float32))
            buf.y32[:rows].copy_(yt.reshape(rows, -1), non_blocking=True)

    def read_stats(self, reset=True):
        """One 64-byte device->host read: ``[sum loss, sum metric_k..., rows]`` -> per-sample means."""
        self._pinned_stats.copy_(self.stats, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        s = self._pinned_stats.clone().numpy()
        if reset:
            self.stats.zero_()
        n = max(s[7], 1.0)
        return [float(s[0] / n)] + [float(s[1 + k] / n) for k in range(len(self.metric_ids))], int(s[7])

    def train_on_batch(self, x, y):
        with torch.cuda.device(self.device):
            rows = len(x)
            buf = self._buf(rows)
            self.stats.zero_()
            self._to_device_xy(x, y, buf, rows)
            self._run(buf, 'train', lambda: self._train_body(buf))
            self._weights_dirty_in_model = True
            return self.read_stats()[0]

    def prepare_train(self, buf):
        """Record the training-step graph for ``buf`` now (no step is taken: the recording runs on a snapshot of the
        engine state that is restored afterwards), so that a timed loop of ``train_step_device`` only replays."""
        if not self.use_graphs:
            return
        key = self._graph_key('train')
        if key not in buf.graphs:
            with _CAPTURE_LOCK:
                self._record(buf, key, lambda: self._train_body(buf), True)

    def train_step_device(self, buf):
        """One step on data already staged in ``buf.x32`` / ``buf.y32`` (no host sync)."""
        self._run(buf, 'train', lambda: self._train_body(buf))
        self._weights_dirty_in_model = True

    def fit(self, x, y, batch_size, epochs, shuffle, history, x_val, y_val, verbose, initial_epoch=0,
            batch_losses=False):
        with torch.cuda.device(self.device):
            n = len(x)
            dev = self.device
            xd = torch.as_tensor(np.ascontiguousarray(x, dtype=np.float32)).reshape(n, -1).to(dev, non_blocking=True)
            yd = torch.as_tensor(np.ascontiguousarray(y, dtype=np.float32)).reshape(n, -1).to(dev, non_blocking=True)
            names = self.model.metrics_names
            for epoch in range(initial_epoch, epochs):
                # CPU generator: the CUDA generator may be in graph-capture mode in another worker thread
                order = torch.randperm(n).to(dev, non_blocking=True) if shuffle else None
                self.stats.zero_()
                for start in range(0, n, batch_size):
                    rows = min(batch_size, n - start)
                    buf = self._buf(rows)
                    if order is not None:
                        idx = order[start:start + rows]
                        torch.index_select(xd, 0, idx, out=buf.x32)
                        torch.index_select(yd, 0, idx, out=buf.y32)
                    else:
                        buf.x32.copy_(xd[start:start + rows])
                        buf.y32.copy_(yd[start:start + rows])
                    self.train_step_device(buf)
                    if batch_losses:
                        vals, _ = self.read_stats(reset=False)
                        history.history.setdefault('batch_loss_running', []).append(vals[0])
                vals, _ = self.read_stats()
                logs = dict(zip(names, vals))
                if x_val is not None and len(x_val):
                    val = self.evaluate(x_val, y_val)
                    logs.update({'val_' + k: v for k, v in zip(names, val)})
                history._append(epoch, logs)
                if verbose:
                    print("Epoch %d/%d - " % (epoch + 1, epochs) + " - ".join("%s: %.4f" % kv for kv in logs.items()))
            self.model.optimizer.iterations = int(self.step.item())

    _CHUNK = 8192

    def predict(self, x):
        with torch.cuda.device(self.device):
            x = np.ascontiguousarray(x, dtype=np.float32) if not isinstance(x, torch.Tensor) else x
            n = len(x)
            x = x.reshape(n, -1)
            outs = []
            for start in range(0, n, self._CHUNK):
                rows = min(self._CHUNK, n - start)
                buf = self._buf(rows)
                self._to_device_xy(x[start:start + rows], None, buf, rows)
                self._run(buf, 'infer', lambda b=buf: (self._forward(b, training=False), self._activate(b)))
                outs.append(buf.out[:rows].clone())
            if not outs:
                return np.zeros((0, self.ops[-1].out_dim), dtype=np.float32)
            return torch.cat(outs, dim=0).cpu().numpy()

    def _activate(self, buf):
        """Output activation only (inference): the loss kernel with no targets-dependent outputs used."""
        op = self.ops[-1]
        native.loss_fwd_bwd(buf.z, buf.y32, buf.B, op.out_dim, native.LOSS_IDS['mean_squared_error']
                            if self.out_act != 'softmax' else native.LOSS_IDS['categorical_crossentropy'],
                            native.ACT_IDS[self.out_act], out=buf.out, dz=None, stats=None,
                            ldy=buf.y32.stride(0))

    def evaluate(self, x, y):
        with torch.cuda.device(self.device):
            saved = self.stats.clone()
            self.stats.zero_()
            x = np.ascontiguousarray(x, dtype=np.float32) if not isinstance(x, torch.Tensor) else x
            y = np.ascontiguousarray(y, dtype=np.float32) if not isinstance(y, torch.Tensor) else y
            n = len(x)
            x, y = x.reshape(n, -1), y.reshape(n, -1)
            for start in range(0, n, self._CHUNK):
                rows = min(self._CHUNK, n - start)
                buf = self._buf(rows)
                self._to_device_xy(x[start:start + rows], y[start:start + rows], buf, rows)
                self._run(buf, 'eval', lambda b=buf: (self._forward(b, training=False),
                                                           self._loss(b, with_grad=False)))
            vals, _ = self.read_stats()
            self.stats.copy_(saved)
            return vals

