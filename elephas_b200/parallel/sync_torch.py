"""Synchronous mode for models the native Dense engine does not cover (CNNs, custom activations), one rank per GPU.

Same semantics as ``sync_fast`` (and as the reference: every worker trains a replica from the same weights, the driver
applies the mean of ``before - after``, ``elephas/spark_model.py:217-231``), same structure: the rank's copy of the
master network *is* the replica. Its trainable weights live in the flat fp32 buffer of the graph-captured torch step
(``ops/torch_step.py``), so a round is

    before <- w32 ; [epoch of CUDA-graph replays] ; delta = before - w32 (K5, straight into peer-mapped memory) ;
    w32 <- before ; fused P2P all-reduce + update (K6)

without a model rebuild from JSON, without host copies of the weights and without re-capturing the step. Non-trainable
state (BatchNormalization moving statistics) is averaged over ranks with one packed all-reduce per round.
``frequency='batch'`` (the DDP-like mode BASELINE.json names for the ResNet-50 config) stays on this path as well: the
optimizer kernel of the captured step emits its update as a bf16 delta into peer-mapped memory and ONE fused kernel
(K6, two-shot over NVLink) averages the deltas of all ranks and applies the mean to the fp32 weights - inside the same
CUDA graph, once per step. The payload of a ResNet-50 (51 MB in bf16) takes about 1-2 % of a step, so it is sent as one
bucket after the backward pass rather than layer by layer. Ranks must have the same number of batches per epoch.
"""
import numpy as np
import torch

from ..models import History
from ..ops import engine as _engine
from ..ops import native
from ..utils import tracing
from ..worker import partition_to_arrays
from .symm import SymmetricMemory

IN_PLACE = object()


def try_fit(spark_model, rt, partitions, mine, train_config, sync_rounds):
    if rt.device.type != 'cuda' or not _engine.native_enabled():
        return None
    if len(partitions) != rt.world_size or len(mine) != 1:
        return None
    if any(train_config.get(k) is not None and train_config.get(k) != [] for k in ('callbacks', 'class_weight', 'sample_weight')):
        return None  # callbacks and row weights belong to Model.fit: the generic worker path calls it with them
    per_batch = rt.distributed and spark_model.frequency == 'batch' and sync_rounds != 'fit'
    if per_batch:
        bs = int(train_config.get('batch_size') or 32)
        rows = len(partition_to_arrays(mine[0][1])[0])
        steps = -(-rows // bs) if rows > bs else 0
        if len(set(rt.all_gather_object((steps, rows % bs)))) != 1 or steps == 0:
            return None  # unequal shards: the generic worker path handles ranks that run out of batches
    master = spark_model.master_network
    if master.device != rt.device:
        master.to(rt.device)
    te = master._torch_engine()
    if not all(rt.all_gather_object(te is not None)):
        return None
    ctx = getattr(spark_model, '_sync_torch_context', None)
    if ctx is None or ctx.te is not te:
        if ctx is not None:
            ctx.close()
        ctx = TorchSyncContext(spark_model, rt, master, te)
        spark_model._sync_torch_context = ctx
    histories = ctx.run(mine[0][1], train_config, sync_rounds)
    if not train_config.get('keep_context'):
        ctx.close()
        spark_model._sync_torch_context = None
    return IN_PLACE, histories


class _StepSync:
    """What the captured step of ``TorchStepEngine`` calls in per-batch mode."""

    def __init__(self, ctx):
        self.delta, self.reduce = ctx.delta_buffer, ctx.reduce_step


class TorchSyncContext:
    def __init__(self, spark_model, rt, model, te):
        self.sm, self.rt, self.model, self.te = spark_model, rt, model, te
        self.mem = None

    def _setup_comm(self):
        if self.mem is not None:
            return
        n = self.te.numel
        # 'delta' / 'avg': fp32 buffers of the per-epoch round; 'd16_*' / 'a16': bf16 buffers of the per-batch mode
        # (delta double buffered: the start barrier of step s + 1 is also "everybody has read the deltas of step s")
        self.mem = SymmetricMemory(self.rt, {'delta': n * 4, 'avg': n * 4, 'pad': 256, 'd16_0': n * 2, 'd16_1': n * 2,
                                             'a16': n * 2, 'pad16': 256})
        self.comm_local = torch.zeros(16, dtype=torch.int32, device=self.rt.device)
        self.comm_local16 = torch.zeros(16, dtype=torch.int32, device=self.rt.device)
        self.delta = self.mem.local('delta', torch.float32, n)
        self.delta16 = [self.mem.local('d16_0', torch.bfloat16, n), self.mem.local('d16_1', torch.bfloat16, n)]

    # -- hooks of the captured step in per-batch mode (ops/torch_step.py: ``te.sync``) --------------------------------
    def delta_buffer(self, flip):
        return self.delta16[flip]

    def reduce_step(self, flip):
        te, rt = self.te, self.rt
        native.allreduce_update(self.mem.ptrs('d16_%d' % flip), self.mem.ptrs('pad16'), self.comm_local16, te.numel,
                                rt.rank, rt.world_size, 1.0 / rt.world_size, True, te.w32, w16=None,
                                avg_ptrs=self.mem.ptrs('a16'), algo=2 if rt.world_size > 2 else 1, device=rt.device)

    def close(self):
        if self.mem is not None:
            torch.cuda.synchronize(self.rt.device)
            self.te.sync = None
            self.te.invalidate_graphs()  # recorded per-batch steps address the buffers that go away now
            self.mem.close()
            self.mem = None

    def _buffers(self):
        return [w for w in self.model.weights if not w.requires_grad]

    def _reduce_round(self, before):
        """Apply the mean delta of all ranks to ``before`` (fused kernel) and average the non-trainable state."""
        te, rt = self.te, self.rt
        native.delta(before, te.w32, out_f32=self.delta)
        te.w32.copy_(before)
        native.allreduce_update(self.mem.ptrs('delta'), self.mem.ptrs('pad'), self.comm_local, te.numel, rt.rank,
                                rt.world_size, 1.0 / rt.world_size, False, te.w32, w16=None,
                                avg_ptrs=self.mem.ptrs('avg'), device=rt.device)
        bufs = self._buffers()
        if bufs:
            packed = self._pack(bufs)
            rt.all_reduce_sum_(packed)
            packed /= rt.world_size
            self._unpack(packed, bufs)

    @staticmethod
    def _pack(bufs):
        return torch.cat([b.reshape(-1).float() for b in bufs])

    @staticmethod
    def _unpack(packed, bufs):
        off = 0
        with torch.no_grad():
            for b in bufs:
                b.copy_(packed[off:off + b.numel()].view(b.shape))
                off += b.numel()

    def run(self, partition, train_config, sync_rounds):
        rt, te, model = self.rt, self.te, self.model
        cfg = dict(train_config)
        with tracing.span('sync_torch/partition_to_arrays'):
            x, y = partition_to_arrays(partition)
        bs = int(cfg.get('batch_size') or 32)
        epochs = int(cfg.get('epochs', 1))
        split = float(cfg.get('validation_split') or 0.0)
        trainable = len(x) > bs  # reference guard (worker.py:41)
        history = History()
        with torch.cuda.device(rt.device):
            distributed = rt.distributed
            # every replica starts from rank 0's weights with a fresh optimizer, like the reference's workers
            te.reset_optimizer()
            model.optimizer.reset_state()
            if distributed:
                with tracing.span('sync_torch/broadcast_start'):
                    rt.broadcast_(te.w32, src=0)
                    bufs = self._buffers()
                    if bufs:
                        packed = self._pack(bufs)
                        rt.broadcast_(packed, src=0)
                        self._unpack(packed, bufs)
                self._setup_comm()
            with tracing.span('sync_torch/h2d'):
                x16 = getattr(partition, 'x16', None)  # pinned bf16 copy staged once by to_sharded_rdd(feature_dtype=...)
                if (trainable and torch.is_tensor(x16) and x16.dtype == torch.bfloat16 and len(x16) == len(x)
                        and x16.shape[1] == int(np.prod(x.shape[1:])) and rt.device.type == 'cuda'
                        and getattr(model, '_compute_dtype', 'float32') == 'bfloat16'):
                    xt = x16.to(rt.device, non_blocking=True).view(x.shape)  # half the bytes, no pageable bounce
                else:
                    xt = torch.as_tensor(np.ascontiguousarray(x, dtype=np.float32)).to(rt.device) if trainable else None
                yt = torch.as_tensor(np.ascontiguousarray(y, dtype=np.float32)).to(rt.device) if trainable else None
            per_batch = distributed and self.sm.frequency == 'batch' and sync_rounds != 'fit'
            if per_batch:
                # first launch of the collective outside of any graph capture (zero deltas: weights unchanged)
                for d in self.delta16:
                    d.zero_()
                torch.cuda.synchronize(rt.device)
                rt.barrier()
                self.reduce_step(0)
                # graphs recorded by an earlier fit of this context stay valid: they are keyed by the flip value and the
                # buffers they address live as long as the context does
                te.sync, te.flip = _StepSync(self), 1
            before = torch.empty_like(te.w32) if (distributed and not per_batch) else None
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record()
            for epoch in range(epochs):
                if before is not None and (epoch == 0 or sync_rounds != 'fit'):
                    before.copy_(te.w32)
                if trainable:
                    with tracing.span('sync_torch/epoch'):
                        h = model.fit(xt, yt, batch_size=bs, epochs=epoch + 1, initial_epoch=epoch, verbose=0,
                                      validation_split=split, shuffle=bool(cfg.get('shuffle', True)))
                    for k, v in h.history.items():
                        history.history.setdefault(k, []).extend(v)
                if before is not None and (sync_rounds != 'fit' or epoch == epochs - 1):
                    with tracing.span('sync_torch/reduce'):
                        self._reduce_round(before)
                elif per_batch:
                    bufs = self._buffers()  # the weights are identical on all ranks already; the moving statistics are not
                    if bufs:
                        packed = self._pack(bufs)
                        rt.all_reduce_sum_(packed)
                        packed /= rt.world_size
                        self._unpack(packed, bufs)
            t1.record()
            if per_batch:
                te.sync = None  # (steps outside of this context run the plain optimizer again; other graph keys)
            torch.cuda.synchronize(rt.device)
            self.sm._last_sync_run = {'device_ms': t0.elapsed_time(t1), 'h2d_bytes_per_step': 0,
                                      'steps_per_epoch': -(-len(x) // bs) if trainable else 0}
            if self.mem is not None:
                status = int(self.comm_local[8].item()) or int(self.comm_local16[8].item())
                if status:
                    raise RuntimeError("rank %d: peer rank %d never reached a fused all-reduce (watchdog fired)"
                                       % (rt.rank, status - 1))
        hist = dict(history.history) if trainable else None
        with tracing.span('sync_torch/gather_histories'):
            gathered = rt.all_gather_object((rt.rank, hist))
        return [h for _, h in sorted(gathered, key=lambda t: t[0])]
