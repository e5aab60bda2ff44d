"""Synchronous mode on the native engine: local steps + ONE fused all-reduce/update kernel per round.

Reference semantics (``elephas/spark_model.py:217-231`` + ``elephas/worker.py:26-49``): every worker
trains a replica from the same weights, reports ``delta = before - after``, the driver subtracts the
mean delta from the master weights. Here one rank = one worker = one B200, and the replica *is* the
rank's copy of the master network (moved to the device once and trained in place - no broadcast, no
collect, no host copy of the weights):

* ``frequency='batch'``: the optimizer kernel runs in *delta mode* - it updates its slots and writes
  the step as a bf16 delta into peer-mapped memory instead of touching the weights - and the
  ``allreduce_update`` kernel (K6) reads all peers' deltas over NVLink, averages them, applies the
  result to the fp32 master weights and refreshes the bf16 compute copy, all in one launch;
* ``frequency='epoch'`` (or ``sync_rounds='fit'``): the replica trains an epoch (or all epochs)
  locally, ``eb_delta`` (K5) forms ``before - after`` in fp32 straight into peer-mapped memory, the
  weights are reset to ``before`` and the same K6 kernel applies the mean delta.

Input pipeline: the partition lives in pinned host memory; ``input_mode='device'`` (default when the
partition fits comfortably) uploads it once and gathers batches on the device, ``input_mode='stream'``
copies every batch host->device on a side stream, double buffered, overlapping the previous step.
``log_batch_loss=True`` additionally reads the running loss back every step (one 64-byte
device->host copy, one step delayed so it never stalls the GPU). ``keep_context=True`` keeps engine,
CUDA graphs and peer-mapped buffers alive on the ``SparkModel`` for the next ``fit`` call.
"""
import os
import time

import numpy as np
import torch

from ..models import History
from ..ops import engine as _engine
from ..ops import native
from ..utils import tracing
from ..worker import partition_to_arrays
from .symm import SymmetricMemory, VmmSymmetricMemory

IN_PLACE = object()  # returned instead of a weight list: the master network already holds the result


def _gather_histories(rt, hist, keys, width):
    """Every rank's history dict (or None), in rank order, on every rank. The values travel as ONE float64 row per rank
    (``[trainable, then per key: count, values...]``, one all-gather) instead of pickled objects: ``keys`` lists every
    name a history of this fit can hold, in the order the dict has them; no list is longer than ``width``."""
    row = torch.zeros(1 + len(keys) * (1 + width), dtype=torch.float64)
    if hist is not None:
        unknown = [k for k in hist if k not in keys]
        assert not unknown and all(len(v) <= width for v in hist.values()), (unknown, width)
        row[0] = 1.0
        for i, k in enumerate(keys):
            vals = hist.get(k)
            if vals is not None:
                at = 1 + i * (1 + width)
                row[at] = len(vals) + 1  # 0: key absent, n + 1: a list of n values
                if len(vals):
                    row[at + 1:at + 1 + len(vals)] = torch.as_tensor([float(v) for v in vals], dtype=torch.float64)
    out = []
    for r in rt.all_gather_rows(row).tolist():
        if r[0] == 0.0:
            out.append(None)
            continue
        h = {}
        for i, k in enumerate(keys):
            at = 1 + i * (1 + width)
            if r[at] > 0:
                h[k] = r[at + 1:at + int(r[at])]
        out.append(h)
    return out


def try_fit(spark_model, rt, partitions, mine, train_config, sync_rounds):
    if rt.device.type != 'cuda' or not _engine.native_enabled():
        return None
    if len(partitions) != rt.world_size or len(mine) != 1:
        return None
    if any(train_config.get(k) is not None and train_config.get(k) != [] for k in ('callbacks', 'class_weight', 'sample_weight')):
        return None  # callbacks and row weights belong to Model.fit: the generic worker path calls it with them
    master = spark_model.master_network
    if master.device != rt.device:
        master.to(rt.device)
    with tracing.span('sync/engine'):
        eng = master._native_engine()
    with tracing.span('sync/agree'):
        if not rt.all_true(eng is not None):
            return None
    ctx = getattr(spark_model, '_sync_context', None)
    if ctx is None or ctx.eng is not eng:
        if ctx is not None:
            ctx.close()
        ctx = SyncContext(spark_model, rt, master, eng)
        spark_model._sync_context = ctx
    histories = ctx.run(mine[0][1], train_config, sync_rounds)
    if not train_config.get('keep_context'):
        ctx.close()
        spark_model._sync_context = None
    return IN_PLACE, histories


_HOST_BF16_CHOICE = {}  # (rows, cols, threads) -> bool: measured once per process


def _host_threads():
    """CPU threads for the host-side input cast: ``ELEPHAS_B200_HOST_THREADS``, else this rank's share of the cores
    (minus one for the Python threads that drive the GPU), between 1 and 16."""
    forced = os.environ.get('ELEPHAS_B200_HOST_THREADS')
    if forced:
        return max(1, int(forced))
    try:
        cores = len(os.sched_getaffinity(0))
    except (AttributeError, OSError):
        cores = os.cpu_count() or 1
    ranks_here = max(1, int(os.environ.get('LOCAL_WORLD_SIZE', '1')))
    return max(1, min(16, cores // ranks_here - 1))


class _InputFeeder:
    """Delivers batches into the engine's static input buffers.

    ``device``: the partition is uploaded once, batches are gathered on the device.
    ``stream``: every batch crosses PCIe while the previous step computes. The features are cast to bf16 on the host
    first (``eb_host_f32_to_bf16``, CPU thread pool, bit-identical to the device cast the engine would do) - half the
    bytes per step - by a helper thread that runs two batches ahead: convert into a pinned slot, enqueue the H2D copy
    on the copy stream, record an event the compute stream waits on.
    """

    SLOTS = 3  # device staging slots of the streamed modes (the host-cast mode keeps two, its helper thread runs ahead)

    def __init__(self, eng, x, y, batch_size, mode, shuffle, x16=None):
        self.eng, self.bs, self.shuffle = eng, batch_size, shuffle
        # features already staged as bf16 in pinned memory by the caller (to_sharded_rdd(feature_dtype='bfloat16')): they
        # are streamed as they are - half the PCIe bytes of the fp32 stream, no per-step CPU work
        self.staged16 = x16 if (x16 is not None and x16.dtype == torch.bfloat16 and len(x16) == len(x)
                                and x16.shape[1] == eng.ops[0].ld_in) else None
        dev = eng.device
        n = len(x)
        self.n = n
        self.x_host = self._pinned(x, n)
        self.y_host = self._pinned(y, n)
        nbytes = self.x_host.numel() * 4 + self.y_host.numel() * 4
        if mode is None:
            mode = 'device' if nbytes < (8 << 30) else 'stream'
        if mode not in ('device', 'stream'):
            raise ValueError("input_mode must be 'device' or 'stream', got %r" % (mode,))
        self.mode = mode
        self.h2d_bytes_per_step = 0
        self.host_bf16 = False
        self.pool = None
        if mode == 'device':
            self.xd = self.x_host.to(dev, non_blocking=True)
            self.yd = self.y_host.to(dev, non_blocking=True)
        else:
            self.copy_stream = torch.cuda.Stream(dev)
            cols, ycols = self.x_host.shape[1], self.y_host.shape[1]
            self.threads = _host_threads()
            self.host_bf16 = self.staged16 is None and self._choose_host_bf16(eng, batch_size)
            if self.staged16 is not None:
                if not self.staged16.is_pinned():
                    self.staged16 = self.staged16.pin_memory()
                ld = eng.ops[0].ld_in
                self.stage = [(torch.empty(batch_size, ld, dtype=torch.bfloat16, device=dev),
                               torch.empty(batch_size, ycols, dtype=torch.float32, device=dev),
                               torch.cuda.Event()) for _ in range(self.SLOTS)]
                self.h2d_bytes_per_step = batch_size * (ld * 2 + ycols * 4)
            elif self.host_bf16:
                ld = eng.ops[0].ld_in
                self.pinned = [torch.empty(batch_size, ld, dtype=torch.bfloat16).pin_memory() for _ in range(3)]
                self.pinned_free = [None, None, None]  # event: the slot's last H2D copy has finished
                self.stage = [(torch.empty(batch_size, ld, dtype=torch.bfloat16, device=dev),
                               torch.empty(batch_size, ycols, dtype=torch.float32, device=dev),
                               torch.cuda.Event()) for _ in range(2)]
                self.h2d_bytes_per_step = batch_size * (ld * 2 + ycols * 4)
                from concurrent.futures import ThreadPoolExecutor
                self.pool = ThreadPoolExecutor(max_workers=1, thread_name_prefix='elephas-feeder')
                self.jobs = {}
            else:
                self.stage = [(torch.empty(batch_size, cols, dtype=torch.float32, device=dev),
                               torch.empty(batch_size, ycols, dtype=torch.float32, device=dev),
                               torch.cuda.Event()) for _ in range(self.SLOTS)]
                self.h2d_bytes_per_step = batch_size * (cols + ycols) * 4
            self.consumed = [None] * self.SLOTS
        self.order = None

    def _choose_host_bf16(self, eng, batch_size):
        """Cast the features to bf16 on the host (half the PCIe bytes) or stream them as fp32 and cast on the device?
        ``ELEPHAS_B200_HOST_BF16=1|0`` decides (default 0); with ``auto`` one batch is timed both ways, once per process and
        input width: with pipelining a step pays max(host cast, bf16 copy) against the fp32 copy. Few cores per GPU favour
        the fp32 copy, many GPUs sharing the host's PCIe switches favour the cast."""
        # default off: on the hosts measured so far (16 vCPUs for one GPU, 32 for eight) the pipelined cast ends up slower
        # than the fp32 copy even where the one-batch calibration says otherwise (2.74 vs 2.65 ms/step end to end)
        forced = os.environ.get('ELEPHAS_B200_HOST_BF16', '0')
        if forced != 'auto':
            return forced == '1'
        rows = min(batch_size, self.n)
        cols = self.x_host.shape[1]
        key = (rows, cols, self.threads)
        if key not in _HOST_BF16_CHOICE:
            dev = eng.device
            dst16 = torch.empty(rows, eng.ops[0].ld_in, dtype=torch.bfloat16).pin_memory()
            d32 = torch.empty(rows, cols, dtype=torch.float32, device=dev)
            src = self.x_host[:rows]
            native.host_f32_to_bf16(src, dst16, rows, cols, threads=self.threads)  # starts the pool, touches the pages
            t0 = time.perf_counter()
            native.host_f32_to_bf16(src, dst16, rows, cols, threads=self.threads)
            t_cast = (time.perf_counter() - t0) * 1e3
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(self.copy_stream):
                d32.copy_(src, non_blocking=True)
                e0.record(self.copy_stream)
                d32.copy_(src, non_blocking=True)
                e1.record(self.copy_stream)
            e1.synchronize()
            t_copy32 = e0.elapsed_time(e1)
            _HOST_BF16_CHOICE[key] = max(t_cast, 0.5 * t_copy32) < 0.9 * t_copy32
            tracing.add_host('sync/input_calibration cast %.2f ms copy32 %.2f ms' % (t_cast, t_copy32), 0.0)
        return _HOST_BF16_CHOICE[key]

    @staticmethod
    def _pinned(a, n):
        """Host tensor view of ``a`` in page-locked memory; arrays that already live in pinned memory
        (e.g. NumPy views of ``torch.empty(..., pin_memory=True)``) are used in place."""
        t = torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32).reshape(n, -1))
        return t if t.is_pinned() else t.pin_memory()

    def close(self):
        if self.pool is not None:
            self.pool.shutdown(wait=True)
            self.pool = None

    def start_epoch(self):
        dev = self.eng.device
        if self.shuffle:
            order = torch.randperm(self.n)  # CPU generator (safe while another thread captures a CUDA graph)
            self.order = order.to(dev, non_blocking=True) if self.mode == 'device' else order
        else:
            self.order = None
        if self.mode == 'stream':
            if self.host_bf16:
                for fut in self.jobs.values():
                    fut.result()
                self.jobs = {}
                self._submit(0, 0)
                self._submit(1, self.bs)
            else:
                for ahead in range(self.SLOTS - 1):  # copies run SLOTS - 1 batches ahead of the step that consumes them
                    self._prefetch(ahead, ahead * self.bs)

    def _host_batch(self, start, rows):
        xsrc = self.staged16 if self.staged16 is not None else self.x_host
        if self.order is None:
            return xsrc[start:start + rows], self.y_host[start:start + rows]
        idx = self.order[start:start + rows]
        return xsrc.index_select(0, idx).pin_memory(), self.y_host.index_select(0, idx).pin_memory()

    # -- fp32 streaming (ELEPHAS_B200_HOST_BF16=0) ---------------------------------------------------------
    def _prefetch(self, step, start):
        rows = min(self.bs, self.n - start)
        if rows <= 0:
            return
        slot = step % self.SLOTS
        xs, ys, ev = self.stage[slot]
        xh, yh = self._host_batch(start, rows)
        if self.consumed[slot] is not None:
            # the slot may only be overwritten once the compute stream has drained its previous contents
            self.copy_stream.wait_event(self.consumed[slot])
        with torch.cuda.stream(self.copy_stream):
            xs[:rows].copy_(xh, non_blocking=True)
            ys[:rows].copy_(yh, non_blocking=True)
            ev.record(self.copy_stream)

    # -- bf16 streaming: helper thread converts and uploads two batches ahead ----------------------------------
    def _submit(self, step, start):
        rows = min(self.bs, self.n - start)
        if rows <= 0:
            return
        consumed = self.consumed[step % 2]
        self.jobs[step] = self.pool.submit(self._job, step, start, rows, consumed)

    def _job(self, step, start, rows, consumed):
        dev = self.eng.device
        with torch.cuda.device(dev):
            pslot, dslot = step % 3, step % 2
            if self.order is None:
                xh, yh = self.x_host[start:start + rows], self.y_host[start:start + rows]
            else:
                idx = self.order[start:start + rows]
                xh, yh = self.x_host.index_select(0, idx), self.y_host.index_select(0, idx).pin_memory()
            free = self.pinned_free[pslot]
            if free is not None:
                free.synchronize()  # the copy that last read this pinned slot is done
            native.host_f32_to_bf16(xh, self.pinned[pslot], rows, xh.shape[1], threads=self.threads)
            xs, ys, ev = self.stage[dslot]
            if consumed is not None:
                self.copy_stream.wait_event(consumed)  # the compute stream has drained this device slot
            with torch.cuda.stream(self.copy_stream):
                xs[:rows].copy_(self.pinned[pslot][:rows], non_blocking=True)
                ys[:rows].copy_(yh, non_blocking=True)
                done = torch.cuda.Event()
                done.record(self.copy_stream)
                ev.record(self.copy_stream)
            self.pinned_free[pslot] = done
        return rows

    def load(self, step, start, buf):
        """Fill the engine's input buffers with batch ``step``; returns the number of rows."""
        rows = min(self.bs, self.n - start)
        if self.mode == 'device':
            if self.order is not None:
                idx = self.order[start:start + rows]
                torch.index_select(self.xd, 0, idx, out=buf.x32)
                torch.index_select(self.yd, 0, idx, out=buf.y32)
            else:
                buf.x32.copy_(self.xd[start:start + rows])
                buf.y32.copy_(self.yd[start:start + rows])
            return rows
        slot = step % (2 if self.host_bf16 else self.SLOTS)
        xs, ys, ev = self.stage[slot]
        cur = torch.cuda.current_stream(self.eng.device)
        if self.host_bf16:
            self.jobs.pop(step).result()  # converted and enqueued (normally long done: it ran during step - 2)
            cur.wait_event(ev)
            buf.x16.copy_(xs[:rows])
        elif self.staged16 is not None:
            cur.wait_event(ev)
            buf.x16.copy_(xs[:rows])
        else:
            cur.wait_event(ev)
            buf.x32.copy_(xs[:rows])
        buf.y32.copy_(ys[:rows])
        done = torch.cuda.Event()
        done.record(cur)
        self.consumed[slot] = done
        if self.host_bf16:
            self._submit(step + 2, start + rows + self.bs)
        else:
            # the batch SLOTS - 1 steps ahead goes into the slot that was consumed longest ago: two copies are in flight
            # behind the running step, so one slow copy (eight GPUs share the host's PCIe switches) does not stall it
            self._prefetch(step + self.SLOTS - 1, start + rows + (self.SLOTS - 2) * self.bs)
        return rows


class SyncContext:
    """Everything a rank keeps between synchronous rounds: engine, peer-mapped buffers, flags."""

    def __init__(self, spark_model, rt, model, eng):
        self.sm, self.rt, self.model, self.eng = spark_model, rt, model, eng
        self.mem = None
        self.dtype_bf16 = None
        self.bs = None
        self.n_sub = rt.world_size
        self.sharded = False
        self.w32_partial = False
        self._private_w16 = None
        self.nvls = False

    # ---------------------------------------------------------------- comm setup
    def _setup_comm(self, dtype_bf16):
        if self.mem is not None and self.dtype_bf16 == dtype_bf16:
            return
        eng = self.eng
        if self.mem is not None:
            self._release_mem()
        n = eng.layout.numel
        item = 2 if dtype_bf16 else 4
        spec = {'delta0': n * item, 'delta1': n * item, 'pad': 256}
        # per-batch mode shards the fp32 master across ranks (each rank updates 1/N of it and stores the refreshed
        # bf16 weights into every peer's compute copy, which therefore lives in peer-mapped memory); the epoch/fit
        # modes keep whole replicas and need the averaged-delta staging buffer of the two-shot kernel instead
        self.sharded = bool(dtype_bf16) and os.environ.get('ELEPHAS_B200_SHARDED_MASTER', '1') != '0'
        spec['w16' if self.sharded else 'avg'] = n * item
        # NVLS (ELEPHAS_B200_NVLS=auto|1|0, default auto = use it where the node supports multicast objects): the buffers
        # come from the VMM allocator and are bound into an NVSwitch multicast object, the sharded kernel then lets the
        # switch add the deltas and replicate the refreshed weights (8 GPUs: 2.56 vs 2.64 ms/step, 2 GPUs: 2.64 vs 2.87)
        self.nvls = False
        nvls_env = os.environ.get('ELEPHAS_B200_NVLS', 'auto')
        want_nvls = self.sharded and nvls_env != '0' and self.rt.distributed
        if want_nvls:
            try:
                # 'stage32': room for the fp32 master, the target of the multimem.st weight broadcast / master gather
                mem = VmmSymmetricMemory(self.rt, dict(spec, stage32=n * 4), multicast=True)
                err = None
            except native.NativeError as exc:
                mem, err = None, str(exc)
            ok = self.rt.all_gather_object(mem is not None and mem.multicast)
            if all(ok):
                self.mem, self.nvls = mem, True
            else:
                if mem is not None:
                    mem.close()
                if self.rt.is_driver and nvls_env == '1':
                    print(">>> NVLS not available on this node (%s): peer-to-peer kernels" % (err or 'no multicast support',))
        if not self.nvls:
            self.mem = SymmetricMemory(self.rt, spec)
        self.comm_local = torch.zeros(16, dtype=torch.int32, device=self.rt.device)
        self.dtype_bf16 = dtype_bf16
        tdt = torch.bfloat16 if dtype_bf16 else torch.float32
        self.delta = [self.mem.local('delta0', tdt, n), self.mem.local('delta1', tdt, n)]
        self.flip = 0
        self.w32_partial = False
        if self.sharded:
            self._private_w16 = eng.w16
            shared = self.mem.local('w16', torch.bfloat16, n)
            shared.copy_(eng.w16)
            eng.w16 = shared
            eng.invalidate_graphs()  # recorded graphs point at the private copy
        # first launch of the collective kernels outside of any graph capture (zero delta: weights unchanged)
        self.delta[0][:1024].zero_()
        torch.cuda.synchronize(self.rt.device)
        self.rt.barrier()
        common = (self.mem.ptrs('delta0'), self.mem.ptrs('pad'), self.comm_local, 1024, self.rt.rank,
                  self.rt.world_size, 1.0 / self.n_sub, dtype_bf16, eng.w32)
        if self.nvls:
            native.allreduce_nvls_update(self.mem.mc_ptr('delta0'), self.mem.mc_ptr('w16'), self.mem.ptrs('pad'),
                                         self.comm_local, 1024, self.rt.rank, self.rt.world_size, 1.0 / self.n_sub, eng.w32,
                                         blocks=2, device=self.rt.device)
        elif self.sharded:
            native.allreduce_shard_update(*common, self.mem.ptrs('w16'), blocks=2, device=self.rt.device)
        else:
            for algo in (1, 2):
                native.allreduce_update(*common, w16=eng.w16, avg_ptrs=self.mem.ptrs('avg'), algo=algo, blocks=2,
                                        device=self.rt.device)
        torch.cuda.synchronize(self.rt.device)

    def _release_mem(self):
        eng = self.eng
        self._gather_master()
        if self.sharded and eng.w16 is not self._private_w16:
            self._private_w16.copy_(eng.w16)
            eng.w16 = self._private_w16
        eng.invalidate_graphs()  # recorded graphs hold pointers into the buffers that go away now
        torch.cuda.synchronize(self.rt.device)
        self.mem.close()
        self.mem = None

    def _owned_ranges(self):
        """Flat element ranges of the fp32 master this rank owns in sharded mode (slice ``rank`` of every bucket,
        the same split as ``allreduce_shard_update_kernel``)."""
        rank, world = self.rt.rank, self.rt.world_size
        out = []
        for lo, n in self.eng.bucket_ranges():
            n8 = n // 8
            per = -(-n8 // world)
            a, b = per * rank, min(n8, per * (rank + 1))
            if b > a:
                out.append((lo + a * 8, lo + b * 8))
        return out

    def _ranges_cover_master(self):
        """True when the buckets tile the whole flat buffer (then the owned ranges of all ranks do, too): the staging
        buffer of the NVLS gather holds no element nobody wrote."""
        at = 0
        for lo, n in sorted(self.eng.bucket_ranges()):
            if lo != at or n % 8:
                return False
            at = lo + n
        return at == self.eng.w32.numel()

    def _gather_master(self):
        """Sharded mode leaves every rank with an up-to-date fp32 master only for the slices it owns: rebuild
        the full replica (once per fit, not per step) so that ``get_weights`` / ``save`` see exact values."""
        if not self.w32_partial:
            return
        eng = self.eng
        if self.nvls and self.mem is not None and self._ranges_cover_master():
            self._nvls_replicate(self._owned_ranges())
        else:
            full = torch.zeros_like(eng.w32)
            for a, b in self._owned_ranges():
                full[a:b].copy_(eng.w32[a:b])
            self.rt.all_reduce_sum_(full)
            eng.w32.copy_(full)
        self.w32_partial = False
        eng._weights_dirty_in_model = True

    def _allreduce_update(self, lo=0, n=None, stream=None, blocks=0):
        """Fused all-reduce + update of flat elements [lo, lo+n) of the current delta buffer."""
        eng = self.eng
        n = eng.layout.numel - lo if n is None else n
        name = 'delta%d' % self.flip
        item = 2 if self.dtype_bf16 else 4
        common = (self.mem.ptrs(name, lo * item), self.mem.ptrs('pad'), self.comm_local, n, self.rt.rank,
                  self.rt.world_size, 1.0 / self.n_sub, self.dtype_bf16, eng.w32[lo:])
        if self.nvls:
            native.allreduce_nvls_update(self.mem.mc_ptr(name, lo * 2), self.mem.mc_ptr('w16', lo * 2), self.mem.ptrs('pad'),
                                         self.comm_local, n, self.rt.rank, self.rt.world_size, 1.0 / self.n_sub,
                                         eng.w32[lo:], blocks=blocks, device=self.rt.device)
        elif self.sharded:
            native.allreduce_shard_update(*common, self.mem.ptrs('w16', lo * 2), blocks=blocks,
                                          device=self.rt.device)
        else:
            native.allreduce_update(*common, w16=eng.w16[lo:], avg_ptrs=self.mem.ptrs('avg', lo * item),
                                    blocks=blocks, device=self.rt.device)
        eng._weights_dirty_in_model = True

    # -- per-bucket communication overlapped with the rest of the backward pass -------------------------
    def _install_hooks(self):
        eng, dev = self.eng, self.rt.device
        self.comm_stream = torch.cuda.Stream(dev)
        # the sharded kernel keeps 2-4 x world 16-byte peer loads in flight per thread: a small grid already
        # saturates NVLink and leaves the SMs to the GEMMs it overlaps with (8 GPUs: 16 blocks 2.68 ms/step,
        # 74 blocks 3.12); the replicated kernel streams the whole fp32 master and wants half the machine
        blocks = int(os.environ.get('ELEPHAS_B200_COMM_BLOCKS', '16' if self.sharded else '74'))

        skip = os.environ.get('ELEPHAS_B200_DEBUG_SKIP_COMM') == '1'  # timing experiments only: wrong results

        # the first layer's bucket and the vector bucket are reduced after the last GEMM of the step: nothing is left to
        # overlap with, so they take the whole machine instead of the small overlap-friendly grid
        tail_from = eng.ops[0].w_seg.offset if self.sharded else None
        tail_blocks = int(os.environ.get('ELEPHAS_B200_COMM_TAIL_BLOCKS', '148'))

        def hook(lo, n):
            if skip:
                return
            cur = torch.cuda.current_stream(dev)
            ready = torch.cuda.Event()
            ready.record(cur)
            self.comm_stream.wait_event(ready)
            tail = tail_from is not None and (lo == eng.ops[0].w_seg.offset or lo >= eng.layout.matrix_end)
            with torch.cuda.stream(self.comm_stream):
                self._allreduce_update(lo, n, blocks=tail_blocks if tail else blocks)

        def join():
            if skip:
                return
            done = torch.cuda.Event()
            done.record(self.comm_stream)
            torch.cuda.current_stream(dev).wait_event(done)

        eng.comm_hook, eng.comm_join = hook, join

    def _remove_hooks(self):
        self.eng.comm_hook = self.eng.comm_join = None

    def _check_watchdog(self):
        """The collective kernels give up on a peer after 20 s and leave a status word instead of hanging."""
        if self.mem is not None:
            status = int(self.comm_local[8].item())
            if status:
                raise RuntimeError("rank %d: peer rank %d never reached a fused all-reduce (watchdog fired); "
                                   "the replicas are no longer consistent" % (self.rt.rank, status - 1))

    def close(self):
        self.eng.delta_target = None
        self._remove_hooks()
        if self.mem is not None:
            self._release_mem()
        else:
            self.eng.invalidate_graphs()

    # -------------------------------------------------------------- one sync step
    def _step(self, buf, per_batch, have_batch):
        eng = self.eng
        if per_batch:
            eng.delta_target = self.delta[self.flip]
            self.w32_partial = self.sharded  # (set here, not in the hook: the hook only runs while capturing)
            if have_batch:
                eng.train_step_device(buf)  # the per-bucket all-reduces are part of the (graph-captured) step
            else:
                # no batch left on this rank: contribute zero deltas through the same sequence of collectives
                self.delta[self.flip].zero_()
                for lo, n in eng.bucket_ranges():
                    eng.comm_hook(lo, n)
                eng.comm_join()
            self.flip ^= 1
            eng._weights_dirty_in_model = True
        elif have_batch:
            eng.delta_target = None
            eng.train_step_device(buf)

    def _sync_start(self):
        """All replicas start a fit from rank 0's weights (the reference broadcasts the driver's)."""
        eng, rt = self.eng, self.rt
        eng.reset_optimizer()
        if rt.distributed:
            if self.nvls and self.mem is not None:
                # rank 0 writes its fp32 master once to the multicast address of the staging buffer; the switch replicates
                # it into every rank's memory (elephas/spark_model.py:206-207: Spark broadcast of the driver's weights)
                self._nvls_replicate([(0, eng.w32.numel())] if rt.rank == 0 else [])
            else:
                rt.broadcast_(eng.w32, src=0)
            eng.refresh_bf16()
            eng._weights_dirty_in_model = True

    def _nvls_replicate(self, ranges):
        """Every rank stores the flat element ``ranges`` of its fp32 master through the multicast address of 'stage32'
        (``multimem.st``: each payload byte leaves the sender once), then all ranks copy the staging buffer - now the
        union of everybody's ranges - over their master. The ranges of different ranks must not overlap and must cover
        the buffer. No barrier is needed BEFORE the stores: the previous reader of 'stage32' on every rank finished
        before a collective every rank has since completed (the agreement all-reduce of ``try_fit`` / the flag barriers
        of the training steps)."""
        eng, rt = self.eng, self.rt
        for a, b in ranges:
            native.nvls_store(eng.w32[a:b], self.mem.mc_ptr('stage32', 4 * a), device=rt.device)
        torch.cuda.synchronize(rt.device)
        rt.barrier()  # every rank's stores have left its GPU and landed
        eng.w32.copy_(self.mem.local('stage32', torch.float32, eng.w32.numel()))

    # ----------------------------------------------------------------------- fit
    def run(self, partition, train_config, sync_rounds):
        rt, eng = self.rt, self.eng
        cfg = dict(train_config)
        with tracing.span('sync/partition_to_arrays'):
            x, y = partition_to_arrays(partition)
        x16 = getattr(partition, 'x16', None)
        bs = self.bs = int(cfg.get('batch_size') or 32)
        epochs = int(cfg.get('epochs', 1))
        split = float(cfg.get('validation_split') or 0.0)
        x_val = y_val = None
        n_all = len(x)
        if 0.0 < split < 1.0:
            at = int(len(x) * (1.0 - split))
            x, x_val, y, y_val = x[:at], x[at:], y[:at], y[at:]
            x16 = x16[:at] if x16 is not None else None
        trainable = n_all > bs and len(x) > 0  # reference guard (worker.py:41)
        log_batch_loss = bool(cfg.get('log_batch_loss', False))
        batch_losses = []
        history = History()
        frequency = self.sm.frequency
        with torch.cuda.device(rt.device):
            distributed = rt.distributed
            per_batch = distributed and frequency == 'batch' and sync_rounds != 'fit'
            if distributed:
                with tracing.span('sync/setup_comm'):
                    self._setup_comm(dtype_bf16=per_batch)
            with tracing.span('sync/broadcast_start'):
                self._sync_start()
            if per_batch:
                self._install_hooks()
            else:
                self._remove_hooks()
            steps_local = -(-len(x) // bs) if trainable else 0
            steps = steps_local
            if distributed:
                steps = int(rt.all_reduce_max_(torch.tensor([steps_local], device=rt.device))[0])
            with tracing.span('sync/input_feeder'):
                feeder = _InputFeeder(eng, x, y, bs, cfg.get('input_mode'), bool(cfg.get('shuffle', True)), x16=x16) \
                    if trainable else None
            eng.input_bf16 = bool(feeder is not None and feeder.mode == 'stream'
                                  and (feeder.host_bf16 or feeder.staged16 is not None))
            names = self.model.metrics_names
            pinned = [torch.zeros(8, dtype=torch.float64).pin_memory() for _ in range(2)]
            events = [torch.cuda.Event() for _ in range(2)]
            before = torch.empty_like(eng.w32) if (distributed and not per_batch) else None
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record()
            try:
                for epoch in range(epochs):
                    if before is not None and (epoch == 0 or sync_rounds != 'fit'):
                        before.copy_(eng.w32)
                    eng.stats.zero_()
                    if feeder:
                        feeder.start_epoch()
                    start = 0
                    t_loop = time.perf_counter()
                    t_step = t_loop
                    for s in range(steps):
                        if tracing.enabled():  # host time between two step submissions: jitter shows up in 'max ms'
                            now = time.perf_counter()
                            if s:
                                tracing.add_host('sync/step_host', now - t_step)
                            t_step = now
                        have = s < steps_local
                        buf = None
                        if have:
                            rows = min(bs, len(x) - start)
                            buf = eng._buf(rows)
                            feeder.load(s, start, buf)
                            start += rows
                        self._step(buf, per_batch, have)
                        if log_batch_loss:
                            slot = s % 2
                            pinned[slot].copy_(eng.stats, non_blocking=True)
                            events[slot].record()
                            if s > 0:
                                events[1 - slot].synchronize()
                                p = pinned[1 - slot]
                                batch_losses.append(float(p[0] / max(float(p[7]), 1.0)))
                    if log_batch_loss and steps > 0:
                        slot = (steps - 1) % 2
                        events[slot].synchronize()
                        batch_losses.append(float(pinned[slot][0] / max(float(pinned[slot][7]), 1.0)))
                    tracing.add_host('sync/step_loop', time.perf_counter() - t_loop)
                    with tracing.span('sync/read_stats'):
                        vals, count = eng.read_stats()
                    logs = dict(zip(names, vals)) if count else {}
                    if before is not None and (sync_rounds != 'fit' or epoch == epochs - 1):
                        # delta = before - after (fp32, straight into peer-mapped memory), reset, fused reduce+update
                        native.delta(before, eng.w32, out_f32=self.delta[self.flip])
                        eng.w32.copy_(before)
                        self._allreduce_update()
                        self.flip ^= 1
                    if x_val is not None and len(x_val) and logs:
                        val = eng.evaluate(x_val, y_val)
                        logs.update({'val_' + k: v for k, v in zip(names, val)})
                    if logs:
                        history._append(epoch, logs)
            except BaseException:
                # leave the engine usable (and the feeder thread stopped) when a step fails
                eng.delta_target = None
                eng.input_bf16 = False
                if feeder is not None:
                    feeder.close()
                raise
            with tracing.device_span('sync/gather_master'):
                self._gather_master()
            t1.record()
            with tracing.span('sync/drain'):
                torch.cuda.synchronize(rt.device)
            self._check_watchdog()
            eng.delta_target = None
            eng.input_bf16 = False
            if feeder is not None:
                feeder.close()
            eng._weights_dirty_in_model = True
            self.model.optimizer.iterations = int(eng.step.item())
        hist = dict(history.history) if trainable else None
        if hist is not None and log_batch_loss:
            hist['batch_loss'] = batch_losses
        self.sm._last_sync_run = {'h2d_bytes_per_step': feeder.h2d_bytes_per_step if feeder else 0,
                                  'steps_per_epoch': steps, 'device_ms': t0.elapsed_time(t1)}
        with tracing.span('sync/gather_histories'):
            keys = list(names) + ['val_' + k for k in names] + ['batch_loss']
            return _gather_histories(rt, hist, keys, max(epochs, epochs * steps if log_batch_loss else 0, 1))

    # ------------------------------------------------------------- kernel timing
    def timed_steps(self, x, y, steps, input_mode='device', on_start=None, on_stop=None):
        """Device time (ms, CUDA events) of exactly ``steps`` synchronous ``frequency='batch'`` steps on data
        that is already resident on the device: the kernel-path number of bench.py."""
        rt, eng, bs = self.rt, self.eng, self.bs
        with torch.cuda.device(rt.device):
            per_batch = rt.distributed
            if per_batch:
                self._setup_comm(dtype_bf16=True)
                self._install_hooks()
            feeder = _InputFeeder(eng, x, y, bs, input_mode, False)
            eng.input_bf16 = feeder.host_bf16
            feeder.start_epoch()
            torch.cuda.synchronize(rt.device)
            rt.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if on_start is not None:
                on_start()  # (e.g. a clock sampler that must cover exactly the timed region)
            e0.record()
            start = 0
            for s in range(steps):
                if start >= len(x):
                    start = 0  # more steps than batches in the data: cycle
                rows = min(bs, len(x) - start)
                buf = eng._buf(rows)
                feeder.load(s, start, buf)
                start += rows
                self._step(buf, per_batch, True)
            e1.record()
            torch.cuda.synchronize(rt.device)
            if on_stop is not None:
                on_stop()
            rt.barrier()
            eng.delta_target = None
            eng.input_bf16 = False
            feeder.close()
            self._gather_master()
            return e0.elapsed_time(e1)
