"""Workers: the code that trains one data partition.

Reference: ``elephas/worker.py:11-131`` (``SparkWorker`` for synchronous mode,
``AsynchronousSparkWorker`` for asynchronous / hogwild). Constructor signatures and the
``train(data_iterator)`` generator protocol are unchanged - ``train`` yields ``[deltas, history]``
(sync) or ``[]`` (async) - so they still work as ``rdd.mapPartitions(worker.train)``.

What changed underneath: a worker is bound to a *device* (one B200 per rank, or the CPU), the
partition can arrive as dense arrays (:class:`~elephas_b200.spark.context.PairBlock`) instead of an
iterator of Python rows, and the synchronous worker additionally exposes a round-based API
(``begin`` / ``train_round`` / ``train_batch``) so ``SparkModel`` can average every epoch or every
batch while the worker keeps its model and optimizer state between rounds.

Reference quirks kept on purpose: partitions with ``rows <= batch_size`` train nothing and yield a
zero delta (``worker.py:41,107,116``, D2); batch-frequency async training never shuffles and
ignores ``validation_split`` (``worker.py:96,114-127``, D4).
"""
import copy

import numpy as np

from .models import model_from_json
from .models import optimizers as _optimizers
from .parameter.client import BaseParameterClient
from .utils.functional_utils import subtract_params


def partition_to_arrays(data_iterator):
    """Materialise a partition as ``(x, y)`` arrays. Accepts a PairBlock, any object with ``x``/``y``
    arrays, or an iterator / list of ``(features, label)`` rows (``worker.py:36-38``)."""
    if hasattr(data_iterator, 'x') and hasattr(data_iterator, 'y'):
        return np.asarray(data_iterator.x), np.asarray(data_iterator.y)
    rows = list(data_iterator)
    if not rows:
        return np.zeros((0,), dtype=np.float32), np.zeros((0,), dtype=np.float32)
    x = np.asarray([np.asarray(r[0]) for r in rows])
    y = np.asarray([np.asarray(r[1]) for r in rows])
    return x, y


def fresh_optimizer(master_optimizer):
    """Every worker gets its own optimizer instance with empty slots (``worker.py:29,88``)."""
    if isinstance(master_optimizer, _optimizers.Optimizer):
        return _optimizers.deserialize(_optimizers.serialize(master_optimizer))
    return _optimizers.get(copy.deepcopy(master_optimizer))


def _weights_of(parameters):
    return parameters.value if hasattr(parameters, 'value') else parameters


# counters for tests / diagnostics: which device paths actually ran in this process
STATS = {'fused_pull_steps': 0, 'device_ps_steps': 0, 'host_store_rounds': 0}

class _WorkerBase:
    def __init__(self, json, parameters, train_config, master_optimizer, master_loss, master_metrics,
                 custom_objects, device=None):
        self.json = json
        self.parameters = parameters
        self.train_config = dict(train_config)
        if self.train_config.get('callbacks'):
            # every worker counts for itself (patience, best value, the model it is attached to), as the pickled copies
            # on Spark executors do under the reference
            self.train_config['callbacks'] = copy.deepcopy(list(self.train_config['callbacks']))
        self.master_optimizer = master_optimizer
        self.master_loss = master_loss
        self.master_metrics = master_metrics
        self.custom_objects = custom_objects
        self.device = device
        self.model = None

    def _build_model(self):
        from .parallel.runtime import default_device
        model = model_from_json(self.json, self.custom_objects)
        model.custom_objects = dict(self.custom_objects or {})
        model.compile(optimizer=fresh_optimizer(self.master_optimizer), loss=self.master_loss,
                      metrics=self.master_metrics)
        model.set_weights(_weights_of(self.parameters))
        model.to(self.device if self.device is not None else default_device())
        dtype = self.train_config.pop('compute_dtype', None)  # extension: bf16 autocast on the generic path
        if dtype:
            model.set_compute_dtype(dtype)
        self.model = model
        return model


class SparkWorker(_WorkerBase):
    """Synchronous worker: trains a replica on its partition and reports ``before - after``."""

    def __init__(self, json, parameters, train_config, master_optimizer, master_loss, master_metrics,
                 custom_objects, device=None):
        super().__init__(json, parameters, train_config, master_optimizer, master_loss, master_metrics,
                         custom_objects, device)
        self.x_train = self.y_train = None
        self.history = {}
        self._epoch_acc = {}

    # ---------------------------------------------------- reference protocol
    def train(self, data_iterator):
        """Train all epochs on the partition, then yield ``[deltas, history_dict_or_None]``."""
        self.begin(data_iterator)
        deltas, history = self.train_round(None, epochs=self.train_config.get('epochs', 1))
        yield [deltas, history]

    # ------------------------------------------------------ round-based API
    def begin(self, data_iterator):
        self._build_model()
        self.x_train, self.y_train = partition_to_arrays(data_iterator)
        self.history = {}
        return self

    @property
    def trainable(self):
        """Reference guard: a partition not larger than one batch is skipped (``worker.py:41``)."""
        return self.x_train is not None and self.x_train.shape[0] > int(self.train_config.get('batch_size') or 32)

    def num_batches(self):
        if not self.trainable:
            return 0
        n = self.x_train.shape[0]
        split = float(self.train_config.get('validation_split') or 0.0)
        n_train = int(n * (1.0 - split)) if 0.0 < split < 1.0 else n
        bs = int(self.train_config.get('batch_size') or 32)
        return -(-n_train // bs)

    def train_round(self, weights, epochs=1, initial_epoch=0):
        """Load ``weights`` (None: keep current), train ``epochs`` epochs, return ``(deltas, history)``."""
        if weights is not None:
            self.model.set_weights(weights)
        before = self.model.get_weights()
        history = None
        if self.trainable:
            cfg = dict(self.train_config)
            cfg['epochs'] = initial_epoch + epochs
            cfg['initial_epoch'] = initial_epoch
            h = self.model.fit(self.x_train, self.y_train, **cfg)
            for k, v in h.history.items():
                self.history.setdefault(k, []).extend(v)
            history = self.history
        after = self.model.get_weights()
        return subtract_params(before, after), history

    def train_batch(self, weights, batch_index):
        """One ``train_on_batch`` on the ``batch_index``-th contiguous batch (no shuffling); returns the
        delta, or a zero delta when this worker has no such batch."""
        if weights is not None:
            self.model.set_weights(weights)
        before = self.model.get_weights()
        bs = int(self.train_config.get('batch_size') or 32)
        start = batch_index * bs
        if self.trainable and start < self.x_train.shape[0]:
            stats = self.model.train_on_batch(self.x_train[start:start + bs], self.y_train[start:start + bs])
            stats = stats if isinstance(stats, list) else [stats]
            n = min(bs, self.x_train.shape[0] - start)
            for k, v in zip(self.model.metrics_names, stats):
                self.history.setdefault('batch_' + k, []).append(v)
                acc = self._epoch_acc.setdefault(k, [0.0, 0])
                acc[0] += v * n
                acc[1] += n
        after = self.model.get_weights()
        return subtract_params(before, after)

    def end_batch_epoch(self):
        """Close an epoch of ``train_batch`` calls: sample-weighted epoch means go to ``history['loss']`` ..."""
        for k, (total, n) in self._epoch_acc.items():
            if n:
                self.history.setdefault(k, []).append(total / n)
        self._epoch_acc = {}


class AsynchronousSparkWorker(_WorkerBase):
    """Asynchronous / hogwild worker: pull -> train -> push against a parameter server, once per
    epoch or once per batch."""

    def __init__(self, json, parameters, client, train_config, frequency, master_optimizer, master_loss,
                 master_metrics, custom_objects, device=None):
        super().__init__(json, parameters, train_config, master_optimizer, master_loss, master_metrics,
                         custom_objects, device)
        if isinstance(client, BaseParameterClient):
            self.client = client
        else:  # a client type name ('http' | 'socket' | 'nvlink'), as the reference allows
            self.client = BaseParameterClient.get_client(client)
        self.frequency = frequency
        self.history = {}
        self.run_info = {}  # device path: device time of the step loop, number of steps, ... (bench.py reads it)

    def train(self, data_iterator):
        x_train, y_train = partition_to_arrays(data_iterator)
        if x_train.size == 0:
            return
        if self.frequency not in ('epoch', 'batch'):
            raise ValueError('frequency parameter can be `epoch` or `batch, got {}'.format(self.frequency))
        # options of the device parameter-server path only; Model.fit must never see them
        self._device_opts = {k: self.train_config.pop(k, None) for k in ('input_mode', 'log_batch_loss', 'timed_from')}
        model = self._build_model()
        epochs = int(self.train_config['epochs'])
        batch_size = int(self.train_config.get('batch_size') or 32)
        n = x_train.shape[0]
        trainable = n > batch_size

        from .parameter.client import DeviceClient
        # callbacks run inside Model.fit: with them the worker takes the generic pull -> fit -> push loop below
        plain = not any(self.train_config.get(k) for k in ('callbacks', 'class_weight'))
        eng = model._native_engine() if model.device.type == 'cuda' and plain else None
        if eng is not None and isinstance(self.client, DeviceClient):
            self._train_against_device_ps(model, eng, x_train, y_train, epochs, batch_size, trainable)
            yield []
            return
        if isinstance(self.client, DeviceClient) and model.device.type == 'cpu' and not self.client._store().on_device:
            self._train_against_host_store(model, self.client._store(), x_train, y_train, epochs, batch_size, trainable)
            yield []
            return

        if self.frequency == 'epoch':
            cfg = dict(self.train_config)
            for epoch in range(epochs):
                before = self.client.get_parameters()
                model.set_weights(before)
                if trainable:
                    cfg['epochs'] = epoch + 1
                    cfg['initial_epoch'] = epoch
                    h = model.fit(x_train, y_train, **cfg)
                    for k, v in h.history.items():
                        self.history.setdefault(k, []).extend(v)
                self.client.update_parameters(subtract_params(before, model.get_weights()))
        else:
            nb_batch = -(-n // batch_size)
            for epoch in range(epochs):
                if not trainable:
                    continue
                for b in range(nb_batch):
                    before = self.client.get_parameters()
                    model.set_weights(before)
                    lo, hi = b * batch_size, min(n, (b + 1) * batch_size)
                    model.train_on_batch(x_train[lo:hi], y_train[lo:hi])
                    self.client.update_parameters(subtract_params(before, model.get_weights()))
        yield []

    def _train_against_host_store(self, model, store, x_train, y_train, epochs, batch_size, trainable):
        """CPU form of the device path (``parameter_server_mode='nvlink'`` on a machine without a GPU: the store is a flat
        tensor in this process behind the host RWLock). The same pull -> train -> push loop as the reference's
        (``elephas/worker.py:102-127``), but the weights travel as ONE flat tensor in the model's own layout: no list of
        NumPy arrays, no Keras-layout transposes, no pickling - two copies and one subtraction per round."""
        import torch
        layout = store.layout
        flat = torch.empty(layout.numel, dtype=torch.float32)
        before = torch.empty_like(flat)
        n = x_train.shape[0]

        def pull():
            store.pull_into(flat)
            before.copy_(flat)
            layout.unpack_to_model(flat, model)

        def push():
            layout.pack_model(model, out=flat)
            store.push_delta(before=before, after=flat)
            STATS['host_store_rounds'] += 1

        if self.frequency == 'epoch':
            cfg = dict(self.train_config)
            for epoch in range(epochs):
                pull()
                if trainable:
                    cfg['epochs'] = epoch + 1
                    cfg['initial_epoch'] = epoch
                    h = model.fit(x_train, y_train, **cfg)
                    for k, v in h.history.items():
                        self.history.setdefault(k, []).extend(v)
                push()
        elif trainable:
            nb_batch = -(-n // batch_size)
            for _ in range(epochs):
                for b in range(nb_batch):
                    pull()
                    lo, hi = b * batch_size, min(n, (b + 1) * batch_size)
                    model.train_on_batch(x_train[lo:hi], y_train[lo:hi])
                    push()

    def _train_against_device_ps(self, model, eng, x_train, y_train, epochs, batch_size, trainable):
        """B200 path of the pull -> train -> push loop: the replica's weights never leave the device.
        Pull is a (peer) copy kernel out of the striped store straight into the engine's flat fp32 / bf16 buffers, push
        is ``red.add(master, -delta)`` into the owners' stripes over NVLink (K7)."""
        import torch
        from .models import History
        store = self.client._store()
        history = History()
        cfg = self.train_config
        if self.frequency == 'epoch':
            before = torch.empty_like(eng.w32)
            split = float(cfg.get('validation_split') or 0.0)
            x, y, x_val, y_val = x_train, y_train, None, None
            if 0.0 < split < 1.0:
                at = int(len(x_train) * (1.0 - split))
                x, x_val, y, y_val = x_train[:at], x_train[at:], y_train[:at], y_train[at:]
            for epoch in range(epochs):
                store.pull_into(eng.w32, eng.w16)
                before.copy_(eng.w32)
                eng._weights_dirty_in_model = True
                if trainable:
                    eng.fit(x, y, batch_size, epoch + 1, bool(cfg.get('shuffle', True)), history, x_val, y_val,
                            int(cfg.get('verbose', 0) or 0), initial_epoch=epoch)
                store.push_delta(before=before, after=eng.w32)
            self.history = history.history
        elif trainable:
            self._train_batches_device(eng, store, x_train, y_train, epochs, batch_size)
        torch.cuda.current_stream(eng.device).synchronize()

    def _train_batches_device(self, eng, store, x_train, y_train, epochs, batch_size):
        """``frequency='batch'`` against the device parameter server: ONE CUDA-graph replay per batch, no host round trip.

        pull   hogwild: the forward GEMMs read their weight tiles from the owner's bf16 mirror (peer memory over NVLink)
               through TMA and side-write them into the local copy the backward pass uses (K2); only the vector tail is
               copied. asynchronous: one seqlock pull kernel (bf16 mirror -> compute copy, fp32 tail) - readers take no
               lock and retry a chunk that a writer touched meanwhile.
        train  forward / loss / backward on the native engine; the gradient stays in the flat fp32 buffer
        push   the optimizer kernel IS the push: the step of every element goes from registers into its owner's master
               (``red.add``; counted in and out of the chunk in asynchronous mode); the delta is never materialised

        Reference loop: ``get_parameters`` / ``set_weights`` / ``train_on_batch`` / ``subtract_params`` /
        ``update_parameters`` per batch (``elephas/worker.py:116-127``). Options (``train_config``): ``input_mode``
        'device' (default: the partition is uploaded once) or 'stream' (every batch is copied from pinned host memory),
        ``log_batch_loss`` (one 64-byte loss read-back per step, one step delayed)."""
        import os
        import torch
        cfg = self.train_config
        dev = eng.device
        lay = eng.layout
        n = x_train.shape[0]
        stream_inputs = self._device_opts.get('input_mode') == 'stream'
        log_loss = bool(self._device_opts.get('log_batch_loss'))
        xh = torch.as_tensor(np.ascontiguousarray(x_train, dtype=np.float32)).reshape(n, -1)
        yh = torch.as_tensor(np.ascontiguousarray(y_train, dtype=np.float32)).reshape(n, -1)
        if stream_inputs:
            xh = xh if xh.is_pinned() else xh.pin_memory()
            yh = yh if yh.is_pinned() else yh.pin_memory()
            copy_stream = torch.cuda.Stream(dev)
            stage = [(torch.empty(batch_size, xh.shape[1], device=dev), torch.empty(batch_size, yh.shape[1], device=dev),
                      torch.cuda.Event()) for _ in range(2)]
            consumed = [None, None]
        else:
            xd, yd = xh.to(dev), yh.to(dev)
        tail = (lay.matrix_end, lay.numel)
        sources = [store.matrix_source(op.w_seg.offset) for op in eng.ops]
        fused_pull = (not store.locked and all(a is not None for a in sources)
                      and os.environ.get('ELEPHAS_B200_FUSED_PULL', '1') != '0')
        saved_pull, saved_sink, saved_delta = eng.pull_source, eng.grad_sink, eng.delta_target
        eng.delta_target = None
        if fused_pull:
            eng.pull_source = sources
            pre = (lambda: store.pull_into(eng.w32, None, f32_range=tail)) if tail[1] > tail[0] else None
        else:
            eng.pull_source = None
            pre = lambda: store.pull_into(eng.w32, eng.w16, f32_range=tail, b16_range=(0, lay.matrix_end))  # noqa: E731
        hyper = eng._opt_kwargs()

        def sink(warmup):
            if warmup:
                # the run that loads the kernels before the capture must not move the master: with a zero gradient and
                # the fresh optimizer slots every worker starts from, the step of every element is zero
                eng.g32.zero_()
            store.optimizer_push(eng.opt_kind, eng.g32, eng.step, eng.slots, advance=not warmup, **hyper)

        eng.grad_sink = sink
        saved_hooks = eng.set_step_hooks(pre=pre)
        pinned = [torch.zeros(8, dtype=torch.float64).pin_memory() for _ in range(2)]
        events = [torch.cuda.Event() for _ in range(2)]
        losses = []
        cur = torch.cuda.current_stream(dev)

        def prefetch(step, lo):
            rows = min(batch_size, n - lo)
            xs, ys, ev = stage[step % 2]
            if consumed[step % 2] is not None:
                copy_stream.wait_event(consumed[step % 2])
            with torch.cuda.stream(copy_stream):
                xs[:rows].copy_(xh[lo:lo + rows], non_blocking=True)
                ys[:rows].copy_(yh[lo:lo + rows], non_blocking=True)
                ev.record(copy_stream)

        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        steps = 0
        timed_from = int(self._device_opts.get('timed_from') or 0)  # steps before this one are warm-up (not device-timed)
        try:
            eng.stats.zero_()
            eng.prepare_train(eng._buf(min(batch_size, n)))  # record the step graph now: the loop below only replays
            if timed_from == 0:
                t0.record(cur)
            for _ in range(epochs):
                if stream_inputs:
                    prefetch(steps, 0)
                for lo in range(0, n, batch_size):
                    rows = min(batch_size, n - lo)
                    buf = eng._buf(rows)
                    if stream_inputs:
                        xs, ys, ev = stage[steps % 2]
                        cur.wait_event(ev)
                        buf.x32.copy_(xs[:rows])
                        buf.y32.copy_(ys[:rows])
                        done = torch.cuda.Event()
                        done.record(cur)
                        consumed[steps % 2] = done
                        if lo + rows < n:
                            prefetch(steps + 1, lo + rows)
                    else:
                        buf.x32.copy_(xd[lo:lo + rows])
                        buf.y32.copy_(yd[lo:lo + rows])
                    eng.train_step_device(buf)
                    if log_loss:
                        slot = steps % 2
                        pinned[slot].copy_(eng.stats, non_blocking=True)
                        events[slot].record(cur)
                        if steps > 0:
                            events[1 - slot].synchronize()
                            losses.append(float(pinned[1 - slot][0] / max(float(pinned[1 - slot][7]), 1.0)))
                    steps += 1
                    if steps == timed_from:
                        t0.record(cur)
                    STATS['fused_pull_steps'] += int(fused_pull)
                    STATS['device_ps_steps'] += 1
            t1.record(cur)
            cur.synchronize()
            if log_loss and steps:
                slot = (steps - 1) % 2
                losses.append(float(pinned[slot][0] / max(float(pinned[slot][7]), 1.0)))
        finally:
            eng.pull_source, eng.grad_sink, eng.delta_target = saved_pull, saved_sink, saved_delta
            eng.set_step_hooks(*saved_hooks)
            eng.invalidate_graphs()  # the recorded steps hold the store's addresses
        vals, count = eng.read_stats()
        if count:
            self.history = dict(zip(eng.model.metrics_names, [[v] for v in vals]))
            if log_loss:
                self.history['batch_loss'] = losses
        self.run_info = {'device_ms': t0.elapsed_time(t1), 'steps': steps, 'timed_steps': steps - timed_from,
                         'fused_pull': fused_pull, 'graph_kernels': eng.last_graph_kernels,
                         'h2d_bytes_per_step': batch_size * (xh.shape[1] + yh.shape[1]) * 4 if stream_inputs else 0}
        store.pull_into(eng.w32, eng.w16)  # leave the replica with a complete, current copy
        eng._weights_dirty_in_model = True
