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
STATS = {'fused_pull_steps': 0}
# largest model (parameters) whose locked per-batch pull across ranks is fused into the forward pass: the lock is then held
# for the forward instead of for one copy kernel, which only pays while the step is launch-bound
FUSED_LOCKED_MAX = 4 << 20

class _WorkerBase:
    def __init__(self, json, parameters, train_config, master_optimizer, master_loss, master_metrics,
                 custom_objects, device=None):
        self.json = json
        self.parameters = parameters
        self.train_config = dict(train_config)
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

    def train(self, data_iterator):
        x_train, y_train = partition_to_arrays(data_iterator)
        if x_train.size == 0:
            return
        if self.frequency not in ('epoch', 'batch'):
            raise ValueError('frequency parameter can be `epoch` or `batch, got {}'.format(self.frequency))
        model = self._build_model()
        epochs = int(self.train_config['epochs'])
        batch_size = int(self.train_config.get('batch_size') or 32)
        n = x_train.shape[0]
        trainable = n > batch_size

        from .parameter.client import DeviceClient
        eng = model._native_engine() if model.device.type == 'cuda' else None
        if eng is not None and isinstance(self.client, DeviceClient):
            self._train_against_device_ps(model, eng, x_train, y_train, epochs, batch_size, trainable)
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

    def _train_against_device_ps(self, model, eng, x_train, y_train, epochs, batch_size, trainable):
        """B200 path of the pull -> train -> push loop: the replica's weights never leave the device.
        Pull is a (peer) copy kernel straight into the engine's flat fp32/bf16 buffers, push is
        ``red.add(master, -(before - after))`` fused with the delta computation (K7)."""
        import torch
        from .models import History
        store = self.client._store()
        before = torch.empty_like(eng.w32)
        history = History()
        cfg = self.train_config
        if self.frequency == 'epoch':
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
        else:
            n = x_train.shape[0]
            dev = eng.device
            xd = torch.as_tensor(np.ascontiguousarray(x_train, dtype=np.float32)).reshape(n, -1).to(dev)
            yd = torch.as_tensor(np.ascontiguousarray(y_train, dtype=np.float32)).reshape(n, -1).to(dev)
            if store.mirror is not None and trainable and (not store.locked or store.stream_exclusive or
                                                           (store.ticket_only and eng.layout.numel <= FUSED_LOCKED_MAX)):
                self._train_fused_pull(eng, store, xd, yd, epochs, batch_size)
                torch.cuda.current_stream(eng.device).synchronize()
                return
            for epoch in range(epochs):
                if not trainable:
                    continue
                for lo in range(0, n, batch_size):
                    rows = min(batch_size, n - lo)
                    buf = eng._buf(rows)
                    store.pull_into(eng.w32, eng.w16)
                    before.copy_(eng.w32)
                    buf.x32.copy_(xd[lo:lo + rows])
                    buf.y32.copy_(yd[lo:lo + rows])
                    eng.train_step_device(buf)
                    store.push_delta(before=before, after=eng.w32)
        torch.cuda.current_stream(eng.device).synchronize()

    @staticmethod
    def _train_fused_pull(eng, store, xd, yd, epochs, batch_size):
        """``frequency='batch'`` against a device parameter server with a bf16 mirror: ONE graph replay per batch.
        (hogwild: always; asynchronous: inside one process - the store's stream serialises whole batches - or, with one
        worker per rank and a small model, with the ticket lock recorded around the pulling forward and the push.)

        pull   = the forward GEMMs read their weight tiles from the store's mirror (peer memory over NVLink) through
                 TMA and side-write them into the local copy the backward pass uses (K2); only the bias tail is copied
        train  = forward / loss / backward on the native engine, the optimizer runs in delta mode
        push   = ``master -= delta`` with ``red.add`` straight into the store + refresh of the mirror, same graph

        Reference loop: ``get_parameters`` / ``set_weights`` / ``train_on_batch`` / ``update_parameters`` per batch
        (``elephas/worker.py:116-127``)."""
        import torch
        lay = eng.layout
        tail_lo, tail_n = lay.matrix_end, lay.numel - lay.matrix_end
        delta16 = torch.zeros(lay.numel, dtype=torch.bfloat16, device=eng.device)
        n = xd.shape[0]
        saved = (eng.pull_source, eng.delta_target, eng.pre_step, eng.post_step, eng.mid_step)
        eng.pull_source = store.mirror.data_ptr()
        eng.delta_target = delta16
        if store.ticket_only:
            # mode='asynchronous' across ranks, small model: the ticket lock is taken inside the graph, once around the
            # pulling forward pass and once around the push (a rank's holders are always on other GPUs)
            def pre():
                store.raw_lock()
                if tail_n > 0:
                    store.raw_pull_range(eng.w32, tail_lo, tail_n)

            def post():
                store.raw_lock()
                store.raw_push(delta_bf16=delta16)
                store.raw_unlock()

            eng.pre_step, eng.mid_step, eng.post_step = pre, store.raw_unlock, post
        else:
            eng.pre_step = (lambda: store.raw_pull_range(eng.w32, tail_lo, tail_n)) if tail_n > 0 else None
            eng.post_step = lambda: store.raw_push(delta_bf16=delta16)
        # mode='asynchronous' inside one process: the store's in-order stream is the lock, and the whole batch (pull,
        # step, push) runs on it as ONE exclusive operation - the strictest form of the reference's locked updates
        exclusive = store.stream_exclusive

        def one_batch(lo, rows):
            buf = eng._buf(rows)
            buf.x32.copy_(xd[lo:lo + rows])
            buf.y32.copy_(yd[lo:lo + rows])
            eng.train_step_device(buf)

        try:
            store.push_delta(delta_bf16=delta16)  # all-zero delta: loads the push kernel outside of graph capture
            for _ in range(epochs):
                for lo in range(0, n, batch_size):
                    rows = min(batch_size, n - lo)
                    if exclusive:
                        store.exclusive(lambda: one_batch(lo, rows))
                    else:
                        one_batch(lo, rows)
                    STATS['fused_pull_steps'] += 1
            torch.cuda.current_stream(eng.device).synchronize()
        finally:
            eng.pull_source, eng.delta_target, eng.pre_step, eng.post_step, eng.mid_step = saved
            eng.invalidate_graphs()  # the recorded steps hold the store's addresses
        store.pull_into(eng.w32, eng.w16)  # leave the replica with a complete, current copy
        eng._weights_dirty_in_model = True
