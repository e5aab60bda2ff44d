"""CPU restatement of the reference's algorithm, in NumPy only.

The reference (arpitjain799/elephas 4.1.0) cannot be installed here (TensorFlow <= 2.10, PySpark <= 3.3, Flask, h5py,
Python < 3.11 - see DESIGN.md). This file restates WHAT IT COMPUTES, independently of the ``elephas_b200`` package, so
that it can serve as (a) a numerical oracle for the package's distributed modes and (b) a timed CPU baseline:

* ``Mlp``                 the Keras model of the reference's tests / examples (Dense + relu ... softmax / linear, categorical
                          cross-entropy or MSE, plain SGD), forward / backward written out in NumPy, Keras conventions
                          (last partial batch kept, loss averaged over the batch)
* ``sync_fit``            ``SparkModel(mode='synchronous')``: every partition trains a replica for all epochs from the same
                          weights, the driver subtracts the mean of ``before - after``
                          (``elephas/worker.py:26-49``, ``elephas/spark_model.py:217-231``)
* ``ParameterServer``     master weights behind the writer-priority reader/writer lock (``elephas/utils/rwlock.py``),
                          ``get`` / ``update`` = ``w <- w - delta`` (``elephas/parameter/server.py:107-132``)
* ``async_fit``           ``mode='asynchronous' | 'hogwild'``: per-epoch or per-batch pull -> train -> push against the server
                          from one thread per partition (``elephas/worker.py:76-131``)

    python baseline/cpu_restatement.py            # timed baselines of BASELINE.json configs 1 and 2 on this CPU (JSON lines)
"""
import json
import threading
import time

import numpy as np


class Mlp:
    """``dims = [in, h1, ..., out]``; hidden activation relu; output 'softmax' (categorical CE) or 'linear' (MSE)."""

    def __init__(self, dims, output='softmax', seed=0):
        rng = np.random.default_rng(seed)
        self.output = output
        self.weights = []
        for fan_in, fan_out in zip(dims[:-1], dims[1:]):
            limit = np.sqrt(6.0 / (fan_in + fan_out))  # Glorot uniform, the Keras default
            self.weights.append(rng.uniform(-limit, limit, size=(fan_in, fan_out)).astype(np.float32))  # Keras (in, out)
            self.weights.append(np.zeros(fan_out, dtype=np.float32))

    def get_weights(self):
        return [w.copy() for w in self.weights]

    def set_weights(self, weights):
        self.weights = [np.array(w, dtype=np.float32, copy=True) for w in weights]

    def forward(self, x):
        acts = [x]
        h = x
        n_layers = len(self.weights) // 2
        for i in range(n_layers):
            z = h @ self.weights[2 * i] + self.weights[2 * i + 1]
            if i < n_layers - 1:
                h = np.maximum(z, 0.0)
            elif self.output == 'softmax':
                z = z - z.max(axis=1, keepdims=True)
                e = np.exp(z)
                h = e / e.sum(axis=1, keepdims=True)
            else:
                h = z
            acts.append(h)
        return acts

    def loss(self, x, y):
        out = self.forward(x)[-1]
        if self.output == 'softmax':
            return float(-(y * np.log(np.clip(out, 1e-7, 1.0))).sum(axis=1).mean())
        return float(((out - y) ** 2).mean())

    def train_on_batch(self, x, y, lr):
        acts = self.forward(x)
        out = acts[-1]
        n = len(x)
        # softmax + categorical CE and linear + MSE share the form of the output gradient
        dz = (out - y) / n if self.output == 'softmax' else 2.0 * (out - y) / (n * y.shape[1])
        n_layers = len(self.weights) // 2
        for i in reversed(range(n_layers)):
            a_prev = acts[i]
            dw = a_prev.T @ dz
            db = dz.sum(axis=0)
            if i > 0:
                dz = (dz @ self.weights[2 * i].T) * (acts[i] > 0)
            self.weights[2 * i] -= lr * dw
            self.weights[2 * i + 1] -= lr * db

    def fit(self, x, y, epochs, batch_size, lr):
        for _ in range(epochs):  # (no shuffling: the oracle comparisons run the package with shuffle=False)
            for lo in range(0, len(x), batch_size):
                self.train_on_batch(x[lo:lo + batch_size], y[lo:lo + batch_size], lr)


def _sub(a, b):
    return [x - y for x, y in zip(a, b)]


def partitions(x, y, n):
    """Contiguous, near-equal split - what ``rdd.repartition(n)`` amounts to for an ordered local collection."""
    base, extra = divmod(len(x), n)  # the first `extra` partitions hold one row more
    out, start = [], 0
    for i in range(n):
        end = start + base + (1 if i < extra else 0)
        out.append((x[start:end], y[start:end]))
        start = end
    return out


def sync_fit(model, parts, epochs, batch_size, lr):
    """elephas/spark_model.py:217-231 + elephas/worker.py:26-49."""
    start = model.get_weights()
    new = [w.copy() for w in start]
    for px, py in parts:
        replica = Mlp([1, 1], model.output)
        replica.set_weights(start)
        if len(px) > batch_size:  # worker.py:41: a partition not larger than one batch trains nothing
            replica.fit(px, py, epochs, batch_size, lr)
        delta = _sub(start, replica.get_weights())
        new = _sub(new, [d / len(parts) for d in delta])
    model.set_weights(new)
    return model


class RWLock:
    """Writer-priority reader/writer lock (elephas/utils/rwlock.py:24-48)."""

    def __init__(self):
        self.cond = threading.Condition()
        self.readers, self.writer, self.waiting_writers = 0, False, 0

    def acquire_read(self):
        with self.cond:
            while self.writer or self.waiting_writers:
                self.cond.wait()
            self.readers += 1

    def acquire_write(self):
        with self.cond:
            self.waiting_writers += 1
            while self.writer or self.readers:
                self.cond.wait()
            self.waiting_writers -= 1
            self.writer = True

    def release(self):
        with self.cond:
            if self.writer:
                self.writer = False
            else:
                self.readers -= 1
            self.cond.notify_all()


class ParameterServer:
    """elephas/parameter/server.py: ``mode='asynchronous'`` locks reads and updates, ``'hogwild'`` does not."""

    def __init__(self, weights, mode):
        self.weights = [w.copy() for w in weights]
        self.locked = mode == 'asynchronous'
        self.lock = RWLock()

    def get(self):
        if self.locked:
            self.lock.acquire_read()
        try:
            return [w.copy() for w in self.weights]
        finally:
            if self.locked:
                self.lock.release()

    def update(self, delta):
        if self.locked:
            self.lock.acquire_write()
        try:
            self.weights = _sub(self.weights, delta)
        finally:
            if self.locked:
                self.lock.release()


def async_fit(model, parts, epochs, batch_size, lr, mode='asynchronous', frequency='epoch'):
    """elephas/worker.py:76-131 with one thread per partition against an in-process parameter server."""
    server = ParameterServer(model.get_weights(), mode)

    def worker(px, py):
        replica = Mlp([1, 1], model.output)
        replica.set_weights(server.get())
        if len(px) <= batch_size:
            return
        if frequency == 'epoch':
            for _ in range(epochs):
                before = server.get()
                replica.set_weights(before)
                replica.fit(px, py, 1, batch_size, lr)
                server.update(_sub(before, replica.get_weights()))
        else:
            for _ in range(epochs):
                for lo in range(0, len(px), batch_size):
                    before = server.get()
                    replica.set_weights(before)
                    replica.train_on_batch(px[lo:lo + batch_size], py[lo:lo + batch_size], lr)
                    server.update(_sub(before, replica.get_weights()))

    threads = [threading.Thread(target=worker, args=p) for p in parts]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    model.set_weights(server.get())
    return model


def synthetic_mnist(n, seed=0):
    """MNIST-shaped synthetic classification data (784 features, 10 classes) with learnable structure."""
    rng = np.random.default_rng(seed)
    x = rng.random((n, 784), dtype=np.float32)
    teacher = np.random.default_rng(99).normal(size=(784, 10)).astype(np.float32)
    labels = (x @ teacher).argmax(axis=1)
    return x, np.eye(10, dtype=np.float32)[labels]


def main():
    n, epochs, bs, lr = 60000, 3, 64, 0.1
    x, y = synthetic_mnist(n)
    rows = []
    for name, run in (
            ('config 1: MNIST MLP synchronous/epoch, 2 workers', lambda m: sync_fit(m, partitions(x, y, 2), epochs, bs, lr)),
            ('config 2: MNIST MLP asynchronous/batch, 2 workers', lambda m: async_fit(m, partitions(x, y, 2), epochs, bs, lr,
                                                                                       'asynchronous', 'batch')),
            ('MNIST MLP hogwild/batch, 2 workers', lambda m: async_fit(m, partitions(x, y, 2), epochs, bs, lr, 'hogwild', 'batch'))):
        model = Mlp([784, 128, 128, 10], 'softmax', seed=1)
        before = model.loss(x[:4096], y[:4096])
        t0 = time.perf_counter()
        run(model)
        dt = time.perf_counter() - t0
        rows.append({'impl': 'cpu restatement of the reference algorithm (NumPy)', 'config': name, 'device': 'cpu',
                     'value': round(n * epochs / dt, 1), 'unit': 'samples/sec', 'seconds': round(dt, 3),
                     'loss_before': round(before, 4), 'final_loss': round(model.loss(x[:4096], y[:4096]), 4)})
        print(json.dumps(rows[-1]), flush=True)
    return rows


if __name__ == '__main__':
    main()
