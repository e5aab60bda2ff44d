"""Intra-op thread count for models that run on the CPU.

The reference's workloads are small MLPs (784-128-128-10 at batch 64): one matrix product of a step is a few hundred
kilo-FLOPs, far below the size at which splitting it over the cores pays for the fork / join of the thread pool - and on
virtual machines whose vCPUs are not scheduled together a pool-wide barrier can cost milliseconds (measured here: a
step of the MNIST MLP takes 87 ms with 8 threads, 2 ms with one). torch picks "all cores" by default. This module times
the model's largest matrix product once per process and shape with a few candidate thread counts and keeps the fastest;
``ELEPHAS_B200_CPU_THREADS=N`` fixes the count, ``=default`` leaves torch alone. An explicit ``OMP_NUM_THREADS`` /
``torch.set_num_threads`` by the user is respected as the upper bound (candidates never exceed the current value).

No counterpart in the reference (TensorFlow makes the same kind of decision inside its own thread pools).
"""
import os
import threading
import time

import torch

_lock = threading.Lock()
_chosen = {}  # (rows, k, n, upper bound) -> thread count


SMALL_FLOPS = 5e7  # below this a matrix product (about a millisecond on one core) is not worth a thread pool


def _time_matmul(rows, k, n, reps=4):
    a, b = torch.randn(rows, k), torch.randn(k, n)
    torch.mm(a, b)
    t0 = time.perf_counter()
    for _ in range(reps):
        torch.mm(a, b)
    return (time.perf_counter() - t0) / reps


def tune_for(model, batch_size):
    """Set torch's intra-op thread count for training / inference of ``model`` on the CPU; returns the count in effect."""
    setting = os.environ.get('ELEPHAS_B200_CPU_THREADS', 'auto')
    if setting == 'default':
        return torch.get_num_threads()
    if setting != 'auto':
        n = max(1, int(setting))
        if torch.get_num_threads() != n:
            torch.set_num_threads(n)
        return n
    shapes = [tuple(w.shape) for w in model.trainable_weights if w.dim() == 2]
    if not shapes:
        return torch.get_num_threads()  # (convolutional / exotic models: leave the default alone)
    n_out, k_in = max(shapes, key=lambda s: s[0] * s[1])
    rows = max(1, min(int(batch_size or 32), 4096))
    with _lock:
        upper = getattr(tune_for, '_upper', None)
        if upper is None:
            upper = tune_for._upper = torch.get_num_threads()  # what the user / torch started with
        key = (rows, k_in, n_out, upper)
        best = _chosen.get(key)
        if best is None:
            if 2.0 * rows * k_in * n_out < SMALL_FLOPS:
                best = 1
            else:
                # ascending, and stop at the first count that does not pay: each step up must win 10 % (fewer threads
                # leave cores to the other workers of the process; a pathological count is never timed at length)
                candidates = sorted({c for c in (1, 2, 4, upper // 2, upper) if 1 <= c <= upper})
                best, best_t = candidates[0], None
                for c in candidates:
                    torch.set_num_threads(c)
                    t = _time_matmul(rows, k_in, n_out)
                    if best_t is None or t < 0.9 * best_t:
                        best, best_t = c, t
                    else:
                        break
            _chosen[key] = best
        if torch.get_num_threads() != best:
            torch.set_num_threads(best)
        return best
