"""Tracing / profiling hooks (the reference has none: SURVEY.md §5.1 - no timers, NVTX or callbacks anywhere
under ``elephas/``).

``span(name)`` brackets a host-side phase: it always opens an NVTX range when CUDA is up (so an external
timeline tool sees ``fit / broadcast / step / reduce / gather`` ranges for free), and - when tracing is switched
on with ``ELEPHAS_B200_TRACE=1`` or :func:`enable` - also accumulates wall-clock time per name.
``device_span(name)`` does the same with a pair of CUDA events on the current stream, resolved lazily in
:func:`summary` so that tracing never adds a synchronisation to the step loop.

    from elephas_b200.utils import tracing
    tracing.enable()
    spark_model.fit(rdd, epochs=1)
    print(tracing.report())
"""
import contextlib
import os
import threading
import time

_lock = threading.Lock()
_enabled = os.environ.get('ELEPHAS_B200_TRACE', '0') not in ('', '0')
_host = {}     # name -> [calls, total seconds, max seconds]
_device = {}   # name -> [calls, total ms, max ms]
_pending = []  # (name, start event, stop event)


def enable(on=True):
    global _enabled
    _enabled = bool(on)


def enabled():
    return _enabled


def reset():
    with _lock:
        _host.clear()
        _device.clear()
        del _pending[:]


def _nvtx():
    try:
        import torch
        if torch.cuda.is_available() and torch.cuda.is_initialized():
            return torch.cuda.nvtx
    except Exception:  # pragma: no cover - torch missing or broken CUDA runtime
        pass
    return None


def add_host(name, seconds):
    """Account ``seconds`` of host time to ``name`` (for loops that are cheaper to time by hand)."""
    if _enabled:
        with _lock:
            rec = _host.setdefault(name, [0, 0.0, 0.0])
            rec[0] += 1
            rec[1] += seconds
            rec[2] = max(rec[2], seconds)


@contextlib.contextmanager
def span(name):
    """Host-side phase: NVTX range + (when tracing is on) wall-clock accounting."""
    nvtx = _nvtx()
    if nvtx is not None:
        nvtx.range_push(name)
    t0 = time.perf_counter() if _enabled else 0.0
    try:
        yield
    finally:
        if _enabled:
            dt = time.perf_counter() - t0
            with _lock:
                rec = _host.setdefault(name, [0, 0.0, 0.0])
                rec[0] += 1
                rec[1] += dt
                rec[2] = max(rec[2], dt)
        if nvtx is not None:
            nvtx.range_pop()


@contextlib.contextmanager
def device_span(name, stream=None):
    """Device-side phase on ``stream`` (default: the current stream), timed with CUDA events."""
    if not _enabled or _nvtx() is None:
        with span(name):
            yield
        return
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(s)
    try:
        with span(name):
            yield
    finally:
        e1.record(s)
        with _lock:
            _pending.append((name, e0, e1))


def _resolve():
    with _lock:
        pending = list(_pending)
        del _pending[:]
    for name, e0, e1 in pending:
        e1.synchronize()
        ms = e0.elapsed_time(e1)
        with _lock:
            rec = _device.setdefault(name, [0, 0.0, 0.0])
            rec[0] += 1
            rec[1] += ms
            rec[2] = max(rec[2], ms)


def summary():
    """``{'host': {name: {'calls', 'total_ms', 'max_ms'}}, 'device': {...}}`` of everything traced so far."""
    _resolve()
    with _lock:
        host = {k: {'calls': v[0], 'total_ms': v[1] * 1e3, 'max_ms': v[2] * 1e3} for k, v in _host.items()}
        dev = {k: {'calls': v[0], 'total_ms': v[1], 'max_ms': v[2]} for k, v in _device.items()}
    return {'host': host, 'device': dev}


def report():
    """Human-readable table of :func:`summary`, longest phases first."""
    s = summary()
    lines = []
    for kind in ('host', 'device'):
        rows = sorted(s[kind].items(), key=lambda kv: -kv[1]['total_ms'])
        if rows:
            lines.append('%-6s %-34s %8s %12s %12s' % (kind, 'phase', 'calls', 'total ms', 'max ms'))
            for name, r in rows:
                lines.append('%-6s %-34s %8d %12.3f %12.3f' % ('', name, r['calls'], r['total_ms'], r['max_ms']))
    return '\n'.join(lines)
