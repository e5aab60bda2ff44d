"""Device-resident parameter server (``parameter_server_mode='nvlink'``).

The reference's parameter server is a Flask process or a socket thread that receives pickled NumPy
lists (``elephas/parameter/server.py:42-233``). Here the master weights are one flat fp32 buffer in
the memory of the owning B200; a *pull* is a vectorised copy kernel out of that buffer (a peer load
over NVLink when the caller is another GPU) and a *push* applies ``master -= delta`` with
``red.global.add.f32`` straight into it - fused with the worker-side delta computation - so no
weight ever crosses the host. ``mode='asynchronous'`` keeps the reference's reader/writer exclusion
(``utils/rwlock.py``) with a device-side sequence lock: writers take a ticket and bump the sequence
word before and after the burst, readers retry while it is odd or changed; ``mode='hogwild'`` skips
both, exactly like ``server.py:109,120``.

A store is addressed by the server's ``port`` number (it plays the role of the URL). Ranks other
than the owner map the buffer through a CUDA IPC handle (``open_remote``). On machines without a
GPU the same class runs on CPU tensors with a host ``RWLock``, which is what the CPU test-suite
exercises.
"""
import threading

import torch

from ..utils.rwlock import RWLock

_stores = {}
_stores_guard = threading.Lock()


class DeviceStore:
    def __init__(self, port, layout_model, arrays, mode, device, remote_handle=None, mirror=False):
        from .flat import FlatLayout
        self.port, self.mode, self.device = port, mode, torch.device(device)
        self.layout = FlatLayout(layout_model)
        self._raw = None
        self._remote_base = None
        self.mirror = None  # bf16 copy of the master, kept current by every push: the operand of the pull-fused GEMM
        n = self.layout.numel
        if self.device.type == 'cuda':
            from ..ops import native
            ctrl_off = n * 4
            mirror_off = ctrl_off + 256
            if remote_handle is None:
                # owner: one IPC-exportable allocation holding the master weights, the control words and (for
                # stores that serve pull-fused GEMMs) the bf16 mirror
                self._raw = native.RawBuffer(mirror_off + (n * 2 if mirror else 0), self.device)
                base = self._raw.ptr
                self.flat = self._raw.tensor(torch.float32, n)
                self.ctrl = self._raw.tensor(torch.int32, 8, offset_bytes=ctrl_off)
                self.layout.pack_arrays(arrays, out=self.flat)
                if mirror:
                    self.mirror = self._raw.tensor(torch.bfloat16, n, offset_bytes=mirror_off)
                    native.cast_f32_to_bf16(self.flat, self.mirror)
            else:
                # another rank: map the owner's allocation; pulls/pushes go over NVLink
                with torch.cuda.device(self.device):
                    base = native.ipc_import(remote_handle)
                self._remote_base = base
                self.flat = _PeerPointer(base, n, self.device)
                self.ctrl = _PeerPointer(base + ctrl_off, 8, self.device)
                if mirror:
                    self.mirror = _PeerPointer(base + mirror_off, n, self.device)
        else:
            self.flat = self.layout.pack_arrays(arrays, device=self.device)
            # ctrl[0]: sequence word (odd while a writer is inside), ctrl[1]: next ticket, ctrl[2]: now serving
            self.ctrl = torch.zeros(8, dtype=torch.int32, device=self.device)
        self._host_lock = RWLock()
        self.pushes = 0
        # Locked operations of all worker threads of this process go through ONE stream, enqueued under a host mutex:
        # a ticket-lock kernel that spins on the GPU must never sit in front of the unlock kernel it waits for, and two
        # streams of one GPU can share a hardware queue. On a single in-order stream the previous holder's unlock always
        # precedes the next lock, so a lock kernel only ever waits for holders on OTHER GPUs (other ranks).
        self._ps_stream = None
        self._enqueue_lock = threading.Lock()
        # how the exclusion of mode='asynchronous' is realised (see configure()): 'ticket+stream' is always safe
        self._exclusion = 'ticket+stream'

    def export_handle(self):
        from ..ops import native
        return native.ipc_export(self._raw.ptr) if self._raw is not None else None

    def close(self):
        from ..ops import native
        if self._remote_base is not None:
            try:
                native.ipc_close(self._remote_base)
            except native.NativeError:
                pass
            self._remote_base = None
        if self._raw is not None:
            torch.cuda.synchronize(self.device)
            self._raw.free()
            self._raw = None

    @property
    def locked(self):
        return self.mode == 'asynchronous'

    def refresh_mirror(self):
        """Rebuild the bf16 mirror from the master (after the owner overwrote the master wholesale)."""
        if self.mirror is not None and self._raw is not None:
            from ..ops import native
            native.cast_f32_to_bf16(self.flat, self.mirror)

    def configure(self, local_workers, distributed):
        """Tell the store who uses it, so that it can pick the cheapest correct exclusion for ``asynchronous`` mode:

        * one process, several worker threads: all locked operations go through the store's in-order stream - the
          stream order IS the lock, no ticket-lock kernels are launched ('stream');
        * several processes, one worker each: ticket-lock kernels on the worker's own stream - every other holder
          runs on another GPU, so a spinning lock kernel never waits for work queued behind it ('ticket');
        * several processes with several workers each: both ('ticket+stream', also the default when unknown)."""
        if not distributed:
            self._exclusion = 'stream'
        elif local_workers <= 1:
            self._exclusion = 'ticket'
        else:
            self._exclusion = 'ticket+stream'

    @property
    def _ticket(self):
        """Launch ticket-lock kernels around an operation?"""
        return self.locked and self._exclusion != 'stream'

    # ---------------------------------------------------------- flat, on device
    def _device_op(self, fn):
        """Run ``fn`` (which enqueues [lock,] kernel [, unlock]) - on the caller's stream, or, where the exclusion
        needs it, on the store's own stream, ordered against the caller's stream with events."""
        if not self.locked or self._exclusion == 'ticket':
            fn()
            return
        cur = torch.cuda.current_stream(self.device)
        with self._enqueue_lock:
            if self._ps_stream is None:
                self._ps_stream = torch.cuda.Stream(self.device)
            ready = torch.cuda.Event()
            ready.record(cur)
            self._ps_stream.wait_event(ready)
            with torch.cuda.stream(self._ps_stream):
                fn()
            done = torch.cuda.Event()
            done.record(self._ps_stream)
        cur.wait_event(done)

    def pull_into(self, dst_flat, dst_bf16=None, stream=None):
        """``dst_flat <- master`` (and optionally the bf16 compute copy in the same pass)."""
        if self.flat.is_cuda:
            from ..ops import native
            self._device_op(lambda: native.ps_pull(self.flat, self.ctrl, dst_flat, dst_bf16, locked=self._ticket))
            return
        if self.locked:
            with self._host_lock.reading():
                dst_flat.copy_(self.flat)
        else:
            dst_flat.copy_(self.flat)
        if dst_bf16 is not None:
            dst_bf16.copy_(dst_flat)

    # -- raw forms: enqueue on the CURRENT stream without any exclusion (the caller provides it, e.g. by running a whole
    #    captured pull+step+push graph inside ``exclusive``) ------------------------------------------------------------
    def raw_pull_range(self, dst_flat, lo, n):
        from ..ops import native
        native.ps_pull(self.flat, self.ctrl, dst_flat[lo:lo + n], None, locked=False, n=n, offset=lo)

    def raw_push(self, delta_flat=None, delta_bf16=None):
        from ..ops import native
        self.pushes += 1
        native.ps_push(self.flat, self.ctrl, delta_flat, None, None, locked=False, mirror=self.mirror,
                       delta_bf16=delta_bf16)

    def raw_lock(self):
        from ..ops import native
        native.ps_lock(self.ctrl, self.device)

    def raw_unlock(self):
        from ..ops import native
        native.ps_unlock(self.ctrl, self.device)

    @property
    def ticket_only(self):
        """Locked store whose exclusion is the ticket lock alone (one worker per rank): lock / unlock kernels may be
        recorded into the worker's own step graph - every other holder runs on another GPU."""
        return self.locked and self._exclusion == 'ticket' and self.flat.is_cuda

    @property
    def stream_exclusive(self):
        """True when locked operations are serialised by the store's own stream (one process, no ticket kernels): a
        worker may then run a whole pull+step+push graph as ONE exclusive operation."""
        return self.locked and self._exclusion == 'stream' and self.flat.is_cuda

    def exclusive(self, fn):
        """Run ``fn`` as one exclusive operation of the store (see ``_device_op``)."""
        self._device_op(fn)

    def pull_range_into(self, dst_flat, lo, n):
        """``dst_flat[lo:lo+n] <- master[lo:lo+n]`` (device stores; used for the bias tail next to pull-fused GEMMs)."""
        from ..ops import native
        self._device_op(lambda: native.ps_pull(self.flat, self.ctrl, dst_flat[lo:lo + n], None, locked=self._ticket,
                                               n=n, offset=lo))

    def push_delta(self, delta_flat=None, before=None, after=None, delta_bf16=None):
        """``master -= delta`` where ``delta`` is given directly (fp32, or bf16 for mirrored stores) or as
        ``before - after`` (fused). A mirrored store refreshes its bf16 mirror in the same kernel."""
        self.pushes += 1
        if self.flat.is_cuda:
            from ..ops import native
            self._device_op(lambda: native.ps_push(self.flat, self.ctrl, delta_flat, before, after, locked=self._ticket,
                                                   mirror=self.mirror, delta_bf16=delta_bf16))
            return
        if delta_bf16 is not None:
            delta_flat = delta_bf16.float()
        delta = delta_flat if delta_flat is not None else (before - after)
        if self.locked:
            with self._host_lock.writing():
                self.flat.sub_(delta)
        else:
            self.flat.sub_(delta)

    # ------------------------------------------------- list-of-arrays contract
    def pull_arrays(self):
        tmp = torch.empty(self.layout.numel, dtype=torch.float32, device=self.device)
        self.pull_into(tmp)
        if tmp.is_cuda:
            torch.cuda.current_stream(tmp.device).synchronize()
            if isinstance(self.ctrl, torch.Tensor) and int(self.ctrl[4].item()):
                raise RuntimeError("device parameter server: a lock holder never released the ticket lock "
                                   "(watchdog fired after 10 s); master weights may have been updated unlocked")
        return self.layout.unpack_arrays(tmp)

    def push_arrays(self, delta):
        self.push_delta(self.layout.pack_arrays(delta, device=self.device))


class _PeerPointer:
    """A raw (peer-mapped) device address that quacks enough like a tensor for the native wrappers."""

    def __init__(self, ptr, numel, device):
        self._ptr, self._numel, self.device, self.is_cuda = ptr, numel, device, True

    def data_ptr(self):
        return self._ptr

    def numel(self):
        return self._numel


def create_store(port, arrays, mode, model=None, device=None, mirror=False):
    from .runtime import default_device
    if model is None:
        raise ValueError("create_store needs the model whose weights it holds")
    store = DeviceStore(port, model, arrays, mode, device or default_device(), mirror=mirror)
    with _stores_guard:
        _stores[port] = store
    return store


def attach_remote_store(port, handle, mode, model, device=None, mirror=False):
    """Register a proxy for a store owned by another rank (``handle``: its CUDA IPC handle)."""
    from .runtime import default_device
    store = DeviceStore(port, model, None, mode, device or default_device(), remote_handle=handle, mirror=mirror)
    with _stores_guard:
        _stores[port] = store
    return store


def get_store(port):
    with _stores_guard:
        store = _stores.get(port)
    if store is None:
        raise RuntimeError("no device parameter server is running on port %s in this process" % (port,))
    return store


def destroy_store(port):
    with _stores_guard:
        store = _stores.pop(port, None)
    if store is not None:
        store.close()
