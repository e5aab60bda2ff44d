"""Device-resident parameter server (``parameter_server_mode='nvlink'``), striped over all GPUs.

The reference's parameter server is a Flask process or a socket thread that holds the master weights and
receives pickled NumPy lists (``elephas/parameter/server.py:42-233``); every worker of every executor
goes through that one process. Here the master weights are one flat fp32 buffer (``parallel/flat.py``)
cut into one **stripe per rank**: stripe *r* lives in the HBM of GPU *r* and is mapped into every other
rank's address space over NVLink (``parallel/symm.py``), together with a bf16 mirror of it and the lock
records of its chunks. Ownership is decided per Dense kernel (whole matrices, largest first onto the
least loaded rank, so the pull-fused GEMM can read a layer through one tensor map) plus an even split of
the vector tail; with N ranks the pushes and pulls of N workers therefore land on N memories, N L2s and N
sets of NVLink ports.

* **push** (``client.update_parameters``, ``server.py:117-132,201-210``): ONE kernel; the block that serves a
  chunk counts itself in (``started += 1``, one 64-bit atomic that also returns the chunk's gate), applies
  ``master -= delta`` with ``red.global.add.f32`` straight into the owner's memory, and counts itself out
  (``finished += 1``). Element-wise atomic additions commute, so writers need no exclusion among themselves:
  every interleaving equals a serial order of the same updates, and no writer ever queues behind another.
  The delta can be given (fp32 / bf16), formed as ``before - after`` on the fly, or - ``optimizer_push`` - be
  the optimizer step itself, computed from gradient and slots in registers and never written to memory.
* **pull** (``client.get_parameters``, ``server.py:107-115,212-218``): readers take **no lock**. In
  ``mode='asynchronous'`` a chunk is copied between two reads of its counters (``finished == started``:
  nobody inside) and copied again when ``started`` moved (seqlock). A reader that keeps losing (8 attempts)
  raises the chunk's gate - writers that see it step back out and wait - copies the quiet chunk and lowers
  the gate, which bounds its wait. Any number of workers read at once - the device-side form of the
  reference's reader/writer lock (``utils/rwlock.py:24-48``) - and none of them ever sees a half-applied push
  inside a chunk. bf16 destinations are converted from the fp32 master on the fly.
* ``mode='hogwild'`` skips the counters (``server.py:109,120``): relaxed ``red.add``, plain
  loads. Only here do pushes also refresh the bf16 mirror (read back, round, store), because only here is
  it read: by bf16 pulls and by the pull-fused GEMM (K2), whose TMA loads fetch weight tiles from the
  owner's mirror. In the locked mode that read-back would only lengthen the window in which readers have to retry.

Consistency unit: the **chunk** (at least 8192 elements, at most 64 chunks per stripe). A seqlock pull
never sees a half-applied push inside a chunk, and pushes are element-wise atomic in every mode, so no
update is ever lost (the reference's socket server can lose one: it reads the weights before it takes the
write lock, ``server.py:204-206``). Pushes of one worker to different chunks are not one transaction.
In hogwild mode a concurrent pusher's older read-back may overwrite a newer bf16 mirror entry; the entry
is corrected the next time the element is pushed (the fp32 master is always exact).

A store is addressed by the server's ``port`` number (it plays the role of the URL). Under ``torchrun``
creation is collective: every rank calls ``create_store`` and contributes its stripe. On machines
without a GPU the same class runs on CPU tensors with a host ``RWLock``, which is what the CPU test-suite
exercises.
"""
import threading

import torch

from ..utils.rwlock import RWLock

_stores = {}
_stores_guard = threading.Lock()

MIN_CHUNK = 8192        # elements; the consistency / locking unit of a stripe
MAX_CHUNKS = 64         # lock records per stripe (EB_PS_MAX_CHUNKS)
MAX_SPANS = 48          # EB_PS_MAX_SPANS
TAIL_PIECE_MIN = 16384  # the vector tail is split over ranks in pieces of at least this many elements


def _round_up(x, m):
    return (x + m - 1) // m * m


class StripePlan:
    """Which rank owns which part of the flat parameter buffer.

    ``spans``: ``(flat_off, n, owner, owner_off)`` - flat elements ``[flat_off, flat_off + n)`` live at
    ``[owner_off, owner_off + n)`` of rank ``owner``'s stripe. Deterministic in (layout, world): every rank
    computes the same plan without talking to anyone."""

    def __init__(self, layout, world):
        self.world = world
        pieces = []  # (flat_off, n, is_matrix)
        mats = sorted((s for s in layout.segments if getattr(s, '_is_matrix', False)), key=lambda s: s.offset)
        if mats and len(mats) + world <= MAX_SPANS:
            for i, seg in enumerate(mats):
                end = mats[i + 1].offset if i + 1 < len(mats) else layout.matrix_end
                pieces.append((seg.offset, end - seg.offset, True))
            tail_lo = layout.matrix_end
        else:
            tail_lo = 0  # no whole-matrix spans (too many matrices, or none): one even split of everything
        tail_n = layout.numel - tail_lo
        if tail_n > 0:
            parts = max(1, min(world, tail_n // TAIL_PIECE_MIN))
            per = _round_up(-(-tail_n // parts), 64)
            lo = tail_lo
            while lo < layout.numel:
                n = min(per, layout.numel - lo)
                pieces.append((lo, n, False))
                lo += n
        load = [0] * world
        owner_of = {}
        for flat_off, n, _ in sorted(pieces, key=lambda p: (-p[1], p[0])):
            r = min(range(world), key=lambda k: (load[k], k))
            owner_of[flat_off] = r
            load[r] += n
        fill = [0] * world
        self.spans = []
        self._matrix_home = {}
        for flat_off, n, is_matrix in sorted(pieces):
            r = owner_of[flat_off]
            self.spans.append((flat_off, n, r, fill[r]))
            if is_matrix:
                self._matrix_home[flat_off] = (r, fill[r])
            fill[r] += n
        self.stripe_len = fill
        longest = max(fill + [64])
        self.chunk_len = max(MIN_CHUNK, _round_up(-(-longest // MAX_CHUNKS), 64))
        self.chunks = max(1, -(-longest // self.chunk_len))
        self.capacity = self.chunks * self.chunk_len  # elements allocated per stripe (same on every rank)
        self.numel = layout.numel

    def matrix_home(self, flat_off):
        """``(owner, owner_off)`` of the Dense kernel that starts at flat element ``flat_off``, or None when kernels are
        not stored whole (then there is no single address a pull-fused GEMM could read the layer from)."""
        return self._matrix_home.get(flat_off)

    def chunk_ranges_flat(self):
        """Flat ``(lo, hi)`` ranges that are each wholly inside one chunk: the granularity at which a seqlock pull is
        consistent (tests compare snapshots range by range)."""
        out = []
        for flat_off, n, _, owner_off in self.spans:
            a = owner_off
            while a < owner_off + n:
                b = min(owner_off + n, (a // self.chunk_len + 1) * self.chunk_len)
                out.append((flat_off + a - owner_off, flat_off + b - owner_off))
                a = b
        return out


class _SoloRuntime:
    """Stand-in for the rank runtime when a store lives in one process (one stripe, no exchange of handles)."""
    world_size, rank, distributed = 1, 0, False

    def __init__(self, device):
        self.device = torch.device(device)

    def barrier(self):
        pass

    def all_gather_object(self, obj):
        return [obj]

    def broadcast_(self, tensor, src=0):
        return tensor


class DeviceStore:
    def __init__(self, port, layout_model, arrays, mode, device, rt=None):
        from .flat import FlatLayout
        self.port, self.mode, self.device = port, mode, torch.device(device)
        self.layout = FlatLayout(layout_model)
        self.mem = None
        self.plan = None
        self.pushes = 0
        n = self.layout.numel
        if self.device.type == 'cuda':
            from ..ops import native
            from .symm import SymmetricMemory
            self.rt = rt if (rt is not None and rt.distributed) else _SoloRuntime(self.device)
            plan = self.plan = StripePlan(self.layout, self.rt.world_size)
            cap = plan.capacity
            with torch.cuda.device(self.device):
                self.mem = SymmetricMemory(self.rt, {'master': cap * 4, 'mirror': cap * 2,
                                                     'ctrl': MAX_CHUNKS * native.PS_LOCK_WORDS * 4})
                self.args = native.ps_store_args(self.mem.ptrs('master'), self.mem.ptrs('mirror'), self.mem.ptrs('ctrl'),
                                                 plan.stripe_len, plan.chunk_len, plan.chunks, self.rt.rank, plan.spans)
                self.stripe32 = self.mem.local('master', torch.float32, cap)
                self.stripe16 = self.mem.local('mirror', torch.bfloat16, cap)
                self.ctrl = self.mem.local('ctrl', torch.int32, MAX_CHUNKS * native.PS_LOCK_WORDS)
                if arrays is not None:
                    self.load_arrays(arrays)
        else:
            self.rt = None
            self.flat = self.layout.pack_arrays(arrays, device=self.device)
        self._host_lock = RWLock()

    # ------------------------------------------------------------------ life cycle
    def load_arrays(self, arrays):
        """(Collective under ``torchrun``.) Replace the master by rank 0's ``arrays`` and rebuild the bf16 mirror."""
        if self.mem is None:
            self.layout.pack_arrays(arrays, out=self.flat)
            return
        from ..ops import native
        with torch.cuda.device(self.device):
            flat = self.layout.pack_arrays(arrays, device=self.device)
            self.rt.broadcast_(flat, src=0)  # plumbing: every stripe is cut from the driver's weights
            self.load_flat(flat)

    def load_flat(self, flat):
        """This rank's stripe <- its spans of ``flat`` (a full flat fp32 buffer on this device); mirror rebuilt."""
        from ..ops import native
        for flat_off, n, owner, owner_off in self.plan.spans:
            if owner == self.rt.rank:
                self.stripe32[owner_off:owner_off + n].copy_(flat[flat_off:flat_off + n])
        native.cast_f32_to_bf16(self.stripe32, self.stripe16)
        torch.cuda.current_stream(self.device).synchronize()
        self.rt.barrier()

    def close(self):
        if self.mem is not None:
            with torch.cuda.device(self.device):
                self.mem.close()
            self.mem = None

    @property
    def locked(self):
        return self.mode == 'asynchronous'

    @property
    def on_device(self):
        return self.mem is not None

    def configure(self, local_workers=1, distributed=False):
        """Kept for callers of the first design (one owner GPU, lock kernels on a dedicated stream): the chunk counters are
        now read and moved inside the push / pull kernels themselves, so the same code is correct for
        worker threads of one process and for one worker per rank. Nothing to choose any more."""

    def matrix_source(self, flat_off):
        """Device address of the bf16 mirror of the Dense kernel stored at flat element ``flat_off`` (in its owner's
        stripe, usually peer memory): the B operand of the pull-fused GEMM (K2). None when kernels are not stored whole."""
        home = self.plan.matrix_home(flat_off) if self.plan is not None else None
        if home is None:
            return None
        owner, owner_off = home
        return self.mem.ptrs('mirror')[owner] + 2 * owner_off

    def watchdog_status(self):
        """Non-zero when a waiter of one of THIS rank's chunk locks gave up (10 s): 1 = a writer, 2 = a reader."""
        if self.mem is None:
            return 0
        from ..ops import native
        words = self.ctrl.view(MAX_CHUNKS, native.PS_LOCK_WORDS)[:, 3]
        return int(words.max().item())

    # ---------------------------------------------------------- flat, on device
    def pull_into(self, dst_flat, dst_bf16=None, f32_range=None, b16_range=None):
        """``dst_flat <- master`` and / or ``dst_bf16 <- bf16(master)`` (flat indexed; ranges default to everything;
        hogwild stores serve the bf16 copy from their mirror)."""
        if self.mem is not None:
            from ..ops import native
            native.ps_striped_pull(self.args, dst_flat, f32_range, dst_bf16, b16_range, seqlock=self.locked,
                                   device=self.device)
            return
        lo, hi = f32_range if f32_range is not None else (0, self.layout.numel)
        if self.locked:
            with self._host_lock.reading():
                dst_flat[lo:hi].copy_(self.flat[lo:hi])
        else:
            dst_flat[lo:hi].copy_(self.flat[lo:hi])
        if dst_bf16 is not None:
            lo, hi = b16_range if b16_range is not None else (0, self.layout.numel)
            dst_bf16[lo:hi].copy_(self.flat[lo:hi])

    def push_delta(self, delta_flat=None, before=None, after=None, delta_bf16=None):
        """``master -= delta`` where ``delta`` is given directly (fp32 or bf16) or as ``before - after`` (fused)."""
        self.pushes += 1
        if self.mem is not None:
            from ..ops import native
            native.ps_striped_push(self.args, delta_flat, delta_bf16, before, after, locked=self.locked, device=self.device)
            return
        if delta_bf16 is not None:
            delta_flat = delta_bf16.float()
        delta = delta_flat if delta_flat is not None else (before - after)
        if self.locked:
            with self._host_lock.writing():
                self.flat.sub_(delta)
        else:
            self.flat.sub_(delta)

    def optimizer_push(self, kind, g32, step, slots, advance=True, **hyper):
        """The optimizer step of one batch fused with its push: slots are updated from ``g32`` and the step of every
        element goes straight into its owner's master (``red.add`` over NVLink); no delta buffer."""
        from ..ops import native
        self.pushes += 1
        native.ps_striped_optimizer_push(self.args, kind, self.layout.numel, g32, step, slots=slots, locked=self.locked,
                                         advance=advance, **hyper)

    # ------------------------------------------------- list-of-arrays contract
    def pull_flat(self):
        tmp = torch.empty(self.layout.numel, dtype=torch.float32, device=self.device)
        self.pull_into(tmp)
        if tmp.is_cuda:
            torch.cuda.current_stream(tmp.device).synchronize()
            if self.watchdog_status():
                raise RuntimeError("device parameter server: a waiter of a chunk lock gave up after 10 s (a lock holder "
                                   "never released it); master weights may have been updated or read without exclusion")
        return tmp

    def pull_arrays(self):
        return self.layout.unpack_arrays(self.pull_flat())

    def push_arrays(self, delta):
        self.push_delta(self.layout.pack_arrays(delta, device=self.device))


def create_store(port, arrays, mode, model=None, device=None, rt=None, **_ignored):
    """Create the store of ``port``. Under ``torchrun`` (``rt.distributed``) this is collective: every rank calls it, each
    allocates its stripe, rank 0's ``arrays`` become the master."""
    from .runtime import default_device
    if model is None:
        raise ValueError("create_store needs the model whose weights it holds")
    store = DeviceStore(port, model, arrays, mode, device or default_device(), rt=rt)
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
