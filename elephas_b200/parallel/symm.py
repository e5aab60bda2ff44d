"""Symmetric (peer-mapped) device memory and the collectives built on it.

Every rank allocates the same set of communication buffers with ``cudaMalloc`` (through the native
library, so they are IPC-exportable), exports a CUDA IPC handle, and maps every peer's allocation
into its own address space (``cudaIpcOpenMemHandle`` enables NVLink peer access lazily). After
that a kernel on any GPU can load from / store to / ``red.add`` into any peer's buffer - the
substrate of the fused all-reduce + update (K6) and of the device-resident parameter server (K7).
``torch.distributed`` is only used to exchange the 64-byte handles (the reference's counterpart is
the Spark broadcast / collect plumbing, ``elephas/spark_model.py:207,220``).
"""
import torch

from ..ops import native

_ALIGN = 256


class SymmetricMemory:
    """``spec``: ordered ``{name: nbytes}``; the same spec must be used on every rank."""

    def __init__(self, rt, spec):
        self.rt = rt
        self.offsets = {}
        total = 0
        for name, nbytes in spec.items():
            self.offsets[name] = (total, int(nbytes))
            total += (int(nbytes) + _ALIGN - 1) // _ALIGN * _ALIGN
        self.total = max(total, _ALIGN)
        self.raw = native.RawBuffer(self.total, rt.device)
        self.base = [0] * rt.world_size
        self.base[rt.rank] = self.raw.ptr
        self._imported = []
        if rt.distributed:
            handles = rt.all_gather_object(native.ipc_export(self.raw.ptr))
            for r, h in enumerate(handles):
                if r != rt.rank:
                    with torch.cuda.device(rt.device):
                        p = native.ipc_import(h)
                    self.base[r] = p
                    self._imported.append(p)
            rt.barrier()

    def ptrs(self, name, extra_offset=0):
        off, _ = self.offsets[name]
        return [b + off + extra_offset for b in self.base]

    def local(self, name, dtype, numel=None):
        off, nbytes = self.offsets[name]
        itemsize = torch.empty((), dtype=dtype).element_size()
        return self.raw.tensor(dtype, numel if numel is not None else nbytes // itemsize, offset_bytes=off)

    def close(self):
        torch.cuda.synchronize(self.rt.device)
        if self.rt.distributed:
            self.rt.barrier()
        for p in self._imported:
            try:
                native.ipc_close(p)
            except native.NativeError:
                pass
        self._imported = []
        if self.rt.distributed:
            self.rt.barrier()
        self.raw.free()


class SumAllReduce:
    """One-shot P2P sum of fp32 buffers (generic delta lists, evaluate statistics)."""

    def __init__(self, rt, numel):
        self.rt = rt
        self.numel = (int(numel) + 7) // 8 * 8
        self.mem = SymmetricMemory(rt, {'d0': self.numel * 4, 'd1': self.numel * 4, 'pad': 256})
        self.local = torch.zeros(16, dtype=torch.int32, device=rt.device)
        self.out = torch.zeros(self.numel, dtype=torch.float32, device=rt.device)
        self.flip = 0

    def __call__(self, flat):
        name = 'd%d' % self.flip
        self.flip ^= 1
        buf = self.mem.local(name, torch.float32, self.numel)
        buf.zero_()
        buf[:flat.numel()].copy_(flat)
        native.allreduce_sum(self.mem.ptrs(name), self.mem.ptrs('pad'), self.local, self.numel, self.rt.rank,
                             self.rt.world_size, False, self.out, device=self.rt.device)
        return self.out[:flat.numel()].clone()


_sum_cache = {}


def all_reduce_sum(rt, flat):
    """Sum ``flat`` (fp32 CUDA tensor) over all ranks with the P2P kernel; falls back to NCCL for
    payloads where a bandwidth-optimal library collective is the better tool (> 64 MB)."""
    if not rt.distributed:
        return flat
    if flat.numel() * 4 > (64 << 20):
        return rt.all_reduce_sum_(flat)
    key = (rt.rank, rt.world_size)
    comm = _sum_cache.get(key)
    if comm is None or comm.numel < flat.numel():
        size = max(flat.numel(), 1 << 20)
        sizes = [int(v) for v in rt.all_gather_object(size)]
        comm = SumAllReduce(rt, max(sizes))
        _sum_cache[key] = comm
    return comm(flat.float())
