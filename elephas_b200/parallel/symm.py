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


# ---------------------------------------------------------------------------------------------------------------------
# Symmetric memory on the VMM API with an NVSwitch multicast view (NVLS)
def exchange_fds(rt, fds, tag):
    """Every rank contributes a list of file descriptors and receives every other rank's list (``result[r]`` = duplicates
    of rank r's descriptors, valid in THIS process). Descriptors cannot travel through torch.distributed: they are sent
    as SCM_RIGHTS messages over abstract Unix sockets (one listener per rank, named after the rendezvous port, the tag
    and the rank). ``tag`` must be the same on all ranks and unique per exchange."""
    import os
    import socket
    import struct
    world, rank = rt.world_size, rt.rank
    out = [None] * world
    out[rank] = list(fds)
    if world == 1:
        return out
    base = '\0elephas_b200_%s_%s_' % (os.environ.get('MASTER_PORT', '0'), tag)
    server = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    server.bind(base + str(rank))
    server.listen(world)
    server.settimeout(120.0)  # a peer that died before it connected must not leave this rank in accept() for ever
    rt.barrier()  # every listener is up
    try:
        for peer in range(world):
            if peer == rank:
                continue
            with socket.socket(socket.AF_UNIX, socket.SOCK_STREAM) as c:
                c.connect(base + str(peer))
                socket.send_fds(c, [struct.pack('<ii', rank, len(fds))], list(fds))
        for _ in range(world - 1):
            conn, _ = server.accept()
            with conn:
                msg, got, _, _ = socket.recv_fds(conn, 8, 64)
                src, count = struct.unpack('<ii', msg)
                assert count == len(got), "rank %d sent %d descriptors, %d arrived" % (src, count, len(got))
                out[src] = list(got)
    finally:
        server.close()
    rt.barrier()
    return out


class VmmSymmetricMemory:
    """Same interface as :class:`SymmetricMemory` (``ptrs`` / ``local`` / ``close``), built on ``cuMemCreate`` allocations
    whose POSIX file descriptors are exchanged between the ranks, plus - where the GPUs sit behind an NVSwitch that
    supports it - ONE multicast object over all ranks' allocations: ``mc_ptr(name)`` is an address at which
    ``multimem.ld_reduce`` reads the SUM of every rank's buffer (added inside the switch) and ``multimem.st`` writes to
    every rank's buffer at once (the NVLS kernels of ``csrc/comm_kernels.cu``).

    Reference mechanism replaced: Spark broadcast of the weights and collect of the deltas through the driver
    (``elephas/spark_model.py:206-207,220-227``)."""

    _serial = [0]

    def __init__(self, rt, spec, multicast=True):
        self.rt = rt
        self.offsets = {}
        total = 0
        for name, nbytes in spec.items():
            self.offsets[name] = (total, int(nbytes))
            total += (int(nbytes) + _ALIGN - 1) // _ALIGN * _ALIGN
        world = rt.world_size
        VmmSymmetricMemory._serial[0] += 1
        tag = 'vmm%d' % VmmSymmetricMemory._serial[0]
        with torch.cuda.device(rt.device):
            vmm_ok, mc_ok = native.vmm_supported()
            if not vmm_ok:
                raise native.NativeError("this device cannot export VMM allocations as file descriptors")
            self.multicast = bool(multicast and mc_ok and world > 1)
            flags = rt.all_gather_object(self.multicast)
            self.multicast = all(flags)
            self.handle, self.total = native.vmm_create(max(total, _ALIGN), world if self.multicast else 0)
            sizes = rt.all_gather_object(self.total)
            assert len(set(sizes)) == 1, "ranks disagree on the allocation size: %r" % (sizes,)
            self.base = [0] * world
            self.base[rt.rank] = native.vmm_map(self.handle, self.total)
            native.tensor_from_ptr(self.base[rt.rank], torch.uint8, self.total, rt.device).zero_()
            torch.cuda.synchronize(rt.device)
            my_fds = [native.vmm_export_fd(self.handle)]
            self.mc_handle, self.mc_base = None, 0
            if self.multicast and rt.rank == 0:
                self.mc_handle = native.mc_create(world, self.total)
                my_fds.append(native.vmm_export_fd(self.mc_handle))
            import os
            all_fds = exchange_fds(rt, my_fds, tag)
            self._peer_handles = {}
            for r in range(world):
                if r == rt.rank:
                    continue
                h = native.vmm_import_fd(all_fds[r][0])
                self._peer_handles[r] = h
                self.base[r] = native.vmm_map(h, self.total)
            if self.multicast:
                if rt.rank != 0:
                    self.mc_handle = native.vmm_import_fd(all_fds[0][1])
                native.mc_add_device(self.mc_handle)
                rt.barrier()  # every device has joined before anybody binds memory
                native.mc_bind(self.mc_handle, self.handle, self.total)
                rt.barrier()
                self.mc_base = native.vmm_map(self.mc_handle, self.total)
            for r in range(world):  # the descriptors have done their job
                for fd in (all_fds[r] if r != rt.rank else my_fds):
                    try:
                        os.close(fd)
                    except OSError:
                        pass
            rt.barrier()

    def ptrs(self, name, extra_offset=0):
        off, _ = self.offsets[name]
        return [b + off + extra_offset for b in self.base]

    def mc_ptr(self, name, extra_offset=0):
        """Multicast address of buffer ``name`` (0 when the node has no multicast support)."""
        return self.mc_base + self.offsets[name][0] + extra_offset if self.mc_base else 0

    def local(self, name, dtype, numel=None):
        off, nbytes = self.offsets[name]
        itemsize = torch.empty((), dtype=dtype).element_size()
        return native.tensor_from_ptr(self.base[self.rt.rank] + off, dtype,
                                      numel if numel is not None else nbytes // itemsize, self.rt.device)

    def close(self):
        rt = self.rt
        torch.cuda.synchronize(rt.device)
        rt.barrier()
        with torch.cuda.device(rt.device):
            try:
                if self.mc_base:
                    native.vmm_unmap(self.mc_base, self.total)
                for r, b in enumerate(self.base):
                    if b:
                        native.vmm_unmap(b, self.total)
                for h in self._peer_handles.values():
                    native.vmm_release(h)
                if self.mc_handle is not None:
                    native.vmm_release(self.mc_handle)
                native.vmm_release(self.handle)
            except native.NativeError:
                pass
        self.mc_base, self.base, self._peer_handles, self.mc_handle = 0, [0] * rt.world_size, {}, None
        rt.barrier()
