"""Rank runtime: who am I, which device do I own, how do ranks talk.

The reference distributes work with Spark executors - one Python worker process per RDD partition
(``elephas/spark_model.py:214,220``) - and moves weights through the driver. Here a *rank* is one
process that owns one B200 (or, on a machine without GPUs, one CPU process / thread):

* launched by ``torchrun`` (``RANK`` / ``WORLD_SIZE`` / ``LOCAL_RANK`` in the environment): SPMD,
  every rank executes the user's script, ``SparkModel`` hands partition *i* to rank ``i % world``;
  ``torch.distributed`` (NCCL on GPUs, gloo on CPUs) is used for rendezvous, object exchange and as
  the collective fallback, the hot collectives are the P2P kernels of ``parallel/collectives.py``;
* plain ``python script.py``: one rank, partitions run as worker threads on the one device.
"""
import datetime
import os
import threading

import torch

_lock = threading.Lock()
_runtime = None


def default_device():
    """``ELEPHAS_B200_DEVICE`` wins; otherwise ``cuda:LOCAL_RANK`` when CUDA is present, else the CPU."""
    forced = os.environ.get('ELEPHAS_B200_DEVICE')
    if forced:
        return torch.device(forced)
    if torch.cuda.is_available():
        return torch.device('cuda', int(os.environ.get('LOCAL_RANK', 0)) % torch.cuda.device_count())
    return torch.device('cpu')


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(','):
        if not part:
            continue
        lo, _, hi = part.partition('-')
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def bind_to_gpu_numa_node(device):
    """Pin this rank's process to the CPUs of the NUMA node its GPU hangs off (sysfs: PCI device -> numa_node ->
    cpulist). Pinned host buffers are then first-touched on that node, so the per-step host->device copies do not
    cross the socket interconnect - with eight ranks streaming inputs at once that link is the first to saturate.
    Returns the node, or None when there is nothing to do (one node, no sysfs, ``ELEPHAS_B200_NUMA_BIND=0``)."""
    if os.environ.get('ELEPHAS_B200_NUMA_BIND', '1') == '0' or not hasattr(os, 'sched_setaffinity'):
        return None
    try:
        props = torch.cuda.get_device_properties(device)
        bdf = '%04x:%02x:%02x.0' % (props.pci_domain_id, props.pci_bus_id, props.pci_device_id)
        with open('/sys/bus/pci/devices/%s/numa_node' % bdf) as fh:
            node = int(fh.read().strip())
        if node < 0:
            return None
        with open('/sys/devices/system/node/node%d/cpulist' % node) as fh:
            cpus = _parse_cpulist(fh.read())
        allowed = os.sched_getaffinity(0)
        cpus &= allowed
        if not cpus or cpus == allowed:
            return node if cpus else None
        os.sched_setaffinity(0, cpus)
        return node
    except (OSError, ValueError, AttributeError, RuntimeError):
        return None


class Runtime:
    def __init__(self):
        self.device = default_device()
        self.world_size = int(os.environ.get('WORLD_SIZE', 1))
        self.rank = int(os.environ.get('RANK', 0))
        self.local_rank = int(os.environ.get('LOCAL_RANK', 0))
        self._group = None
        self.numa_node = None
        if self.device.type == 'cuda':
            torch.cuda.set_device(self.device)
            if self.world_size > 1:
                self.numa_node = bind_to_gpu_numa_node(self.device)
        if self.world_size > 1:
            self._init_process_group()

    def _init_process_group(self):
        import torch.distributed as dist
        if not dist.is_initialized():
            backend = 'nccl' if self.device.type == 'cuda' else 'gloo'
            os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
            os.environ.setdefault('MASTER_PORT', '29500')
            kwargs = {}
            if backend == 'nccl':
                kwargs['device_id'] = self.device
            dist.init_process_group(backend=backend, rank=self.rank, world_size=self.world_size,
                                    timeout=datetime.timedelta(seconds=600), **kwargs)
        self._dist = dist

    @property
    def distributed(self):
        return self.world_size > 1

    @property
    def is_driver(self):
        return self.rank == 0

    # ------------------------------------------------------------ plumbing
    def barrier(self):
        if self.distributed:
            if self.device.type == 'cuda':
                self._dist.barrier(device_ids=[self.device.index])
            else:
                self._dist.barrier()

    def all_gather_object(self, obj):
        if not self.distributed:
            return [obj]
        out = [None] * self.world_size
        self._dist.all_gather_object(out, obj)
        return out

    def all_true(self, flag):
        """``all(flag over ranks)`` with one small all-reduce (no pickling, one device->host read)."""
        if not self.distributed:
            return bool(flag)
        import torch
        t = torch.tensor([1 if flag else 0], dtype=torch.int32, device=self.device)
        self._dist.all_reduce(t, op=self._dist.ReduceOp.MIN)
        return bool(int(t.item()))

    def all_gather_rows(self, row):
        """``row`` (1-D tensor of the same length on every rank) from all ranks as ``[world, len]`` on the host."""
        import torch
        if not self.distributed:
            return row.detach().cpu().unsqueeze(0)
        src = row.to(self.device).contiguous().view(-1)
        out = torch.empty(self.world_size * src.numel(), dtype=src.dtype, device=self.device)  # (flat: every backend takes it)
        self._dist.all_gather_into_tensor(out, src)
        return out.view(self.world_size, -1).cpu()

    def broadcast_object(self, obj, src=0):
        if not self.distributed:
            return obj
        box = [obj]
        self._dist.broadcast_object_list(box, src=src)
        return box[0]

    def all_reduce_sum_(self, tensor):
        """In-place sum over ranks (library fallback; the fused kernels live in ``collectives.py``)."""
        if self.distributed:
            self._dist.all_reduce(tensor, op=self._dist.ReduceOp.SUM)
        return tensor

    def all_reduce_max_(self, tensor):
        if self.distributed:
            self._dist.all_reduce(tensor, op=self._dist.ReduceOp.MAX)
        return tensor

    def broadcast_(self, tensor, src=0):
        if self.distributed:
            self._dist.broadcast(tensor, src=src)
        return tensor

    def shutdown(self):
        if self.distributed and self._dist.is_initialized():
            self._dist.destroy_process_group()


def get_runtime() -> Runtime:
    global _runtime
    with _lock:
        if _runtime is None:
            _runtime = Runtime()
        return _runtime


def reset_runtime():
    """Drop the cached runtime (tests that change WORLD_SIZE / device between cases)."""
    global _runtime
    with _lock:
        if _runtime is not None:
            try:
                _runtime.shutdown()
            except Exception:
                pass
        _runtime = None
