"""Standalone timing of the fused all-reduce + update kernels on the headline payload (one rank per GPU, torchrun):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 scripts/bench_collective.py

Payload = the wide MLP's 134,254,592 parameters as bf16 deltas (268.5 MB). Per call every rank reduces its 1/N slice
of all peers' deltas over NVLink, updates its fp32 master slice and stores the refreshed bf16 weights into every
peer's compute copy (``allreduce_shard_update``, K6b; ``nvls_update`` = the multimem variant K6c) - nothing else runs on the
GPU. ``nccl_all_reduce_only`` times torch.distributed's all-reduce of the same payload (no master update, no bf16 refresh). CUDA events around K calls, max
over ranks. Bytes that must cross NVLink per rank: (N-1)/N x payload inbound (delta loads) and the same outbound (bf16
weight stores); the two directions are independent links. Reported against the measured peer-copy rate of this pool
(770 GB/s per direction per GPU, B200_PROFILING.md) and the nominal 900 GB/s. Prints one JSON line per variant."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
import torch  # noqa: E402

from elephas_b200.ops import native  # noqa: E402
from elephas_b200.parallel import get_runtime  # noqa: E402
from elephas_b200.parallel.symm import SymmetricMemory, VmmSymmetricMemory  # noqa: E402

N_ELEMS = int(os.environ.get('COLL_ELEMS', 134254592))
STEPS, WARM = int(os.environ.get('COLL_STEPS', 20)), 3
MEASURED_GBS, NOMINAL_GBS = 770.0, 900.0


def main():
    rt = get_runtime()
    W, R = rt.world_size, rt.rank
    n = N_ELEMS // 64 * 64
    mem = SymmetricMemory(rt, {'delta0': n * 2, 'delta1': n * 2, 'w16': n * 2, 'avg': n * 2, 'pad': 256})
    local = torch.zeros(16, dtype=torch.int32, device=rt.device)
    w32 = torch.zeros(n, dtype=torch.float32, device=rt.device)
    for name in ('delta0', 'delta1'):
        mem.local(name, torch.bfloat16, n).normal_()
    variants = [('shard_update', blocks) for blocks in (16, 32, 74, 148)] + [('twoshot_update', 148)]
    if W > 1:
        variants.append(('nccl_all_reduce_only', 0))  # the library's all-reduce of the same bf16 payload, WITHOUT the update
        nccl_buf = torch.randn(n, device=rt.device).to(torch.bfloat16)
    vm = None
    if os.environ.get('COLL_NVLS', '1') != '0' and W > 1:
        try:
            vm = VmmSymmetricMemory(rt, {'delta0': n * 2, 'delta1': n * 2, 'w16': n * 2, 'pad': 256}, multicast=True)
            if not vm.multicast:
                vm.close()
                vm = None
        except native.NativeError as exc:
            if rt.is_driver:
                print('no NVLS arm: %s' % exc, file=sys.stderr)
        if not all(rt.all_gather_object(vm is not None)):
            if vm is not None:
                vm.close()
            vm = None
    if vm is not None:
        for name in ('delta0', 'delta1'):
            vm.local(name, torch.bfloat16, n).normal_()
        vlocal = torch.zeros(16, dtype=torch.int32, device=rt.device)
        variants += [('nvls_update', blocks) for blocks in (8, 16, 32, 74)]
    for kind, blocks in variants:
        def call(flip):
            common = (mem.ptrs('delta%d' % flip), mem.ptrs('pad'), local, n, R, W, 1.0 / W, True, w32)
            if kind == 'nccl_all_reduce_only':
                rt._dist.all_reduce(nccl_buf)
            elif kind == 'nvls_update':
                native.allreduce_nvls_update(vm.mc_ptr('delta%d' % flip), vm.mc_ptr('w16'), vm.ptrs('pad'), vlocal, n, R, W,
                                             1.0 / W, w32, blocks=blocks, device=rt.device)
            elif kind == 'shard_update':
                native.allreduce_shard_update(*common, mem.ptrs('w16'), blocks=blocks, device=rt.device)
            else:
                native.allreduce_update(*common, w16=mem.local('w16', torch.bfloat16, n), avg_ptrs=mem.ptrs('avg'), algo=2,
                                        blocks=blocks, device=rt.device)
        torch.cuda.synchronize()
        rt.barrier()
        for i in range(WARM):
            call(i & 1)
        torch.cuda.synchronize()
        rt.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(STEPS):
            call((WARM + i) & 1)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / STEPS], device=rt.device, dtype=torch.float64)
        rt.all_reduce_max_(t)
        ms = float(t[0])
        one_way = (W - 1) / W * n * 2  # bytes in (and, independently, out) per rank
        bound_us = one_way / (MEASURED_GBS * 1e9) * 1e6
        if rt.is_driver:
            print(json.dumps({'kernel': kind, 'blocks': blocks, 'n_gpus': W, 'payload_mb': round(n * 2 / 1e6, 1),
                              'us_per_call': round(ms * 1e3, 1), 'nvlink_gbs_per_direction': round(one_way / (ms * 1e-3) / 1e9, 1),
                              'fraction_of_measured_770': round(one_way / (ms * 1e-3) / 1e9 / MEASURED_GBS, 3),
                              'fraction_of_nominal_900': round(one_way / (ms * 1e-3) / 1e9 / NOMINAL_GBS, 3),
                              'bound_us_at_770': round(bound_us, 1),
                              'serial_both_directions_bound_us_at_900': round(2 * one_way / (NOMINAL_GBS * 1e9) * 1e6, 1)}),
                  flush=True)
    status = int(local[8].item())
    assert status == 0, "watchdog fired: peer %d" % (status - 1)
    if vm is not None:
        vm.close()
    mem.close()
    rt.shutdown()


if __name__ == '__main__':
    main()
