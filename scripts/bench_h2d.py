"""Ceiling of the host link: every rank copies a pinned 67 MB buffer host -> device back to back, all ranks at once.
    python -m torch.distributed.run --nproc-per-node N scripts/bench_h2d.py
Prints the per-GPU rate (min / mean over ranks): what an end-to-end step that streams its inputs can reach at best."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
import torch  # noqa: E402

from elephas_b200.parallel import get_runtime  # noqa: E402

rt = get_runtime()
for label, nbytes in (('fp32 batch 4096x4096', 4096 * 4096 * 4), ('bf16 batch 4096x4096', 4096 * 4096 * 2)):
    host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    dev = torch.empty(nbytes, dtype=torch.uint8, device=rt.device)
    for _ in range(3):
        dev.copy_(host, non_blocking=True)
    torch.cuda.synchronize()
    rt.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 30
    e0.record()
    for _ in range(reps):
        dev.copy_(host, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    rates = rt.all_gather_object(nbytes / (ms * 1e-3) / 1e9)
    rt.barrier()
    if rt.is_driver:
        print(json.dumps({'what': 'concurrent pinned H2D, ' + label, 'n_gpus': rt.world_size, 'ms_per_copy_rank0': round(ms, 3),
                          'gbs_per_gpu_min': round(min(rates), 2), 'gbs_per_gpu_mean': round(sum(rates) / len(rates), 2),
                          'samples_per_sec_ceiling_all_gpus': round(rt.world_size * 4096 / (nbytes / (min(rates) * 1e9)), 0)}),
              flush=True)
rt.shutdown()
