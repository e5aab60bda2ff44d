"""Every non-GEMM kernel of the library once at the wide-MLP shapes (B=4096, 4096 features, 134 M parameters):
device time (CUDA events, after warm-up) against the bytes it has to move, i.e. its fraction of the measured HBM
copy bandwidth. Also the target of the ncu captures (scripts/gpu_round3.sh). One GPU, world=1 for collectives."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from elephas_b200.ops import native  # noqa: E402

dev = torch.device('cuda', 0)
torch.cuda.set_device(dev)
ONCE = os.environ.get('TOUR_ONCE') == '1'  # profiler mode: one launch per kernel, smaller parameter vector
B, D, P = 4096, 4096, (16 << 20) if ONCE else 134254592 // 64 * 64
peaks = {}
try:
    peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'MEASURED_PEAKS.json')))
except Exception:
    pass
HBM = float(peaks.get('hbm_gbs', 6457.1))
rows = []


def timed(name, nbytes, fn, reps=10):
    if ONCE:
        fn()
        torch.cuda.synchronize()
        return
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        flush.add_(1.0)  # > L2: the next launch reads its operands from HBM
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    gbs = nbytes / best / 1e6
    rows.append({'kernel': name, 'us': round(best * 1e3, 2), 'MB': round(nbytes / 1e6, 1), 'GB/s': round(gbs, 1),
                 'frac_of_hbm_copy': round(gbs / HBM, 3)})
    print('%-34s %9.2f us %9.1f MB %9.1f GB/s  %5.1f%% of measured copy bandwidth' % (name, best * 1e3, nbytes / 1e6, gbs,
                                                                                   100 * gbs / HBM), flush=True)


flush = torch.zeros(80 << 20, device=dev)  # 320 MB
x32 = torch.randn(B, D, device=dev)
x16 = torch.empty(B, D, dtype=torch.bfloat16, device=dev)
y16 = torch.empty_like(x16)
z16 = torch.empty_like(x16)
step = torch.zeros(1, dtype=torch.int64, device=dev)
timed('stage_f32_to_bf16', B * D * 6, lambda: native.stage_f32_to_bf16(x32, x16, B, D))
w32 = torch.randn(P, device=dev) * 0.01
w16 = torch.empty(P, dtype=torch.bfloat16, device=dev)
timed('cast_f32_to_bf16 (134M)', P * 6, lambda: native.cast_f32_to_bf16(w32, w16))
bias = torch.zeros(D, device=dev)
timed('colsum_bf16', B * D * 2, lambda: native.colsum_bf16(x16, B, D, bias))
timed('act_dropout_fwd relu p=.2', B * D * 4, lambda: native.act_dropout_fwd(x16, y16, B, D, act=1, drop_p=0.2, seed=1,
                                                                             step_ptr=step))
timed('act_dropout_bwd relu p=.2', B * D * 6, lambda: native.act_dropout_bwd(x16, y16, z16, B, D, act=1, drop_p=0.2,
                                                                             seed=1, step_ptr=step))
after = w32 + 0.001
d32 = torch.empty(P, device=dev)
timed('delta f32 (before-after)', P * 12, lambda: native.delta(w32, after, out_f32=d32))
# loss: softmax + categorical CE (10 classes, MNIST batch) and mse on the regression head
R, C = 65536, 10
zl = torch.randn(R, 12, device=dev)
yl = torch.zeros(R, C, device=dev)
yl[:, 3] = 1.0
dzl = torch.empty(R, 16, dtype=torch.bfloat16, device=dev)
outl = torch.empty(R, C, device=dev)
stats = torch.zeros(8, dtype=torch.float64, device=dev)
timed('loss softmax+CE 65536x10', R * (C * 4 * 3 + C * 2),
      lambda: native.loss_fwd_bwd(zl, yl, R, C, native.LOSS_IDS['categorical_crossentropy'], native.ACT_IDS['softmax'],
                                  out=outl, dz=dzl, stats=stats, metrics=(1, 0, 0, 0), grad_scale=1.0 / R))
g32 = torch.randn(P, device=dev) * 1e-3
m = torch.zeros(P, device=dev)
v = torch.zeros(P, device=dev)
timed('optimizer SGD apply', P * (4 + 4 + 4 + 2),
      lambda: native.optimizer_step(native.OPT_IDS['SGD'], P, g32, step, w32=w32, w16=w16, lr=1e-3, advance=False))
timed('optimizer SGD+momentum', P * (4 + 8 + 8 + 2),
      lambda: native.optimizer_step(native.OPT_IDS['SGD'], P, g32, step, w32=w32, w16=w16, slots=(m,), lr=1e-3,
                                    momentum=0.9, advance=False))
timed('optimizer Adam', P * (4 + 8 + 8 + 8 + 2),
      lambda: native.optimizer_step(native.OPT_IDS['Adam'], P, g32, step, w32=w32, w16=w16, slots=(m, v), lr=1e-3,
                                    advance=False))
dl16 = torch.empty(P, dtype=torch.bfloat16, device=dev)
timed('optimizer SGD delta-mode (bf16)', P * (4 + 2),
      lambda: native.optimizer_step(native.OPT_IDS['SGD'], P, g32, step, w32=w32, w16=w16, delta_bf16=dl16, apply=False,
                                    lr=1e-3, advance=False))
# parameter server: push (red.add into the master) and pull (+ bf16 cast), same GPU = HBM roofline
master = torch.zeros(P, device=dev)
ctrl = torch.zeros(16, dtype=torch.int32, device=dev)
timed('ps_push delta (red.add.f32 v4)', P * 12, lambda: native.ps_push(master, ctrl, delta=g32))
timed('ps_push before-after (locked)', P * 16, lambda: native.ps_push(master, ctrl, before=w32, after=after, locked=True))
timed('ps_pull f32 + bf16 cast', P * 10, lambda: native.ps_pull(master, ctrl, d32, dst_bf16=w16))
# fused all-reduce + update, world = 1 (own buffer): the HBM side of the multi-GPU kernels
pad = torch.zeros(64, dtype=torch.int32, device=dev)
local = torch.zeros(16, dtype=torch.int32, device=dev)
n1 = min(P, 4096 * 4096)
timed('allreduce_update one-shot bf16 W=1', n1 * (2 + 8 + 2),
      lambda: native.allreduce_update([dl16.data_ptr()], [pad.data_ptr()], local, n1, 0, 1, 1.0, True, w32, w16=w16,
                                      avg_ptrs=[z16.data_ptr()], algo=1, device=dev))
timed('allreduce_shard_update bf16 W=1', n1 * (2 + 8 + 2),
      lambda: native.allreduce_shard_update([dl16.data_ptr()], [pad.data_ptr()], local, n1, 0, 1, 1.0, True, w32,
                                            [w16.data_ptr()], device=dev))
os.makedirs('gpurun_out', exist_ok=True)
json.dump({'hbm_copy_gbs': HBM, 'rows': rows}, open('gpurun_out/kernel_tour.json', 'w'), indent=1)
