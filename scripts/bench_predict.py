"""Forward pass of the reference's MNIST MLP (784-128-128-10; MODEL=otto: the Otto MLP 93-512-512-512-9 on the streamed-
weights variant): the one-kernel K8 forward against the per-layer K1 GEMM chain, device time (CUDA events, best of 10) for a
block of rows already on the device, plus Model.predict end to end."""
import json
import os
import sys
import time

os.environ.setdefault('ELEPHAS_B200_FUSED_MLP_WIDE', '1')

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from elephas_b200.models import SGD, zoo  # noqa: E402
from elephas_b200.ops import engine as engine_mod  # noqa: E402
from elephas_b200.ops import native  # noqa: E402

rows = int(os.environ.get('ROWS', 65536))
torch.manual_seed(0)
which = os.environ.get('MODEL', 'mnist')
m = zoo.otto_mlp() if which == 'otto' else zoo.mnist_mlp()
in_dim = 93 if which == 'otto' else 784
m.compile(SGD(learning_rate=0.05), 'categorical_crossentropy', ['acc'])
m.to('cuda')
eng = m._native_engine()
x = np.random.default_rng(0).random((rows, in_dim), dtype=np.float32)
res = {}
flush = torch.zeros(64 << 20, device='cuda')
for fused in (True, False):
    engine_mod._FUSED_MLP = fused
    eng._fused_mlp = None
    eng.invalidate_graphs()
    buf = eng._buf(min(rows, 8192))
    buf.x32.copy_(torch.from_numpy(x[:buf.B]).cuda())
    for _ in range(3):
        eng._forward(buf, training=False)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(10):
        flush.add_(1.0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        eng._forward(buf, training=False)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    m.predict(x[:1024])
    t0 = time.perf_counter()
    out = m.predict(x)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    res['fused' if fused else 'gemm_chain'] = {'forward_us_per_%d_rows' % buf.B: round(best * 1e3, 1),
                                               'predict_rows_per_s': round(rows / dt)}
    res.setdefault('check', []).append(float(out.sum()))
res['model'] = which
print(json.dumps(res))
os.makedirs('gpurun_out', exist_ok=True)
json.dump(res, open('gpurun_out/bench_predict_%s.json' % which, 'w'))
