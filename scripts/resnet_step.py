"""A few ResNet-50 training steps (batch 64, 224x224, bf16) on one GPU, eagerly (no CUDA graph), for launch lists:
    python scripts/resnet_step.py [steps]       ELEPHAS_B200_NATIVE_CONV=0 for the cuDNN arm"""
import os
import sys
import time

os.environ.setdefault('ELEPHAS_B200_TORCH_GRAPHS', os.environ.get('RESNET_GRAPHS', '0'))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from elephas_b200.models import SGD, zoo  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
bs = int(os.environ.get('RESNET_BATCH', '64'))
rng = np.random.default_rng(0)
x = torch.from_numpy(rng.standard_normal((bs, 224, 224, 3), dtype=np.float32)).cuda()
y = torch.from_numpy(np.eye(1000, dtype=np.float32)[rng.integers(0, 1000, bs)]).cuda()
model = zoo.resnet50()
model.compile(SGD(learning_rate=0.01, momentum=0.9), 'categorical_crossentropy', ['acc'])
model.set_compute_dtype('bfloat16')
model.to('cuda')
for i in range(steps):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = model.train_on_batch(x, y)
    torch.cuda.synchronize()
    print('step %d: %.2f ms, loss %.4f' % (i, (time.perf_counter() - t0) * 1e3, out[0]), flush=True)
