"""The kernels added after the first tour, once each at representative shapes (ncu target): the one-kernel MLP forward
(K8), the pull-fused GEMM with its TMA side-write (K2), the mirror-refreshing parameter-server push and the narrow-head
kernels. One GPU."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from elephas_b200.ops import native  # noqa: E402

dev = torch.device('cuda', 0)
torch.cuda.set_device(dev)
torch.manual_seed(0)


def bf(t):
    return t.to(torch.bfloat16)


# K8: MNIST MLP 784-128-128-10 on 16384 rows
rows = 16384
x = bf(torch.randn(rows, 784, device=dev))
layers, k = [], 784
for units in (128, 128, 10):
    layers.append((bf(torch.randn(units, k, device=dev) / k ** 0.5), units, torch.zeros(units, device=dev), native.ACT_IDS['relu']))
    k = units
z = torch.empty(rows, 12, device=dev)
native.mlp_forward(x, rows, 784, layers, z)
# K2: pull-fused GEMM 4096^3 (B read from a "server" buffer, side-written to the local copy)
a = bf(torch.randn(4096, 4096, device=dev))
server = bf(torch.randn(4096, 4096, device=dev) / 64)
local = torch.zeros_like(server)
out = torch.empty(4096, 4096, dtype=torch.bfloat16, device=dev)
native.gemm(a, server.data_ptr(), 4096, 4096, 4096, ldb=4096, out_bf16=out, b_copy=local)
# mirror-refreshing push, 16 M parameters, bf16 delta
n = 16 << 20
master = torch.zeros(n, device=dev)
mirror = torch.zeros(n, dtype=torch.bfloat16, device=dev)
ctrl = torch.zeros(16, dtype=torch.int32, device=dev)
native.ps_push(master, ctrl, delta_bf16=bf(torch.randn(n, device=dev)), mirror=mirror)
# narrow head: 4096 x 4096 activations, one output unit
w1 = bf(torch.randn(1, 4096, device=dev) / 64)
z1 = torch.empty(4096, 4, device=dev)
native.head_fwd(a, w1, 4096, 4096, 1, z1)
dz = bf(torch.randn(4096, 8, device=dev))
dw = torch.empty(1, 4096, device=dev)
native.head_wgrad(dz, a, 4096, 4096, 1, dw)
dzp = torch.empty(4096, 4096, dtype=torch.bfloat16, device=dev)
native.head_dgrad(dz, w1, a, dzp, 4096, 4096, 1, act=native.ACT_IDS['relu'])
torch.cuda.synchronize()
assert torch.equal(local, server)
print('tour2 ok')
