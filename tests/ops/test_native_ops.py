"""Numerics of every hand-written sm_100a kernel against a plain PyTorch fp32 reference of the same op."""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from elephas_b200.models import optimizers  # noqa: E402


@pytest.fixture(scope='module')
def native():
    from elephas_b200.ops import native as n
    n.require()
    return n


def _bf(t):
    return t.to(torch.bfloat16)


def _pad(t, ld):
    out = torch.zeros(t.shape[0], ld, dtype=t.dtype, device=t.device)
    out[:, :t.shape[1]] = t
    return out


ACT = {0: lambda x: x, 1: torch.relu, 2: torch.sigmoid, 3: torch.tanh, 4: torch.nn.functional.softplus,
       5: torch.nn.functional.softsign, 6: torch.nn.functional.elu, 7: torch.nn.functional.selu, 8: torch.exp,
       9: lambda x: torch.clamp(0.2 * x + 0.5, 0, 1)}


@pytest.mark.parametrize('M,N,K', [(64, 128, 784), (64, 10, 128), (128, 512, 93), (300, 200, 520), (4096, 1024, 512)])
def test_dense_forward(native, M, N, K):
    torch.manual_seed(0)
    dev = 'cuda'
    x, w, b = torch.randn(M, K, device=dev), torch.randn(N, K, device=dev) / math.sqrt(K), torch.randn(N, device=dev)
    ldk, ldn = (K + 7) // 8 * 8, (N + 7) // 8 * 8
    x16, w16 = _pad(_bf(x), ldk), _pad(_bf(w), ldk)
    out = torch.zeros(M, ldn, dtype=torch.bfloat16, device=dev)
    native.gemm(x16, w16, M, N, K, out_bf16=out, bias=b, act=1)
    ref = torch.relu(x16[:, :K].float() @ w16[:, :K].float().t() + b)
    assert torch.allclose(out[:, :N].float(), ref, atol=2e-2, rtol=2e-2)
    assert torch.all(out[:, N:] == 0)


@pytest.mark.parametrize('act', sorted(ACT))
def test_activations_in_epilogue(native, act):
    torch.manual_seed(act)
    M, N, K = 128, 64, 64
    x, w = torch.randn(M, K, device='cuda') * 0.5, torch.randn(N, K, device='cuda') * 0.2
    x16, w16 = _bf(x), _bf(w)
    z32 = torch.zeros(M, N, device='cuda')
    native.gemm(x16, w16, M, N, K, out_f32=z32, act=act)
    ref = ACT[act](x16.float() @ w16.float().t())
    assert torch.allclose(z32, ref, atol=1e-4, rtol=1e-4)


def test_dgrad_and_wgrad_match_autograd(native):
    torch.manual_seed(1)
    B, IN, OUT = 256, 200, 136
    ld_in, ld_out = 200, 136
    a_prev = torch.rand(B, IN, device='cuda')          # stored activation (relu output, > 0 mostly)
    a_prev[a_prev < 0.3] = 0.0
    w = torch.randn(OUT, IN, device='cuda') / math.sqrt(IN)
    dz = torch.randn(B, OUT, device='cuda') * 0.1
    a16, w16, dz16 = _bf(a_prev), _bf(w), _bf(dz)
    # wgrad: dW = dZ^T A_prev (fp32 out)
    dw = torch.zeros(OUT, ld_in, device='cuda')
    native.gemm(dz16, a16, OUT, IN, B, a_mn=True, b_mn=True, out_f32=dw)
    assert torch.allclose(dw, dz16.float().t() @ a16.float(), atol=1e-3, rtol=1e-3)
    # dgrad with relu derivative taken from the stored output
    dprev = torch.zeros(B, ld_in, dtype=torch.bfloat16, device='cuda')
    native.gemm(dz16, w16, B, IN, OUT, b_mn=True, out_bf16=dprev, epi=native.EPI_ACTGRAD, act=1, aux=a16)
    ref = (dz16.float() @ w16.float()) * (a16.float() > 0)
    assert torch.allclose(dprev.float(), ref, atol=2e-2, rtol=2e-2)
    # bias gradient
    db = torch.zeros(OUT, device='cuda')
    native.colsum_bf16(dz16, B, OUT, db)
    assert torch.allclose(db, dz16.float().sum(0), atol=1e-3, rtol=1e-3)


def test_dropout_epilogue_statistics_and_backward_mask(native):
    torch.manual_seed(2)
    M, N, K, p = 512, 256, 64, 0.2
    x, w = torch.rand(M, K, device='cuda') + 0.1, torch.rand(N, K, device='cuda') + 0.1
    out = torch.zeros(M, N, dtype=torch.bfloat16, device='cuda')
    nodrop = torch.zeros_like(out)
    native.gemm(_bf(x), _bf(w), M, N, K, out_bf16=nodrop, act=1)
    native.gemm(_bf(x), _bf(w), M, N, K, out_bf16=out, act=1, drop_p=p, seed=123, drop_stream=5)
    kept = out != 0
    assert abs(kept.float().mean().item() - (1 - p)) < 0.01
    assert torch.allclose(out[kept].float(), nodrop[kept].float() / (1 - p), rtol=2e-2)
    # same seed/stream reproduces the mask, another stream does not
    again = torch.zeros_like(out)
    native.gemm(_bf(x), _bf(w), M, N, K, out_bf16=again, act=1, drop_p=p, seed=123, drop_stream=5)
    assert torch.equal(again, out)
    other = torch.zeros_like(out)
    native.gemm(_bf(x), _bf(w), M, N, K, out_bf16=other, act=1, drop_p=p, seed=123, drop_stream=6)
    assert not torch.equal(other != 0, kept)
    # backward regenerates the same mask: gradient is zero exactly where the forward output was dropped
    dz = torch.ones(M, 64, dtype=torch.bfloat16, device='cuda')
    w2 = _bf(torch.ones(64, N, device='cuda'))
    dprev = torch.zeros(M, N, dtype=torch.bfloat16, device='cuda')
    native.gemm(dz, w2, M, N, 64, b_mn=True, out_bf16=dprev, epi=native.EPI_ACTGRAD, act=1, aux=out, drop_p=p,
                seed=123, drop_stream=5)
    assert torch.equal(dprev != 0, kept)
    assert torch.allclose(dprev[kept].float(), torch.full_like(dprev[kept].float(), 64 / (1 - p)), rtol=1e-2)


def test_stage_cast_delta_actdrop(native):
    torch.manual_seed(3)
    x = torch.randn(37, 93, device='cuda')
    dst = torch.full((37, 96), 7.0, dtype=torch.bfloat16, device='cuda')
    native.stage_f32_to_bf16(x, dst, 37, 93)
    assert torch.equal(dst[:, :93], _bf(x)) and torch.all(dst[:, 93:] == 0)
    flat = torch.randn(100003 // 8 * 8 + 8, device='cuda')
    out = torch.zeros_like(flat, dtype=torch.bfloat16)
    native.cast_f32_to_bf16(flat, out)
    assert torch.equal(out, _bf(flat))
    a, b = torch.randn(4099, device='cuda'), torch.randn(4099, device='cuda')
    d32, d16 = torch.zeros(4099, device='cuda'), torch.zeros(4099, dtype=torch.bfloat16, device='cuda')
    native.delta(a, b, out_f32=d32, out_bf16=d16)
    assert torch.equal(d32, a - b) and torch.equal(d16, _bf(a - b))
    h = _bf(torch.randn(64, 40, device='cuda'))
    y = torch.zeros_like(h)
    native.act_dropout_fwd(h, y, 64, 40, act=3, drop_p=0.5, seed=9, drop_stream=1)
    keep = y != 0
    assert 0.4 < keep.float().mean().item() < 0.6
    assert torch.allclose(y[keep].float(), 2 * torch.tanh(h.float())[keep], atol=2e-2, rtol=2e-2)
    dy, dx = _bf(torch.ones(64, 40, device='cuda')), torch.zeros_like(h)
    native.act_dropout_bwd(dy, y, dx, 64, 40, act=3, drop_p=0.5, seed=9, drop_stream=1)
    ref = torch.where(keep, 2 * (1 - (y.float() * 0.5) ** 2), torch.zeros_like(y.float()))
    assert torch.allclose(dx.float(), ref, atol=2e-2, rtol=2e-2)


def test_vector_and_scalar_paths_of_elementwise_kernels(native):
    """Every elementwise kernel has a 16-byte vector path (cols, ld multiples of 8) and a scalar one."""
    torch.manual_seed(11)
    for rows, cols, ld in ((64, 256, 264), (5, 4096, 4096), (300, 784, 784), (33, 93, 96), (7, 12, 16)):
        x = torch.randn(rows, cols, device='cuda')
        dst = torch.full((rows, ld), 7.0, dtype=torch.bfloat16, device='cuda')
        native.stage_f32_to_bf16(x, dst, rows, cols)
        assert torch.equal(dst[:, :cols], _bf(x)) and torch.all(dst[:, cols:] == 0), (rows, cols, ld)
        h = _bf(torch.randn(rows, ld, device='cuda'))
        out = torch.full((cols,), 3.0, device='cuda')
        native.colsum_bf16(h, rows, cols, out, ld=ld)
        assert torch.allclose(out, h[:, :cols].float().sum(0), atol=1e-3, rtol=1e-4), (rows, cols, ld)
        native.colsum_bf16(h, rows, cols, out, accumulate=True, ld=ld)
        assert torch.allclose(out, 2 * h[:, :cols].float().sum(0), atol=2e-3, rtol=1e-4)
        # activation + dropout: forward statistics, backward mask identical to the forward one
        y = torch.zeros_like(h)
        native.act_dropout_fwd(h, y, rows, cols, act=2, drop_p=0.25, seed=5, drop_stream=3, ld=ld)
        keep = y[:, :cols] != 0
        if rows * cols > 2000:
            assert 0.70 < keep.float().mean().item() < 0.80
        ref = torch.sigmoid(h[:, :cols].float()) / 0.75
        assert torch.allclose(y[:, :cols][keep].float(), ref[keep], atol=2e-2, rtol=2e-2)
        dy, dx = _bf(torch.randn(rows, ld, device='cuda')), torch.zeros_like(h)
        native.act_dropout_bwd(dy, y, dx, rows, cols, act=2, drop_p=0.25, seed=5, drop_stream=3, ld=ld)
        a = y[:, :cols].float() * 0.75
        ref = torch.where(keep, dy[:, :cols].float() / 0.75 * a * (1 - a), torch.zeros_like(a))
        assert torch.allclose(dx[:, :cols].float(), ref, atol=2e-2, rtol=3e-2), (rows, cols, ld)
        # no dropout: plain activation
        native.act_dropout_fwd(h, y, rows, cols, act=1, ld=ld)
        assert torch.equal(y[:, :cols], _bf(torch.relu(h[:, :cols].float())))


LOSS_CASES = [('categorical_crossentropy', 'softmax', 10), ('binary_crossentropy', 'sigmoid', 1),
              ('categorical_crossentropy', 'softmax', 24), ('categorical_crossentropy', 'softmax', 1000),
              ('mean_squared_error', 'linear', 40), ('mean_squared_error', 'sigmoid', 2),
              ('mean_squared_error', 'linear', 1), ('mean_absolute_error', 'linear', 3),
              ('mean_squared_error', 'relu', 7), ('mean_absolute_percentage_error', 'linear', 1),
              ('logcosh', 'tanh', 5), ('mean_squared_logarithmic_error', 'softplus', 4)]


@pytest.mark.parametrize('loss_name,act_name,n_out', LOSS_CASES)
def test_fused_loss_matches_autograd(native, loss_name, act_name, n_out):
    from elephas_b200.models import activations, losses
    torch.manual_seed(4)
    B = 333
    z = (torch.randn(B, n_out, device='cuda') * 1.5).requires_grad_(True)
    if loss_name == 'categorical_crossentropy':
        y = torch.eye(n_out, device='cuda')[torch.randint(0, n_out, (B,), device='cuda')]
    elif loss_name == 'binary_crossentropy':
        y = torch.randint(0, 2, (B, n_out), device='cuda').float()
    else:
        y = torch.rand(B, n_out, device='cuda') * 2 + 0.5
    act_fn = activations.get(act_name)[1]
    a = act_fn(z)
    per_sample = losses.get(loss_name)(y, a)
    per_sample.mean().backward()
    ldz = (n_out + 3) // 4 * 4
    zbuf = torch.zeros(B, ldz, device='cuda')
    zbuf[:, :n_out] = z.detach()
    out = torch.zeros(B, n_out, device='cuda')
    ldd = (n_out + 7) // 8 * 8
    dz = torch.zeros(B, ldd, dtype=torch.bfloat16, device='cuda')
    stats = torch.zeros(8, dtype=torch.float64, device='cuda')
    mets = [native.METRIC_IDS['categorical_accuracy' if act_name == 'softmax' else 'binary_accuracy'],
            native.METRIC_IDS['mean_absolute_error'], native.METRIC_IDS['mean_squared_error'],
            native.METRIC_IDS['mean_absolute_percentage_error']]
    native.loss_fwd_bwd(zbuf, y, B, n_out, native.LOSS_IDS[loss_name], native.ACT_IDS[act_name], out=out, dz=dz,
                        stats=stats, metrics=mets, grad_scale=1.0 / B)
    s = stats.cpu().numpy()
    assert s[7] == B
    assert torch.allclose(out, a.detach(), atol=1e-5, rtol=1e-5)
    assert abs(s[0] / B - per_sample.mean().item()) < 1e-4 * max(1.0, abs(per_sample.mean().item()))
    assert torch.allclose(dz[:, :n_out].float(), z.grad, atol=2e-4, rtol=2e-2)
    if act_name == 'softmax':
        assert abs(s[1] / B - (a.argmax(1) == y.argmax(1)).float().mean().item()) < 1e-6
    else:
        assert abs(s[1] / B - ((a > 0.5).float() == y).float().mean().item()) < 1e-6
    assert abs(s[2] / B - (a - y).abs().mean().item()) < 1e-4
    assert abs(s[3] / B - ((a - y) ** 2).mean().item()) < 1e-3


OPT_CASES = [('SGD', {}), ('SGD', {'momentum': 0.9}), ('SGD', {'momentum': 0.9, 'nesterov': True, 'decay': 1e-2}),
             ('RMSprop', {}), ('RMSprop', {'momentum': 0.5, 'centered': True}), ('Adam', {}),
             ('Adam', {'amsgrad': True, 'decay': 1e-3}), ('Adagrad', {'learning_rate': 0.1}),
             ('Adadelta', {'learning_rate': 1.0}), ('Adamax', {})]


@pytest.mark.parametrize('name,cfg', OPT_CASES)
def test_fused_optimizers_match_python(native, name, cfg):
    torch.manual_seed(5)
    n = 4096 + 24
    opt = optimizers.deserialize({'class_name': name, 'config': dict(cfg)})
    w_ref = torch.randn(n, dtype=torch.float64)
    w32 = w_ref.float().cuda()
    w16 = torch.zeros(n, dtype=torch.bfloat16, device='cuda')
    c = opt._config
    init0 = c.get('initial_accumulator_value', 0.0) if name == 'Adagrad' else 0.0
    slots = [torch.full((n,), init0 if i == 0 else 0.0, device='cuda') for i in range(3)]
    step = torch.zeros(1, dtype=torch.int64, device='cuda')
    for it in range(5):
        g = torch.randn(n, dtype=torch.float64)
        opt.apply_gradients([w_ref], [g])
        native.optimizer_step(native.OPT_IDS[name], n, g.float().cuda(), step, w32=w32, w16=w16, slots=slots,
                              lr=c['learning_rate'], decay=c.get('decay', 0.0), momentum=c.get('momentum', 0.0),
                              rho=c.get('rho', 0.9), eps=c.get('epsilon', 1e-7), beta1=c.get('beta_1', 0.9),
                              beta2=c.get('beta_2', 0.999), nesterov=c.get('nesterov', False),
                              centered=c.get('centered', False), amsgrad=c.get('amsgrad', False))
    assert int(step.item()) == 5
    assert torch.allclose(w32.cpu().double(), w_ref, atol=2e-5, rtol=2e-4)
    assert torch.equal(w16, w32.to(torch.bfloat16))


def test_optimizer_delta_mode_leaves_weights(native):
    n = 1024
    w32 = torch.randn(n, device='cuda')
    w0 = w32.clone()
    g = torch.randn(n, device='cuda')
    d16 = torch.zeros(n, dtype=torch.bfloat16, device='cuda')
    step = torch.zeros(1, dtype=torch.int64, device='cuda')
    native.optimizer_step(native.OPT_IDS['SGD'], n, g, step, w32=w32, delta_bf16=d16, apply=False, lr=0.5)
    assert torch.equal(w32, w0) and torch.equal(d16, (0.5 * g).to(torch.bfloat16))


def test_allreduce_update_single_rank(native):
    """world=1 exercises the kernel's barrier (trivially), reduction, scaling, update and bf16 refresh."""
    torch.manual_seed(6)
    n = 1 << 16
    w32 = torch.randn(n, device='cuda')
    w0 = w32.clone()
    w16 = torch.zeros(n, dtype=torch.bfloat16, device='cuda')
    pad = torch.zeros(64, dtype=torch.int32, device='cuda')
    local = torch.zeros(16, dtype=torch.int32, device='cuda')
    for it, dtype in enumerate((torch.bfloat16, torch.float32, torch.bfloat16)):
        d = torch.randn(n, device='cuda').to(dtype)
        avg = torch.zeros(n, dtype=dtype, device='cuda')
        native.allreduce_update([d.data_ptr()], [pad.data_ptr()], local, n, 0, 1, 1.0, dtype == torch.bfloat16, w32,
                                w16=w16, avg_ptrs=[avg.data_ptr()], algo=1 if it < 2 else 2)
        w0 = w0 - d.float()
        torch.cuda.synchronize()
        assert torch.allclose(w32, w0, atol=1e-6)
        assert torch.equal(w16, w32.to(torch.bfloat16))
    assert int(local[0].item()) == 3


def test_emulated_four_rank_reduction(native):
    """The reduction/scale/update arithmetic for N=4 'ranks' whose buffers live on one GPU: one launch per
    emulated rank, each seeing all four delta buffers (flags pre-published so that no launch waits on another)."""
    torch.manual_seed(7)
    n, world = 1 << 14, 4
    deltas = [torch.randn(n, device='cuda').to(torch.bfloat16) for _ in range(world)]
    pads = [torch.full((64,), 1 << 30, dtype=torch.int32, device='cuda') for _ in range(world)]
    mean = sum(d.float() for d in deltas) / world
    for r in range(world):
        w32 = torch.ones(n, device='cuda')
        local = torch.zeros(16, dtype=torch.int32, device='cuda')
        native.allreduce_update([d.data_ptr() for d in deltas], [p.data_ptr() for p in pads], local, n, r, world,
                                1.0 / world, True, w32, algo=1)
        torch.cuda.synchronize()
        assert torch.allclose(w32, 1.0 - mean, atol=1e-6)


def _emulated_ranks(n, world, dtype, seed):
    torch.manual_seed(seed)
    deltas = [torch.randn(n, device='cuda').to(dtype) for _ in range(world)]
    # flags far in the future: no launch ever waits for another one (ranks emulated on one GPU run one after another)
    pads = [torch.full((64,), 1 << 30, dtype=torch.int32, device='cuda') for _ in range(world)]
    for r, pad in enumerate(pads):
        pad[8 + r] = 0  # ... except a rank's own phase flag: its blocks really wait for one another's stores
    mean = sum(d.float() for d in deltas) / world
    return deltas, pads, mean


@pytest.mark.parametrize('dtype', [torch.bfloat16, torch.float32])
def test_emulated_eight_rank_twoshot_reduction(native, dtype):
    """Two-shot kernel (reduce-scatter of slice r, all-gather through every rank's averaged-delta buffer, local apply)
    with 8 ranks emulated on one GPU, against a PyTorch fp32 mean. The rank under test is launched last, when all the
    other ranks' slices have already landed in its buffer (the kernel's own barrier does this between real GPUs)."""
    n, world = 1 << 15, 8
    deltas, pads, mean = _emulated_ranks(n, world, dtype, 21)
    if dtype == torch.bfloat16:
        mean = mean.to(torch.bfloat16).float()  # the averaged slice travels in the delta's own type
    for target in (0, 3, 7):
        avgs = [torch.full((n,), float('nan'), dtype=dtype, device='cuda') for _ in range(world)]
        for r, pad in enumerate(pads):  # fresh flags for this round (the previous one left every own flag published)
            pad.fill_(1 << 30)
            pad[8 + r] = 0
        order = [r for r in range(world) if r != target] + [target]
        for r in order:
            w32 = torch.ones(n, device='cuda')
            w16 = torch.zeros(n, dtype=torch.bfloat16, device='cuda')
            local = torch.zeros(16, dtype=torch.int32, device='cuda')
            native.allreduce_update([d.data_ptr() for d in deltas], [p.data_ptr() for p in pads], local, n, r, world,
                                    1.0 / world, dtype == torch.bfloat16, w32, w16=w16,
                                    avg_ptrs=[a.data_ptr() for a in avgs], algo=2, blocks=4)
        torch.cuda.synchronize()
        assert torch.allclose(w32, 1.0 - mean, atol=1e-6), (w32 - (1.0 - mean)).abs().max()
        assert torch.equal(w16, w32.to(torch.bfloat16))
        for a in avgs:  # every rank's buffer holds the complete averaged delta
            assert torch.allclose(a.float(), mean, atol=1e-6)
        assert int(local[8].item()) == 0 and int(local[0].item()) == 1


def test_emulated_eight_rank_sharded_master_reduction(native):
    """Sharded-master kernel (K6b): emulated rank r reduces slice r of all eight delta buffers, updates slice r of ITS
    fp32 master and stores the refreshed bf16 weights into every rank's compute copy."""
    n, world = 1 << 15, 8
    deltas, pads, mean = _emulated_ranks(n, world, torch.bfloat16, 22)
    w16s = [torch.full((n,), float('nan'), dtype=torch.bfloat16, device='cuda') for _ in range(world)]
    masters = []
    for r in range(world):
        w32 = torch.ones(n, device='cuda')
        local = torch.zeros(16, dtype=torch.int32, device='cuda')
        native.allreduce_shard_update([d.data_ptr() for d in deltas], [p.data_ptr() for p in pads], local, n, r, world,
                                      1.0 / world, True, w32, [w.data_ptr() for w in w16s], blocks=4)
        masters.append((w32, local))
    torch.cuda.synchronize()
    per = -(-(n // 8) // world) * 8
    expect = 1.0 - mean
    for r, (w32, local) in enumerate(masters):
        a, b = per * r, min(n, per * (r + 1))
        assert torch.allclose(w32[a:b], expect[a:b], atol=1e-6)           # the slice this rank owns
        untouched = torch.ones(n, device='cuda')
        untouched[a:b] = w32[a:b]
        assert torch.equal(untouched, w32)                                 # nothing else was written
        assert int(local[8].item()) == 0
    for w16 in w16s:                                                       # all-gather: every copy complete, bf16(master)
        assert torch.allclose(w16.float(), expect, atol=2e-2)
        for r, (w32, _) in enumerate(masters):
            a, b = per * r, min(n, per * (r + 1))
            assert torch.equal(w16[a:b], w32[a:b].to(torch.bfloat16))


@pytest.mark.parametrize('locked', [True, False])
def test_parameter_server_kernels(native, locked):
    torch.manual_seed(8)
    n = 50000
    master = torch.randn(n, device='cuda')
    m0 = master.clone()
    ctrl = torch.zeros(8, dtype=torch.int32, device='cuda')
    streams = [torch.cuda.Stream() for _ in range(4)]
    deltas = [torch.randn(n, device='cuda') for _ in range(4)]
    torch.cuda.synchronize()
    for s, d in zip(streams, deltas):
        with torch.cuda.stream(s):
            for _ in range(5):
                before = torch.zeros(n, device='cuda')
                native.ps_push(master, ctrl, before=before, after=-d, locked=locked)  # delta = before - after = d
    torch.cuda.synchronize()
    expect = m0 - 5 * sum(deltas)
    assert torch.allclose(master, expect, atol=1e-4)   # red.add is atomic per element: exact in both modes
    dst32, dst16 = torch.zeros(n, device='cuda'), torch.zeros(n, dtype=torch.bfloat16, device='cuda')
    native.ps_pull(master, ctrl, dst32, dst16, locked=locked)
    torch.cuda.synchronize()
    assert torch.equal(dst32, master) and torch.equal(dst16, master.to(torch.bfloat16))
    if locked:
        assert int(ctrl[1].item()) == 21 and int(ctrl[2].item()) == 21 and int(ctrl[0].item()) % 2 == 0


@pytest.mark.parametrize('M,N,K', [(64, 128, 784), (64, 10, 128), (300, 200, 520), (4096, 1024, 512), (512, 4096, 4096)])
def test_gemm_fused_pull_side_writes_the_weights(native, M, N, K):
    """K2: the B operand is read from a 'remote' buffer and every tile is also stored into the local copy."""
    torch.manual_seed(21)
    ld = (K + 7) // 8 * 8
    x = _bf(torch.randn(M, ld, device='cuda'))
    x[:, K:] = 0
    remote = _bf(torch.randn(N, ld, device='cuda') * 0.1)
    local = torch.full((N, ld), 7.0, dtype=torch.bfloat16, device='cuda')
    bias = torch.randn(N, device='cuda')
    ldo = (N + 7) // 8 * 8
    out = torch.zeros(M, ldo, dtype=torch.bfloat16, device='cuda')
    ref_out = torch.zeros_like(out)
    native.gemm(x, remote, M, N, K, out_bf16=ref_out, bias=bias, act=1)            # plain GEMM on the remote weights
    native.gemm(x, remote.data_ptr(), M, N, K, ldb=ld, out_bf16=out, bias=bias, act=1, b_copy=local)  # raw address + copy
    assert torch.equal(out, ref_out)
    assert torch.equal(local[:, :K], remote[:, :K])
    assert torch.all(local[:, K:] == 7.0)  # padding columns are outside the tensor map: untouched
    ref = torch.relu(x[:, :K].float() @ remote[:, :K].float().t() + bias)
    assert torch.allclose(out[:, :N].float(), ref, atol=3e-2, rtol=3e-2)


@pytest.mark.parametrize('kind', ['f32', 'bf16', 'before_after'])
def test_parameter_server_push_refreshes_the_bf16_mirror(native, kind):
    torch.manual_seed(22)
    n = 4096 * 3 + 5
    master = torch.randn(n, device='cuda')
    mirror = master.to(torch.bfloat16)
    ctrl = torch.zeros(16, dtype=torch.int32, device='cuda')
    expect = master.clone()
    for it in range(3):
        d = torch.randn(n, device='cuda') * 0.1
        d[::7] = 0.0  # untouched elements keep their (already current) mirror entry
        if kind == 'f32':
            native.ps_push(master, ctrl, delta=d, mirror=mirror)
            expect -= d
        elif kind == 'bf16':
            d16 = d.to(torch.bfloat16)
            native.ps_push(master, ctrl, delta_bf16=d16, mirror=mirror, locked=True)
            expect -= d16.float()
        else:
            after = expect - d
            native.ps_push(master, ctrl, before=expect.clone(), after=after, mirror=mirror)
            expect = expect - (expect - after)
    torch.cuda.synchronize()
    assert torch.allclose(master, expect, atol=1e-5)
    assert torch.equal(mirror, master.to(torch.bfloat16))
    # ranges: push into a slice of the store
    native.ps_push(master, ctrl, delta=torch.ones(64, device='cuda'), mirror=mirror, offset=128, n=64)
    torch.cuda.synchronize()
    expect[128:192] -= 1.0
    assert torch.allclose(master, expect, atol=1e-5) and torch.equal(mirror, master.to(torch.bfloat16))
    tail = torch.zeros(n, device='cuda')
    native.ps_pull(master, ctrl, tail[100:164], n=64, offset=100)
    assert torch.equal(tail[100:164], master[100:164]) and float(tail[:100].abs().sum()) == 0.0


@pytest.mark.parametrize('rows,K,out', [(64, 128, 1), (300, 520, 3), (4096, 4096, 1), (1000, 512, 8), (5, 8, 2)])
def test_narrow_head_kernels_match_torch(native, rows, K, out):
    """Dense layers with <= 8 output units: forward, weight gradient and input gradient at HBM speed."""
    torch.manual_seed(31)
    ld = K + 8
    A = _bf(torch.randn(rows, ld, device='cuda'))
    W = _bf(torch.randn(out, ld, device='cuda') * 0.1)
    bias = torch.randn(out, device='cuda')
    z = torch.full((rows, max(4, (out + 3) // 4 * 4)), 9.0, device='cuda')
    native.head_fwd(A, W, rows, K, out, z, bias=bias)
    ref = A[:, :K].float() @ W[:, :K].float().t() + bias
    assert torch.allclose(z[:, :out], ref, atol=2e-3, rtol=2e-3)
    assert torch.all(z[:, out:] == 9.0)
    ldz = 8
    dz = _bf(torch.randn(rows, ldz, device='cuda') * 0.05)
    dw = torch.full((out, ld), 3.0, device='cuda')
    native.head_wgrad(dz, A, rows, K, out, dw, lddw=ld)
    ref_dw = dz[:, :out].float().t() @ A[:, :K].float()
    assert torch.allclose(dw[:, :K], ref_dw, atol=3e-3, rtol=3e-3), (dw[:, :K] - ref_dw).abs().max()
    for act_name in ('relu', 'tanh', 'linear'):
        aux = _bf(torch.randn(rows, ld, device='cuda'))
        if act_name == 'relu':
            aux = torch.relu(aux)
        dzp = torch.full((rows, ld), 7.0, dtype=torch.bfloat16, device='cuda')
        native.head_dgrad(dz, W, aux, dzp, rows, K, out, act=native.ACT_IDS[act_name])
        g = dz[:, :out].float() @ W[:, :K].float()
        a = aux[:, :K].float()
        grad = {'relu': (a > 0).float(), 'tanh': 1 - a * a, 'linear': torch.ones_like(a)}[act_name]
        assert torch.allclose(dzp[:, :K].float(), g * grad, atol=2e-2, rtol=2e-2)
        assert torch.all(dzp[:, K:] == 7.0)


@pytest.mark.parametrize('rows,dims,acts', [
    (300, (784, 128, 128, 10), ('relu', 'relu')),        # the reference's MNIST MLP
    (128, (64, 64, 16), ('tanh',)),
    (1000, (200, 32, 48, 128, 3), ('relu', 'sigmoid', 'elu')),
    (20000, (784, 128, 128, 10), ('relu', 'relu')),      # several tiles per CTA
    (77, (96, 16, 1), ('linear',)),
    (700, (96, 512, 512, 512, 9), ('relu', 'relu', 'relu')),   # the Otto MLP: weights streamed, activations resident
    (129, (200, 144, 320, 48), ('tanh', 'relu')),              # widths that are no multiples of 64 / 128
    (5000, (512, 512, 512), ('relu',)),                        # full-width input, several tiles per CTA
])
def test_fused_mlp_forward_matches_layer_by_layer(native, rows, dims, acts):
    """K8: one persistent kernel carries a row tile through every layer (activations stay on chip)."""
    torch.manual_seed(41)
    in_dim = dims[0]
    ldx = in_dim + 8
    x = _bf(torch.randn(rows, ldx, device='cuda'))
    layers, ref = [], x[:, :in_dim].float()
    k = in_dim
    for i, units in enumerate(dims[1:]):
        ld = (k + 7) // 8 * 8 + 8
        w = _bf(torch.randn(units, ld, device='cuda') / (k ** 0.5))
        b = torch.randn(units, device='cuda') * 0.1
        last = i == len(dims) - 2
        act = 'linear' if last else acts[i]
        layers.append((w, units, b, native.ACT_IDS[act]))
        ref = ref @ w[:, :k].float().t() + b
        if not last:
            ref = _bf({'relu': torch.relu, 'tanh': torch.tanh, 'sigmoid': torch.sigmoid, 'linear': lambda t: t,
                       'elu': torch.nn.functional.elu}[act](ref)).float()   # the kernel keeps hidden activations in bf16
        k = units
    assert native.mlp_forward_supported(in_dim, dims[1:])
    ldz = max(4, (dims[-1] + 3) // 4 * 4)
    z = torch.full((rows, ldz), 9.0, device='cuda')
    native.mlp_forward(x, rows, in_dim, layers, z)
    torch.cuda.synchronize()
    assert torch.allclose(z[:, :dims[-1]], ref, atol=3e-2, rtol=3e-2), (z[:, :dims[-1]] - ref).abs().max()
    assert torch.all(z[:, dims[-1]:] == 9.0)
    assert not native.mlp_forward_supported(in_dim, (1024, 10)) and not native.mlp_forward_supported(in_dim, (10,))
    assert not native.mlp_forward_supported(784, (512, 10))  # wide layers need the input tile on chip (<= 512 columns)


def test_launch_counter_moves(native):
    before = native.launch_count()
    native.cast_f32_to_bf16(torch.zeros(64, device='cuda'), torch.zeros(64, dtype=torch.bfloat16, device='cuda'))
    assert native.launch_count() == before + 1
