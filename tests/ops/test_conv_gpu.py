"""K9 on a B200: the implicit-GEMM convolution kernels (forward, input gradient, split-K weight gradient), the
BatchNormalization kernels and the Dense GEMM wrapper against plain PyTorch fp32 references of the same op, on odd
shapes, strides and paddings; then a small residual CNN end to end through the public Model API."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from elephas_b200.ops import cnn, native  # noqa: E402


def _rounded(*shape, scale=1.0, seed=0):
    g = torch.Generator(device='cuda').manual_seed(seed)
    return (torch.randn(*shape, device='cuda', generator=g) * scale).to(torch.bfloat16).float()


def _close(got, want, rel=2e-2):
    """bf16 outputs of fp32 accumulations: compare against the magnitude of the reference tensor."""
    err = (got.float() - want).abs().max().item()
    ref = want.abs().max().item()
    assert err <= rel * max(ref, 1e-3), "max |err| %.4g vs max |ref| %.4g" % (err, ref)


CONV_CASES = [
    # N, H, W, C, K, R, stride, pad
    (2, 56, 56, 64, 64, 3, 1, 1),      # ResNet stage 2, 3x3
    (3, 14, 14, 256, 256, 3, 1, 1),    # several images per tile
    (5, 7, 7, 512, 512, 3, 1, 1),      # tiny feature map
    (2, 13, 9, 24, 40, 3, 1, 1),       # odd sizes, channel tails (24 < 64, 40 < 64)
    (2, 28, 28, 128, 512, 1, 1, 0),    # 1x1
    (2, 56, 56, 256, 128, 1, 2, 0),    # 1x1 stride 2 (element strides of the tensor map)
    (1, 17, 11, 72, 136, 1, 2, 0),     # 1x1 stride 2, odd
    (2, 12, 12, 32, 48, 3, 1, 0),      # valid padding
    (1, 9, 9, 16, 24, 5, 1, 2),        # 5x5
]


@pytest.mark.parametrize('N,H,W,C,K,R,stride,pad', CONV_CASES)
def test_conv2d_forward_and_gradients_match_pytorch(N, H, W, C, K, R, stride, pad):
    x = _rounded(N, H, W, C, seed=1).requires_grad_(True)
    w = _rounded(K, R, R, C, scale=(R * R * C) ** -0.5, seed=2).requires_grad_(True)
    b = _rounded(K, seed=3).requires_grad_(True)
    with torch.autocast('cuda', dtype=torch.bfloat16):
        y = cnn.conv2d(x, w, b, stride, (pad, pad, pad, pad))
    ref = F.conv2d(x.permute(0, 3, 1, 2), w.permute(0, 3, 1, 2), b, stride=stride, padding=pad).permute(0, 2, 3, 1)
    assert y.dtype == torch.bfloat16 and tuple(y.shape) == tuple(ref.shape)
    _close(y, ref.detach())
    g = _rounded(*ref.shape, seed=4)
    dx, dw, db = torch.autograd.grad(y, (x, w, b), g.to(torch.bfloat16))
    rx, rw, rb = torch.autograd.grad(ref, (x, w, b), g)
    _close(dx, rx)
    _close(dw, rw, rel=1e-2)  # fp32 output of an fp32 accumulation (bf16 products): only summation order differs
    _close(db, rb, rel=1e-2)


def test_weight_gradient_accumulates_into_an_existing_grad_buffer():
    """The training engine pre-assigns ``kernel.grad`` (a view of the flat gradient buffer, zeroed every step): the
    weight-gradient kernel adds into it directly and autograd gets no tensor to add."""
    x = _rounded(2, 10, 10, 32, seed=5)
    w = _rounded(64, 3, 3, 32, scale=0.1, seed=6).requires_grad_(True)
    w.grad = torch.full_like(w, 0.5)
    w._eb_flat_grad = True  # what ops/torch_step.py sets on the parameters whose gradient views it owns
    with torch.autocast('cuda', dtype=torch.bfloat16):
        y = cnn.conv2d(x, w, None, 1, (1, 1, 1, 1))
    g = _rounded(*y.shape, seed=7)
    y.backward(g.to(torch.bfloat16))
    ref = F.conv2d(x.permute(0, 3, 1, 2), w.detach().permute(0, 3, 1, 2).requires_grad_(True), padding=1)
    wr = w.detach().clone().requires_grad_(True)
    F.conv2d(x.permute(0, 3, 1, 2), wr.permute(0, 3, 1, 2), padding=1).backward(g.permute(0, 3, 1, 2))
    _close(w.grad - 0.5, wr.grad, rel=1e-2)
    assert ref.shape[1] == 64


def test_input_layer_goes_through_the_patch_matrix():
    """7x7 / stride 2 over 3 channels (fp32 images): patch matrix + the same GEMM kernels."""
    x = _rounded(2, 38, 38, 3, seed=8)
    w = _rounded(64, 7, 7, 3, scale=0.1, seed=9).requires_grad_(True)
    b = _rounded(64, seed=10).requires_grad_(True)
    assert cnn.conv2d_supported(x, w, (2, 2), (1, 1), 1, (0, 0, 0, 0))
    with torch.autocast('cuda', dtype=torch.bfloat16):
        y = cnn.conv2d(x, w, b, 2, (0, 0, 0, 0))
    ref = F.conv2d(x.permute(0, 3, 1, 2), w.permute(0, 3, 1, 2), b, stride=2).permute(0, 2, 3, 1)
    _close(y, ref.detach())
    g = _rounded(*ref.shape, seed=11)
    dw, db = torch.autograd.grad(y, (w, b), g.to(torch.bfloat16))
    rw, rb = torch.autograd.grad(ref, (w, b), g)
    _close(dw, rw, rel=1e-2)
    _close(db, rb, rel=1e-2)


@pytest.mark.parametrize('M,C', [(2 * 56 * 56, 64), (3 * 7 * 7, 2048), (101, 24), (5000, 136)])
@pytest.mark.parametrize('relu,residual', [(False, False), (True, False), (True, True)])
def test_batch_norm_matches_pytorch(M, C, relu, residual):
    x = _rounded(M, C, scale=2.0, seed=12) + 0.5
    x = x.to(torch.bfloat16).float().requires_grad_(True)
    gamma = (_rounded(C, seed=13) * 0.2 + 1.0).requires_grad_(True)
    beta = (_rounded(C, seed=14) * 0.3).requires_grad_(True)
    res = _rounded(M, C, seed=15).requires_grad_(True) if residual else None
    rm, rv = torch.zeros(C, device='cuda'), torch.ones(C, device='cuda')
    rm_ref, rv_ref = rm.clone(), rv.clone()
    with torch.autocast('cuda', dtype=torch.bfloat16):
        y = cnn.batch_norm(x, gamma, beta, rm, rv, 1e-3, 0.9, True, residual=res, relu=relu)
    ref = F.batch_norm(x, rm_ref, rv_ref, gamma, beta, training=True, momentum=0.1, eps=1e-3)
    if residual:
        ref = ref + res
    if relu:
        ref = torch.relu(ref)
    _close(y, ref.detach(), rel=1e-2)
    assert torch.allclose(rm, rm_ref, atol=1e-3) and torch.allclose(rv, rv_ref, rtol=1e-2, atol=1e-3)
    g = _rounded(M, C, seed=16)
    # the ReLU mask is taken from each implementation's own output; compare where both agree on the sign
    inputs = (x, gamma, beta) + ((res,) if residual else ())
    got = torch.autograd.grad(y, inputs, g.to(torch.bfloat16))
    want = torch.autograd.grad(ref, inputs, g)
    agree = ((y.float() > 0) == (ref > 0)).float().mean().item() if relu else 1.0
    assert agree > 0.995
    _close(got[1], want[1], rel=3e-2)
    _close(got[2], want[2], rel=3e-2)
    err = (got[0].float() - want[0]).abs()
    assert err.mean().item() <= 2e-2 * want[0].abs().mean().item() + 1e-4
    if residual:
        err = (got[3].float() - want[3]).abs()
        assert err.mean().item() <= 2e-2 * want[3].abs().mean().item() + 1e-4
    # inference mode uses the moving statistics
    with torch.no_grad(), torch.autocast('cuda', dtype=torch.bfloat16):
        yi = cnn.batch_norm(x.detach(), gamma.detach(), beta.detach(), rm, rv, 1e-3, 0.9, False, relu=relu)
    ri = F.batch_norm(x.detach(), rm.clone(), rv.clone(), gamma.detach(), beta.detach(), training=False, eps=1e-3)
    _close(yi, torch.relu(ri) if relu else ri, rel=1e-2)


POOL_CASES = [
    # N, H, W, C, window, stride, pads (top, bottom, left, right), zero padding?
    (2, 112, 112, 64, 3, 2, (1, 1, 1, 1), True),    # ResNet pool1 with its ZeroPadding2D folded in
    (3, 13, 9, 24, 3, 2, (0, 1, 0, 1), False),      # Keras padding='same' on odd sizes (-inf outside)
    (2, 8, 8, 16, 2, 2, (0, 0, 0, 0), False),       # plain 2x2
    (1, 11, 7, 40, 3, 1, (1, 1, 1, 1), True),       # overlapping windows, stride 1
    (2, 10, 10, 8, 5, 3, (2, 2, 2, 2), False),      # wide window, stride 3
]


@pytest.mark.parametrize('N,H,W,C,k,stride,pads,zero', POOL_CASES)
def test_max_pool_matches_pytorch(N, H, W, C, k, stride, pads, zero):
    """Forward values and the routing of the gradient (bf16 inputs have many exact ties: the first maximum in scan order
    must win, as in the library kernel; a zero of the padding takes part like an element)."""
    x = _rounded(N, H, W, C, seed=11)
    if zero:
        x = x - 1.0  # mostly negative: the padding zeros win many border windows
    x = x.to(torch.bfloat16).float().requires_grad_(True)
    pt, pb, pl, pr = pads
    with torch.autocast('cuda', dtype=torch.bfloat16):
        y = cnn.max_pool2d(x, (k, k), stride, pads, pad_is_zero=zero)
    xp = F.pad(x.permute(0, 3, 1, 2), (pl, pr, pt, pb), value=0.0 if zero else float('-inf'))
    ref = F.max_pool2d(xp, k, stride).permute(0, 2, 3, 1)
    assert y.dtype == torch.bfloat16 and tuple(y.shape) == tuple(ref.shape)
    assert torch.equal(y.float(), ref.detach())  # a maximum is one of the inputs: exact
    g = _rounded(*ref.shape, seed=12)
    (dx,) = torch.autograd.grad(y, x, g.to(torch.bfloat16))
    (rx,) = torch.autograd.grad(ref, x, g)
    _close(dx, rx, rel=1e-2)  # sums of up to four bf16 gradients, rounded once


def test_global_average_pool_matches_pytorch():
    x = _rounded(5, 7, 7, 136, seed=13).requires_grad_(True)
    with torch.autocast('cuda', dtype=torch.bfloat16):
        y = cnn.global_avg_pool(x)
    ref = x.mean(dim=(1, 2))
    assert y.dtype == torch.bfloat16 and tuple(y.shape) == (5, 136)
    _close(y, ref.detach(), rel=1e-2)
    g = _rounded(5, 136, seed=14)
    (dx,) = torch.autograd.grad(y, x, g.to(torch.bfloat16))
    (rx,) = torch.autograd.grad(ref, x, g)
    _close(dx, rx, rel=1e-2)


def test_linear_matches_pytorch():
    x = _rounded(70, 2048, seed=17).requires_grad_(True)
    w = _rounded(1000, 2048, scale=0.02, seed=18).requires_grad_(True)
    b = _rounded(1000, seed=19).requires_grad_(True)
    with torch.autocast('cuda', dtype=torch.bfloat16):
        y = cnn.linear(x, w, b)
    ref = F.linear(x, w, b)
    _close(y, ref.detach())
    g = _rounded(70, 1000, seed=20)
    got = torch.autograd.grad(y, (x, w, b), g.to(torch.bfloat16))
    want = torch.autograd.grad(ref, (x, w, b), g)
    for a, r in zip(got, want):
        _close(a, r)


def _small_resnet():
    from elephas_b200.models import Model
    from elephas_b200.models.layers import (Activation, Add, BatchNormalization, Conv2D, Dense, GlobalAveragePooling2D,
                                            Input, MaxPooling2D, ZeroPadding2D)
    inp = Input(shape=(32, 32, 3))
    x = ZeroPadding2D(3)(inp)
    x = Conv2D(32, 7, strides=2, name='stem')(x)
    x = Activation('relu')(BatchNormalization(epsilon=1e-5)(x))
    x = MaxPooling2D(3, strides=2)(ZeroPadding2D(1)(x))
    for i, (f, s) in enumerate(((16, 1), (32, 2))):
        short = BatchNormalization(epsilon=1e-5)(Conv2D(4 * f, 1, strides=s)(x))
        y = Activation('relu')(BatchNormalization(epsilon=1e-5)(Conv2D(f, 1, strides=s)(x)))
        y = Activation('relu')(BatchNormalization(epsilon=1e-5)(Conv2D(f, 3, padding='same')(y)))
        y = BatchNormalization(epsilon=1e-5)(Conv2D(4 * f, 1)(y))
        x = Activation('relu')(Add()([short, y]))
    x = GlobalAveragePooling2D()(x)
    return Model(inputs=inp, outputs=Dense(10, activation='softmax')(x))


def test_residual_cnn_trains_on_native_kernels_and_matches_the_torch_path(monkeypatch):
    from elephas_b200.models import SGD
    rng = np.random.default_rng(0)
    x = rng.normal(size=(64, 32, 32, 3)).astype(np.float32)
    labels = (x.mean(axis=(1, 2, 3)) > 0).astype(int) + 2 * (x[:, :16].mean(axis=(1, 2, 3)) > 0).astype(int)
    y = np.eye(10, dtype=np.float32)[labels]
    losses = {}
    for flag in ('1', '0'):
        monkeypatch.setenv('ELEPHAS_B200_NATIVE_CONV', flag)
        torch.manual_seed(3)
        model = _small_resnet()
        model.compile(SGD(learning_rate=0.05, momentum=0.9), 'categorical_crossentropy', ['acc'])
        model.set_compute_dtype('bfloat16')
        model.to('cuda')
        before = native.launch_count()
        losses[flag] = [float(model.train_on_batch(x, y)[0]) for _ in range(12)]
        launched = native.launch_count() - before
        preds = model.predict(x[:8])
        assert preds.shape == (8, 10) and np.allclose(preds.sum(1), 1.0, atol=2e-2)
        if flag == '1':
            assert launched > 12 * 40, "the native convolution kernels did not run (%d launches)" % launched
    assert losses['1'][-1] < 0.7 * losses['1'][0], losses['1']
    assert abs(losses['1'][0] - losses['0'][0]) < 0.05 * max(1.0, losses['0'][0])   # same first loss as cuDNN
    assert abs(losses['1'][-1] - losses['0'][-1]) < 0.35 * max(losses['0'][0], 1e-3)  # and a comparable trajectory
