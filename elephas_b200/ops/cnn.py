"""Native convolutional layers (K9): autograd functions around the sm_100a kernels of ``csrc/conv_sm100.cu`` /
``csrc/conv_aux.cu``.

The reference trains CNNs through ``keras.Model.fit`` on the workers (``elephas/worker.py:42,108``). Here, for a model
that computes in bf16 on a B200, every Conv2D / BatchNormalization / ReLU / residual Add / Dense of the forward and
backward pass is one of this repository's kernels:

* ``conv2d``      implicit GEMM on tcgen05: forward, input gradient (the same OHWI kernel array read N-major, no
                  transposed copy) and split-K weight gradient; the 3-channel input layer goes through a small patch
                  matrix (``eb_conv_im2col``) and then the same GEMM kernels
* ``batch_norm``  statistics -> normalise (+ residual) (+ ReLU) in one pass; backward = one reduction pass + one apply
                  pass that also emits the gradient of the residual branch
* ``linear``      the tcgen05 GEMM of the Dense engine (K1)
* ``max_pool2d`` / ``global_avg_pool``  ``csrc/pool_kernels.cu``: the forward keeps a byte map of the winning taps, the
                  backward gathers per input pixel (no atomics); a ZeroPadding2D in front of a pool is folded into it

torch autograd only orders the calls (and the whole step is replayed from a CUDA graph by ``ops/torch_step.py``);
activations travel channels-last in bf16, master weights stay fp32. ``ELEPHAS_B200_NATIVE_CONV=0`` switches the layers
back to the torch / cuDNN ops (the A/B arm of the benchmark).
"""
import os
import threading

import torch

from . import native


_scope = threading.local()  # .fresh: inside a training step whose engine has just refreshed the flat bf16 weight copy


def weights_bf16(p):
    """bf16 values of parameter ``p``: a view of the engine's flat bf16 copy inside a captured training step (ONE cast
    kernel per step over all weights, ``TorchStepEngine._body``), a fresh cast everywhere else."""
    w16 = getattr(p, '_eb_w16', None)
    if w16 is not None and getattr(_scope, 'fresh', False) and w16.shape == p.shape:
        return w16
    return p.detach().to(torch.bfloat16)


class fresh_weight_copy:
    """Context of ``TorchStepEngine._body``: the flat bf16 copy matches the fp32 weights for the duration of the forward
    pass (per thread: worker threads of one process train replicas side by side)."""

    def __enter__(self):
        self.prev = getattr(_scope, 'fresh', False)
        _scope.fresh = True

    def __exit__(self, *exc):
        _scope.fresh = self.prev


def enabled():
    return os.environ.get('ELEPHAS_B200_NATIVE_CONV', '1') != '0' and native.available()


def usable(x):
    """bf16 compute on a CUDA device requested (autocast region of ``Model.set_compute_dtype('bfloat16')``)."""
    return x.is_cuda and torch.is_autocast_enabled('cuda') and torch.get_autocast_dtype('cuda') == torch.bfloat16 and enabled()


def _bf16(t):
    t = t if t.dtype == torch.bfloat16 else t.to(torch.bfloat16)
    return t if t.is_contiguous() else t.contiguous()


def _out_size(size, k, stride, pad_lo, pad_hi):
    return (size + pad_lo + pad_hi - k) // stride + 1


class _Conv2d(torch.autograd.Function):
    """``x`` NHWC, ``kernel`` fp32 OHWI ``[K, R, S, C]``; ``pads`` = (top, bottom, left, right) zeros (handled by the TMA
    out-of-bounds fill: no padded copy of the activation is made)."""

    @staticmethod
    def forward(ctx, x, kernel, bias, stride, pads):
        wb = weights_bf16(kernel)
        N, H, W, C = x.shape
        K, R, S, _ = wb.shape
        xb = _bf16(x) if C % 8 == 0 or not (x.dtype == torch.float32 and x.is_contiguous()) else None
        pt, pb, pl, pr = pads
        P, Q = _out_size(H, R, stride, pt, pb), _out_size(W, S, stride, pl, pr)
        y = torch.empty(N, P, Q, K, dtype=torch.bfloat16, device=x.device)
        b32 = bias.detach().float() if bias is not None else None
        small_c = C % 8 != 0
        cols = None
        if small_c:
            # input layer (3 channels): explicit patch matrix, then a 1x1 convolution over it
            ld = (R * S * C + 31) // 32 * 32
            cols = torch.empty(N, P, Q, ld, dtype=torch.bfloat16, device=x.device)
            src = x if (x.dtype == torch.float32 and x.is_contiguous()) else xb  # the kernel casts while it gathers
            native.conv_im2col(src, R, S, stride, pt, P, Q, cols, ld)
            wpad = torch.zeros(K, 1, 1, ld, dtype=torch.bfloat16, device=x.device)
            wpad.view(K, ld)[:, :R * S * C] = wb.view(K, R * S * C)
            native.conv_igemm(cols, wpad, y, P, Q, 1, 1, bias=b32)
            ctx.save_for_backward(cols, wpad)
        elif R == 1 and S == 1 and stride == 1 and not any(pads):
            # a 1x1 / stride-1 convolution IS the GEMM [pixels, C] x [K, C]^T: the clustered CTA-pair kernel of the Dense
            # engine (K1, B tiles multicast across the cluster) serves it without any patch logic
            native.gemm(xb.view(-1, C), wb.view(K, C), N * H * W, K, C, out_bf16=y.view(-1, K), bias=b32)
            ctx.save_for_backward(xb, wb)
        else:
            native.conv_igemm(xb, wb, y, P, Q, R, S, stride=stride, pad_h=pt, pad_w=pl, bias=b32)
            ctx.save_for_backward(xb, wb)
        ctx.geom = (x.shape, x.dtype, tuple(kernel.shape), stride, pads, P, Q, small_c, bias is not None)
        ctx.kernel_param, ctx.bias_param = kernel, bias
        return y

    @staticmethod
    def backward(ctx, gy):
        saved_x, wb = ctx.saved_tensors
        x_shape, x_dtype, k_shape, stride, pads, P, Q, small_c, has_bias = ctx.geom
        kernel = ctx.kernel_param
        K, R, S, C = k_shape
        N, H, W, _ = x_shape
        pt, pb, pl, pr = pads
        g = _bf16(gy)
        torch.cuda.set_device(g.device)  # autograd's worker thread: the TMA descriptor encoder needs the context
        dx = dw = db = None
        if ctx.needs_input_grad[0] and not small_c:
            if stride == 1 and R == 1 and S == 1 and not any(pads):
                dxb = torch.empty(N, H, W, C, dtype=torch.bfloat16, device=g.device)
                native.gemm(g.view(-1, K), wb.view(K, C), N * H * W, C, K, b_mn=True, out_bf16=dxb.view(-1, C))
            elif stride == 1:
                dxb = torch.empty(N, H, W, C, dtype=torch.bfloat16, device=g.device)
                # dx[h] = sum_r dy[h + pad - r] w[r]: a stride-1 convolution of dy with the flipped taps, padding R-1-pad
                native.conv_igemm(g, wb, dxb, H, W, R, S, stride=1, pad_h=R - 1 - pt, pad_w=S - 1 - pl, w_mn=True,
                                  flip_taps=True)
            elif R == 1 and S == 1:
                dxb = torch.zeros(N, H, W, C, dtype=torch.bfloat16, device=g.device)
                native.conv_igemm(g, wb, dxb, P, Q, 1, 1, stride=1, os=stride, w_mn=True)  # scatter to the sampled pixels
            else:
                raise NotImplementedError("input gradient of a strided %dx%d convolution" % (R, S))
            dx = dxb if x_dtype == torch.bfloat16 else dxb.to(x_dtype)
        if ctx.needs_input_grad[1]:
            target = None if small_c else _flat_grad(kernel)
            if small_c:
                ld = saved_x.shape[3]
                dwp = torch.zeros(K, 1, 1, ld, dtype=torch.float32, device=g.device)
                native.conv_wgrad(saved_x, g, dwp, 1, 1)
                dw = dwp.view(K, ld)[:, :R * S * C].reshape(K, R, S, C)
            elif target is not None:
                # accumulate straight into the (pre-zeroed) flat gradient buffer: no temporary, no add kernel
                native.conv_wgrad(saved_x, g, target, R, S, stride=stride, pad_h=pt, pad_w=pl)
            else:
                dw = torch.zeros(K, R, S, C, dtype=torch.float32, device=g.device)
                native.conv_wgrad(saved_x, g, dw, R, S, stride=stride, pad_h=pt, pad_w=pl)
        if has_bias and ctx.needs_input_grad[2]:
            into = _flat_grad(ctx.bias_param)
            if into is not None:
                native.colsum_bf16(g.view(-1, K), N * P * Q, K, into, accumulate=True)
            else:
                db = torch.empty(K, dtype=torch.float32, device=g.device)
                native.colsum_bf16(g.view(-1, K), N * P * Q, K, db)
        return dx, dw, db, None, None


def conv2d(x, kernel, bias, stride, pads):
    return _Conv2d.apply(x, kernel, bias, int(stride), tuple(int(v) for v in pads))


def conv2d_supported(x, kernel, strides, dilation, groups, pads=(0, 0, 0, 0)):
    K, R, S, C = kernel.shape
    if x.dim() != 4 or tuple(dilation) != (1, 1) or groups != 1 or strides[0] != strides[1] or K % 8:
        return False
    if C % 8:
        # the input layer: goes through the patch matrix (one symmetric padding value), never needs an input gradient
        return C * R * S <= 512 and len(set(pads)) == 1 and not x.requires_grad
    return strides[0] == 1 or (R == 1 and S == 1)


_PARTIALS = {}


def _partials(device, M, C):
    """fp32 scratch for the per-block partial sums of the two BatchNormalization reductions, per device / stream / width
    (everything on it is ordered by the stream: one layer at a time). Never needs clearing."""
    rows = max(1, min(native.BN_MAX_BLOCKS, -(-int(M) // 8)))
    key = (device.index, C, torch.cuda.current_stream(device).cuda_stream)
    buf = _PARTIALS.get(key)
    if buf is None or buf.numel() < rows * 2 * C:
        buf = _PARTIALS[key] = torch.empty(rows * 2 * C, dtype=torch.float32, device=device)
    return buf


def _flat_grad(p):
    """The pre-zeroed slice of the training engine's flat gradient buffer that belongs to ``p`` (``ops/torch_step.py``
    marks the parameters it manages), or None: then gradients are returned to autograd as tensors."""
    g = p.grad if p is not None and getattr(p, '_eb_flat_grad', False) else None
    return g if (g is not None and g.dtype == torch.float32 and g.is_contiguous()) else None


class _BatchNorm(torch.autograd.Function):
    """Training-mode BatchNormalization over the last axis (+ residual) (+ ReLU) of a channels-last bf16 tensor."""

    @staticmethod
    def forward(ctx, x, gamma, beta, running_mean, running_var, residual, relu, eps, momentum, training):
        xb = _bf16(x)
        C = xb.shape[-1]
        M = xb.numel() // C
        dev = x.device
        res = _bf16(residual) if residual is not None else None
        y = torch.empty_like(xb)
        g32 = gamma.detach() if gamma is not None else None
        b32 = beta.detach() if beta is not None else None
        if training:
            mean = torch.empty(C, dtype=torch.float32, device=dev)
            rstd = torch.empty(C, dtype=torch.float32, device=dev)
            native.bn_stats(xb, M, C, _partials(dev, M, C), eps, momentum, mean, rstd, running_mean, running_var)
        else:
            mean = running_mean.detach().float()
            rstd = torch.rsqrt(running_var.detach().float() + eps)
        native.bn_apply(xb, M, C, mean, rstd, g32, b32, res, relu, y)
        ctx.save_for_backward(xb, y if relu else None, mean, rstd, g32)
        ctx.meta = (x.dtype, relu, residual is not None, M, C, training)
        ctx.params = (gamma, beta)
        return y

    @staticmethod
    def backward(ctx, gy):
        xb, y, mean, rstd, g32 = ctx.saved_tensors
        x_dtype, relu, has_res, M, C, training = ctx.meta
        gamma, beta = ctx.params
        if not training:
            raise NotImplementedError("gradient through inference-mode BatchNormalization")
        g = _bf16(gy)
        torch.cuda.set_device(g.device)
        # d beta = sum dz and d gamma = sum dz * xhat are exactly the two reductions the input gradient needs: when the
        # training engine owns the (zeroed) gradient buffers they are accumulated there directly and read back from there
        dgamma_buf, dbeta_buf = _flat_grad(gamma), _flat_grad(beta)
        direct = dgamma_buf is not None and dbeta_buf is not None
        if direct:
            sum_dz, sum_dzx = dbeta_buf, dgamma_buf
        else:
            both = torch.empty(2, C, dtype=torch.float32, device=g.device)
            sum_dz, sum_dzx = both[0], both[1]
        native.bn_bwd_reduce(g, y, xb, M, C, mean, rstd, relu, _partials(g.device, M, C), sum_dz, sum_dzx, accumulate=direct)
        dx = torch.empty_like(xb)
        dres = torch.empty_like(xb) if has_res else None
        native.bn_bwd_apply(g, y, xb, M, C, mean, rstd, g32, sum_dz, sum_dzx, relu, dx, dres)
        dgamma = None if (direct or gamma is None) else sum_dzx
        dbeta = None if (direct or beta is None) else sum_dz
        return (dx if x_dtype == torch.bfloat16 else dx.to(x_dtype)), dgamma, dbeta, None, None, dres, None, None, None, None


def batch_norm(x, gamma, beta, running_mean, running_var, eps, momentum, training, residual=None, relu=False):
    return _BatchNorm.apply(x, gamma, beta, running_mean, running_var, residual, bool(relu), float(eps), float(momentum),
                            bool(training))


class _Linear(torch.autograd.Function):
    """``y = x W^T + b`` on the tcgen05 GEMM: ``x`` ``[..., in]``, ``weight`` fp32 ``[out, in]`` (the Dense layer's own layout)."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        out_dim, in_dim = weight.shape
        x2 = _bf16(x.reshape(-1, in_dim))
        wb = weights_bf16(weight)
        m = x2.shape[0]
        ldo = (out_dim + 7) // 8 * 8
        y = torch.empty(m, ldo, dtype=torch.bfloat16, device=x.device)
        native.gemm(x2, wb, m, out_dim, in_dim, out_bf16=y, ldo=ldo, bias=bias.detach().float() if bias is not None else None)
        ctx.save_for_backward(x2, wb)
        ctx.meta = (x.shape, x.dtype, bias is not None, ldo)
        return y[:, :out_dim].reshape(*x.shape[:-1], out_dim)

    @staticmethod
    def backward(ctx, gy):
        x2, wb = ctx.saved_tensors
        x_shape, x_dtype, has_bias, ldo = ctx.meta
        out_dim, in_dim = wb.shape
        m = x2.shape[0]
        torch.cuda.set_device(gy.device)
        g = torch.zeros(m, ldo, dtype=torch.bfloat16, device=gy.device)
        g[:, :out_dim] = gy.reshape(m, out_dim)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dxb = torch.empty(m, in_dim, dtype=torch.bfloat16, device=gy.device)
            native.gemm(g, wb, m, in_dim, out_dim, b_mn=True, lda=ldo, out_bf16=dxb)
            dx = dxb.reshape(x_shape).to(x_dtype)
        if ctx.needs_input_grad[1]:
            dw = torch.empty(out_dim, in_dim, dtype=torch.float32, device=gy.device)
            native.gemm(g, x2, out_dim, in_dim, m, a_mn=True, b_mn=True, lda=ldo, out_f32=dw)
        if has_bias and ctx.needs_input_grad[2]:
            db = torch.empty(out_dim, dtype=torch.float32, device=gy.device)
            native.colsum_bf16(g, m, out_dim, db, ld=ldo)
        return dx, dw, db


def linear(x, weight, bias):
    return _Linear.apply(x, weight, bias)


def linear_supported(weight):
    return weight.shape[1] % 8 == 0


class _MaxPool2d(torch.autograd.Function):
    """``x`` NHWC bf16; ``pads`` = (top, bottom, left, right); elements outside the image count as zeros
    (``pad_is_zero``: a ZeroPadding2D folded into the pool) or as -inf (Keras ``padding='same'``)."""

    @staticmethod
    def forward(ctx, x, window, stride, pads, pad_is_zero):
        xb = _bf16(x)
        N, H, W, C = xb.shape
        R, S = window
        pt, pb, pl, pr = pads
        P, Q = _out_size(H, R, stride, pt, pb), _out_size(W, S, stride, pl, pr)
        y = torch.empty(N, P, Q, C, dtype=torch.bfloat16, device=x.device)
        idx = torch.empty(N, P, Q, C, dtype=torch.uint8, device=x.device)
        native.pool_max_fwd(xb, R, S, stride, pt, pl, P, Q, pad_is_zero, y, idx)
        ctx.save_for_backward(idx)
        ctx.geo = (tuple(xb.shape), R, S, stride, pt, pl, x.dtype)
        return y

    @staticmethod
    def backward(ctx, gy):
        (idx,) = ctx.saved_tensors
        shape, R, S, stride, pt, pl, dtype = ctx.geo
        dx = torch.empty(shape, dtype=torch.bfloat16, device=gy.device)
        native.pool_max_bwd(_bf16(gy), idx, shape, R, S, stride, pt, pl, dx)
        return (dx if dtype == torch.bfloat16 else dx.to(dtype)), None, None, None, None


def max_pool2d(x, window, stride, pads=(0, 0, 0, 0), pad_is_zero=False):
    return _MaxPool2d.apply(x, tuple(window), int(stride), tuple(pads), bool(pad_is_zero))


def max_pool_supported(x, window, strides, pads=(0, 0, 0, 0)):
    R, S = window
    pt, pb, pl, pr = pads
    return (x.dim() == 4 and x.shape[-1] % 8 == 0 and strides[0] == strides[1] and R * S < 255
            and max(pt, pb) < R and max(pl, pr) < S
            and _out_size(x.shape[1], R, strides[0], pt, pb) >= 1 and _out_size(x.shape[2], S, strides[0], pl, pr) >= 1)


class _GlobalAvgPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        xb = _bf16(x)
        y = torch.empty(xb.shape[0], xb.shape[-1], dtype=torch.bfloat16, device=x.device)
        native.pool_gap_fwd(xb, y)
        ctx.geo = (tuple(xb.shape), x.dtype)
        return y

    @staticmethod
    def backward(ctx, gy):
        shape, dtype = ctx.geo
        dx = torch.empty(shape, dtype=torch.bfloat16, device=gy.device)
        native.pool_gap_bwd(_bf16(gy), dx)
        return dx if dtype == torch.bfloat16 else dx.to(dtype)


def global_avg_pool(x):
    return _GlobalAvgPool.apply(x)
