"""1x1 convolutions on the tcgen05 GEMM (first piece of K9).

A 1x1, stride-1 convolution over a channels-last activation ``[N, H, W, Cin]`` is the GEMM
``[N*H*W, Cin] x [Cout, Cin]^T`` - exactly the operand layouts of K1 - and two thirds of the convolutions of a
ResNet-50 are of this kind. ``conv1x1`` is an autograd function around three launches of ``eb_gemm_bf16``
(forward with the bias in the epilogue, input gradient with the MN-major weight operand, weight gradient with both
operands MN-major into fp32) plus the column-sum kernel for the bias gradient; activations travel in bf16, the master
weights stay fp32. All launches are graph-capturable, so the layer drops into the captured torch step
(``ops/torch_step.py``). Everything else (3x3, 7x7, strided convolutions) stays on cuDNN.

Opt-in: ``ELEPHAS_B200_NATIVE_CONV1X1=1``. Measured on the ResNet-50 shape (batch 64, bf16, one B200): 1,934 samples/s with
the 1x1 layers here against 2,208 with cuDNN for everything - the GEMMs are fine, the extra bf16 casts and the
separate bias-gradient pass around them are not fused the way cuDNN's are - so cuDNN stays the default.
"""
import os

import torch

from . import native


def enabled():
    return os.environ.get('ELEPHAS_B200_NATIVE_CONV1X1', '0') == '1'


def eligible(x, kernel, strides, dilation, groups, padding_applied):
    """CUDA tensor, bf16 compute requested by the caller, 1x1 / stride 1 / dense, channel counts multiples of 8."""
    return (x.is_cuda and kernel.shape[1] == 1 and kernel.shape[2] == 1 and tuple(strides) == (1, 1)
            and tuple(dilation) == (1, 1) and groups == 1 and not padding_applied
            and kernel.shape[0] % 8 == 0 and kernel.shape[3] % 8 == 0 and native.available())


class _Conv1x1(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, kernel, bias):
        cout, cin = kernel.shape[0], kernel.shape[3]
        x2 = x.reshape(-1, cin)
        xb = x2 if x2.dtype == torch.bfloat16 else x2.to(torch.bfloat16)
        xb = xb.contiguous()
        wb = kernel.detach().reshape(cout, cin).to(torch.bfloat16)
        m = xb.shape[0]
        out = torch.empty(m, cout, dtype=torch.bfloat16, device=x.device)
        native.gemm(xb, wb, m, cout, cin, out_bf16=out, bias=bias.detach().float() if bias is not None else None, act=0)
        ctx.save_for_backward(xb, wb)
        ctx.meta = (x.shape, x.dtype, kernel.shape, bias is not None)
        return out.view(*x.shape[:-1], cout)

    @staticmethod
    def backward(ctx, gy):
        xb, wb = ctx.saved_tensors
        x_shape, x_dtype, k_shape, has_bias = ctx.meta
        cout, cin = wb.shape
        m = xb.shape[0]
        g = gy.reshape(m, cout)
        g = (g if g.dtype == torch.bfloat16 else g.to(torch.bfloat16)).contiguous()
        dx = dw = db = None
        # autograd runs backward on its own thread: make the device (and with it the CUDA context the TMA descriptor
        # encoder needs) current there
        torch.cuda.set_device(g.device)
        if ctx.needs_input_grad[0]:
            dxb = torch.empty(m, cin, dtype=torch.bfloat16, device=g.device)
            native.gemm(g, wb, m, cin, cout, b_mn=True, out_bf16=dxb)          # dX = dY x W
            dx = dxb.view(x_shape).to(x_dtype)
        if ctx.needs_input_grad[1]:
            dw32 = torch.empty(cout, cin, dtype=torch.float32, device=g.device)
            native.gemm(g, xb, cout, cin, m, a_mn=True, b_mn=True, out_f32=dw32)  # dW = dY^T x X
            dw = dw32.view(k_shape)
        if has_bias and ctx.needs_input_grad[2]:
            db = torch.empty(cout, dtype=torch.float32, device=g.device)
            native.colsum_bf16(g, m, cout, db)
        return dx, dw, db


def conv1x1(x, kernel, bias):
    """``x`` ``[..., Cin]`` channels-last, ``kernel`` ``[Cout, Cin, 1, 1]`` (fp32 master), ``bias`` ``[Cout]`` or None;
    returns bf16 ``[..., Cout]``."""
    return _Conv1x1.apply(x, kernel, bias)
