"""eb_allreduce_update_noalloc: the single-rank fused "apply averaged delta" step for hosts that may not
allocate device memory when they need it - e.g. a process whose GPU memory is entirely owned by its own
pool, or a loop that cannot tolerate the device-wide stall of cudaMalloc/cudaFree. Its pad / counter /
averaged-delta scratch lives in an array built into the library instead of caller-allocated buffers.

The entry point is called through the C ABI exactly as a C host would call it and must reproduce the
regular entry point (eb_allreduce_update, world = 1, two-shot) on two consecutive calls - the second
call sees the scratch contents the first one left behind. This file sorts last on purpose: a run that
stops at the first failure has then already executed every other test.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_ELEMS = 1 << 20


class _CommArgs(ctypes.Structure):
    _fields_ = [('delta', ctypes.c_void_p * 8), ('avg', ctypes.c_void_p * 8), ('pad', ctypes.c_void_p * 8),
                ('local', ctypes.c_void_p), ('w32', ctypes.c_void_p), ('w16', ctypes.c_void_p),
                ('out32', ctypes.c_void_p), ('n', ctypes.c_int64), ('rank', ctypes.c_int32),
                ('world', ctypes.c_int32), ('scale', ctypes.c_float), ('dtype_bf16', ctypes.c_int32),
                ('algo', ctypes.c_int32), ('blocks', ctypes.c_int32)]


@pytest.fixture(scope='module')
def lib():
    from elephas_b200.ops import build
    so = ctypes.CDLL(build.build())
    so.eb_last_error.restype = ctypes.c_char_p
    so.eb_allreduce_update.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    so.eb_allreduce_update_noalloc.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32, ctypes.c_float,
                                               ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    return so


def _reference_update(lib, delta, w32, w16, scale, state):
    """The existing entry point with caller-allocated scratch (kept across calls, like a real caller would)."""
    n = delta.numel()
    if 'pad' not in state:
        state['pad'] = torch.zeros(64, dtype=torch.int32, device='cuda')
        state['local'] = torch.zeros(16, dtype=torch.int32, device='cuda')
        state['avg'] = torch.zeros(n, dtype=delta.dtype, device='cuda')
    a = _CommArgs()
    a.delta[0], a.avg[0], a.pad[0] = delta.data_ptr(), state['avg'].data_ptr(), state['pad'].data_ptr()
    a.local, a.w32, a.w16 = state['local'].data_ptr(), w32.data_ptr(), w16.data_ptr()
    a.n, a.rank, a.world, a.scale = n, 0, 1, scale
    a.dtype_bf16, a.algo, a.blocks = int(delta.dtype == torch.bfloat16), 2, 0
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    assert lib.eb_allreduce_update(ctypes.byref(a), stream) == 0, lib.eb_last_error()


@pytest.mark.parametrize('dtype', [torch.bfloat16, torch.float32])
def test_matches_regular_entry_point_on_two_calls_in_a_row(lib, dtype):
    torch.manual_seed(11)
    n, scale = 300000 // 8 * 8, 0.25
    w_ref = torch.randn(n, device='cuda')
    w_new = w_ref.clone()
    h_ref = torch.zeros(n, dtype=torch.bfloat16, device='cuda')
    h_new = torch.zeros_like(h_ref)
    state = {}
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    for call in range(2):
        delta = torch.randn(n, device='cuda').to(dtype)
        _reference_update(lib, delta, w_ref, h_ref, scale, state)
        rc = lib.eb_allreduce_update_noalloc(delta.data_ptr(), n, int(dtype == torch.bfloat16), scale,
                                             w_new.data_ptr(), h_new.data_ptr(), stream)
        assert rc == 0, lib.eb_last_error()
        torch.cuda.synchronize()
        assert torch.allclose(w_new, w_ref, atol=1e-6, rtol=0), "call %d: fp32 weights differ" % call
        assert torch.equal(h_new, h_ref), "call %d: bf16 copy differs" % call
    # and both agree with the plain formula
    assert torch.equal(h_new, w_new.to(torch.bfloat16))


@pytest.mark.parametrize('dtype', [torch.bfloat16, torch.float32])
def test_matches_plain_pytorch_arithmetic_with_a_scale_that_is_not_a_power_of_two(lib, dtype):
    """Independent oracle: the documented formula evaluated with PyTorch in fp32, not another entry point of this
    library (a mistake shared by both entry points would pass the comparison above). scale = 1/3 is not exact in
    bf16, so this also pins down WHERE the bf16 path rounds: the averaged delta is rounded to bf16 before it is
    subtracted."""
    torch.manual_seed(5)
    n, scale = 200000 // 8 * 8, 1.0 / 3.0
    w = torch.randn(n, device='cuda')
    w_oracle = w.clone()
    h = torch.zeros(n, dtype=torch.bfloat16, device='cuda')
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    scale32 = torch.tensor(scale, dtype=torch.float32, device='cuda')
    for call in range(2):
        delta = torch.randn(n, device='cuda').to(dtype)
        step = delta.float() * scale32
        if dtype == torch.bfloat16:
            step = step.to(torch.bfloat16).float()
        w_oracle = w_oracle - step
        rc = lib.eb_allreduce_update_noalloc(delta.data_ptr(), n, int(dtype == torch.bfloat16), scale,
                                             w.data_ptr(), h.data_ptr(), stream)
        assert rc == 0, lib.eb_last_error()
        torch.cuda.synchronize()
        # (a bf16 delta rounded in the wrong place would be off by up to 2^-9 * |delta|, thousands of times this bound)
        assert torch.allclose(w, w_oracle, atol=1e-6, rtol=0), \
            "call %d: max |diff| %g" % (call, (w - w_oracle).abs().max().item())
        assert torch.equal(h, w.to(torch.bfloat16))


def test_rejects_payloads_beyond_the_built_in_scratch(lib):
    n = MAX_ELEMS + 8
    delta = torch.zeros(n, device='cuda')
    w32 = torch.ones(n, device='cuda')
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.eb_allreduce_update_noalloc(delta.data_ptr(), n, 0, 1.0, w32.data_ptr(), None, stream)
    assert rc != 0
    message = lib.eb_last_error().decode()
    assert 'exceeds the built-in scratch capacity' in message and str(MAX_ELEMS) in message
    torch.cuda.synchronize()
    assert torch.all(w32 == 1.0)  # nothing was launched
    # the largest accepted size works
    n = MAX_ELEMS
    delta, w32 = torch.full((n,), 2.0, device='cuda'), torch.ones(n, device='cuda')
    assert lib.eb_allreduce_update_noalloc(delta.data_ptr(), n, 0, 0.5, w32.data_ptr(), None, stream) == 0
    torch.cuda.synchronize()
    assert torch.all(w32 == 0.0)
