"""ctypes bindings of ``libelephas_b200.so`` (C ABI declared in ``include/elephas_b200.h``).

Every wrapper takes torch CUDA tensors, passes raw device pointers and the *current* torch CUDA
stream, and raises ``NativeError`` with the library's message on failure - there is no silent
fallback: on a machine with a GPU a missing library is an error (``require()``), so a test that
passes has run the sm_100a kernels.
"""
import ctypes
import os
import threading

import torch

from . import build as _build

c_void_p, c_int32, c_int64, c_float, c_uint64 = (ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64, ctypes.c_float,
                                                   ctypes.c_uint64)

ACT_IDS = {'linear': 0, 'relu': 1, 'sigmoid': 2, 'tanh': 3, 'softplus': 4, 'softsign': 5, 'elu': 6, 'selu': 7,
           'exponential': 8, 'hard_sigmoid': 9, 'softmax': 100}
EPI_FWD, EPI_ACTGRAD = 0, 1
LOSS_IDS = {'mean_squared_error': 0, 'mean_absolute_error': 1, 'mean_absolute_percentage_error': 2,
            'mean_squared_logarithmic_error': 3, 'categorical_crossentropy': 4, 'binary_crossentropy': 5,
            'logcosh': 6}
METRIC_IDS = {'none': 0, 'categorical_accuracy': 1, 'binary_accuracy': 2, 'mean_absolute_error': 3,
              'mean_squared_error': 4, 'mean_absolute_percentage_error': 5}
OPT_IDS = {'SGD': 1, 'RMSprop': 2, 'Adam': 3, 'Adagrad': 4, 'Adadelta': 5, 'Adamax': 6}


class NativeError(RuntimeError):
    pass


class GemmArgs(ctypes.Structure):
    _fields_ = [('M', c_int32), ('N', c_int32), ('K', c_int32),
                ('A', c_void_p), ('lda', c_int32), ('a_mn', c_int32),
                ('B', c_void_p), ('ldb', c_int32), ('b_mn', c_int32),
                ('alpha', c_float), ('epi', c_int32), ('act', c_int32),
                ('bias', c_void_p), ('aux', c_void_p), ('ldaux', c_int32),
                ('drop_p', c_float), ('seed', c_uint64), ('drop_stream', c_uint64), ('step_ptr', c_void_p),
                ('out_bf16', c_void_p), ('ldo', c_int32),
                ('out_f32', c_void_p), ('ldo32', c_int32), ('beta', c_float), ('bf16_from_f32', c_int32),
                ('block_n', c_int32), ('cta_group', c_int32), ('multicast', c_int32), ('no_tail_split', c_int32),
                ('b_copy', c_void_p), ('operands_local', c_int32)]


class HeadArgs(ctypes.Structure):
    _fields_ = [('rows', c_int32), ('K', c_int32), ('out', c_int32),
                ('A', c_void_p), ('lda', c_int64), ('W', c_void_p), ('ldw', c_int64), ('bias', c_void_p),
                ('z', c_void_p), ('ldz', c_int64), ('dz', c_void_p), ('lddz', c_int64),
                ('dw', c_void_p), ('lddw', c_int64), ('aux', c_void_p), ('ldaux', c_int64),
                ('dzp', c_void_p), ('lddzp', c_int64), ('act', c_int32)]


MLP_MAX_LAYERS = 4


class MlpArgs(ctypes.Structure):
    _fields_ = [('rows', c_int32), ('layers', c_int32), ('dims', c_int32 * (MLP_MAX_LAYERS + 1)),
                ('x', c_void_p), ('ldx', c_int64), ('w', c_void_p * MLP_MAX_LAYERS), ('ldw', c_int64 * MLP_MAX_LAYERS),
                ('bias', c_void_p * MLP_MAX_LAYERS), ('act', c_int32 * MLP_MAX_LAYERS), ('z', c_void_p),
                ('ldz', c_int64)]


class LossArgs(ctypes.Structure):
    _fields_ = [('z', c_void_p), ('ldz', c_int64), ('y', c_void_p), ('ldy', c_int64),
                ('out', c_void_p), ('ldo', c_int64), ('dz', c_void_p), ('lddz', c_int64),
                ('stats', c_void_p), ('rows', c_int32), ('cols', c_int32), ('loss', c_int32), ('act', c_int32),
                ('metrics', c_int32 * 4), ('grad_scale', c_float)]


class OptimArgs(ctypes.Structure):
    _fields_ = [('kind', c_int32), ('n', c_int64), ('w32', c_void_p), ('g32', c_void_p),
                ('slot0', c_void_p), ('slot1', c_void_p), ('slot2', c_void_p), ('w16', c_void_p),
                ('delta_bf16', c_void_p), ('delta_f32', c_void_p), ('apply', c_int32), ('advance', c_int32),
                ('lr', c_float), ('decay', c_float), ('momentum', c_float), ('rho', c_float), ('eps', c_float),
                ('beta1', c_float), ('beta2', c_float), ('nesterov', c_int32), ('centered', c_int32),
                ('amsgrad', c_int32), ('step', c_void_p)]


class CommArgs(ctypes.Structure):
    _fields_ = [('delta', c_void_p * 8), ('avg', c_void_p * 8), ('pad', c_void_p * 8), ('local', c_void_p),
                ('w32', c_void_p), ('w16', c_void_p), ('out32', c_void_p), ('n', c_int64),
                ('rank', c_int32), ('world', c_int32), ('scale', c_float), ('dtype_bf16', c_int32),
                ('algo', c_int32), ('blocks', c_int32)]


PS_MAX_RANKS, PS_MAX_SPANS, PS_MAX_CHUNKS, PS_LOCK_WORDS = 8, 48, 64, 8


class PsSpan(ctypes.Structure):
    _fields_ = [('flat_off', c_int64), ('n', c_int64), ('owner_off', c_int64), ('owner', c_int32), ('reserved', c_int32)]


class PsStore(ctypes.Structure):
    _fields_ = [('master', c_void_p * PS_MAX_RANKS), ('mirror', c_void_p * PS_MAX_RANKS), ('ctrl', c_void_p * PS_MAX_RANKS),
                ('stripe_len', c_int64 * PS_MAX_RANKS), ('chunk_len', c_int64), ('chunks', c_int32),
                ('world', c_int32), ('rank', c_int32), ('nspans', c_int32), ('spans', PsSpan * PS_MAX_SPANS)]


class ConvArgs(ctypes.Structure):
    _fields_ = [('inp', c_void_p), ('N', c_int32), ('H', c_int32), ('W', c_int32), ('C', c_int32),
                ('w', c_void_p), ('K', c_int32), ('R', c_int32), ('S', c_int32),
                ('w_mn', c_int32), ('flip_taps', c_int32), ('stride', c_int32), ('pad_h', c_int32), ('pad_w', c_int32),
                ('P', c_int32), ('Q', c_int32), ('pn', c_int32), ('ph', c_int32), ('pw', c_int32),
                ('out', c_void_p), ('OH', c_int32), ('OW', c_int32), ('os', c_int32), ('bias', c_void_p),
                ('block_n', c_int32)]


class ConvWgradArgs(ctypes.Structure):
    _fields_ = [('x', c_void_p), ('N', c_int32), ('H', c_int32), ('W', c_int32), ('C', c_int32),
                ('dy', c_void_p), ('P', c_int32), ('Q', c_int32), ('K', c_int32),
                ('R', c_int32), ('S', c_int32), ('stride', c_int32), ('pad_h', c_int32), ('pad_w', c_int32),
                ('pn', c_int32), ('ph', c_int32), ('pw', c_int32), ('dw', c_void_p), ('splits', c_int32)]


_lib = None
_lib_lock = threading.Lock()


def lib_path():
    return _build.LIB_PATH


def available() -> bool:
    return os.path.exists(_build.LIB_PATH)


def load():
    """Load (building first if the sources are newer and nvcc is present) and return the library."""
    global _lib
    with _lib_lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(_build.LIB_PATH):
            _build.build()
        lib = ctypes.CDLL(_build.LIB_PATH, mode=ctypes.RTLD_GLOBAL)
        lib.eb_last_error.restype = ctypes.c_char_p
        lib.eb_launch_count.restype = c_uint64
        lib.eb_version.restype = c_int32
        for name in ('eb_gemm_bf16', 'eb_loss_fwd_bwd', 'eb_optimizer_step', 'eb_allreduce_update',
                     'eb_allreduce_sum', 'eb_allreduce_shard_update', 'eb_head_fwd', 'eb_head_wgrad', 'eb_head_dgrad',
                     'eb_mlp_forward'):
            getattr(lib, name).argtypes = [c_void_p, c_void_p]
        lib.eb_stage_f32_to_bf16.argtypes = [c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int32, c_void_p]
        lib.eb_cast_f32_to_bf16.argtypes = [c_void_p, c_void_p, c_int64, c_void_p]
        lib.eb_colsum_bf16.argtypes = [c_void_p, c_int64, c_int32, c_int32, c_void_p, c_int32, c_void_p]
        lib.eb_act_dropout_fwd.argtypes = [c_void_p, c_void_p, c_int64, c_int32, c_int32, c_int32, c_float, c_uint64,
                                           c_uint64, c_void_p, c_void_p]
        lib.eb_act_dropout_bwd.argtypes = [c_void_p, c_void_p, c_void_p, c_int64, c_int32, c_int32, c_int32, c_float,
                                           c_uint64, c_uint64, c_void_p, c_void_p]
        lib.eb_delta.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p]
        lib.eb_ps_lock.argtypes = [c_void_p, c_void_p]
        lib.eb_ps_unlock.argtypes = [c_void_p, c_void_p]
        lib.eb_ps_push.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p]
        lib.eb_ps_pull.argtypes = [c_void_p, c_void_p, c_void_p, c_int64, c_void_p]
        lib.eb_ps_push_mirror.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p]
        lib.eb_ps_striped_pull.argtypes = [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_int64, c_int64, c_int32, c_void_p]
        lib.eb_ps_striped_push.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int32, c_void_p]
        lib.eb_ps_striped_optimizer_push.argtypes = [c_void_p, c_void_p, c_int32, c_void_p]
        lib.eb_conv_igemm.argtypes = [c_void_p, c_void_p]
        lib.eb_conv_wgrad.argtypes = [c_void_p, c_void_p]
        lib.eb_conv_im2col.argtypes = [c_void_p, c_int32] + [c_int32] * 10 + [c_void_p, c_int32, c_void_p]
        lib.eb_nvls_store.argtypes = [c_void_p, c_void_p, c_int64, c_int32, c_void_p]
        lib.eb_pool_max_fwd.argtypes = [c_void_p] + [c_int32] * 12 + [c_void_p, c_void_p, c_void_p]
        lib.eb_pool_max_bwd.argtypes = [c_void_p, c_void_p] + [c_int32] * 11 + [c_void_p, c_void_p]
        lib.eb_pool_gap_fwd.argtypes = [c_void_p, c_int32, c_int32, c_int32, c_void_p, c_void_p]
        lib.eb_pool_gap_bwd.argtypes = [c_void_p, c_int32, c_int32, c_int32, c_void_p, c_void_p]
        lib.eb_bn_stats.argtypes = [c_void_p, c_int64, c_int32, c_void_p, c_int32, c_float, c_float, c_void_p, c_void_p,
                                    c_void_p, c_void_p, c_void_p]
        lib.eb_bn_apply.argtypes = [c_void_p, c_int64, c_int32, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                    c_int32, c_void_p, c_void_p]
        lib.eb_bn_bwd_reduce.argtypes = [c_void_p, c_void_p, c_void_p, c_int64, c_int32, c_void_p, c_void_p, c_int32,
                                         c_void_p, c_int32, c_void_p, c_void_p, c_int32, c_void_p]
        lib.eb_bn_bwd_apply.argtypes = [c_void_p, c_void_p, c_void_p, c_int64, c_int32, c_void_p, c_void_p, c_void_p,
                                        c_void_p, c_void_p, c_int32, c_void_p, c_void_p, c_void_p]
        lib.eb_host_cast_isa.restype = ctypes.c_char_p
        lib.eb_host_f32_to_bf16.argtypes = [c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int32, c_int32]
        lib.eb_allreduce_nvls_update.argtypes = [c_void_p, c_void_p]
        lib.eb_vmm_supported.argtypes = [ctypes.POINTER(c_int32), ctypes.POINTER(c_int32)]
        lib.eb_vmm_create.argtypes = [c_int64, c_int32, ctypes.POINTER(c_uint64), ctypes.POINTER(c_int64)]
        lib.eb_vmm_export_fd.argtypes = [c_uint64, ctypes.POINTER(c_int32)]
        lib.eb_vmm_import_fd.argtypes = [c_int32, ctypes.POINTER(c_uint64)]
        lib.eb_vmm_map.argtypes = [c_uint64, c_int64, ctypes.POINTER(c_void_p)]
        lib.eb_vmm_unmap.argtypes = [c_void_p, c_int64]
        lib.eb_vmm_release.argtypes = [c_uint64]
        lib.eb_mc_create.argtypes = [c_int32, c_int64, ctypes.POINTER(c_uint64)]
        lib.eb_mc_add_device.argtypes = [c_uint64]
        lib.eb_mc_bind.argtypes = [c_uint64, c_uint64, c_int64]
        lib.eb_device_malloc.argtypes = [ctypes.POINTER(c_void_p), c_int64]
        lib.eb_device_free.argtypes = [c_void_p]
        lib.eb_ipc_export.argtypes = [c_void_p, c_void_p]
        lib.eb_ipc_import.argtypes = [c_void_p, ctypes.POINTER(c_void_p)]
        lib.eb_ipc_close.argtypes = [c_void_p]
        lib.eb_enable_peer_access.argtypes = [c_int32]
        _lib = lib
        return lib


def require():
    """The library, or a loud failure: used on every GPU code path."""
    try:
        return load()
    except Exception as exc:
        raise NativeError("libelephas_b200.so is required on a CUDA device but could not be loaded/built: %s" % exc)


def _check(rc, what):
    if rc != 0:
        raise NativeError("%s failed: %s" % (what, load().eb_last_error().decode('utf8', 'replace')))


def _ptr(t):
    return None if t is None else c_void_p(t.data_ptr())


def _stream(device=None):
    return c_void_p(torch.cuda.current_stream(device).cuda_stream)


_replayed = [0]


def count_replay(n: int):
    """Kernels launched by a CUDA-graph replay do not pass through the library's launch wrappers;
    the engine reports them here so that ``launch_count`` stays the number of native kernels run."""
    _replayed[0] += int(n)


def launch_count() -> int:
    base = int(load().eb_launch_count()) if _lib is not None or available() else 0
    return base + _replayed[0]


# ----------------------------------------------------------------------- GEMM
def gemm(A, B, M, N, K, a_mn=False, b_mn=False, lda=None, ldb=None, out_bf16=None, ldo=None, out_f32=None,
         ldo32=None, beta=0.0, alpha=1.0, epi=EPI_FWD, act=0, bias=None, aux=None, ldaux=None, drop_p=0.0, seed=0,
         drop_stream=0, step_ptr=None, block_n=0, cta_group=0, bf16_from_f32=False, b_copy=None):
    """``C[M,N] = epilogue(A x B^T)`` on the tcgen05 GEMM (see ``EbGemmArgs`` for the operand layouts).

    ``B`` may be a raw device address (``int``, e.g. a peer-mapped parameter-server buffer; ``ldb`` is then
    required); with ``b_copy`` (a local bf16 tensor of B's layout) the GEMM doubles as the weight pull (K2)."""
    g = GemmArgs()
    g.M, g.N, g.K = int(M), int(N), int(K)
    g.A, g.lda, g.a_mn = A.data_ptr(), int(lda if lda is not None else A.stride(0)), int(bool(a_mn))
    if isinstance(B, int):
        g.B, g.ldb, g.b_mn = B, int(ldb), int(bool(b_mn))
    else:
        g.B, g.ldb, g.b_mn = B.data_ptr(), int(ldb if ldb is not None else B.stride(0)), int(bool(b_mn))
    g.b_copy = b_copy.data_ptr() if b_copy is not None else None
    g.alpha, g.epi, g.act = float(alpha), int(epi), int(act)
    g.bias = bias.data_ptr() if bias is not None else None
    g.aux = aux.data_ptr() if aux is not None else None
    g.ldaux = int(ldaux if ldaux is not None else (aux.stride(0) if aux is not None else 0))
    g.drop_p, g.seed, g.drop_stream = float(drop_p), int(seed) & (2 ** 64 - 1), int(drop_stream) & (2 ** 64 - 1)
    g.step_ptr = step_ptr.data_ptr() if step_ptr is not None else None
    g.out_bf16 = out_bf16.data_ptr() if out_bf16 is not None else None
    g.ldo = int(ldo if ldo is not None else (out_bf16.stride(0) if out_bf16 is not None else 0))
    g.out_f32 = out_f32.data_ptr() if out_f32 is not None else None
    g.ldo32 = int(ldo32 if ldo32 is not None else (out_f32.stride(0) if out_f32 is not None else 0))
    g.beta = float(beta)
    g.bf16_from_f32 = int(bool(bf16_from_f32))
    g.block_n, g.cta_group = int(block_n), int(cta_group)
    # torch tensors on the launching device are this GPU's memory; a raw address (a parameter server's stripe, usually peer
    # memory) is not declared local, which keeps the kernel on single-CTA, non-multicast TMA loads (TMA multicast loads
    # here read only this GPU's memory)
    g.operands_local = int(not isinstance(B, int) and b_copy is None and A.device == B.device)
    _check(require().eb_gemm_bf16(ctypes.byref(g), _stream(A.device)), 'eb_gemm_bf16')


# ------------------------------------------------------------- fused MLP forward (K8)
def mlp_forward_supported(in_dim, units):
    """Can ``eb_mlp_forward`` run a Dense chain with these widths? 2-4 layers, hidden widths multiples of 16, 8-aligned
    input width; every layer up to 128 units (weights resident in shared memory), or up to 512 units with an input width
    of at most 512 (weights streamed through the TMA ring, activations resident)."""
    units = list(units)
    if not (2 <= len(units) <= MLP_MAX_LAYERS and in_dim % 8 == 0 and all(u % 16 == 0 and u > 0 for u in units[:-1])
            and units[-1] > 0):
        return False
    if max(units) <= 128:
        return True
    return max(units) <= 512 and in_dim <= 512


def mlp_forward(x16, rows, in_dim, layers, z, ldx=None):
    """Whole-MLP forward in one persistent kernel. ``layers``: list of ``(w16 [units, ld], units, bias32, act_id)``;
    ``z``: fp32 ``[rows, >= units_last]`` receives the last layer's pre-activations."""
    a = MlpArgs()
    a.rows, a.layers = int(rows), len(layers)
    a.dims[0] = int(in_dim)
    a.x, a.ldx = x16.data_ptr(), int(ldx if ldx is not None else x16.stride(0))
    for i, (w, units, bias, act) in enumerate(layers):
        a.dims[i + 1] = int(units)
        a.w[i], a.ldw[i] = w.data_ptr(), int(w.stride(0))
        a.bias[i] = bias.data_ptr() if bias is not None else None
        a.act[i] = int(act)
    a.z, a.ldz = z.data_ptr(), z.stride(0)
    _check(require().eb_mlp_forward(ctypes.byref(a), _stream(x16.device)), 'eb_mlp_forward')


# ------------------------------------------------------------- narrow Dense layers
HEAD_MAX_OUT = 8


def _head_args(rows, K, out):
    a = HeadArgs()
    a.rows, a.K, a.out = int(rows), int(K), int(out)
    return a


def head_fwd(A, W, rows, K, out, z, bias=None, lda=None, ldw=None):
    """``z[:rows, :out] = A[:rows, :K] @ W[:out, :K].T + bias`` for ``out <= 8`` (fp32 ``z``)."""
    a = _head_args(rows, K, out)
    a.A, a.lda = A.data_ptr(), int(lda if lda is not None else A.stride(0))
    a.W, a.ldw = W.data_ptr(), int(ldw if ldw is not None else W.stride(0))
    a.bias = bias.data_ptr() if bias is not None else None
    a.z, a.ldz = z.data_ptr(), z.stride(0)
    _check(require().eb_head_fwd(ctypes.byref(a), _stream(A.device)), 'eb_head_fwd')


def head_wgrad(dz, A, rows, K, out, dw, lddz=None, lda=None, lddw=None):
    """``dw[:out, :K] = dz[:rows, :out].T @ A[:rows, :K]`` (fp32 ``dw``, overwritten)."""
    a = _head_args(rows, K, out)
    a.A, a.lda = A.data_ptr(), int(lda if lda is not None else A.stride(0))
    a.W, a.ldw = A.data_ptr(), 8  # unused by this entry point (kept valid for the shared argument check)
    a.dz, a.lddz = dz.data_ptr(), int(lddz if lddz is not None else dz.stride(0))
    a.dw, a.lddw = dw.data_ptr(), int(lddw if lddw is not None else K)
    _check(require().eb_head_wgrad(ctypes.byref(a), _stream(A.device)), 'eb_head_wgrad')


def head_dgrad(dz, W, aux, dzp, rows, K, out, act=0, lddz=None, ldw=None, ldaux=None, lddzp=None):
    """``dzp[:rows, :K] = (dz[:rows, :out] @ W[:out, :K]) * act'(aux)`` (bf16 ``dzp``)."""
    a = _head_args(rows, K, out)
    a.A, a.lda = aux.data_ptr(), 8  # unused here
    a.W, a.ldw = W.data_ptr(), int(ldw if ldw is not None else W.stride(0))
    a.dz, a.lddz = dz.data_ptr(), int(lddz if lddz is not None else dz.stride(0))
    a.aux, a.ldaux = aux.data_ptr(), int(ldaux if ldaux is not None else aux.stride(0))
    a.dzp, a.lddzp = dzp.data_ptr(), int(lddzp if lddzp is not None else dzp.stride(0))
    a.act = int(act)
    _check(require().eb_head_dgrad(ctypes.byref(a), _stream(W.device)), 'eb_head_dgrad')


# ---------------------------------------------------------------- elementwise
def stage_f32_to_bf16(src, dst, rows, cols):
    _check(require().eb_stage_f32_to_bf16(_ptr(src), src.stride(0), _ptr(dst), dst.stride(0), int(rows), int(cols),
                                          _stream(src.device)), 'eb_stage_f32_to_bf16')


def cast_f32_to_bf16(src, dst, n=None):
    _check(require().eb_cast_f32_to_bf16(_ptr(src), _ptr(dst), int(n if n is not None else src.numel()),
                                         _stream(src.device)), 'eb_cast_f32_to_bf16')


def colsum_bf16(x, rows, cols, out, accumulate=False, ld=None):
    _check(require().eb_colsum_bf16(_ptr(x), int(ld if ld is not None else x.stride(0)), int(rows), int(cols),
                                    _ptr(out), int(bool(accumulate)), _stream(x.device)), 'eb_colsum_bf16')


def act_dropout_fwd(x, y, rows, cols, act=0, drop_p=0.0, seed=0, drop_stream=0, ld=None, step_ptr=None):
    _check(require().eb_act_dropout_fwd(_ptr(x), _ptr(y), int(ld if ld is not None else x.stride(0)), int(rows),
                                        int(cols), int(act), float(drop_p), int(seed), int(drop_stream),
                                        _ptr(step_ptr), _stream(x.device)), 'eb_act_dropout_fwd')


def act_dropout_bwd(dy, y, dx, rows, cols, act=0, drop_p=0.0, seed=0, drop_stream=0, ld=None, step_ptr=None):
    _check(require().eb_act_dropout_bwd(_ptr(dy), _ptr(y), _ptr(dx), int(ld if ld is not None else y.stride(0)),
                                        int(rows), int(cols), int(act), float(drop_p), int(seed), int(drop_stream),
                                        _ptr(step_ptr), _stream(y.device)), 'eb_act_dropout_bwd')


def delta(before, after, out_f32=None, out_bf16=None, n=None):
    _check(require().eb_delta(_ptr(before), _ptr(after), _ptr(out_f32), _ptr(out_bf16),
                              int(n if n is not None else before.numel()), _stream(before.device)), 'eb_delta')


# ----------------------------------------------------------------------- loss
def loss_fwd_bwd(z, y, rows, cols, loss, act, out=None, dz=None, stats=None, metrics=(0, 0, 0, 0), grad_scale=1.0,
                 ldz=None, ldy=None, ldo=None, lddz=None):
    a = LossArgs()
    a.z, a.ldz = z.data_ptr(), int(ldz if ldz is not None else z.stride(0))
    a.y = y.data_ptr()
    a.ldy = int(ldy if ldy is not None else (y.stride(0) if y.dim() > 1 else 0))
    a.out = out.data_ptr() if out is not None else None
    a.ldo = int(ldo if ldo is not None else (out.stride(0) if out is not None else 0))
    a.dz = dz.data_ptr() if dz is not None else None
    a.lddz = int(lddz if lddz is not None else (dz.stride(0) if dz is not None else 0))
    a.stats = stats.data_ptr() if stats is not None else None
    a.rows, a.cols, a.loss, a.act = int(rows), int(cols), int(loss), int(act)
    m = list(metrics) + [0] * (4 - len(metrics))
    a.metrics = (c_int32 * 4)(*m[:4])
    a.grad_scale = float(grad_scale)
    _check(require().eb_loss_fwd_bwd(ctypes.byref(a), _stream(z.device)), 'eb_loss_fwd_bwd')


# ------------------------------------------------------------------ optimizer
def optimizer_step(kind, n, g32, step, w32=None, w16=None, slots=(None, None, None), delta_bf16=None, delta_f32=None,
                   apply=True, advance=True, lr=0.01, decay=0.0, momentum=0.0, rho=0.9, eps=1e-7, beta1=0.9,
                   beta2=0.999, nesterov=False, centered=False, amsgrad=False):
    a = _optim_args(kind, n, g32, step, w32, w16, slots, delta_bf16, delta_f32, apply, advance, lr, decay, momentum, rho,
                    eps, beta1, beta2, nesterov, centered, amsgrad)
    _check(require().eb_optimizer_step(ctypes.byref(a), _stream(g32.device)), 'eb_optimizer_step')


def _optim_args(kind, n, g32, step, w32=None, w16=None, slots=(None, None, None), delta_bf16=None, delta_f32=None,
                apply=True, advance=True, lr=0.01, decay=0.0, momentum=0.0, rho=0.9, eps=1e-7, beta1=0.9,
                beta2=0.999, nesterov=False, centered=False, amsgrad=False):
    a = OptimArgs()
    a.kind, a.n = int(kind), int(n)
    a.w32, a.g32 = (w32.data_ptr() if w32 is not None else None), g32.data_ptr()
    s = list(slots) + [None] * (3 - len(slots))
    a.slot0, a.slot1, a.slot2 = [(t.data_ptr() if t is not None else None) for t in s[:3]]
    a.w16 = w16.data_ptr() if w16 is not None else None
    a.delta_bf16 = delta_bf16.data_ptr() if delta_bf16 is not None else None
    a.delta_f32 = delta_f32.data_ptr() if delta_f32 is not None else None
    a.apply, a.advance = int(bool(apply)), int(bool(advance))
    a.lr, a.decay, a.momentum, a.rho, a.eps = float(lr), float(decay), float(momentum), float(rho), float(eps)
    a.beta1, a.beta2 = float(beta1), float(beta2)
    a.nesterov, a.centered, a.amsgrad = int(bool(nesterov)), int(bool(centered)), int(bool(amsgrad))
    a.step = step.data_ptr()
    return a


# ------------------------------------------------------------------ collectives
def _comm_args(delta_ptrs, pad_ptrs, local, n, rank, world, scale, dtype_bf16, avg_ptrs=None, w32=None, w16=None,
               out32=None, algo=0, blocks=0):
    a = CommArgs()
    for i in range(8):
        a.delta[i] = delta_ptrs[i] if i < world else None
        a.pad[i] = pad_ptrs[i] if i < world else None
        a.avg[i] = avg_ptrs[i] if (avg_ptrs is not None and i < world) else None
    a.local = local.data_ptr() if hasattr(local, 'data_ptr') else local
    a.w32 = w32.data_ptr() if w32 is not None else None
    a.w16 = w16.data_ptr() if w16 is not None else None
    a.out32 = out32.data_ptr() if out32 is not None else None
    a.n, a.rank, a.world, a.scale = int(n), int(rank), int(world), float(scale)
    a.dtype_bf16, a.algo, a.blocks = int(bool(dtype_bf16)), int(algo), int(blocks)
    return a


def allreduce_update(delta_ptrs, pad_ptrs, local, n, rank, world, scale, dtype_bf16, w32, w16=None, avg_ptrs=None,
                     algo=0, blocks=0, device=None):
    a = _comm_args(delta_ptrs, pad_ptrs, local, n, rank, world, scale, dtype_bf16, avg_ptrs, w32, w16, None, algo,
                   blocks)
    _check(require().eb_allreduce_update(ctypes.byref(a), _stream(device)), 'eb_allreduce_update')


def allreduce_shard_update(delta_ptrs, pad_ptrs, local, n, rank, world, scale, dtype_bf16, w32, w16_ptrs, blocks=0,
                           device=None):
    """Sharded-master fused all-reduce + update: ``w16_ptrs`` are every rank's peer-mapped bf16 weight buffers."""
    a = _comm_args(delta_ptrs, pad_ptrs, local, n, rank, world, scale, dtype_bf16, w16_ptrs, w32, None, None, 0,
                   blocks)
    _check(require().eb_allreduce_shard_update(ctypes.byref(a), _stream(device)), 'eb_allreduce_shard_update')


def allreduce_nvls_update(mc_delta, mc_w16, pad_ptrs, local, n, rank, world, scale, w32, blocks=0, device=None):
    """NVLS fused all-reduce + update: ``mc_delta`` / ``mc_w16`` are MULTICAST addresses (``symm.VmmSymmetricMemory``)."""
    a = _comm_args([mc_delta] + [None] * 7, pad_ptrs, local, n, rank, world, scale, True, [mc_w16] + [None] * 7, w32, None,
                   None, 0, blocks)
    a.delta[0], a.avg[0] = mc_delta, mc_w16
    _check(require().eb_allreduce_nvls_update(ctypes.byref(a), _stream(device)), 'eb_allreduce_nvls_update')


def nvls_store(src, mc_dst, nbytes=None, blocks=0, device=None):
    """Replicate ``src`` (tensor) into every rank's symmetric buffer through its multicast address ``mc_dst`` (int)."""
    nbytes = src.numel() * src.element_size() if nbytes is None else int(nbytes)
    _check(require().eb_nvls_store(src.data_ptr(), mc_dst, nbytes, int(blocks), _stream(device if device is not None else src.device)),
           'eb_nvls_store')


def allreduce_sum(delta_ptrs, pad_ptrs, local, n, rank, world, dtype_bf16, out32, scale=1.0, blocks=0, device=None):
    a = _comm_args(delta_ptrs, pad_ptrs, local, n, rank, world, scale, dtype_bf16, None, None, None, out32, 1, blocks)
    _check(require().eb_allreduce_sum(ctypes.byref(a), _stream(device)), 'eb_allreduce_sum')


# ------------------------------------------------------------ parameter server
def ps_lock(ctrl, device=None):
    """Enqueue the ticket-lock kernel of a device parameter server (pair every call with :func:`ps_unlock`)."""
    cptr = ctrl.data_ptr() if hasattr(ctrl, 'data_ptr') else ctrl
    _check(require().eb_ps_lock(cptr, _stream(device if device is not None else ctrl.device)), 'eb_ps_lock')


def ps_unlock(ctrl, device=None):
    cptr = ctrl.data_ptr() if hasattr(ctrl, 'data_ptr') else ctrl
    _check(require().eb_ps_unlock(cptr, _stream(device if device is not None else ctrl.device)), 'eb_ps_unlock')


def ps_pull(master, ctrl, dst_f32, dst_bf16=None, locked=False, n=None, offset=0):
    lib = require()
    s = _stream(dst_f32.device if dst_f32 is not None else dst_bf16.device)
    n = int(n if n is not None else (dst_f32.numel() if dst_f32 is not None else dst_bf16.numel()))
    mptr = (master.data_ptr() if hasattr(master, 'data_ptr') else master) + 4 * int(offset)
    cptr = ctrl.data_ptr() if hasattr(ctrl, 'data_ptr') else ctrl
    if locked:
        _check(lib.eb_ps_lock(cptr, s), 'eb_ps_lock')
    try:
        _check(lib.eb_ps_pull(mptr, _ptr(dst_f32), _ptr(dst_bf16), n, s), 'eb_ps_pull')
    finally:
        if locked:  # a lock that was enqueued is always followed by its unlock, whatever happens in between
            _check(lib.eb_ps_unlock(cptr, s), 'eb_ps_unlock')


def ps_push(master, ctrl, delta=None, before=None, after=None, locked=False, n=None, mirror=None, delta_bf16=None,
            offset=0):
    """``master -= delta``. ``mirror`` (address of the store's bf16 mirror) switches to the kernel that also refreshes
    the mirror; ``delta_bf16`` is accepted only there. ``offset``: first element of master / mirror this push covers."""
    lib = require()
    ref = delta if delta is not None else (delta_bf16 if delta_bf16 is not None else before)
    s = _stream(ref.device)
    n = int(n if n is not None else ref.numel())
    mptr = (master.data_ptr() if hasattr(master, 'data_ptr') else master) + 4 * int(offset)
    cptr = ctrl.data_ptr() if hasattr(ctrl, 'data_ptr') else ctrl
    if delta_bf16 is not None and mirror is None:
        raise NativeError("a bf16 delta can only be pushed into a store with a bf16 mirror")
    if locked:
        _check(lib.eb_ps_lock(cptr, s), 'eb_ps_lock')
    try:
        if mirror is not None:
            mir = (mirror.data_ptr() if hasattr(mirror, 'data_ptr') else mirror) + 2 * int(offset)
            _check(lib.eb_ps_push_mirror(mptr, mir, _ptr(delta), _ptr(delta_bf16), _ptr(before), _ptr(after), n, s),
                   'eb_ps_push_mirror')
        else:
            _check(lib.eb_ps_push(mptr, _ptr(delta), _ptr(before), _ptr(after), n, s), 'eb_ps_push')
    finally:
        if locked:
            _check(lib.eb_ps_unlock(cptr, s), 'eb_ps_unlock')


# ---------------------------------------------- parameter server, striped over all ranks
def ps_store_args(master_ptrs, mirror_ptrs, ctrl_ptrs, stripe_len, chunk_len, chunks, rank, spans):
    """``EbPsStore`` from Python values. ``spans``: iterable of ``(flat_off, n, owner, owner_off)``."""
    world = len(master_ptrs)
    spans = list(spans)
    if world > PS_MAX_RANKS or len(spans) > PS_MAX_SPANS or chunks > PS_MAX_CHUNKS:
        raise NativeError("striped parameter server: %d ranks / %d spans / %d chunks exceed the kernel limits (%d / %d / %d)"
                          % (world, len(spans), chunks, PS_MAX_RANKS, PS_MAX_SPANS, PS_MAX_CHUNKS))
    a = PsStore()
    for r in range(world):
        a.master[r], a.mirror[r], a.ctrl[r] = master_ptrs[r], mirror_ptrs[r], ctrl_ptrs[r]
        a.stripe_len[r] = int(stripe_len[r])
    a.chunk_len, a.chunks, a.world, a.rank, a.nspans = int(chunk_len), int(chunks), world, int(rank), len(spans)
    for i, (flat_off, n, owner, owner_off) in enumerate(spans):
        a.spans[i].flat_off, a.spans[i].n, a.spans[i].owner_off, a.spans[i].owner = int(flat_off), int(n), int(owner_off), int(owner)
    return a


def ps_striped_pull(store, dst_f32=None, f32_range=None, dst_bf16=None, b16_range=None, seqlock=False, device=None):
    """Flat-indexed pull: ``dst_f32[i] = master[i]`` for ``i`` in ``f32_range``, ``dst_bf16[i] = mirror[i]`` for ``i`` in
    ``b16_range`` (ranges default to the whole destination)."""
    ref = dst_f32 if dst_f32 is not None else dst_bf16
    f_lo, f_hi = f32_range if f32_range is not None else (0, dst_f32.numel() if dst_f32 is not None else 0)
    b_lo, b_hi = b16_range if b16_range is not None else (0, dst_bf16.numel() if dst_bf16 is not None else 0)
    _check(require().eb_ps_striped_pull(ctypes.byref(store), _ptr(dst_f32), int(f_lo), int(f_hi), _ptr(dst_bf16), int(b_lo),
                                        int(b_hi), int(bool(seqlock)), _stream(device if device is not None else ref.device)),
           'eb_ps_striped_pull')


def ps_striped_push(store, delta_f32=None, delta_bf16=None, before=None, after=None, locked=False, device=None):
    ref = delta_f32 if delta_f32 is not None else (delta_bf16 if delta_bf16 is not None else before)
    _check(require().eb_ps_striped_push(ctypes.byref(store), _ptr(delta_f32), _ptr(delta_bf16), _ptr(before), _ptr(after),
                                        int(bool(locked)), _stream(device if device is not None else ref.device)),
           'eb_ps_striped_push')


def ps_striped_optimizer_push(store, kind, n, g32, step, slots=(None, None, None), locked=False, advance=True, **hyper):
    """Optimizer step fused with the push into the striped store (``master -= step`` as red.add into peer memory)."""
    a = _optim_args(kind, n, g32, step, slots=slots, apply=False, advance=advance, **hyper)
    _check(require().eb_ps_striped_optimizer_push(ctypes.byref(store), ctypes.byref(a), int(bool(locked)), _stream(g32.device)),
           'eb_ps_striped_optimizer_push')


# ------------------------------------------------------------------ convolutions (K9)
def conv_patch(n, p, q, limit, stride=1):
    """Output pixels per tile ``(pn, ph, pw)`` with ``pn * ph * pw <= limit`` (and every TMA box edge <= 256 input
    elements) that wastes the fewest tile rows; among equals the widest rows (longest contiguous TMA runs)."""
    edge = max(1, 256 // int(stride))
    best = None
    for pw in range(min(q, limit, edge), 0, -1):
        for ph in range(min(p, limit // pw, edge), 0, -1):
            pn = min(n, limit // (pw * ph), 256)
            tiles = -(-n // pn) * -(-p // ph) * -(-q // pw)
            eff = (n * p * q) / float(tiles * limit)
            if best is None or eff > best[0] + 0.02:
                best = (eff, pn, ph, pw)
    return best[1:]


def conv_igemm(inp, w, out, P, Q, R, S, stride=1, pad_h=0, pad_w=0, os=1, bias=None, w_mn=False, flip_taps=False,
               patch=None, block_n=0):
    """``eb_conv_igemm`` (see ``EbConvArgs``): ``inp`` bf16 ``[N,H,W,C]``, ``out`` bf16 ``[N,OH,OW,K]``; ``w`` is the OHWI
    kernel ``[K,R,S,C]`` (forward) or, with ``w_mn``, the forward kernel ``[C,R,S,K]`` read N-major (input gradient)."""
    a = ConvArgs()
    N, H, W, C = inp.shape
    K = out.shape[3]
    a.inp, a.N, a.H, a.W, a.C = inp.data_ptr(), N, H, W, C
    a.w, a.K, a.R, a.S = w.data_ptr(), K, int(R), int(S)
    a.w_mn, a.flip_taps, a.stride, a.pad_h, a.pad_w = int(bool(w_mn)), int(bool(flip_taps)), int(stride), int(pad_h), int(pad_w)
    a.P, a.Q = int(P), int(Q)
    a.pn, a.ph, a.pw = patch if patch is not None else conv_patch(N, int(P), int(Q), 128, stride)
    a.out, a.OH, a.OW, a.os = out.data_ptr(), out.shape[1], out.shape[2], int(os)
    a.bias = bias.data_ptr() if bias is not None else None
    a.block_n = int(block_n)
    _check(require().eb_conv_igemm(ctypes.byref(a), _stream(inp.device)), 'eb_conv_igemm')


def conv_wgrad(x, dy, dw, R, S, stride=1, pad_h=0, pad_w=0, patch=None, splits=0):
    """``dw[K,R,S,C] (fp32) += sum over pixels dy[..., K] (x) x[shifted, C]`` - zero ``dw`` first (``EbConvWgradArgs``)."""
    a = ConvWgradArgs()
    N, H, W, C = x.shape
    _, P, Q, K = dy.shape
    a.x, a.N, a.H, a.W, a.C = x.data_ptr(), N, H, W, C
    a.dy, a.P, a.Q, a.K = dy.data_ptr(), P, Q, K
    a.R, a.S, a.stride, a.pad_h, a.pad_w = int(R), int(S), int(stride), int(pad_h), int(pad_w)
    a.pn, a.ph, a.pw = patch if patch is not None else conv_patch(N, P, Q, 64, stride)
    a.dw, a.splits = dw.data_ptr(), int(splits)
    _check(require().eb_conv_wgrad(ctypes.byref(a), _stream(x.device)), 'eb_conv_wgrad')


def conv_im2col(x, R, S, stride, pad, P, Q, out, ld):
    N, H, W, C = x.shape
    _check(require().eb_conv_im2col(x.data_ptr(), int(x.dtype == torch.float32), N, H, W, C, int(R), int(S), int(stride),
                                    int(pad), int(P), int(Q), out.data_ptr(), int(ld), _stream(x.device)), 'eb_conv_im2col')


def pool_max_fwd(x, R, S, stride, pad_t, pad_l, P, Q, pad_is_zero, y, idx):
    """Max pooling of channels-last bf16 ``x`` [N, H, W, C] into ``y`` [N, P, Q, C]; ``idx`` (uint8, shape of y) = winning tap."""
    N, H, W, C = x.shape
    _check(require().eb_pool_max_fwd(x.data_ptr(), N, H, W, C, int(R), int(S), int(stride), int(pad_t), int(pad_l), int(P),
                                     int(Q), int(bool(pad_is_zero)), y.data_ptr(), idx.data_ptr(), _stream(x.device)),
           'eb_pool_max_fwd')


def pool_max_bwd(dy, idx, x_shape, R, S, stride, pad_t, pad_l, dx):
    N, H, W, C = x_shape
    P, Q = dy.shape[1], dy.shape[2]
    _check(require().eb_pool_max_bwd(dy.data_ptr(), idx.data_ptr(), N, H, W, C, int(R), int(S), int(stride), int(pad_t),
                                     int(pad_l), P, Q, dx.data_ptr(), _stream(dy.device)), 'eb_pool_max_bwd')


def pool_gap_fwd(x, y):
    """Global average pooling of channels-last bf16 ``x`` [N, ..., C] into ``y`` [N, C]."""
    N, C = x.shape[0], x.shape[-1]
    _check(require().eb_pool_gap_fwd(x.data_ptr(), N, x.numel() // (N * C), C, y.data_ptr(), _stream(x.device)),
           'eb_pool_gap_fwd')


def pool_gap_bwd(dy, dx):
    N, C = dx.shape[0], dx.shape[-1]
    _check(require().eb_pool_gap_bwd(dy.data_ptr(), N, dx.numel() // (N * C), C, dx.data_ptr(), _stream(dy.device)),
           'eb_pool_gap_bwd')


BN_MAX_BLOCKS = 592  # rows of the partial-sum scratch: 4 blocks of 256 threads per SM


def bn_stats(x, M, C, partials, eps, momentum, mean, rstd, running_mean=None, running_var=None):
    """Batch statistics of ``x`` ``[M, C]`` (two kernels: per-block partial sums into ``partials``, then their total)."""
    _check(require().eb_bn_stats(x.data_ptr(), int(M), int(C), partials.data_ptr(), int(partials.numel() // (2 * C)),
                                 float(eps), float(momentum), mean.data_ptr(), rstd.data_ptr(), _ptr(running_mean),
                                 _ptr(running_var), _stream(x.device)), 'eb_bn_stats')


def bn_apply(x, M, C, mean, rstd, gamma, beta, residual, relu, y):
    _check(require().eb_bn_apply(x.data_ptr(), int(M), int(C), mean.data_ptr(), rstd.data_ptr(), _ptr(gamma), _ptr(beta),
                                 _ptr(residual), int(bool(relu)), y.data_ptr(), _stream(x.device)), 'eb_bn_apply')


def bn_bwd_reduce(dy, y, x, M, C, mean, rstd, relu, partials, sum_dz, sum_dzx, accumulate=False):
    _check(require().eb_bn_bwd_reduce(dy.data_ptr(), _ptr(y), x.data_ptr(), int(M), int(C), mean.data_ptr(),
                                      rstd.data_ptr(), int(bool(relu)), partials.data_ptr(),
                                      int(partials.numel() // (2 * C)), sum_dz.data_ptr(), sum_dzx.data_ptr(),
                                      int(bool(accumulate)), _stream(x.device)), 'eb_bn_bwd_reduce')


def bn_bwd_apply(dy, y, x, M, C, mean, rstd, gamma, sum_dz, sum_dzx, relu, dx, dres=None):
    _check(require().eb_bn_bwd_apply(dy.data_ptr(), _ptr(y), x.data_ptr(), int(M), int(C), mean.data_ptr(), rstd.data_ptr(),
                                     _ptr(gamma), sum_dz.data_ptr(), sum_dzx.data_ptr(), int(bool(relu)), dx.data_ptr(),
                                     _ptr(dres), _stream(x.device)), 'eb_bn_bwd_apply')


# ------------------------------------------------------------------ host staging
def host_f32_to_bf16(src, dst, rows=None, cols=None, threads=4):
    """``dst[:rows, :cols] = bf16(src[:rows, :cols])`` on CPU threads (host tensors; ``dst`` usually pinned, bf16 or
    int16 storage), padding columns of ``dst`` zeroed. The GIL is released for the duration of the call."""
    rows = int(src.shape[0] if rows is None else rows)
    cols = int(src.shape[1] if cols is None else cols)
    if src.is_cuda or dst.is_cuda or src.dtype != torch.float32 or dst.element_size() != 2:
        raise NativeError("host_f32_to_bf16 needs a host fp32 source and a host 16-bit destination")
    _check(load().eb_host_f32_to_bf16(src.data_ptr(), src.stride(0), dst.data_ptr(), dst.stride(0), rows, cols,
                                      int(threads)), 'eb_host_f32_to_bf16')


# ------------------------------------------------- VMM allocations and multicast objects
def vmm_supported():
    """``(fd-exportable VMM allocations supported, NVSwitch multicast supported)`` on the current device."""
    v, m = c_int32(0), c_int32(0)
    _check(require().eb_vmm_supported(ctypes.byref(v), ctypes.byref(m)), 'eb_vmm_supported')
    return bool(v.value), bool(m.value)


def vmm_create(nbytes, multicast_ranks=0):
    h, size = c_uint64(0), c_int64(0)
    _check(require().eb_vmm_create(int(nbytes), int(multicast_ranks), ctypes.byref(h), ctypes.byref(size)), 'eb_vmm_create')
    return h.value, size.value


def vmm_export_fd(handle) -> int:
    fd = c_int32(-1)
    _check(require().eb_vmm_export_fd(c_uint64(handle), ctypes.byref(fd)), 'eb_vmm_export_fd')
    return fd.value


def vmm_import_fd(fd) -> int:
    h = c_uint64(0)
    _check(require().eb_vmm_import_fd(int(fd), ctypes.byref(h)), 'eb_vmm_import_fd')
    return h.value


def vmm_map(handle, size) -> int:
    p = c_void_p()
    _check(require().eb_vmm_map(c_uint64(handle), int(size), ctypes.byref(p)), 'eb_vmm_map')
    return p.value


def vmm_unmap(ptr, size):
    _check(require().eb_vmm_unmap(c_void_p(ptr), int(size)), 'eb_vmm_unmap')


def vmm_release(handle):
    _check(require().eb_vmm_release(c_uint64(handle)), 'eb_vmm_release')


def mc_create(ranks, size) -> int:
    h = c_uint64(0)
    _check(require().eb_mc_create(int(ranks), int(size), ctypes.byref(h)), 'eb_mc_create')
    return h.value


def mc_add_device(mc_handle):
    _check(require().eb_mc_add_device(c_uint64(mc_handle)), 'eb_mc_add_device')


def mc_bind(mc_handle, mem_handle, size):
    _check(require().eb_mc_bind(c_uint64(mc_handle), c_uint64(mem_handle), int(size)), 'eb_mc_bind')


def tensor_from_ptr(ptr, dtype, numel, device):
    """A torch tensor over raw device memory owned by somebody else (VMM mappings, library allocations)."""
    typestr = {torch.float32: '<f4', torch.bfloat16: '<u2', torch.int32: '<i4', torch.uint8: '|u1', torch.int64: '<i8',
               torch.float64: '<f8'}[dtype]
    iface = {'shape': (int(numel),), 'typestr': typestr, 'data': (int(ptr), False), 'version': 3}
    holder = type('_CudaArray', (), {'__cuda_array_interface__': iface})()
    t = torch.as_tensor(holder, device=device)
    return t.view(torch.bfloat16) if dtype == torch.bfloat16 else t


# --------------------------------------------------------------- memory / IPC
def device_malloc(nbytes) -> int:
    p = c_void_p()
    _check(require().eb_device_malloc(ctypes.byref(p), int(nbytes)), 'eb_device_malloc')
    return p.value


def device_free(ptr):
    _check(require().eb_device_free(c_void_p(ptr)), 'eb_device_free')


def ipc_export(ptr) -> bytes:
    buf = ctypes.create_string_buffer(64)
    _check(require().eb_ipc_export(c_void_p(ptr), buf), 'eb_ipc_export')
    return buf.raw


def ipc_import(handle: bytes) -> int:
    p = c_void_p()
    buf = ctypes.create_string_buffer(handle, 64)
    _check(require().eb_ipc_import(buf, ctypes.byref(p)), 'eb_ipc_import')
    return p.value


def ipc_close(ptr):
    _check(require().eb_ipc_close(c_void_p(ptr)), 'eb_ipc_close')


class RawBuffer:
    """A device allocation made by the library (cudaMalloc, IPC-exportable) viewed as torch tensors."""

    def __init__(self, nbytes, device):
        self.nbytes = int(nbytes)
        self.device = torch.device(device)
        with torch.cuda.device(self.device):
            self.ptr = device_malloc(self.nbytes)
        self._owned = True

    def tensor(self, dtype, numel=None, offset_bytes=0):
        itemsize = torch.empty((), dtype=dtype).element_size()
        numel = int(numel if numel is not None else (self.nbytes - offset_bytes) // itemsize)
        typestr = {torch.float32: '<f4', torch.bfloat16: '<u2', torch.int32: '<i4', torch.uint8: '|u1',
                   torch.int64: '<i8', torch.float64: '<f8'}[dtype]
        iface = {'shape': (numel,), 'typestr': typestr, 'data': (self.ptr + offset_bytes, False), 'version': 3}
        holder = type('_CudaArray', (), {'__cuda_array_interface__': iface})()
        t = torch.as_tensor(holder, device=self.device)
        return t.view(torch.bfloat16) if dtype == torch.bfloat16 else t

    def free(self):
        if self._owned and self.ptr:
            with torch.cuda.device(self.device):
                device_free(self.ptr)
            self.ptr = 0
