"""Flat parameter layout shared by the native engine, the fused all-reduce and the device PS.

All weights of a model live in ONE contiguous fp32 buffer (plus an index-identical bf16 compute
copy): collectives, optimizer steps, delta computation (``subtract_params`` in the reference,
``elephas/worker.py:44-45``) and parameter-server pushes/pulls become single kernels over one
pointer instead of a Python loop over a list of arrays.

Segment rules: weights appear in Keras order; a Dense kernel is stored ``[out, ld]`` with
``ld = round_up(in, 8)`` (zero padding, so rows satisfy the 16-byte TMA stride rule in bf16); every
segment starts on a 64-element boundary; Dense kernels
are stored first, all other weights after them (``matrix_end`` marks the boundary).
"""
import numpy as np
import torch

ALIGN = 64


def _round_up(x, m):
    return (x + m - 1) // m * m


class Segment:
    __slots__ = ('offset', 'rows', 'cols', 'ld', 'layer', 'index', 'shape', '_is_matrix')

    def __init__(self, offset, rows, cols, ld, layer, index, shape):
        self.offset, self.rows, self.cols, self.ld = offset, rows, cols, ld
        self.layer, self.index, self.shape = layer, index, shape

    @property
    def span(self):
        return self.rows * self.ld


class FlatLayout:
    def __init__(self, model):
        from ..models.layers import Dense
        entries = []
        for layer, idx, w in model._weight_entries():
            shape = tuple(w.shape)
            if isinstance(layer, Dense) and idx == 0:
                rows, cols = shape
                ld = _round_up(cols, 8)
                is_matrix = True
            else:
                rows, cols = 1, int(np.prod(shape)) if shape else 1
                ld = cols
                is_matrix = False
            entries.append(Segment(0, rows, cols, ld, layer, idx, shape))
            entries[-1]._is_matrix = is_matrix
        # storage order: all Dense kernels first, then everything else (biases, BN statistics ...), so that
        # "the matrices" and "the vectors" are two contiguous ranges (the fused SGD path updates the matrices
        # inside the wgrad GEMM epilogue and runs the optimizer kernel over the vector range only).
        offset = 0
        for seg in [e for e in entries if e._is_matrix] + [e for e in entries if not e._is_matrix]:
            seg.offset = offset
            offset = _round_up(offset + seg.rows * seg.ld, ALIGN)
        self.segments = entries  # Keras order
        self.matrix_end = max([_round_up(e.offset + e.span, ALIGN) for e in entries if e._is_matrix] + [0])
        self.numel = max(offset, ALIGN)

    def view(self, flat, seg):
        """2-D ``[rows, cols]`` strided view of one segment inside ``flat``."""
        return flat[seg.offset:seg.offset + seg.span].view(seg.rows, seg.ld)[:, :seg.cols]

    # -- internal-layout tensors (the model's own weight tensors) ----------------
    @torch.no_grad()
    def pack_model(self, model, out=None, device=None, dtype=torch.float32):
        device = device or model.device
        flat = out if out is not None else torch.zeros(self.numel, dtype=dtype, device=device)
        for seg, (_, _, w) in zip(self.segments, model._weight_entries()):
            self.view(flat, seg).copy_(w.detach().reshape(seg.rows, seg.cols))
        return flat

    @torch.no_grad()
    def unpack_to_model(self, flat, model):
        for seg, (_, _, w) in zip(self.segments, model._weight_entries()):
            w.copy_(self.view(flat, seg).reshape(seg.shape))

    # -- Keras-layout arrays (the list-of-ndarrays the reference moves around) ----
    def pack_arrays(self, arrays, device='cpu', out=None):
        flat = out if out is not None else torch.zeros(self.numel, dtype=torch.float32, device=device)
        if out is not None:
            flat.zero_()
        for seg, a in zip(self.segments, arrays):
            if hasattr(a, 'detach'):
                a = a.detach().cpu().numpy()
            src = np.ascontiguousarray(seg.layer.from_keras_layout(seg.index, np.asarray(a, dtype=np.float32)))
            self.view(flat, seg).copy_(torch.from_numpy(src).reshape(seg.rows, seg.cols))
        return flat

    def unpack_arrays(self, flat):
        host = flat.detach().float().cpu()
        out = []
        for seg in self.segments:
            t = self.view(host, seg).reshape(seg.shape)
            out.append(seg.layer.to_keras_layout(seg.index, t).contiguous().numpy().copy())
        return out
