"""Collectives used by the driver-side API.

Reference mechanism replaced: Spark ``collect`` of every worker's delta followed by a sequential
NumPy average on the driver (``elephas/spark_model.py:220-227``, O(N*P) on one core). Here every
rank packs its partial sum into one flat buffer and the ranks all-reduce it: on B200s through the
fused one-shot / two-shot P2P kernels over NVLink peer memory (``csrc/comm_kernels.cu``,
bound in ``elephas_b200/parallel/symm.py``), on CPU ranks through gloo.
"""
import numpy as np
import torch


def all_reduce_arrays_sum(rt, arrays):
    """Sum a list of equally-shaped arrays over all ranks; returns a list of NumPy arrays."""
    shapes = [np.asarray(a).shape for a in arrays]
    sizes = [int(np.prod(s)) if s else 1 for s in shapes]
    flat = torch.from_numpy(np.concatenate([np.asarray(a, dtype=np.float32).ravel() for a in arrays])
                            if arrays else np.zeros(0, dtype=np.float32))
    if rt.device.type == 'cuda':
        from . import symm
        flat = symm.all_reduce_sum(rt, flat.to(rt.device)).cpu()
    else:
        rt.all_reduce_sum_(flat)
    out, pos = [], 0
    host = flat.numpy()
    for shape, size in zip(shapes, sizes):
        out.append(host[pos:pos + size].reshape(shape).copy())
        pos += size
    return out
