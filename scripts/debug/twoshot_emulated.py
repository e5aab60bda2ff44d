"""Where do the NaNs of the emulated 8-rank two-shot run come from? Prints, per variant, the NaN ranges of the target's
weights and of every averaged-delta buffer."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch  # noqa: E402

from elephas_b200.ops import native  # noqa: E402


def ranges(mask):
    idx = torch.nonzero(mask).flatten().tolist()
    out, start, prev = [], None, None
    for i in idx:
        if start is None:
            start = prev = i
        elif i == prev + 1:
            prev = i
        else:
            out.append((start, prev))
            start = prev = i
    if start is not None:
        out.append((start, prev))
    return out[:12], len(idx)


n, world = 1 << 15, 8
for dtype in (torch.bfloat16, torch.float32):
    for blocks in (1, 4):
        for own_zero in (True, False):
            torch.manual_seed(21)
            deltas = [torch.randn(n, device='cuda').to(dtype) for _ in range(world)]
            pads = [torch.full((64,), 1 << 30, dtype=torch.int32, device='cuda') for _ in range(world)]
            if own_zero:
                for r, pad in enumerate(pads):
                    pad[8 + r] = 0
            target = 3
            avgs = [torch.full((n,), float('nan'), dtype=dtype, device='cuda') for _ in range(world)]
            order = [r for r in range(world) if r != target] + [target]
            keep = []
            for r in order:
                w32 = torch.ones(n, device='cuda')
                w16 = torch.zeros(n, dtype=torch.bfloat16, device='cuda')
                local = torch.zeros(16, dtype=torch.int32, device='cuda')
                keep.append((w32, w16, local))
                native.allreduce_update([d.data_ptr() for d in deltas], [p.data_ptr() for p in pads], local, n, r, world,
                                        1.0 / world, dtype == torch.bfloat16, w32, w16=w16,
                                        avg_ptrs=[a.data_ptr() for a in avgs], algo=2, blocks=blocks)
                torch.cuda.synchronize()
                print(dtype, 'blocks', blocks, 'own_zero', own_zero, 'after rank', r, 'avg[target] NaN:',
                      ranges(torch.isnan(avgs[target].float())), 'local', local[:9].tolist(), flush=True)
            print('   target w32 NaN:', ranges(torch.isnan(w32)), 'pads[target]', pads[target][:16].tolist(), flush=True)
