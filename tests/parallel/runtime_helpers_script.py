"""Two CPU ranks: the small tensor collectives of the runtime (agreement flag, one row per rank)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
os.environ['ELEPHAS_B200_DEVICE'] = 'cpu'
import torch  # noqa: E402

from elephas_b200.parallel import get_runtime  # noqa: E402

rt = get_runtime()
assert rt.all_true(True) is True
assert rt.all_true(rt.rank == 0) is False
rows = rt.all_gather_rows(torch.tensor([float(rt.rank), 7.0], dtype=torch.float64))
assert rows.tolist() == [[0.0, 7.0], [1.0, 7.0]], rows
print('RT HELPERS OK', rt.rank, flush=True)
rt.shutdown()
