"""Two or three CPU ranks exchange file descriptors (the transport of the VMM symmetric-memory bootstrap)."""
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
os.environ['ELEPHAS_B200_DEVICE'] = 'cpu'
from elephas_b200.parallel import get_runtime  # noqa: E402
from elephas_b200.parallel.symm import exchange_fds  # noqa: E402

rt = get_runtime()
mine = []
for k in range(2):
    f = tempfile.TemporaryFile()
    f.write(('rank %d file %d' % (rt.rank, k)).encode())
    f.flush()
    mine.append(os.dup(f.fileno()))
    f.close()
for rnd in range(2):  # two exchanges in a row: the socket names must not collide
    got = exchange_fds(rt, mine, 'test%d' % rnd)
    assert len(got) == rt.world_size and got[rt.rank] == mine
    for r in range(rt.world_size):
        assert len(got[r]) == 2
        for k, fd in enumerate(got[r]):
            # (pread: duplicates in other processes share the file OFFSET with this one, a seek + read pair would race)
            assert os.pread(fd, 64, 0) == ('rank %d file %d' % (r, k)).encode(), (r, k)
            if r != rt.rank:
                os.close(fd)
print('FD EXCHANGE OK rank %d' % rt.rank, flush=True)
rt.shutdown()
