"""Multi-process path on the CPU: two ranks over gloo (the host-side logic of the one-rank-per-GPU design)."""
import json
import os
import socket
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _launch(tmp_path, mode, ps, freq, *extra):
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node=2',
           '--master-addr', '127.0.0.1', '--master-port', str(_free_port()),
           os.path.join(HERE, 'spmd_script.py'), mode, ps, freq, str(_free_port()), str(tmp_path)] + list(extra)
    env = dict(os.environ, SPARK_LOCAL_IP='127.0.0.1', OMP_NUM_THREADS='2')
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    return [json.load(open(os.path.join(str(tmp_path), 'rank%d.json' % r))) for r in range(2)]


@pytest.mark.parametrize('mode,ps,freq', [('synchronous', 'none', 'epoch'), ('synchronous', 'none', 'batch'),
                                          ('asynchronous', 'socket', 'epoch'), ('hogwild', 'http', 'epoch')])
def test_two_ranks(tmp_path, mode, ps, freq):
    r0, r1 = _launch(tmp_path, mode, ps, freq)
    # every rank ends with the same master weights, predictions and evaluation
    assert abs(r0['digest'] - r1['digest']) < 1e-3 * max(1.0, abs(r0['digest']))
    assert r0['argmax'] == r1['argmax'] and r0['n_pred'] == 200
    assert abs(r0['eval'][0] - r1['eval'][0]) < 1e-6
    # distributed evaluate == local evaluate, and training made progress
    assert abs(r0['eval'][0] - r0['local'][0]) < 0.01 and abs(r0['eval'][1] - r0['local'][1]) < 0.01
    assert r0['eval'][0] < r0['before'][0]
    if mode == 'synchronous':
        assert r0['hist'] == 2


@pytest.mark.parametrize('mode,ps', [('synchronous', 'none'), ('asynchronous', 'socket')])
def test_two_ranks_with_penalties_constraints_callbacks_and_class_weights(tmp_path, mode, ps):
    """The Keras features that the fused paths do not implement take the generic worker loop on every rank: the ranks
    still agree on the result, each worker's callbacks saw its epochs, the constraint holds on the master weights."""
    r0, r1 = _launch(tmp_path, mode, ps, 'epoch', 'features')
    assert abs(r0['digest'] - r1['digest']) < 1e-3 * max(1.0, abs(r0['digest']))
    assert r0['argmax'] == r1['argmax'] and r0['eval'][0] < r0['before'][0]
    assert r0['epochs_seen'] == [0, 1] and r1['epochs_seen'] == [0, 1]   # one worker per rank, both epochs each
    assert r0['max_col_norm'] <= 4.0 + 1e-3


def test_cpulist_parser_and_numa_binding_is_harmless_without_a_gpu():
    from elephas_b200.parallel import runtime
    assert runtime._parse_cpulist('0-3,8,10-11\n') == {0, 1, 2, 3, 8, 10, 11}
    assert runtime._parse_cpulist('') == set()
    import torch
    assert runtime.bind_to_gpu_numa_node(torch.device('cpu')) is None  # no PCI device behind it: nothing happens


def test_ranks_exchange_file_descriptors_over_unix_sockets():
    """``symm.exchange_fds`` (SCM_RIGHTS over abstract Unix sockets) is how the exported handles of cuMemCreate
    allocations and multicast objects reach the other ranks; three CPU ranks send each other open files."""
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node=3',
           '--master-addr', '127.0.0.1', '--master-port', str(_free_port()),
           os.path.join(HERE, 'fd_exchange_script.py')]
    res = subprocess.run(cmd, env=dict(os.environ, OMP_NUM_THREADS='1'), capture_output=True, text=True, timeout=300)
    assert res.returncode == 0 and res.stdout.count('FD EXCHANGE OK') == 3, res.stdout[-2000:] + res.stderr[-3000:]


def test_small_tensor_collectives_of_the_runtime():
    """``Runtime.all_true`` / ``all_gather_rows`` (the agreement flag and the packed histories of the synchronous fast
    path) over gloo with two CPU ranks."""
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node=2',
           '--master-addr', '127.0.0.1', '--master-port', str(_free_port()),
           os.path.join(HERE, 'runtime_helpers_script.py')]
    res = subprocess.run(cmd, env=dict(os.environ, OMP_NUM_THREADS='1'), capture_output=True, text=True, timeout=300)
    assert res.returncode == 0 and res.stdout.count('RT HELPERS OK') == 2, res.stdout[-2000:] + res.stderr[-3000:]
