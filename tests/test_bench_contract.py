"""bench.py contract pieces that can be checked without a GPU: stdout carries exactly one JSON line, the reference arm
reports itself unavailable, the clock sampler and the data tiler behave."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args):
    return subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py')] + list(args), capture_output=True, text=True,
                          timeout=300, cwd=ROOT)


def test_reference_arm_prints_one_json_line_and_exits_zero():
    res = _run('--impl', 'reference', '--gpus', '1', '--steps', '2', '--warmup', '1')
    assert res.returncode == 0
    lines = [l for l in res.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d['impl'] == 'reference' and 'unavailable' in d and '\n' not in d['unavailable']


def test_stdout_is_one_line_even_when_something_prints_to_fd_1():
    import torch
    if torch.cuda.is_available():
        pytest.skip('CPU-only check of the no-device path')
    res = _run()
    assert res.returncode == 0
    lines = [l for l in res.stdout.splitlines() if l.strip()]
    assert len(lines) == 1 and json.loads(lines[0])['unavailable'] == 'no CUDA device'


def test_clock_sampler_without_a_device_and_data_tiler():
    sys.path.insert(0, ROOT)
    import bench
    sampler = bench.ClockSampler(0)
    sampler.start()
    summary = sampler.summary()
    assert set(summary) >= {'sm_mhz', 'sm_max_mhz', 'reasons'}
    a = np.arange(12, dtype=np.float32).reshape(6, 2)
    assert bench.tile_pinned(a, 6) is a
    t = bench.tile_pinned(a, 15)
    assert t.shape == (15, 2) and np.array_equal(t[:6], a) and np.array_equal(t[6:12], a) and np.array_equal(t[12:], a[:3])
