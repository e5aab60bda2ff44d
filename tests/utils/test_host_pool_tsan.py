"""ThreadSanitizer run of the host thread pool + fp32 -> bf16 row conversion (csrc/host_stage.cu, csrc/host_cast.cpp):
four caller threads convert batches of changing shapes / thread counts concurrently (the pool grows under them), every
result is compared with a scalar conversion, and the sanitizer must report nothing."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


@pytest.mark.skipif(shutil.which('g++') is None, reason='needs g++')
def test_host_pool_is_clean_under_thread_sanitizer(tmp_path):
    exe = str(tmp_path / 'tsan_host_pool')
    cuda_inc = os.path.join(os.environ.get('CUDA_HOME', '/usr/local/cuda'), 'include')
    build = subprocess.run(['g++', '-std=c++17', '-O1', '-g', '-fsanitize=thread', '-I', cuda_inc, '-x', 'c++',
                            os.path.join(ROOT, 'csrc', 'host_stage.cu'), os.path.join(ROOT, 'csrc', 'host_cast.cpp'),
                            os.path.join(ROOT, 'tests', 'native', 'tsan_host_pool.cpp'), '-o', exe, '-lpthread'],
                           capture_output=True, text=True, timeout=600)
    if build.returncode != 0 and ('tsan' in build.stderr or 'sanitize' in build.stderr):
        pytest.skip('this toolchain has no ThreadSanitizer runtime')
    assert build.returncode == 0, build.stderr[-3000:]
    run = subprocess.run([exe], capture_output=True, text=True, timeout=900,
                         env=dict(os.environ, TSAN_OPTIONS='halt_on_error=0 exitcode=66'))
    if 'FATAL: ThreadSanitizer' in run.stderr and 'unexpected memory mapping' in run.stderr:
        pytest.skip('ThreadSanitizer cannot map its shadow memory in this container')
    assert run.returncode == 0 and 'mismatches: 0' in run.stdout, run.stdout[-1000:] + run.stderr[-3000:]
    assert 'WARNING: ThreadSanitizer' not in run.stderr, run.stderr[-3000:]
