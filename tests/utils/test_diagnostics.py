"""``python -m elephas_b200``: the report names the path this machine will take."""
import json
import subprocess
import sys


def test_report_runs_without_a_gpu_and_lists_the_switches():
    env_run = subprocess.run([sys.executable, '-m', 'elephas_b200', '--json'], capture_output=True, text=True, timeout=300,
                             env={**__import__('os').environ, 'ELEPHAS_B200_NVLS': '0'})
    assert env_run.returncode == 0, env_run.stderr[-2000:]
    report = json.loads(env_run.stdout)
    assert {'version', 'torch', 'cuda_available', 'native_library', 'knobs', 'execution', 'spark_backend'} <= set(report)
    assert report['knobs']['ELEPHAS_B200_NVLS'] == '0' and report['knobs']['ELEPHAS_B200_GRAPHS'] is None
    assert report['native_library']['path'].endswith('libelephas_b200.so')
    if not report['cuda_available']:
        assert report['execution'].startswith('CPU') and report['devices'] == []
    # the table form, in this process
    from elephas_b200 import __main__ as tool
    import contextlib
    import io
    out = io.StringIO()
    with contextlib.redirect_stdout(out):
        assert tool.main([]) == 0
    assert 'native library' in out.getvalue() and 'switches set' in out.getvalue()
