"""The docs generator runs and the shipped examples execute (CPU, small data)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_autogen_writes_pages(tmp_path, monkeypatch):
    sys.path.insert(0, os.path.join(ROOT, 'docs'))
    import autogen
    monkeypatch.setattr(autogen, 'OUT', str(tmp_path))
    pages = autogen.main()
    assert 'models/spark-model.md' in pages
    text = (tmp_path / 'models' / 'spark-model.md').read_text()
    assert 'class `SparkModel' in text and 'load_spark_model' in text
    assert 'eb_gemm_bf16' in (tmp_path / 'c-api.md').read_text()


@pytest.mark.parametrize('script', ['mnist_mlp_spark_synchronous.py', 'ml_mlp_regression.py', 'wide_mlp_sync.py', 'cnn_sync.py',
                                    'lstm_callbacks.py'])
def test_example_runs(script):
    env = dict(os.environ, ELEPHAS_B200_DEVICE='cpu', SPARK_LOCAL_IP='127.0.0.1')
    res = subprocess.run([sys.executable, script], cwd=os.path.join(ROOT, 'examples'), env=env, capture_output=True,
                         text=True, timeout=300)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]


def test_pipeline_notebook_cells_run(tmp_path):
    """examples/Spark_ML_Pipeline.ipynb (the reference ships the same notebook): its code cells execute in order."""
    import json
    nb = json.load(open(os.path.join(ROOT, 'examples', 'Spark_ML_Pipeline.ipynb')))
    assert nb['nbformat'] == 4 and nb['cells'][0]['cell_type'] == 'markdown'
    code = '\n\n'.join(''.join(c['source']) for c in nb['cells'] if c['cell_type'] == 'code')
    script = os.path.join(ROOT, 'examples', '_nb_cells_check.py')
    with open(script, 'w') as fh:
        fh.write(code)
    try:
        env = dict(os.environ, ELEPHAS_B200_DEVICE='cpu', SPARK_LOCAL_IP='127.0.0.1')
        res = subprocess.run([sys.executable, script], cwd=os.path.join(ROOT, 'examples'), env=env, capture_output=True,
                             text=True, timeout=600)
    finally:
        os.remove(script)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert 'train precision' in res.stdout
