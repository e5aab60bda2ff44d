"""Generates the API reference (Markdown, mkdocs-ready) from the package's docstrings.

Counterpart of the reference's ``docs/autogen.py`` (which introspects docstrings into mkdocs pages, PAGES table
at ``docs/autogen.py:17-81``, and copies the README into the index, ``:340-345``). Run
``python docs/autogen.py`` to (re)write ``docs/sources/*.md``; ``mkdocs build -f docs/mkdocs.yml`` renders them.
"""
import importlib
import inspect
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, 'docs', 'sources')

PAGES = [
    ('models/spark-model.md', 'elephas_b200.spark_model', ['SparkModel', 'SparkMLlibModel', 'load_spark_model']),
    ('models/spark-ml-model.md', 'elephas_b200.ml_model',
     ['ElephasEstimator', 'ElephasTransformer', 'load_ml_estimator', 'load_ml_transformer']),
    ('models/workers.md', 'elephas_b200.worker', ['SparkWorker', 'AsynchronousSparkWorker']),
    ('models/keras-layer.md', 'elephas_b200.models.model', ['Model', 'Sequential', 'model_from_json', 'load_model']),
    ('models/layers.md', 'elephas_b200.models.layers', None),
    ('models/layers-extra.md', 'elephas_b200.models.layers_extra', None),
    ('models/keras-namespace.md', 'elephas_b200.keras', None),
    ('models/layers-more.md', 'elephas_b200.models.layers_more', None),
    ('models/callbacks.md', 'elephas_b200.models.callbacks', None),
    ('models/regularizers.md', 'elephas_b200.models.regularizers', None),
    ('models/constraints.md', 'elephas_b200.models.constraints', None),
    ('models/schedules.md', 'elephas_b200.models.schedules', None),
    ('models/optimizers.md', 'elephas_b200.models.optimizers', None),
    ('models/zoo.md', 'elephas_b200.models.zoo', None),
    ('parameter/server.md', 'elephas_b200.parameter.server', None),
    ('parameter/client.md', 'elephas_b200.parameter.client', None),
    ('parameter/factory.md', 'elephas_b200.parameter.factory', None),
    ('parameter/device-ps.md', 'elephas_b200.parallel.device_ps', None),
    ('runtime/runtime.md', 'elephas_b200.parallel.runtime', None),
    ('runtime/sync-fast-path.md', 'elephas_b200.parallel.sync_fast', None),
    ('runtime/symmetric-memory.md', 'elephas_b200.parallel.symm', None),
    ('runtime/sync-torch-path.md', 'elephas_b200.parallel.sync_torch', None),
    ('runtime/torch-step.md', 'elephas_b200.ops.torch_step', None),
    ('runtime/native-cnn-ops.md', 'elephas_b200.ops.cnn', None),
    ('runtime/cpu-mlp-step.md', 'elephas_b200.ops.cpu_mlp', None),
    ('utils/cpu_threads.md', 'elephas_b200.utils.cpu_threads', None),
    ('utils/tracing.md', 'elephas_b200.utils.tracing', None),
    ('runtime/engine.md', 'elephas_b200.ops.engine', None),
    ('runtime/native-bindings.md', 'elephas_b200.ops.native', None),
    ('utils/functional_utils.md', 'elephas_b200.utils.functional_utils', None),
    ('utils/rdd_utils.md', 'elephas_b200.utils.rdd_utils', None),
    ('utils/serialization.md', 'elephas_b200.utils.serialization', None),
    ('utils/sockets.md', 'elephas_b200.utils.sockets', None),
    ('utils/rwlock.md', 'elephas_b200.utils.rwlock', None),
    ('utils/model_utils.md', 'elephas_b200.utils.model_utils', None),
    ('ml/params.md', 'elephas_b200.ml.params', None),
    ('ml/adapter.md', 'elephas_b200.ml.adapter', None),
    ('mllib/adapter.md', 'elephas_b200.mllib.adapter', None),
    ('io/hdf5.md', 'elephas_b200.io.hdf5', None),
    ('io/keras-h5.md', 'elephas_b200.io.keras_h5', None),
    ('spark/stand-ins.md', 'elephas_b200.spark.context', None),
    ('spark/feature-stages.md', 'elephas_b200.spark.feature', None),
    ('spark/evaluators.md', 'elephas_b200.spark.ml_evaluation', None),
]


def _signature(obj):
    try:
        return str(inspect.signature(obj))
    except (TypeError, ValueError):
        return '(...)'


def _doc(obj):
    return inspect.getdoc(obj) or '*(undocumented)*'


def render_module(module_name, names=None):
    mod = importlib.import_module(module_name)
    lines = ['# `%s`' % module_name, '', _doc(mod), '']
    if names is None:
        names = [n for n, o in vars(mod).items()
                 if not n.startswith('_') and getattr(o, '__module__', None) == module_name
                 and (inspect.isclass(o) or inspect.isfunction(o))]
    for name in names:
        obj = getattr(mod, name)
        kind = 'class' if inspect.isclass(obj) else 'function'
        lines += ['## %s `%s%s`' % (kind, name, _signature(obj)), '', _doc(obj), '']
        if inspect.isclass(obj):
            for mname, member in vars(obj).items():
                if mname.startswith('_') or not (inspect.isfunction(member) or isinstance(member, property)):
                    continue
                target = member.fget if isinstance(member, property) else member
                lines += ['### `%s.%s%s`' % (name, mname, '' if isinstance(member, property) else _signature(target)),
                          '', _doc(target), '']
    return '\n'.join(lines)


def render_c_api():
    header = open(os.path.join(ROOT, 'include', 'elephas_b200.h')).read()
    return '# C ABI (`include/elephas_b200.h`)\n\n```c\n%s\n```\n' % header


def main():
    written = []
    for rel, module_name, names in PAGES:
        path = os.path.join(OUT, rel)
        os.makedirs(os.path.dirname(path), exist_ok=True)
        with open(path, 'w') as fh:
            fh.write(render_module(module_name, names))
        written.append(rel)
    with open(os.path.join(OUT, 'c-api.md'), 'w') as fh:
        fh.write(render_c_api())
    with open(os.path.join(OUT, 'index.md'), 'w') as fh:
        fh.write(open(os.path.join(ROOT, 'README.md')).read())
    for name in ('DESIGN.md', 'INTEGRATION.md', 'BASELINE.md'):
        with open(os.path.join(OUT, name.lower()), 'w') as fh:
            fh.write(open(os.path.join(ROOT, name)).read())
    print("wrote %d pages to %s" % (len(written) + 5, OUT))
    return written


if __name__ == '__main__':
    main()
