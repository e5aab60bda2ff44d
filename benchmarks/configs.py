"""The BASELINE.json configs other than the headline one (config 4 is ``bench.py``): each prints ONE JSON line.

    python benchmarks/configs.py --config 1                    # MNIST MLP, synchronous/epoch, CPU, 2 workers
    torchrun ... benchmarks/configs.py --config 2 --gpus N     # MNIST MLP, asynchronous/batch, NVLink parameter server
    torchrun ... benchmarks/configs.py --config 3 --gpus N     # Otto pipeline (ElephasEstimator, hogwild) fit+transform
    torchrun ... benchmarks/configs.py --config 5 --gpus N     # ResNet-50-shaped CNN, synchronous, bf16

Everything goes through the public API (SparkModel / ElephasEstimator / Pipeline) on synthetic data of the named
shapes; times are wall clock around the call with a device synchronize + barrier on both sides, max over ranks.
The reference (TF/Keras + PySpark + Flask) cannot be installed in this image, so ``vs_baseline``
is null and the CPU number of config 1 doubles as the "CPU restatement" baseline for the GPU configs.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'examples'))
os.environ.setdefault('SPARK_LOCAL_IP', '127.0.0.1')
os.environ.setdefault('MASTER_ADDR', '127.0.0.1')


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--config', type=int, required=True, choices=[1, 2, 3, 5])
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--epochs', type=int, default=None)
    ap.add_argument('--samples', type=int, default=None, help='training samples (whole job)')
    ap.add_argument('--batch-size', type=int, default=None)
    ap.add_argument('--cpu', action='store_true', help='force the CPU runtime (config 1 always does)')
    return ap.parse_args()


def timed(rt, fn):
    from elephas_b200.utils import tracing
    tracing.reset()
    """Wall-clock seconds of ``fn()`` bracketed by synchronize + barrier, max over ranks."""
    import torch
    cuda = rt.device.type == 'cuda'
    if cuda:
        torch.cuda.synchronize(rt.device)
    rt.barrier()
    t0 = time.perf_counter()
    out = fn()
    if cuda:
        torch.cuda.synchronize(rt.device)
    rt.barrier()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device=rt.device)
    rt.all_reduce_max_(t)
    if tracing.enabled() and rt.is_driver:
        print(tracing.report(), file=sys.stderr)
    return float(t[0]), out


def emit(rt, **kw):
    base = {'unit': 'samples/sec', 'higher_is_better': True, 'vs_baseline': None, 'impl': 'ours',
            'data': 'synthetic (shapes of the named dataset, random-init weights)', 'n_gpus': 0}
    base.update(kw)
    if rt.is_driver:
        print(json.dumps(base), flush=True)


def config_mnist(args, rt, mode, frequency, ps_mode):
    """Configs 1 and 2: the reference's MNIST MLP through SparkModel.fit / evaluate."""
    from _data import mnist_like
    from elephas.spark_model import SparkModel
    from elephas.utils.rdd_utils import to_simple_rdd
    from elephas_b200.models import SGD, zoo
    from elephas_b200.spark import SparkContext
    workers = max(rt.world_size, 2)
    n = args.samples or 60000
    epochs = args.epochs or 3
    bs = args.batch_size or 64
    x, y, xt, yt = mnist_like(n, 2000)
    sc = SparkContext.getOrCreate()
    model = zoo.mnist_mlp()
    model.compile(SGD(learning_rate=0.1), 'categorical_crossentropy', ['acc'])
    sm = SparkModel(model, mode=mode, frequency=frequency, parameter_server_mode=ps_mode, num_workers=workers)
    rdd = to_simple_rdd(sc, x, y)
    sm.fit(to_simple_rdd(sc, x[:workers * bs * 4], y[:workers * bs * 4]), epochs=1, batch_size=bs, verbose=0)  # warm-up
    dt, _ = timed(rt, lambda: sm.fit(rdd, epochs=epochs, batch_size=bs, verbose=0, validation_split=0.1))
    loss, acc = sm.evaluate(xt, yt)
    return dict(value=round(n * 0.9 * epochs / dt, 1), seconds=round(dt, 4), epochs=epochs, final_loss=round(loss, 4),
                final_acc=round(acc, 4), workers=workers, batch_size=bs, train_samples=n)


def config_otto(args, rt):
    """Config 3: StringIndexer -> StandardScaler -> ElephasEstimator (hogwild) pipeline: fit, then distributed
    transform (predict) over the same frame."""
    from _data import otto_like
    from elephas.ml_model import ElephasEstimator
    from elephas_b200.models import optimizers, zoo
    from elephas_b200.spark import (DoubleType, MulticlassMetrics, Pipeline, SparkSession, StandardScaler,
                                    StringIndexer, Vectors, functions)
    n = args.samples or 60000
    epochs = args.epochs or 3
    bs = args.batch_size or 128
    workers = max(rt.world_size, 2)
    spark = SparkSession.builder.appName('otto-bench').getOrCreate()
    feats, cats = otto_like(n)
    df = spark.createDataFrame([(Vectors.dense(f), c) for f, c in zip(feats, cats)], ['features', 'category'])
    classes = df.select('category').distinct().count()

    def pipeline(n_epochs):
        est = ElephasEstimator(featuresCol='scaled_features', labelCol='index_category')
        est.set_keras_model_config(zoo.otto_mlp(input_dim=93, classes=classes).to_json())
        est.set_optimizer_config(optimizers.serialize(optimizers.Adam(learning_rate=0.002)))
        est.set_mode('hogwild')
        est.set_loss('categorical_crossentropy')
        est.set_metrics(['acc'])
        est.set_epochs(n_epochs)
        est.set_batch_size(bs)
        est.set_num_workers(workers)
        est.set_verbosity(0)
        est.set_validation_split(0.1)
        est.set_categorical_labels(True)
        est.set_nb_classes(classes)
        return Pipeline(stages=[StringIndexer(inputCol='category', outputCol='index_category'),
                                StandardScaler(inputCol='features', outputCol='scaled_features', withStd=True,
                                               withMean=True), est])

    pipeline(1).fit(df.limit(workers * bs * 4))  # warm-up: engines, kernels, parameter server
    fit_s, fitted = timed(rt, lambda: pipeline(epochs).fit(df))
    pred_s, pred = timed(rt, lambda: fitted.transform(df))     # first pass: column blocks are built from the rows
    warm_s, _ = timed(rt, lambda: fitted.transform(df))        # same frame again: blocks and label factors are cached
    pnl = pred.select('index_category', functions.array_argmax('prediction').astype(DoubleType()).alias('p'))
    precision = MulticlassMetrics(pnl.rdd.map(lambda r: (r.p, r.index_category))).precision()
    return dict(value=round(n * 0.9 * epochs / fit_s, 1), seconds=round(fit_s, 4), epochs=epochs, workers=workers,
                batch_size=bs, train_samples=n, predict_seconds=round(pred_s, 4),
                predict_samples_per_sec=round(n / pred_s, 1), predict_warm_seconds=round(warm_s, 4),
                predict_warm_samples_per_sec=round(n / warm_s, 1), train_precision=round(float(precision), 4))


def config_resnet(args, rt):
    """Config 5: ResNet-50-shaped CNN, SparkModel synchronous (frequency='batch': replicas averaged every step), bf16.
    Convolutions, BatchNormalization (+ residual + ReLU) and the classifier run on this repository's tcgen05 / sm_100a
    kernels (``ops/cnn.py``; ``ELEPHAS_B200_NATIVE_CONV=0`` = the cuDNN arm), the step is replayed from a CUDA graph, the
    weight deltas go through the fused NVLink all-reduce inside that graph. Every rank generates only its own shard."""
    import numpy as np
    from elephas.spark_model import SparkModel
    from elephas_b200.models import SGD, zoo
    from elephas_b200.ops import cnn
    from elephas_b200.spark import SparkContext
    from elephas_b200.utils.rdd_utils import to_sharded_rdd
    bs = args.batch_size or 64
    steps = 20
    epochs = args.epochs or 3
    side, classes = 224, 1000
    rng = np.random.default_rng(7 + rt.rank)
    n_rank = bs * steps
    x = rng.standard_normal((n_rank, side, side, 3), dtype=np.float32)
    y = np.eye(classes, dtype=np.float32)[rng.integers(0, classes, n_rank)]
    model = zoo.resnet50(input_shape=(side, side, 3), classes=classes)
    model.compile(SGD(learning_rate=0.01, momentum=0.9), 'categorical_crossentropy', ['acc'])
    model.set_compute_dtype('bfloat16')
    frequency = getattr(args, 'frequency', None) or 'batch'
    sm = SparkModel(model, mode='synchronous', frequency=frequency, num_workers=None)
    sc = SparkContext.getOrCreate()
    host_dtype = os.environ.get('BENCH_RESNET_FEATURES', 'bfloat16')  # staged once, outside of fit (like the RDD itself)
    rdd = to_sharded_rdd(sc, x, y, feature_dtype=host_dtype if host_dtype == 'bfloat16' else None)
    sm.fit(rdd, epochs=1, batch_size=bs, verbose=0, keep_context=True)  # warm-up (graph capture, both delta buffers)
    # the phase tracer is on (a perf_counter pair per phase): where the wall clock of a fit goes besides the steps
    from elephas_b200.utils import tracing

    def traced_fit():
        tracing.enable(True)
        out = timed(rt, lambda: sm.fit(rdd, epochs=epochs, batch_size=bs, verbose=0, keep_context=True))
        tracing.enable(False)
        top = sorted(tracing.summary()['host'].items(), key=lambda kv: -kv[1]['total_ms'])[:10]
        return out[0], {k: '%d x, %.1f ms' % (v['calls'], v['total_ms']) for k, v in top}

    dt, phases = traced_fit()
    hist = sm.training_histories[-1] or {}
    import torch
    dev_ms = torch.tensor([float((getattr(sm, '_last_sync_run', None) or {}).get('device_ms', float('nan')))],
                          dtype=torch.float64, device=rt.device)
    rt.all_reduce_max_(dev_ms)  # CUDA events around the epoch loops inside fit (steps + per-epoch reads), max over ranks
    t_dt, phases_again = traced_fit()  # the same call once more
    sm.release_context()
    n = n_rank * max(rt.world_size, 1)
    return dict(value=round(n * epochs / dt, 1), seconds=round(dt, 4), ms_per_step=round(dt * 1e3 / (epochs * steps), 3),
                epochs=epochs, steps_per_epoch=steps, batch_size=bs, train_samples=n, image=side, frequency=frequency,
                final_loss=round(float(hist.get('loss', [float('nan')])[-1]), 4),
                conv_backend='native tcgen05 implicit GEMM + sm_100a BatchNorm kernels' if cnn.enabled()
                else 'cuDNN (library)', dtype='bf16 (fp32 master weights)', host_features=host_dtype,
                device_ms_per_step_inside_fit=round(float(dev_ms[0]) / (epochs * steps), 3),
                device_timed_samples_per_sec=round(n * epochs / (float(dev_ms[0]) / 1e3), 1),
                host_phases_ms=phases, repeat_fit_seconds=round(t_dt, 4), repeat_host_phases_ms=phases_again)


def main():
    args = parse()
    if args.config == 1 or args.cpu:
        os.environ['ELEPHAS_B200_DEVICE'] = 'cpu'
    from elephas_b200.parallel import get_runtime
    rt = get_runtime()
    names = {1: 'MNIST MLP 784-128-128-10 synchronous/epoch CPU local[2]',
             2: 'MNIST MLP asynchronous/batch, NVLink parameter server',
             3: 'Otto (93 features, 9 classes) ElephasEstimator pipeline, hogwild, fit + transform',
             5: 'ResNet-50-shaped CNN synchronous bf16'}
    if args.config == 1:
        res = config_mnist(args, rt, 'synchronous', 'epoch', None)
    elif args.config == 2:
        res = config_mnist(args, rt, 'asynchronous', 'batch', 'nvlink' if rt.device.type == 'cuda' else 'socket')
    elif args.config == 3:
        res = config_otto(args, rt)
    else:
        res = config_resnet(args, rt)
    emit(rt, metric='samples/sec (wall clock around the public API call, max over ranks)', config=names[args.config],
         n_gpus=rt.world_size if rt.device.type == 'cuda' else 0, device=str(rt.device.type), **res)
    rt.shutdown()


if __name__ == '__main__':
    main()
