#!/usr/bin/env python
"""Benchmarks of the BASELINE.json configurations: samples/sec, device-timed, max over ranks.

Default (what the driver runs) = config 4: wide MLP regression, 4096 hidden x 8 layers (134,254,593 parameters), bf16
compute / fp32 master weights, ``SparkModel(mode='synchronous', frequency='batch')``, SGD, synthetic data, random-init
weights, 4096 samples per GPU per step (weak scaling).

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Other rows of the matrix (same launch line plus):
    --config mnist_mlp --mode asynchronous --frequency batch --steps 2000    config 2: NVSwitch parameter server
    --config mnist_mlp --mode hogwild      --frequency batch --steps 2000
    --config otto_pipeline                                                   config 3: ElephasEstimator pipeline, hogwild
    --config resnet50                                                        config 5: ResNet-50-shaped CNN, synchronous bf16
    --impl ddp | localsgd                                                    torch + NCCL baselines for the same model

Prints ONE JSON line on rank 0. ``value`` is the kernel-path throughput (inputs resident on the device, CUDA events
around exactly K steps); ``e2e`` is the same metric through the public API (``SparkModel.fit``) with every step's
inputs copied host->device from pinned memory and the running loss read back device->host every step; ``sustained`` is
the kernel-path number over a run long enough to hit the board's power limit; ``final_loss`` is the loss on a fixed
batch after the timed steps (both arms start from the same weights and see the same data); ``replicas_identical``
says whether all ranks ended with bit-identical weights. ``--impl reference`` reports the reference arm (not
installable in this image: poetry build backend, tensorflow/pyspark/flask/h5py and Python < 3.11 are all missing).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "samples/sec (device-timed, max over ranks)"


_RESULT_FD = os.dup(1)  # the real stdout, kept aside before main() points fd 1 at stderr


def print_result(line):
    """The ONE JSON line of the contract goes to the real stdout; everything else that would land there while the run
    is in progress - the reference's '>>> Fit model' progress lines, but also what C libraries write to fd 1 directly
    (NCCL prints its version banner there when the first communicator comes up) - is routed to stderr by main()."""
    os.write(_RESULT_FD, (line + '\n').encode())


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference', 'ddp', 'localsgd'])
    ap.add_argument('--config', default='wide_mlp',
                    choices=['wide_mlp', 'mnist_mlp', 'otto_mlp', 'otto_pipeline', 'resnet50'])
    ap.add_argument('--mode', default=None, choices=['synchronous', 'asynchronous', 'hogwild'])
    ap.add_argument('--frequency', default='batch', choices=['batch', 'epoch'])
    ap.add_argument('--batch-per-gpu', type=int, default=None)
    ap.add_argument('--hidden', type=int, default=4096)
    ap.add_argument('--depth', type=int, default=8)
    ap.add_argument('--sustained', type=int, default=200, help='steps of the long kernel-path run (0: skip)')
    ap.add_argument('--no-e2e', action='store_true')
    return ap.parse_args()


class ClockSampler(threading.Thread):
    """Samples SM clocks and throttle reasons while the timed region runs: through NVML (a query takes microseconds, so
    a 50 ms region yields dozens of samples), falling back to ``nvidia-smi`` (one sample per ~100 ms) without it."""

    FIELDS = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
              'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')
    NAMES = ('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap')

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], threading.Event()
        self.nvml = self.handle = None
        try:
            import pynvml
            pynvml.nvmlInit()
            visible = os.environ.get('CUDA_VISIBLE_DEVICES')
            physical = int(visible.split(',')[index]) if visible and visible.split(',')[index].isdigit() else index
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(physical)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM)
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def _nvml_sample(self):
        n = self.nvml
        mhz = n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)
        get = getattr(n, 'nvmlDeviceGetCurrentClocksEventReasons', None) or n.nvmlDeviceGetCurrentClocksThrottleReasons
        mask = int(get(self.handle))
        bits = {'hw_slowdown': 0x8, 'hw_thermal_slowdown': 0x40, 'sw_thermal_slowdown': 0x20, 'sw_power_cap': 0x4}
        flags = ['Active' if mask & bits[k] else 'Not Active' for k in self.NAMES]
        return [str(mhz), str(self.max_mhz)] + flags

    def run(self):
        while not self.stop_flag.is_set():
            try:
                if self.nvml is not None:
                    self.samples.append(self._nvml_sample())
                    self.stop_flag.wait(0.002)
                    continue
                out = subprocess.run(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.FIELDS,
                                      '--format=csv,noheader,nounits'], capture_output=True, text=True, timeout=5).stdout
                parts = [p.strip() for p in out.strip().split(',')]
                if len(parts) >= 6:
                    self.samples.append(parts)
            except Exception:
                if self.nvml is not None:
                    self.nvml = None  # NVML refused: fall back to nvidia-smi for the rest of the run
            self.stop_flag.wait(0.2 if self.nvml is None else 0.002)

    def summary(self):
        self.stop_flag.set()
        self.join(timeout=6)
        if not self.samples:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': []}
        clocks = sorted(int(float(s[0])) for s in self.samples)
        reasons = set()
        for s in self.samples:
            for name, val in zip(self.NAMES, s[2:6]):
                if val.lower().startswith('active'):
                    reasons.add(name)
        return {'sm_mhz': clocks[len(clocks) // 2], 'sm_max_mhz': int(float(self.samples[0][1])),
                'reasons': sorted(reasons), 'samples': len(clocks),
                'source': 'nvml' if self.nvml is not None else 'nvidia-smi'}


def build_model(args):
    from elephas_b200.models import SGD, zoo
    if args.config == 'wide_mlp':
        model = zoo.wide_mlp(input_dim=args.hidden, hidden=args.hidden, depth=args.depth, out_dim=1)
        model.compile(SGD(learning_rate=1e-3), 'mse', ['mae'])
        in_dim, out_dim, default_batch = args.hidden, 1, 4096
    elif args.config == 'mnist_mlp':
        model = zoo.mnist_mlp()
        model.compile(SGD(learning_rate=0.1), 'categorical_crossentropy', ['acc'])
        in_dim, out_dim, default_batch = 784, 10, 64
    else:
        model = zoo.otto_mlp()
        model.compile(SGD(learning_rate=0.01), 'categorical_crossentropy', ['acc'])
        in_dim, out_dim, default_batch = 93, 9, 128
    return model, in_dim, out_dim, default_batch


def synthetic(rows, in_dim, out_dim, seed, classification):
    """Synthetic shard in pinned host memory (so the timed H2D copies are real pinned-memory DMA). The targets follow a
    fixed random linear teacher of the features (the same one for every seed / rank) plus noise, so that a training run
    has something to learn and "the loss went down" is a meaningful check; the features are random normal."""
    import torch
    g = torch.Generator().manual_seed(seed)
    x = torch.empty(rows, in_dim, dtype=torch.float32, pin_memory=torch.cuda.is_available())
    x.normal_(generator=g)
    teacher = torch.randn(in_dim, out_dim, generator=torch.Generator().manual_seed(4242)) / in_dim ** 0.5
    y = torch.empty(rows, out_dim, dtype=torch.float32, pin_memory=torch.cuda.is_available())
    t = x @ teacher + 0.1 * torch.randn(rows, out_dim, generator=g)
    if classification:
        y.zero_()
        y[torch.arange(rows), t.argmax(dim=1)] = 1.0
    else:
        y.copy_(t)
    return x.numpy(), y.numpy()


def tile_pinned(a, rows):
    """``rows`` rows in pinned host memory, cycling through the rows of ``a`` (plain memcpy: no new random numbers)."""
    import numpy as np
    import torch
    if len(a) == rows:
        return a
    out = torch.empty((rows,) + tuple(a.shape[1:]), dtype=torch.float32, pin_memory=torch.cuda.is_available()).numpy()
    for lo in range(0, rows, len(a)):
        n = min(len(a), rows - lo)
        np.copyto(out[lo:lo + n], a[:n])
    return out


def weights_digest(rt, flat):
    """(all replicas bit-identical?, digest of this rank): exact integer sum of the weights' bit patterns + abs-sum."""
    import torch
    bits = int(flat.view(torch.int32).to(torch.int64).sum().item())
    mag = float(flat.double().abs().sum().item())
    gathered = rt.all_gather_object((bits, mag))
    return all(g == gathered[0] for g in gathered), gathered[0]


def max_over_ranks(rt, value):
    import torch
    t = torch.tensor([value], device=rt.device, dtype=torch.float64)
    rt.all_reduce_max_(t)
    return float(t[0])


def reference_arm():
    """The reference cannot be installed here; say why in one line (see DESIGN.md for the attempt)."""
    why = None
    ref = os.path.join(ROOT, 'baseline', '_ref')
    try:
        if not os.path.isdir(os.path.join(ref, 'elephas')):
            raise ImportError("baseline/_ref/elephas absent: pip install of /root/reference fails offline "
                              "(build backend poetry-core missing)")
        sys.path.insert(0, ref)
        import pyspark  # noqa: F401
        import tensorflow  # noqa: F401
    except Exception as exc:
        why = ("reference (elephas 4.1.0) not installable in this image: %s; it also needs tensorflow<=2.10, "
               "pyspark<=3.3, flask, h5py and python<3.11, none of which are available offline" % exc)
    if os.environ.get('RANK', '0') == '0':  # one line per job, also under torchrun
        print_result(json.dumps({'impl': 'reference', 'unavailable': why or 'reference stack import failed'}))


def run_sync(args, rt):
    """Configs 4 (default) and the synchronous rows of the small MLPs: SparkModel(mode='synchronous', frequency='batch')."""
    import numpy as np
    import torch
    from elephas_b200.ops import native
    from elephas_b200.spark import SparkContext
    from elephas_b200.spark_model import SparkModel
    from elephas_b200.utils import tracing
    from elephas_b200.utils.rdd_utils import to_sharded_rdd

    model, in_dim, out_dim, default_batch = build_model(args)
    B = args.batch_per_gpu or default_batch
    K, W, N = args.steps, args.warmup, rt.world_size
    classification = args.config != 'wide_mlp'
    sc = SparkContext.getOrCreate()

    # ------------------------------------------------------------------ kernel path
    spark_model = SparkModel(model, mode='synchronous', frequency='batch', num_workers=None)
    xw, yw = synthetic(max(W, 3) * B, in_dim, out_dim, 1000 + rt.rank, classification)  # never fewer than 3 warm-up steps
    # distinct timed batches: at most DISTINCT of them (1 GB of inputs per GPU at the default shape, far beyond the L2);
    # longer runs cycle through them, so host and device memory stay bounded whatever K the caller asks for
    DISTINCT, E2E_MAX = 16, 64
    xk, yk = synthetic(min(K, DISTINCT) * B, in_dim, out_dim, 2000 + rt.rank, classification)
    x_eval, y_eval = synthetic(B, in_dim, out_dim, 3000, classification)  # the same held-out batch on every rank / arm
    model.to(rt.device)
    loss_before = float(model.evaluate(x_eval, y_eval)[0])  # before any training step, like the torch arms
    # warm-up fit: builds the engine, symmetric memory, CUDA graphs (W untimed steps)
    spark_model.fit(to_sharded_rdd(sc, xw, yw), epochs=1, batch_size=B, shuffle=False, input_mode='device',
                    keep_context=True)
    ctx = spark_model._sync_context
    launches_before = native.launch_count()
    sampler = ClockSampler(rt.local_rank)
    # the sampler covers exactly the timed steps (not the upload of the data set that precedes them)
    ms_kernel = ctx.timed_steps(xk, yk, K, input_mode='device', on_start=sampler.start, on_stop=sampler.stop_flag.set)
    launches = native.launch_count() - launches_before
    clocks = sampler.summary()
    # correctness evidence for the multi-GPU kernels, outside of the timed region
    identical, digest = weights_digest(rt, ctx.eng.w32)
    final_loss = float(model.evaluate(x_eval, y_eval)[0])

    # ------------------------------------------------------------------ long run (power-limited regime)
    sustained = None
    if args.sustained and args.sustained > K:
        S = args.sustained
        s_sampler = ClockSampler(rt.local_rank)
        ms_s = max_over_ranks(rt, ctx.timed_steps(xk, yk, S, input_mode='device', on_start=s_sampler.start,
                                                   on_stop=s_sampler.stop_flag.set))
        sustained = {'steps': S, 'ms_per_step': round(ms_s / S, 4), 'value': round(N * B * S / (ms_s / 1e3), 2),
                     'clocks': s_sampler.summary(), 'final_loss': round(float(model.evaluate(x_eval, y_eval)[0]), 6)}

    # ------------------------------------------------------------------ end to end through the public API
    # Two data layouts of the same pinned host shard: 'bf16' = the shard was staged as bf16 when the RDD was built
    # (to_sharded_rdd(feature_dtype='bfloat16'): what a bf16 model is meant to be fed with, 2 bytes per feature and
    # step over PCIe) and 'fp32' (4 bytes per feature, cast on the device). Building the RDD is data preparation and
    # happens before the clock starts; every step of the timed fit copies its own batch host -> device.
    e2e = e2e_fp32 = None
    if not args.no_e2e:
        Ke = max(2, min(K, E2E_MAX))
        xe, ye = tile_pinned(xk, Ke * B), tile_pinned(yk, Ke * B)

        def timed_fit(feature_dtype):
            warm = to_sharded_rdd(sc, xw, yw, feature_dtype=feature_dtype)
            spark_model.fit(warm, epochs=1, batch_size=B, shuffle=False, input_mode='stream', log_batch_loss=True,
                            keep_context=True)  # warm-up of this streaming path
            rdd = to_sharded_rdd(sc, xe, ye, feature_dtype=feature_dtype)
            rt.barrier()
            torch.cuda.synchronize()
            tracing.reset()
            rt.barrier()
            t0 = time.perf_counter()
            spark_model.fit(rdd, epochs=1, batch_size=B, shuffle=False, input_mode='stream', log_batch_loss=True,
                            keep_context=True)
            torch.cuda.synchronize()
            rt.barrier()
            wall_ms = (time.perf_counter() - t0) * 1e3
            dev_ms = spark_model._last_sync_run.get('device_ms', wall_ms)
            if tracing.enabled() and rt.is_driver:
                print('--- phases of the timed fit, host features %s (wall %.3f ms, device %.3f ms)\n%s'
                      % (feature_dtype or 'float32', wall_ms, dev_ms, tracing.report()), file=sys.stderr)
            ms = max_over_ranks(rt, max(wall_ms, dev_ms))
            hist = spark_model.training_histories[-1]
            assert len(hist['batch_loss']) == Ke, "one loss read-back per step expected"
            return {'value': round(N * B * Ke / (ms / 1e3), 2), 'unit': 'samples/sec', 'ms_per_step': round(ms / Ke, 4),
                    'steps': Ke, 'h2d_bytes_per_step': int(spark_model._last_sync_run['h2d_bytes_per_step']),
                    'd2h_bytes_per_step': 64, 'host_features': feature_dtype or 'float32',
                    'device_ms_per_step_inside_fit': round(max_over_ranks(rt, dev_ms) / Ke, 4),
                    'timing': 'wall clock around SparkModel.fit incl. setup, max over ranks'}

        e2e_fp32 = timed_fit(None)
        e2e = timed_fit('bfloat16') if os.environ.get('BENCH_E2E_BF16', '1') != '0' else e2e_fp32
    ms_total = max_over_ranks(rt, ms_kernel)
    value = N * B * K / (ms_total / 1e3)
    params = model.count_params()
    flops_step = 6.0 * params * B  # fwd 2PB + bwd 4PB (dgrad of the first layer is skipped: slight over-count)
    result = {
        'metric': METRIC, 'value': round(value, 2), 'unit': 'samples/sec', 'n_gpus': N, 'steps': K, 'warmup': W,
        'ms_per_step': round(ms_total / K, 4), 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': 'bf16', 'data': 'synthetic (random normal features, linear-teacher targets, random-init weights)',
        'impl': 'ours',
        'config': {'model': 'wide MLP regression %dx%d (%d params)' % (args.hidden, args.depth, params)
                   if args.config == 'wide_mlp' else args.config,
                   'global_batch': N * B, 'seq_len': None, 'parallelism': 'dp%d' % N,
                   'mode': 'synchronous', 'frequency': 'batch', 'optimizer': 'SGD',
                   'collective': ('none (one GPU)' if N == 1 else
                                  'fused all-reduce + update kernel, ' + ('NVLS multimem' if getattr(ctx, 'nvls', False)
                                                                         else 'peer-to-peer loads / stores')),
                   'l2': 'working set (weights+grads+activations, >1 GB) exceeds the 126 MB L2; no flush needed'},
        'clocks': clocks, 'e2e': e2e, 'e2e_fp32_host_features': e2e_fp32, 'gpu_launches': int(launches),
        'tflops_per_gpu': round(flops_step / (ms_total / K / 1e3) / 1e12, 1),
        'sustained': sustained, 'loss_before': round(loss_before, 6), 'final_loss': round(final_loss, 6),
        'loss_decreased': bool(final_loss < loss_before), 'replicas_identical': bool(identical),
        'weights_digest': list(digest),
    }
    if N > 1:
        assert identical, "replicas diverged: %r" % (rt.all_gather_object(digest),)
    if rt.is_driver:
        print_result(json.dumps(result))
    spark_model.release_context()


def run_ps(args, rt):
    """Config 2 (and its hogwild sibling): SparkModel(mode='asynchronous' | 'hogwild', frequency='batch',
    parameter_server_mode='nvlink') - every rank one worker against the parameter server striped over all GPUs.
    Device-timed: CUDA events around exactly K pull -> step -> optimizer+push graph replays per rank, after W untimed
    ones in the same fit (graphs recorded before the first step), max over ranks."""
    import torch
    from elephas_b200.ops import native
    from elephas_b200.spark import SparkContext
    from elephas_b200.spark_model import SparkModel
    from elephas_b200.utils.rdd_utils import to_sharded_rdd

    model, in_dim, out_dim, default_batch = build_model(args)
    B = args.batch_per_gpu or default_batch
    K, W, N = args.steps, max(args.warmup, 3), rt.world_size
    sc = SparkContext.getOrCreate()
    mode = args.mode or 'asynchronous'
    x, y = synthetic((W + K) * B, in_dim, out_dim, 2000 + rt.rank, True)
    x_eval, y_eval = synthetic(4096, in_dim, out_dim, 3000, True)
    port = 4700 + (os.getpid() % 200)
    port = int(rt.broadcast_object(port, src=0))
    sm = SparkModel(model, mode=mode, frequency=args.frequency, parameter_server_mode='nvlink', num_workers=None, port=port)
    init = [w.copy() for w in model.get_weights()]
    loss_before = float(sm.master_network.evaluate(x_eval, y_eval)[0])
    # untimed fit: loads every kernel, NCCL, IPC mappings
    sm.fit(to_sharded_rdd(sc, x[:4 * B], y[:4 * B]), epochs=1, batch_size=B)
    sm.master_network.set_weights(init)
    rt.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(rt.local_rank)
    launches_before = native.launch_count()
    sampler.start()
    sm.fit(to_sharded_rdd(sc, x, y), epochs=1, batch_size=B, timed_from=W)
    torch.cuda.synchronize()
    sampler.stop_flag.set()
    launches = native.launch_count() - launches_before
    info = sm._last_async_run[0]
    assert info['timed_steps'] == K, info
    ms_total = max_over_ranks(rt, info['device_ms'])
    clocks = sampler.summary()
    final = sm.master_network.evaluate(x_eval, y_eval)
    identical, digest = weights_digest(rt, torch.cat([torch.as_tensor(w).reshape(-1) for w in sm.master_network.get_weights()])
                                       .to(rt.device))
    e2e = None
    if not args.no_e2e:
        rt.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        sm.fit(to_sharded_rdd(sc, x, y), epochs=1, batch_size=B, input_mode='stream', log_batch_loss=True)
        torch.cuda.synchronize()
        rt.barrier()
        wall_ms = max_over_ranks(rt, (time.perf_counter() - t0) * 1e3)
        run = sm._last_async_run[0]
        steps = run['steps']
        assert len(sm.training_histories[-1]['batch_loss']) == steps
        e2e = {'value': round(N * B * steps / (wall_ms / 1e3), 2), 'unit': 'samples/sec', 'steps': steps,
               'ms_per_step': round(wall_ms / steps, 4), 'h2d_bytes_per_step': int(run['h2d_bytes_per_step']),
               'd2h_bytes_per_step': 64, 'step_loop_device_ms_per_step': round(max_over_ranks(rt, run['device_ms']) / steps, 4),
               'timing': 'wall clock around SparkModel.fit (server start, worker + engine build, graph capture, steps, '
                         'final pull), max over ranks'}
    result = {
        'metric': METRIC, 'value': round(N * B * K / (ms_total / 1e3), 2), 'unit': 'samples/sec', 'n_gpus': N, 'steps': K,
        'warmup': W, 'ms_per_step': round(ms_total / K, 4), 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'bf16', 'impl': 'ours',
        'data': 'synthetic (random normal features, linear-teacher classes, random-init weights)',
        'config': {'model': args.config, 'global_batch': N * B, 'seq_len': None, 'parallelism': 'dp%d' % N, 'mode': mode,
                   'frequency': args.frequency, 'parameter_server': 'nvlink (striped over %d GPUs)' % N,
                   'optimizer': 'SGD', 'pull': 'fused into the forward GEMMs (K2)' if info.get('fused_pull') else
                   'seqlock pull kernel', 'l2': 'latency-bound config: each step touches < 2 MB; inputs cycle through '
                   '%d distinct batches' % (W + K)},
        'clocks': clocks, 'e2e': e2e, 'gpu_launches': int(launches), 'kernels_per_step': int(info.get('graph_kernels', 0)),
        'loss_before': round(loss_before, 6), 'final_loss': round(float(final[0]), 6), 'final_acc': round(float(final[1]), 4),
        'loss_decreased': bool(final[0] < loss_before), 'replicas_identical': bool(identical),
    }
    if rt.is_driver:
        print_result(json.dumps(result))


def run_config_script(args, rt):
    """Configs 3 and 5 go through benchmarks/configs.py (wall clock around the public API call)."""
    sys.path.insert(0, os.path.join(ROOT, 'benchmarks'))
    import configs as _configs
    ns = argparse.Namespace(config=3 if args.config == 'otto_pipeline' else 5, gpus=rt.world_size, epochs=None,
                            samples=None, batch_size=args.batch_per_gpu, cpu=False, frequency=args.frequency)
    res = _configs.config_otto(ns, rt) if ns.config == 3 else _configs.config_resnet(ns, rt)
    out = {'metric': 'samples/sec (wall clock around the public API call, max over ranks)', 'unit': 'samples/sec',
           'n_gpus': rt.world_size, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'impl': 'ours',
           'dtype': 'bf16', 'data': 'synthetic', 'config': {'model': args.config, 'parallelism': 'dp%d' % rt.world_size}}
    out.update(res)
    if rt.is_driver:
        print_result(json.dumps(out))


def _torch_twin(args, model_k, dev):
    """A plain torch.nn twin of the Keras-style model, starting from the SAME initial weights."""
    import torch
    import torch.nn as nn
    from elephas_b200.models import layers as L
    mods = []
    for layer in model_k.layers:
        if isinstance(layer, L.Dense):
            lin = nn.Linear(int(layer.kernel.shape[1]), int(layer.units))
            with torch.no_grad():
                lin.weight.copy_(layer.kernel.detach().cpu())
                lin.bias.copy_(layer.bias.detach().cpu())
            mods.append(lin)
            if layer.activation_name == 'relu':
                mods.append(nn.ReLU())
        elif isinstance(layer, L.Activation) and layer.activation_name == 'relu':
            mods.append(nn.ReLU())
        elif isinstance(layer, L.Dropout):
            mods.append(nn.Dropout(layer.rate))
    return nn.Sequential(*mods).to(dev)  # (a trailing softmax is folded into CrossEntropyLoss)


def run_torch(args, rt):
    """Library baselines on the same model, initial weights and data:
    ``ddp``      torch DistributedDataParallel over NCCL (bucketed gradient all-reduce overlapped with backward), bf16
                 autocast, fp32 master weights, fused SGD, fresh inputs every step, the step replayed from a CUDA graph
                 where capture succeeds (always on one GPU);
    ``localsgd`` the reference's synchronous semantics on torch + NCCL: every rank takes a local step, the weight DELTAS
                 are all-reduced and averaged (elephas/spark_model.py:217-231 per batch)."""
    import torch
    import torch.nn as nn
    import torch.distributed as dist
    model_k, in_dim, out_dim, default_batch = build_model(args)
    B = args.batch_per_gpu or default_batch
    K, W, N = args.steps, max(args.warmup, 3), rt.world_size
    dev = rt.device
    classification = args.config != 'wide_mlp'
    lr = float(model_k.optimizer._config['learning_rate'])
    net = _torch_twin(args, model_k, dev)
    if N > 1:
        for p_ in net.parameters():
            dist.broadcast(p_.data, src=0)
    loss_fn = nn.CrossEntropyLoss() if classification else nn.MSELoss()
    # the same data policy as the ours arm: 16 distinct 67 MB batches for the wide MLP, a fresh batch at every step
    # for the small models (2000 steps over 16 batches would just memorise them)
    DISTINCT = 16 if args.config == 'wide_mlp' else K
    xw, yw = synthetic(W * B, in_dim, out_dim, 1000 + rt.rank, classification)
    xk, yk = synthetic(min(K, DISTINCT) * B, in_dim, out_dim, 2000 + rt.rank, classification)
    x_eval, y_eval = synthetic(B if args.config == 'wide_mlp' else 4096, in_dim, out_dim, 3000, classification)
    to_dev = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
    xs, ys = torch.cat([to_dev(xw), to_dev(xk)]), torch.cat([to_dev(yw), to_dev(yk)])
    if classification:
        ys = ys.argmax(dim=1)
        y_eval_t = to_dev(y_eval).argmax(dim=1)
    else:
        y_eval_t = to_dev(y_eval)
    x_eval_t = to_dev(x_eval)
    xb, yb = torch.empty_like(xs[:B]), torch.empty_like(ys[:B])  # static inputs of the (graphed) step
    ddp = args.impl == 'ddp'
    wrapped = nn.parallel.DistributedDataParallel(net, device_ids=[dev.index], gradient_as_bucket_view=True,
                                                  static_graph=True) if (ddp and N > 1) else net
    params = list(net.parameters())
    try:
        opt = torch.optim.SGD(params, lr=lr, fused=True)
    except (TypeError, RuntimeError):
        opt = torch.optim.SGD(params, lr=lr, foreach=True)
    before_w = [torch.empty_like(p_) for p_ in params] if not ddp else None

    def evaluate():
        net.eval()
        with torch.no_grad(), torch.autocast('cuda', dtype=torch.bfloat16):
            out = net(x_eval_t)
        net.train()
        return float(loss_fn(out.float(), y_eval_t))

    def step():
        if not ddp and N > 1:
            torch._foreach_copy_(before_w, params)
        opt.zero_grad(set_to_none=False)
        with torch.autocast('cuda', dtype=torch.bfloat16):
            out = wrapped(xb)
        loss = loss_fn(out.float(), yb)
        loss.backward()
        opt.step()
        if not ddp and N > 1:
            # delta = before - after, averaged over ranks, applied to the common starting point
            deltas = torch._foreach_sub(before_w, params)
            flat = torch.cat([d.reshape(-1) for d in deltas])
            dist.all_reduce(flat)
            flat.div_(N)
            off = 0
            for p_, b_ in zip(params, before_w):
                n_ = p_.numel()
                p_.data.copy_(b_ - flat[off:off + n_].view_as(p_))
                off += n_

    loss_before = evaluate()
    n_w, n_k = W, min(K, DISTINCT)

    def load(i):
        lo = i * B
        xb.copy_(xs[lo:lo + B])
        yb.copy_(ys[lo:lo + B])

    graph = None
    warm = max(W, 11 if (ddp and N > 1) else 3)  # DDP wants ~11 eager iterations before a capture
    for i in range(warm):
        load(i % n_w)
        step()
    torch.cuda.synchronize()
    # whole-step capture only on one GPU: DDP's bucket hooks and an NCCL all-reduce issued from Python do not survive a
    # capture here (the attempt fails and leaves the dropout generator in capture mode), so N > 1 runs eagerly
    if N == 1 and os.environ.get('BENCH_TORCH_GRAPH', '1') != '0':
        try:
            g = torch.cuda.CUDAGraph()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            snapshot = [p_.detach().clone() for p_ in params]
            with torch.cuda.stream(side):
                with torch.cuda.graph(g, stream=side):
                    step()
            torch.cuda.current_stream().wait_stream(side)
            graph = g
            for p_, s_ in zip(params, snapshot):
                p_.data.copy_(s_)
        except Exception as exc:  # capture not supported for this arm / world size: run eagerly
            print('torch arm: CUDA-graph capture failed (%s); running eagerly' % (str(exc).splitlines()[0],), file=sys.stderr)
            graph = None
            torch.cuda.synchronize()
    rt.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(rt.local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(K):
        load(n_w + i % n_k)
        if graph is not None:
            graph.replay()
        else:
            step()
    e1.record()
    torch.cuda.synchronize()
    sampler.stop_flag.set()
    rt.barrier()
    ms = max_over_ranks(rt, e0.elapsed_time(e1))
    final_loss = evaluate()
    identical, _ = weights_digest(rt, torch.cat([p_.detach().reshape(-1) for p_ in params]).float().contiguous())
    if rt.is_driver:
        print_result(json.dumps({'metric': METRIC, 'value': round(N * B * K / (ms / 1e3), 2), 'unit': 'samples/sec',
                                 'n_gpus': N, 'steps': K, 'warmup': W, 'ms_per_step': round(ms / K, 4),
                                 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'bf16',
                                 'data': 'synthetic (same generator, same initial weights as the ours arm)',
                                 'impl': args.impl,
                                 'config': {'model': args.config, 'global_batch': N * B, 'parallelism': 'dp%d' % N,
                                            'cuda_graph': graph is not None, 'optimizer': 'SGD fused',
                                            'inputs': 'fresh batch every step (device resident)'},
                                 'clocks': sampler.summary(), 'gpu_launches': 0, 'loss_before': round(loss_before, 6),
                                 'final_loss': round(final_loss, 6), 'replicas_identical': bool(identical)}))


def run_torch_resnet(args, rt):
    """Library baseline for config 5: torchvision's ResNet-50 (the same architecture, 25.6 M parameters), channels-last,
    bf16 autocast with fp32 master weights, fused SGD with momentum, cuDNN convolutions, DistributedDataParallel over NCCL
    at N > 1 (gradient buckets overlapped with backward), the step replayed from a CUDA graph on one GPU. Batch 64 per
    GPU, 224 x 224, a fresh device-resident batch every step; device-timed with CUDA events, max over ranks."""
    import torch
    import torch.nn as nn
    from torchvision.models import resnet50
    N, dev = rt.world_size, rt.device
    B = args.batch_per_gpu or 64
    K, W = args.steps, max(args.warmup, 3)
    torch.manual_seed(5)
    net = resnet50(num_classes=1000).to(dev).to(memory_format=torch.channels_last)
    if N > 1:
        import torch.distributed as dist
        for t in list(net.parameters()) + list(net.buffers()):
            dist.broadcast(t.data, src=0)
    wrapped = nn.parallel.DistributedDataParallel(net, device_ids=[dev.index], gradient_as_bucket_view=True) \
        if N > 1 else net
    opt = torch.optim.SGD(net.parameters(), lr=0.01, momentum=0.9, fused=True)
    loss_fn = nn.CrossEntropyLoss()
    gen = torch.Generator(device=dev).manual_seed(7 + rt.rank)
    DISTINCT = 8
    xs = torch.randn(DISTINCT * B, 3, 224, 224, device=dev, generator=gen).to(memory_format=torch.channels_last)
    ys = torch.randint(0, 1000, (DISTINCT * B,), device=dev, generator=gen)
    xb, yb = torch.empty_like(xs[:B]), torch.empty_like(ys[:B])
    loss_box = torch.zeros((), device=dev)

    def step():
        opt.zero_grad(set_to_none=False)
        with torch.autocast('cuda', dtype=torch.bfloat16):
            out = wrapped(xb)
        loss = loss_fn(out.float(), yb)
        loss.backward()
        opt.step()
        loss_box.copy_(loss.detach())

    def load(i):
        lo = (i % DISTINCT) * B
        xb.copy_(xs[lo:lo + B])
        yb.copy_(ys[lo:lo + B])

    for i in range(max(W, 11)):
        load(i)
        step()
    torch.cuda.synchronize()
    graph = None
    if N == 1 and os.environ.get('BENCH_TORCH_GRAPH', '1') != '0':
        try:
            g, side = torch.cuda.CUDAGraph(), torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                with torch.cuda.graph(g, stream=side):
                    step()
            torch.cuda.current_stream().wait_stream(side)
            graph = g
        except Exception as exc:
            print('torch ResNet arm: CUDA-graph capture failed (%s); running eagerly' % (str(exc).splitlines()[0],),
                  file=sys.stderr)
            torch.cuda.synchronize()
    rt.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(rt.local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(K):
        load(i)
        if graph is not None:
            graph.replay()
        else:
            step()
    e1.record()
    torch.cuda.synchronize()
    sampler.stop_flag.set()
    rt.barrier()
    ms = max_over_ranks(rt, e0.elapsed_time(e1))
    if rt.is_driver:
        print_result(json.dumps({'metric': METRIC, 'value': round(N * B * K / (ms / 1e3), 2), 'unit': 'samples/sec',
                                 'n_gpus': N, 'steps': K, 'warmup': W, 'ms_per_step': round(ms / K, 4),
                                 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'bf16',
                                 'data': 'synthetic', 'impl': args.impl,
                                 'config': {'model': 'resnet50 (torchvision, channels-last, cuDNN)', 'global_batch': N * B,
                                            'parallelism': 'dp%d' % N, 'cuda_graph': graph is not None,
                                            'optimizer': 'SGD momentum fused'},
                                 'clocks': sampler.summary(), 'gpu_launches': 0, 'final_loss': round(float(loss_box), 4)}))


def main():
    args = parse()
    sys.stdout.flush()
    os.dup2(2, 1)            # fd level: C libraries (NCCL) write to fd 1 without going through sys.stdout
    sys.stdout = sys.stderr  # keep stdout clean for the result line
    if args.impl == 'reference':
        reference_arm()
        return
    import torch
    if not torch.cuda.is_available():
        if os.environ.get('RANK', '0') != '0':
            return
        print_result(json.dumps({'metric': METRIC, 'value': None, 'unavailable': 'no CUDA device', 'impl': args.impl}))
        return
    os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
    from elephas_b200.parallel import get_runtime
    rt = get_runtime()
    if rt.world_size != args.gpus and rt.is_driver:
        print("warning: --gpus %d but WORLD_SIZE=%d; launch with torch.distributed.run" % (args.gpus, rt.world_size),
              file=sys.stderr)
    mode = args.mode or {'mnist_mlp': 'asynchronous'}.get(args.config, 'synchronous')
    if args.impl == 'ddp' and args.config == 'resnet50':
        run_torch_resnet(args, rt)
    elif args.impl in ('ddp', 'localsgd'):
        run_torch(args, rt)
    elif args.config in ('otto_pipeline', 'resnet50'):
        run_config_script(args, rt)
    elif mode == 'synchronous':
        run_sync(args, rt)
    else:
        args.mode = mode
        run_ps(args, rt)
    rt.shutdown()


if __name__ == '__main__':
    main()
