"""``python -m elephas_b200 [--json]``: what this installation will do on this machine.

Device(s), the native library (path, whether it is built and loads), the multi-GPU facilities (peer access, NVSwitch
multicast), the Spark backend in use and the value of every ``ELEPHAS_B200_*`` switch - the first thing to look at when a
run takes a path other than the expected one (the reference has no such tool: SURVEY.md §5.1).
"""
import json
import os
import sys

KNOBS = ('DEVICE', 'NATIVE', 'GRAPHS', 'TORCH_GRAPHS', 'NATIVE_CONV', 'FUSE_BN', 'FUSED_MLP', 'FUSED_MLP_WIDE', 'FUSED_PULL',
         'NVLS', 'SHARDED_MASTER', 'COMM_BLOCKS', 'COMM_TAIL_BLOCKS', 'SIDE_STREAM', 'HOST_BF16', 'HOST_THREADS', 'NUMA_BIND',
         'CPU_ENGINE', 'CPU_THREADS', 'CPU_PARALLEL_WORKERS', 'PS_TIMEOUT', 'FORCE_LOCAL_SPARK', 'TRACE')


def info():
    import torch

    from . import __version__, spark
    from .ops import native
    report = {'version': __version__, 'python': sys.version.split()[0], 'torch': torch.__version__,
              'cuda_available': bool(torch.cuda.is_available()), 'devices': [], 'spark_backend':
              'pyspark' if spark.HAVE_PYSPARK else 'local stand-ins (elephas_b200.spark)',
              'distributed_backends': [b for b in ('nccl', 'gloo') if getattr(torch.distributed, 'is_%s_available' % b)()],
              'launched_by_torchrun': 'RANK' in os.environ and 'WORLD_SIZE' in os.environ,
              'knobs': {'ELEPHAS_B200_' + k: os.environ.get('ELEPHAS_B200_' + k) for k in KNOBS}}
    path = native.lib_path()
    lib = {'path': path, 'built': os.path.exists(path)}
    if lib['built']:
        lib['bytes'] = os.path.getsize(path)
    if report['cuda_available']:
        for i in range(torch.cuda.device_count()):
            p = torch.cuda.get_device_properties(i)
            report['devices'].append({'index': i, 'name': p.name, 'sm': '%d.%d' % (p.major, p.minor),
                                      'multiprocessors': p.multi_processor_count, 'memory_gb': round(p.total_memory / 2 ** 30, 1)})
        n = torch.cuda.device_count()
        report['peer_access'] = [[i == j or bool(torch.cuda.can_device_access_peer(i, j)) for j in range(n)] for i in range(n)]
        try:
            loaded = native.load()
            lib['loads'], lib['abi_version'] = True, int(loaded.eb_version())
        except Exception as exc:  # noqa: BLE001 - a diagnostics tool reports, it does not raise
            lib['loads'], lib['error'] = False, '%s: %s' % (type(exc).__name__, exc)
        if lib.get('loads'):
            try:
                vmm, multicast = native.vmm_supported()
                report['vmm_allocations'], report['nvswitch_multicast'] = vmm, multicast
            except Exception as exc:  # noqa: BLE001
                report['nvswitch_multicast'] = '%s: %s' % (type(exc).__name__, exc)
    report['native_library'] = lib
    native_on = os.environ.get('ELEPHAS_B200_NATIVE', '1') != '0'
    if not report['cuda_available']:
        report['execution'] = 'CPU: torch operators, hand-written step for Dense chains (ops/cpu_mlp.py), gloo between ranks'
    elif not native_on:
        report['execution'] = 'GPU, torch operators only (ELEPHAS_B200_NATIVE=0)'
    elif not lib.get('loads'):
        report['execution'] = 'GPU, but the native library does not load: fits will raise (run python -m elephas_b200.ops.build)'
    else:
        report['execution'] = 'GPU: sm_100a kernels, CUDA-graph replay, fused collectives between ranks'
    return report


def main(argv=None):
    argv = sys.argv[1:] if argv is None else argv
    report = info()
    if '--json' in argv:
        print(json.dumps(report, indent=2))
        return 0
    print("elephas_b200 %s  (python %s, torch %s)" % (report['version'], report['python'], report['torch']))
    print("execution      : %s" % report['execution'])
    for d in report['devices']:
        print("device %d       : %s, sm_%s, %d SMs, %.1f GiB" % (d['index'], d['name'], d['sm'].replace('.', ''),
                                                                d['multiprocessors'], d['memory_gb']))
    lib = report['native_library']
    state = 'not built' if not lib['built'] else ('%d bytes' % lib['bytes'] + ('' if 'loads' not in lib else
                                                                             (', loads' if lib['loads'] else ', DOES NOT LOAD: ' + lib['error'])))
    print("native library : %s (%s)" % (lib['path'], state))
    if 'nvswitch_multicast' in report:
        print("multicast      : %s" % report['nvswitch_multicast'])
    print("spark backend  : %s" % report['spark_backend'])
    print("collectives    : %s%s" % (', '.join(report['distributed_backends']),
                                     ' (launched by torchrun)' if report['launched_by_torchrun'] else ''))
    chosen = {k: v for k, v in report['knobs'].items() if v is not None}
    print("switches set   : %s" % (', '.join('%s=%s' % kv for kv in chosen.items()) or 'none (defaults)'))
    return 0


if __name__ == '__main__':
    sys.exit(main())
