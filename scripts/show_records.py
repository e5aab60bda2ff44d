"""Print the key fields of bench JSON records: python scripts/show_records.py gpurun_out/m4_*.json"""
import json
import sys

KEYS = ('value', 'ms_per_step', 'final_loss', 'loss_before', 'replicas_identical', 'seconds', 'device_ms_per_step_inside_fit',
        'device_timed_samples_per_sec', 'host_phases_ms', 'repeat_fit_seconds', 'predict_samples_per_sec',
        'predict_warm_samples_per_sec')
for path in sys.argv[1:]:
    try:
        lines = [l for l in open(path).read().strip().splitlines() if l.startswith('{')]
        for line in lines:
            d = json.loads(line)
            if 'kernel' in d or 'what' in d:
                print(path, line[:260])
                continue
            row = {k: d[k] for k in KEYS if d.get(k) is not None}
            cfg = d.get('config') or {}
            if isinstance(cfg, dict):
                for k in ('collective', 'cuda_graph'):
                    if k in cfg:
                        row[k] = cfg[k]
            print(path, d.get('impl'), d.get('n_gpus'), row)
            for k in ('e2e', 'e2e_fp32_host_features', 'sustained'):
                if d.get(k):
                    print('    ', k, {a: b for a, b in d[k].items() if a in ('value', 'ms_per_step', 'steps', 'device_ms_per_step_inside_fit', 'h2d_bytes_per_step', 'final_loss')})
            if d.get('clocks'):
                print('     clocks', d['clocks'])
    except Exception as exc:
        print(path, 'ERR', exc)
        try:
            print(open(path.replace('.json', '.err')).read()[-1200:])
        except OSError:
            pass
