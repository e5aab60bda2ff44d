"""Markdown table of the metrics that matter from an .ncu-rep (one row per captured kernel)."""
import csv
import io
import subprocess
import sys

WANT = [('gpu__time_duration.sum', 'time'),
        ('dram__bytes_read.sum', 'dram rd'), ('dram__bytes_write.sum', 'dram wr'),
        ('dram__bytes.sum.per_second', 'dram B/s'),
        ('FBSP.TriageCompute.dram__throughput.avg.pct_of_peak_sustained_elapsed', 'dram %peak'),
        ('lts__throughput.avg.pct_of_peak_sustained_elapsed', 'L2 %peak'),
        ('sm__throughput.avg.pct_of_peak_sustained_elapsed', 'SM %peak'),
        ('sm__warps_active.avg.pct_of_peak_sustained_active', 'occupancy %'),
        ('launch__registers_per_thread', 'regs'), ('launch__grid_size', 'grid'), ('launch__block_size', 'block'),
        ('sm__inst_executed_pipe_tc.sum', 'tensor-core inst'),
        ('smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'stall long-scoreboard')]


def main(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h = next(i for i, r in enumerate(rows) if r and r[0] == 'ID')
    names, units = rows[h], rows[h + 1]
    idx = {n: i for i, n in enumerate(names)}
    cols = [(m, label) for m, label in WANT if m in idx]
    print('| kernel | ' + ' | '.join('%s [%s]' % (label, units[idx[m]] or '-') for m, label in cols) + ' |')
    print('|---|' + '---:|' * len(cols))
    for r in rows[h + 2:]:
        if len(r) < len(names):
            continue
        k = r[idx['Kernel Name']].split('(')[0][:60]
        vals = []
        for m, _ in cols:
            v = r[idx[m]]
            try:
                f = float(v.replace(',', ''))
                v = '%.3g' % f if abs(f) < 1e5 else '%.4g' % f
            except ValueError:
                pass
            vals.append(v)
        print('| `%s` | ' % k + ' | '.join(vals) + ' |')


if __name__ == '__main__':
    main(sys.argv[1])
