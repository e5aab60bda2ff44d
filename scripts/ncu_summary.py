"""Markdown summary of .ncu-rep files (read on a machine without a GPU): python scripts/ncu_summary.py a.ncu-rep b.ncu-rep"""
import csv
import io
import subprocess
import sys

COLS = [('gpu__time_duration.sum', 'time us', 1.0), ('launch__grid_size', 'grid', 1), ('launch__block_size', 'block', 1),
        ('launch__registers_per_thread', 'regs', 1), ('launch__shared_mem_per_block_dynamic', 'dyn smem KB', 1.0),
        ('sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active', 'tensor pipe %', 1.0),
        ('sm__throughput.avg.pct_of_peak_sustained_elapsed', 'SM %', 1.0),
        ('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'DRAM %', 1.0),
        ('dram__bytes_read.sum', 'DRAM read MB', 1.0), ('dram__bytes_write.sum', 'DRAM write MB', 1.0),
        ('lts__t_sector_hit_rate.pct', 'L2 hit %', 1.0), ('sm__warps_active.avg.pct_of_peak_sustained_active', 'occupancy %', 1.0)]
STALL = 'smsp__average_warps_issue_stalled_%s_per_issue_active.ratio'
STALLS = ('long_scoreboard', 'barrier', 'sleeping', 'wait', 'short_scoreboard', 'lg_throttle', 'membar', 'mio_throttle',
          'math_pipe_throttle', 'branch_resolving', 'no_instruction', 'dispatch_stall')


def main():
    for path in sys.argv[1:]:
        text = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(text)))
        hdr, units = rows[0], rows[1]
        print('### %s\n' % path.split('/')[-1])
        print('| kernel | ' + ' | '.join(c[1] for c in COLS) + ' | top stalls (warps per issue) |')
        print('|---|' + '---:|' * len(COLS) + '---|')
        for row in rows[2:]:
            get = lambda name: row[hdr.index(name)] if name in hdr else ''  # noqa: E731
            cells = []
            for name, _, kind in COLS:
                v = get(name)
                try:
                    cells.append('%.1f' % float(v.replace(',', '')) if isinstance(kind, float) else v)
                except ValueError:
                    cells.append(v)
            stalls = []
            for s in STALLS:
                try:
                    stalls.append((float(get(STALL % s)), s))
                except ValueError:
                    pass
            top = ', '.join('%s %.2f' % (s, v) for v, s in sorted(stalls, reverse=True)[:3])
            print('| `%s` | %s | %s |' % (get('Kernel Name').split('(')[0][:60], ' | '.join(cells), top))
        print('\nunits: ' + ', '.join('%s [%s]' % (c[1], units[hdr.index(c[0])]) for c in COLS if c[0] in hdr) + '\n')


if __name__ == '__main__':
    main()
