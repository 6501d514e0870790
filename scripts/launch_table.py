"""Print a per-launch table from an `ncu --csv --metrics ...` log."""
import csv, sys
rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
h = rows[0]; ki = h.index('Kernel Name'); mi = h.index('Metric Name'); vi = h.index('Metric Value')
d = {}
for r in rows[1:]:
    d.setdefault(int(r[0]), {'k': r[ki].split('(')[0]})[r[mi]] = r[vi]
short = {'gpu__time_duration.sum': 'ns', 'smsp__thread_inst_executed_per_inst_executed.ratio': 'lanes', 'smsp__inst_executed.sum': 'winst',
         'smsp__issue_active.avg.pct_of_peak_sustained_active': 'issue%', 'launch__registers_per_thread': 'regs',
         'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed': 'l1wf%', 'lts__t_sector_hit_rate.pct': 'l2hit%',
         'dram__bytes_read.sum': 'dramR', 'dram__bytes_write.sum': 'dramW', 'l1tex__t_sector_hit_rate.pct': 'l1hit%'}
lim = int(sys.argv[2]) if len(sys.argv) > 2 else 60
for i in sorted(d)[:lim]:
    x = d[i]
    print(i, x['k'][-28:], ' '.join('%s=%s' % (short.get(k, k), v) for k, v in x.items() if k != 'k'))
