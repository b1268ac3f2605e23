"""One line per profiled launch of an .ncu-rep (read here): duration, tensor-pipe activity, DRAM traffic, L2 hit rate.
  python tools/ncu_table.py rep [rep ...]"""
import csv, io, subprocess, sys
KEYS = [('gpu__time_duration.sum', 'us'), ('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'tensor%'),
        ('dram__bytes_read.sum', 'dramR'), ('dram__bytes_write.sum', 'dramW'), ('lts__t_sector_hit_rate.pct', 'L2hit%'),
        ('smsp__inst_executed.sum', 'warp-inst'), ('launch__grid_size', 'grid'),
        ('sm__cycles_elapsed.avg.per_second', 'GHz'), ('launch__shared_mem_per_block_dynamic', 'smemKB'),
        ('launch__registers_per_thread', 'regs')]
for path in sys.argv[1:]:
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h, u = rows[0], rows[1]
    units = dict(zip(h, u))
    print('== ' + path)
    print('id  kernel                      ' + ' '.join('%10s' % ('%s[%s]' % (n, units.get(k, '')))[:10] for k, n in KEYS))
    for r in rows[2:]:
        d = dict(zip(h, r))
        name = d['Kernel Name'].replace('void ', '').replace('apvae::', '').split('(')[0]
        print('%-3s %-27s ' % (d['ID'], name[:27]) + ' '.join('%10s' % d.get(k, '?')[:10] for k, _ in KEYS))
