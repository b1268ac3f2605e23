"""Summaries of the ncu CSV logs tools/r1_profile.sh brings back (read here, no GPU needed):
  python tools/ncu_launch_summary.py list gpurun_out/r1_launch_list.csv        # per-kernel totals of the launch list
  python tools/ncu_launch_summary.py chunk gpurun_out/r1_chunk_metrics.csv     # one line per launch of the metric window"""
import csv
import sys
from collections import OrderedDict, defaultdict


def read(path):
    rows = list(csv.reader(open(path)))
    h = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
    return rows[h], [r for r in rows[h + 1:] if r and r[0].isdigit()]


def short(name):
    return name.replace('void ', '').replace('apvae::', '').replace('<unnamed>::', '').split('(')[0][:58]


def main():
    mode, path = sys.argv[1], sys.argv[2]
    hdr, rows = read(path)
    if mode == 'list':
        if 'Metric Value' in hdr:      # long format
            ix = {h: i for i, h in enumerate(hdr)}
            rows = [(r[ix['ID']], r[ix['Kernel Name']], r[ix['Metric Value']]) for r in rows if r[ix['Metric Name']] == 'gpu__time_duration.sum']
        agg = defaultdict(lambda: [0, 0.0])
        for _, name, ns in rows:
            a = agg[short(name)]
            a[0] += 1
            a[1] += float(ns.replace(',', '')) / 1e6
        tot = sum(a[1] for a in agg.values())
        print('launches %d, total %.2f ms' % (sum(a[0] for a in agg.values()), tot))
        for name, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            print('%-60s n=%5d %10.2f ms %6.1f%%  avg %8.1f us' % (name, n, ms, 100 * ms / tot, 1e3 * ms / n))
    else:
        ix = {h: i for i, h in enumerate(hdr)}
        per = OrderedDict()
        for r in rows:
            d = per.setdefault(int(r[ix['ID']]), {'name': short(r[ix['Kernel Name']]), 'grid': r[ix['Grid Size']]})
            d[r[ix['Metric Name']]] = float(r[ix['Metric Value']].replace(',', ''))
        print('id  kernel                                   grid        us  tensor%%  dramR MB  dramW MB  dram%%  L2hit%%  smemTC%%   GHz')
        for i, d in per.items():
            print('%-3d %-40s %-9s %8.1f %7.1f %9.1f %9.1f %6.1f %7.1f %8.1f %5.2f' % (
                i, d['name'][:40], d['grid'].replace(', 1, 1', '').strip('()'), d.get('gpu__time_duration.sum', 0) / 1e3,
                d.get('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 0),
                d.get('dram__bytes_read.sum', 0) / 1e6, d.get('dram__bytes_write.sum', 0) / 1e6,
                d.get('dram__throughput.avg.pct_of_peak_sustained_elapsed', 0), d.get('lts__t_sector_hit_rate.pct', 0),
                d.get('l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed', 0),
                d.get('sm__cycles_elapsed.avg.per_second', 0) / 1e9))


if __name__ == '__main__':
    main()
