"""Condenses an .ncu-rep (read here, no GPU needed) into the handful of numbers the roofline uses.
  python tools/ncu_summary.py gpurun_out/prof.ncu-rep [more.ncu-rep ...]"""
import csv
import io
import subprocess
import sys

WANT = [
    'gpu__time_duration.sum', 'sm__cycles_elapsed.avg', 'sm__cycles_active.avg', 'sm__cycles_elapsed.avg.per_second',
    'sm__throughput.avg.pct_of_peak_sustained_elapsed',
    'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
    'sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed',
    'sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg',
    'sm__inst_executed_pipe_uniform.sum', 'sm__inst_executed_pipe_tc.sum', 'sm__inst_executed_pipe_tmem.sum',
    'l1tex__data_pipe_tc_wavefronts_mem_shared.sum', 'l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed',
    'lts__t_bytes.sum', 'lts__t_bytes.sum.per_second', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
    'lts__t_sectors_op_read.sum', 'lts__t_sectors_op_write.sum', 'lts__t_sector_hit_rate.pct',
    'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
    'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
    'l1tex__m_xbar2l1tex_read_bytes.sum', 'l1tex__m_l1tex2xbar_write_bytes.sum',
    'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'launch__cluster_size',
    'launch__shared_mem_per_block_dynamic', 'sm__warps_active.avg.pct_of_peak_sustained_active',
    'smsp__cycles_active.avg', 'smsp__inst_executed.sum',
]


def main():
    for path in sys.argv[1:]:
        out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(out)))
        hdr, units = rows[0], rows[1]
        for vals in rows[2:]:
            d = dict(zip(hdr, vals))
            u = dict(zip(hdr, units))
            print('== %s | %s' % (path, d.get('Kernel Name', '?')[:90]))
            for k in WANT:
                if k in d:
                    print('  %-78s %s %s' % (k, d[k], u[k]))
            for k in hdr:                                 # stall reasons, whatever this ncu calls them
                if 'warp_issue_stalled' in k and k.endswith('_per_warp_active.pct') and float(d[k] or 0) > 2:
                    print('  %-78s %s' % (k, d[k]))


if __name__ == '__main__':
    main()
