#!/bin/bash
# Round-1 profile refresh at the bench defaults.  Only CSV / text comes back (an .ncu-rep of a whole chunk is > 64 MiB).
#  1. plain run (must exit 0)   2. ncu launch list of the same command (gpu__time_duration)
#  3. per-launch metric list (DRAM bytes, tensor pipe, L2 hit, clock) over a window covering one whole chunk
#  4. one --set full capture of the dominant op (G_theta/level_9: patchgemm_kernel<32, 1>), exported to CSV on the box
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-op-times"
O=gpurun_out
timeout 600 $CMD > $O/prof_plain.log 2>&1 || { echo "plain run failed"; tail -5 $O/prof_plain.log; exit 1; }
tail -1 $O/prof_plain.log | cut -c1-200
L=$(tail -1 $O/prof_plain.log | python -c "import json,sys; print(json.loads(sys.stdin.read())['gpu_launches'])")
echo "launches per step: $L"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file $O/r1_launch_list.csv $CMD > $O/prof_ncu_list.log 2>&1
echo "launch list rc=$? lines=$(wc -l < $O/r1_launch_list.csv)"
SKIP=$((3 * L + 60))
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,lts__t_sector_hit_rate.pct,sm__cycles_elapsed.avg.per_second,launch__grid_size,smsp__inst_executed.sum,l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed,dram__throughput.avg.pct_of_peak_sustained_elapsed
timeout 900 ncu --metrics $M --clock-control none -s $SKIP -c 60 --csv --log-file $O/r1_chunk_metrics.csv $CMD > $O/prof_ncu_metrics.log 2>&1
echo "metrics rc=$? lines=$(wc -l < $O/r1_chunk_metrics.csv)"
# first level_9 launch (pair-mode kc = 32 patch kernel) inside the window
ID=$(python - <<'PY'
import csv
rows = list(csv.reader(open('gpurun_out/r1_chunk_metrics.csv')))
hdr = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
ids = sorted({int(r[0]) for r in rows[hdr + 2:] if len(r) > 4 and 'patchgemm_kernel<32, 1>' in r[4]})
print(ids[0] if ids else -1)
PY
)
echo "first level_9 launch in window: $ID"
if [ "$ID" -ge 0 ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -s $((SKIP + ID)) -c 2 -f -o /tmp/r1_dom $CMD > $O/prof_ncu_full.log 2>&1
  echo "full rc=$?"
  ncu -i /tmp/r1_dom.ncu-rep --page raw --csv > $O/r1_dom_raw.csv 2>/dev/null
  ncu -i /tmp/r1_dom.ncu-rep --page details --csv > $O/r1_dom_details.csv 2>/dev/null
  ls -la /tmp/r1_dom.ncu-rep $O/r1_dom_raw.csv
fi
