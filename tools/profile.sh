#!/bin/bash
# Profile refresh at the bench defaults:  bash tools/profile.sh r2   (one GPU; only CSV / text comes back).
#  1. plain run (must exit 0)   2. ncu launch list of the same command (gpu__time_duration)
#  3. per-launch metric list (DRAM bytes, tensor pipe, L2 hit, clock) over a window covering one whole chunk
#  4. one --set full capture of the dominant op (G_theta/level_9: pair-mode kc = 32 patch kernel), exported to CSV on the box
R=${1:-r2}
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-op-times --no-mala"
O=gpurun_out
mkdir -p $O
# ONLY_FULL=1: just the plain run and step 4 (re-uses gpurun_out/${R}_chunk_metrics.csv when present, else profiles/)
timeout 600 $CMD > $O/prof_plain.log 2>&1 || { echo "plain run failed"; tail -5 $O/prof_plain.log; exit 1; }
tail -1 $O/prof_plain.log | cut -c1-200
L=$(tail -1 $O/prof_plain.log | python -c "import json,sys; print(json.loads(sys.stdin.read())['gpu_launches'])")
echo "launches per step: $L"
if [ -z "$ONLY_FULL" ]; then
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file $O/${R}_launch_list.csv $CMD > $O/prof_ncu_list.log 2>&1
echo "launch list rc=$? lines=$(wc -l < $O/${R}_launch_list.csv)"
SKIP=$((3 * L + 60))
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,lts__t_sector_hit_rate.pct,sm__cycles_elapsed.avg.per_second,launch__grid_size,smsp__inst_executed.sum,l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed,dram__throughput.avg.pct_of_peak_sustained_elapsed
timeout 900 ncu --metrics $M --clock-control none -s $SKIP -c 60 --csv --log-file $O/${R}_chunk_metrics.csv $CMD > $O/prof_ncu_metrics.log 2>&1
echo "metrics rc=$? lines=$(wc -l < $O/${R}_chunk_metrics.csv)"
else
SKIP=$((3 * L + 60)); [ -f $O/${R}_chunk_metrics.csv ] || cp profiles/${R}_chunk_metrics.csv $O/
fi
ID=$(python - $O/${R}_chunk_metrics.csv <<'PY'
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
ids = sorted({int(r[0]) for r in rows[hdr + 2:] if len(r) > 4 and 'patchgemm_kernel<32, 1' in r[4]})
print(ids[2] if len(ids) > 2 else -1)     # <32, pair> launches of a chunk in program order: level_8/conv_1, level_9/conv_0, level_9/conv_1
PY
)
echo "G_theta/level_9/conv_1 launch in window: $ID"
if [ "$ID" -ge 0 ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -s $((SKIP + ID)) -c 1 -f -o /tmp/${R}_dom $CMD > $O/prof_ncu_full.log 2>&1
  echo "full rc=$?"
  ncu -i /tmp/${R}_dom.ncu-rep --page raw --csv > $O/${R}_dom_raw.csv 2>/dev/null
  ncu -i /tmp/${R}_dom.ncu-rep --page details --csv > $O/${R}_dom_details.csv 2>/dev/null
  ls -la /tmp/${R}_dom.ncu-rep $O/${R}_dom_raw.csv
fi
