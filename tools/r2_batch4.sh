#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 1100 python -m pytest tests -q -m gpu > $O/r2_gputest.log 2>&1; echo "gpu tests rc=$?"; tail -6 $O/r2_gputest.log | cut -c1-300
for P in 2 1 0; do
  timeout 300 python tools/bench_mala.py --precise $P --op-times > $O/r2_mala_blk_p$P.log 2> $O/r2_mala_blk_p$P.ops; tail -1 $O/r2_mala_blk_p$P.log | cut -c1-330; tail -1 $O/r2_mala_blk_p$P.ops
done
bash tools/sweep_zdim.sh 1 $O/r2_sweep_zdim_1gpu.jsonl
