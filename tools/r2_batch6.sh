#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_model_gpu.py tests/test_evaluator_gpu.py -q -m gpu > $O/r2_gputest_model.log 2>&1; echo "model tests rc=$?"; tail -6 $O/r2_gputest_model.log | cut -c1-300
for P in 2 1 0; do
  timeout 300 python tools/bench_mala.py --precise $P --op-times > $O/r2_mala_adj_p$P.log 2> $O/r2_mala_adj_p$P.ops; tail -1 $O/r2_mala_adj_p$P.log | cut -c1-330; tail -1 $O/r2_mala_adj_p$P.ops
done
grep "level_0" $O/r2_mala_adj_p2.ops | cut -c1-120
