#!/bin/bash
# two epilogue warpgroups: layer tests, then A/B of the whole step and per-op tables (auto vs off), sampler
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_layers_gpu.py -q -m gpu -x > $O/r2_gputest_layers.log 2>&1; echo "layer tests rc=$?"; tail -4 $O/r2_gputest_layers.log | cut -c1-300
for E in 0 -1 0 -1; do
  timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-mala --tuning patch_epi2=$E > $O/r2_epi2_${E}_$RANDOM.log 2>/dev/null
done
for f in $O/r2_epi2_*.log; do python tools/show_bench.py $f; done
timeout 300 python tools/bench_mala.py --precise 0 --op-times > $O/r2_mala_e2_p0.log 2> $O/r2_mala_e2_p0.ops; tail -1 $O/r2_mala_e2_p0.log | cut -c1-300; tail -1 $O/r2_mala_e2_p0.ops
timeout 300 python tools/bench_mala.py --precise 2 > $O/r2_mala_e2_p2.log 2>/dev/null; tail -1 $O/r2_mala_e2_p2.log | cut -c1-300
