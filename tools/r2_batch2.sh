#!/bin/bash
# Round-2 second one-GPU batch: full GPU tests on the split-epilogue build, sampler modes, A-ring depth A/B, --set full of the dominant op.
O=gpurun_out; mkdir -p $O
timeout 1100 python -m pytest tests -q -m gpu -x > $O/r2_gputest.log 2>&1; echo "gpu tests rc=$?"; tail -3 $O/r2_gputest.log
for P in 2 1; do
  timeout 300 python tools/bench_mala.py --precise $P --op-times > $O/r2_mala_patch_p$P.log 2> $O/r2_mala_patch_p$P.ops; tail -1 $O/r2_mala_patch_p$P.log | cut -c1-330; tail -1 $O/r2_mala_patch_p$P.ops
done
for A in 0 6 8; do
  timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-mala --tuning patch_a_cap=$A > $O/r2_acap$A.log 2> $O/r2_acap$A.err; python tools/show_bench.py $O/r2_acap$A.log
done
ONLY_FULL=1 bash tools/profile.sh r2
