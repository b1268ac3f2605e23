#!/bin/bash
# Round-2 one-GPU batch: GPU tests, bench line, reference arm, sampler op table, profiles.
O=gpurun_out; mkdir -p $O
timeout 1100 python -m pytest tests -q -m gpu -x > $O/r2_gputest.log 2>&1; echo "gpu tests rc=$?"; tail -3 $O/r2_gputest.log
timeout 600 python bench.py > $O/r2_bench_default.log 2> $O/r2_bench_default.err; echo "bench rc=$?"; python tools/show_bench.py $O/r2_bench_default.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/r2_bench_reference.log 2>&1; echo "ref rc=$?"; tail -1 $O/r2_bench_reference.log | cut -c1-300
for P in 0 1 2; do
  timeout 300 python tools/bench_mala.py --precise $P --op-times > $O/r2_mala_p$P.log 2> $O/r2_mala_p$P.ops; tail -1 $O/r2_mala_p$P.log | cut -c1-400
done
bash tools/profile.sh r2
timeout 600 python bench.py --images 10000 --warmup 3 --no-cpu-baseline --no-op-times --no-mala > $O/r2_strong_1gpu.log 2> $O/r2_strong_1gpu.err; echo "strong rc=$?"; python tools/show_bench.py $O/r2_strong_1gpu.log
