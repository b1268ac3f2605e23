#!/bin/bash
# One-GPU validation batch (tools/gpu_retry.sh 2400 "bash tools/one_gpu_batch.sh"): GPU tests, sampler modes, bench line, reference arm, smoke, profiles (tools/profile.sh r2)
O=gpurun_out; mkdir -p $O
timeout 1100 python -m pytest tests -q -m gpu > $O/r2_gputest.log 2>&1; echo "gpu tests rc=$?"; tail -4 $O/r2_gputest.log | cut -c1-300
for P in 2 1 0; do timeout 300 python tools/bench_mala.py --precise $P --op-times > $O/r2_mala_fin_p$P.log 2> $O/r2_mala_fin_p$P.ops; tail -1 $O/r2_mala_fin_p$P.log | cut -c1-300; tail -1 $O/r2_mala_fin_p$P.ops; done
timeout 600 python bench.py > $O/r2_bench_default.log 2> $O/r2_bench_default.err; echo "bench rc=$?"; python tools/show_bench.py $O/r2_bench_default.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/r2_bench_reference.log 2>&1; echo "ref rc=$?"
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
bash tools/profile.sh r2
