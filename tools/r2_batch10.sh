#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 1100 python -m pytest tests -q -m gpu > $O/r2_gputest.log 2>&1; echo "gpu tests rc=$?"; tail -4 $O/r2_gputest.log | cut -c1-300
rm -f $O/r2_epi2_*.log
for E in 0 -1 0 -1; do
  timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-mala --tuning patch_epi2=$E > $O/r2_epi2_${E}_$RANDOM.log 2>/dev/null
done
for f in $O/r2_epi2_*.log; do python tools/show_bench.py $f; done
for P in 2 0; do timeout 300 python tools/bench_mala.py --precise $P --op-times > $O/r2_mala_e2_p$P.log 2> $O/r2_mala_e2_p$P.ops; tail -1 $O/r2_mala_e2_p$P.log | cut -c1-300; tail -1 $O/r2_mala_e2_p$P.ops; done
timeout 300 python bench.py --arch B --dataset cifar10 --z-dim 128 --steps 3 --warmup 3 --no-cpu-baseline --no-mala > $O/r2_archB_epi2.log 2>/dev/null; python tools/show_bench.py $O/r2_archB_epi2.log
timeout 300 python bench.py --arch B --dataset cifar10 --z-dim 128 --steps 3 --warmup 3 --no-cpu-baseline --no-mala --tuning patch_epi2=-1 > $O/r2_archB_noepi2.log 2>/dev/null; python tools/show_bench.py $O/r2_archB_noepi2.log
