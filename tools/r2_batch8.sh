#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 1100 python -m pytest tests -q -m gpu > $O/r2_gputest.log 2>&1; echo "gpu tests rc=$?"; tail -4 $O/r2_gputest.log | cut -c1-300
timeout 600 python bench.py > $O/r2_bench_default.log 2> $O/r2_bench_default.err; echo "bench rc=$?"; python tools/show_bench.py $O/r2_bench_default.log
timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-mala > $O/r2_bench_again.log 2>/dev/null; python tools/show_bench.py $O/r2_bench_again.log
bash tools/profile.sh r2
