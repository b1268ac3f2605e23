#!/bin/bash
# weight-item depth A/B for the 192-wide layers: taps per item 3 (default) / 1 / 2, per-op tables in one call
O=gpurun_out; mkdir -p $O
for T in 0 1 2; do
  timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-mala --tuning patch_tpi=$T > $O/r2_tpi$T.log 2> $O/r2_tpi$T.err; python tools/show_bench.py $O/r2_tpi$T.log
done
