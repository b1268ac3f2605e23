#!/bin/bash
for r in 1 2; do
for k in 1500 500 250 100 1; do
  APVAE_PATCH_2CTA_MINK=$k timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-op-times > gpurun_out/bench_mink$k.log 2>&1
  echo "mink $k: $(tail -1 gpurun_out/bench_mink$k.log | cut -c60-130)"
done
done
APVAE_PATCH_2CTA_MINK=1 timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/bench_mink1_ops.log 2>&1
