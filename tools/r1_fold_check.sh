#!/bin/bash
# 8-wide (folded) patch ops: parity first, then an A/B of the bench with and without them in one call.
set -x
timeout 1200 python -m pytest tests/test_layers_gpu.py -x -q -m gpu > gpurun_out/fold_layers.log 2>&1; echo "layers rc=$?" 
tail -5 gpurun_out/fold_layers.log
timeout 1500 python -m pytest tests/test_model_gpu.py -x -q -m gpu > gpurun_out/fold_model.log 2>&1; echo "model rc=$?"
tail -5 gpurun_out/fold_model.log
APVAE_PATCH_FOLD=0 timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/bench_nofold.log 2>&1
timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/bench_fold.log 2>&1
APVAE_PATCH_FOLD=0 timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-op-times > gpurun_out/bench_nofold2.log 2>&1
timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-op-times > gpurun_out/bench_fold2.log 2>&1
for f in nofold fold nofold2 fold2; do tail -1 gpurun_out/bench_$f.log | cut -c1-200; done
