#!/bin/bash
# fused CUDA-core D_psi/level_0 block: model parity, A/B of the bench, pair-mode probes.
timeout 1500 python -m pytest tests/test_model_gpu.py -x -q -m gpu > gpurun_out/b0_model.log 2>&1; echo "model rc=$?"
tail -5 gpurun_out/b0_model.log
APVAE_SMALL_BLOCK=0 timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/bench_nob0.log 2>&1
timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/bench_b0.log 2>&1
APVAE_SMALL_BLOCK=0 timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-op-times > gpurun_out/bench_nob0b.log 2>&1
timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-op-times > gpurun_out/bench_b0b.log 2>&1
for f in nob0 b0 nob0b b0b; do tail -1 gpurun_out/bench_$f.log | cut -c1-200; done
python tools/layer_suite.py --only pair_ 2>&1 | tail -8
python tools/layer_suite.py --only pair_ 2>&1 | tail -8
