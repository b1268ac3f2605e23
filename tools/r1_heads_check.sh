#!/bin/bash
# CUDA-core heads kernel: model parity, then an A/B of the bench against the 1x1 tensor-core op in one call.
timeout 1500 python -m pytest tests/test_model_gpu.py -x -q -m gpu > gpurun_out/heads_model.log 2>&1; echo "model rc=$?"
tail -5 gpurun_out/heads_model.log
APVAE_HEADS_GEMM=1 timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/bench_headsgemm.log 2>&1
timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/bench_heads.log 2>&1
APVAE_HEADS_GEMM=1 timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-op-times > gpurun_out/bench_headsgemm2.log 2>&1
timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-op-times > gpurun_out/bench_heads2.log 2>&1
for f in headsgemm heads headsgemm2 heads2; do tail -1 gpurun_out/bench_$f.log | cut -c1-200; done
