#!/bin/bash
# stride-2 convs on the patch kernel + CUDA-core heads: parity, then an A/B of the bench in one call.
timeout 1200 python -m pytest tests/test_layers_gpu.py -x -q -m gpu > gpurun_out/s2_layers.log 2>&1; echo "layers rc=$?"
tail -5 gpurun_out/s2_layers.log
timeout 1500 python -m pytest tests/test_model_gpu.py -x -q -m gpu > gpurun_out/s2_model.log 2>&1; echo "model rc=$?"
tail -5 gpurun_out/s2_model.log
APVAE_PATCH_S2=0 timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/bench_nos2.log 2>&1
timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/bench_s2.log 2>&1
APVAE_PATCH_S2=0 timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-op-times > gpurun_out/bench_nos2b.log 2>&1
timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-op-times > gpurun_out/bench_s2b.log 2>&1
for f in nos2 s2 nos2b s2b; do tail -1 gpurun_out/bench_$f.log | cut -c1-200; done
