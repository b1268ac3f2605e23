#!/bin/bash
timeout 1800 python -m pytest tests/test_model_gpu.py tests/test_evaluator_gpu.py -x -q -m gpu > gpurun_out/tiny_model.log 2>&1; echo "model rc=$?"; tail -4 gpurun_out/tiny_model.log
for r in 1 2; do
  APVAE_TINY=0 python tools/bench_mala.py --chunk 1024 2>/dev/null | tail -1 | cut -c1-140
  python tools/bench_mala.py --chunk 1024 2>/dev/null | tail -1 | cut -c1-140
done
python tools/bench_mala.py --chunk 1024 --op-times 2>&1 | grep "tiny\|sum of"
python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-op-times 2>/dev/null | tail -1 | cut -c60-130
