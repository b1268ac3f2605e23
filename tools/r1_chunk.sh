#!/bin/bash
for r in 1 2; do
for c in 7400 18500 37000; do
  timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-op-times --chunk $c > gpurun_out/bench_chunk$c.log 2>&1
  echo "chunk $c: $(tail -1 gpurun_out/bench_chunk$c.log | cut -c60-130) $(tail -1 gpurun_out/bench_chunk$c.log | grep -o '"nll": {[^}]*}')"
  nvidia-smi --query-gpu=memory.used --format=csv,noheader
done
done
