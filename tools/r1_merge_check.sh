#!/bin/bash
timeout 1200 python -m pytest tests/test_layers_gpu.py -x -q -m gpu -k "phases or conv_layers" > gpurun_out/mg_layers.log 2>&1; echo "layers rc=$?"
tail -5 gpurun_out/mg_layers.log
timeout 1500 python -m pytest tests/test_model_gpu.py -x -q -m gpu > gpurun_out/mg_model.log 2>&1; echo "model rc=$?"
tail -5 gpurun_out/mg_model.log
for r in 1 2; do
APVAE_PATCH_MERGE=0 timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/bench_nomg$r.log 2>&1
timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/bench_mg$r.log 2>&1
done
for f in nomg1 mg1 nomg2 mg2; do tail -1 gpurun_out/bench_$f.log | cut -c1-200; done
