#!/bin/bash
timeout 900 python -m pytest tests/test_layers_gpu.py -x -q -m gpu -k "mask or phases" > gpurun_out/pm_layers.log 2>&1; echo "layers rc=$?"; tail -4 gpurun_out/pm_layers.log
timeout 1500 python -m pytest tests/test_model_gpu.py -x -q -m gpu -k "gradient or mala or energy_distribution or image_energy" > gpurun_out/pm_model.log 2>&1; echo "model rc=$?"; tail -4 gpurun_out/pm_model.log
python tools/bench_mala.py 2>/dev/null | tail -1 | cut -c1-140
python tools/bench_mala.py --op-times 2>&1 | grep "\^T\|sum of" | grep "conv_1\|sum of"
