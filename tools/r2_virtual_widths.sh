#!/bin/bash
# One-GPU validation of the virtual-width (arch M) patch ops: the arch-M layer cases and model tests, then configs[0]
# with the planner's accumulator count, with one accumulator per tile (the seven-tile plan), and on the tap-GEMM kernel.
# FULL=1 runs the whole GPU suite as well.  Usage (on the GPU box): tools/r2_virtual_widths.sh
mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_layers_gpu.py -x -q --timeout 60 -k "-28-28- or -14-14- or -7-7- or -14-28-" > gpurun_out/r2_virt_layers.log 2>&1
rc=$?; echo "layers rc=$rc"; tail -4 gpurun_out/r2_virt_layers.log | cut -c1-300
[ $rc -eq 0 ] || exit $rc
timeout 150 python -m pytest tests/test_model_gpu.py -x -q --timeout 100 -k "M40 or arch_m or errors_are_loud or energy_grad_requires" > gpurun_out/r2_virt_model.log 2>&1
rc=$?; echo "model rc=$rc"; tail -4 gpurun_out/r2_virt_model.log | cut -c1-300
[ $rc -eq 0 ] || exit $rc
if [ "${FULL:-0}" = 1 ]; then
  timeout 420 python -m pytest tests -q -m gpu -x --timeout 120 > gpurun_out/r2_gputest_virt.log 2>&1
  rc=$?; echo "suite rc=$rc"; tail -4 gpurun_out/r2_gputest_virt.log | cut -c1-300
fi
M="--arch M --z-dim 40 --n-z 100 --batch 64 --chunk 6400 --steps 20 --warmup 5 --no-mala --no-cpu-baseline"
timeout 90 python bench.py $M > gpurun_out/r2_bench_c0_archM_virtual.log 2>/dev/null
timeout 90 python bench.py $M --tuning patch_macc=1 > gpurun_out/r2_bench_c0_archM_macc1.log 2>/dev/null
[ "${FULL:-0}" = 1 ] && timeout 90 python bench.py $M --tuning patch_virtual=-1 > gpurun_out/r2_bench_c0_archM_tapkernel.log 2>/dev/null
python tools/show_bench.py gpurun_out/r2_bench_c0_archM_*.log
exit $rc
