#!/bin/bash
P=adversarial-prior-vae_b200
timeout 900 python -m pytest tests/test_layers_gpu.py -x -q -m gpu -k "mask or tma_epilogue" > gpurun_out/mask_layers.log 2>&1; echo "layers rc=$?"; tail -4 gpurun_out/mask_layers.log
timeout 1500 python -m pytest tests/test_model_gpu.py -x -q -m gpu -k "gradient or mala or energy_distribution or image_energy" > gpurun_out/mask_model.log 2>&1; echo "model rc=$?"; tail -4 gpurun_out/mask_model.log
for r in 1 2; do
  APVAE_FAST_MASK=0 python tools/bench_mala.py --chunk 1024 2>/dev/null | tail -1 | cut -c1-140
  python tools/bench_mala.py --chunk 1024 2>/dev/null | tail -1 | cut -c1-140
done
cp $P/libapvae_b200.so /tmp/new.so
for r in 1 2; do
  cp $P/libapvae_prev.so $P/libapvae_b200.so; echo "prev: $(python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-op-times 2>/dev/null | tail -1 | cut -c60-130)"
  cp /tmp/new.so $P/libapvae_b200.so; echo "new:  $(python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-op-times 2>/dev/null | tail -1 | cut -c60-130)"
done
python tools/bench_mala.py --chunk 1024 --op-times > gpurun_out/mala_ops_mask.log 2>&1
