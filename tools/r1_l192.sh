#!/bin/bash
L192="--k 3 --h 16 --w 16 --cin 192 --cout 192 --n 2000 --res same"
for d in 0 1 2 3 4 5; do echo "debug $d: $(APVAE_PATCH_DEBUG=$d python tools/bench_layer.py $L192 2>&1 | tail -1)"; done
for b in 2 3 4 6; do echo "bstages $b: $(APVAE_PATCH_BSTAGES=$b python tools/bench_layer.py $L192 2>&1 | tail -1)"; done
for a in 2 3; do echo "astages $a: $(APVAE_PATCH_ASTAGES=$a python tools/bench_layer.py $L192 2>&1 | tail -1)"; done
echo "acap 6: $(APVAE_PATCH_ACAP=6 python tools/bench_layer.py $L192 2>&1 | tail -1)"
export APVAE_PATCH_TRACE=1
python tools/bench_layer.py --k 3 --h 16 --w 16 --cin 192 --cout 192 --n 2000 --res same --reps 2 2>&1 | head -30
