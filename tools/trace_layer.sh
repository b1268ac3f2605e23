#!/bin/bash
# Clock-stamp traces of CTA 0 (APVAE_PATCH_TRACE) for a few layer shapes; slots per tile:
#  0 MMA warp before tempty wait, 1 after, 2 first patch landed, 3 all MMAs of the tile issued,
#  4 epilogue warp 4 before tfull wait, 5 after; block 0: 6 accumulator in registers, 7 math done, 8 staging free,
#  9 staged, 10 fenced, 11 stores issued; 12 block 1 done, 13 last block done,
#  14 patch producer issues the tile's first load, 15 its last
export APVAE_PATCH_TRACE=1
D10="--k 3 --h 32 --w 32 --cin 16 --cout 16 --n 1000 --outs act --reps 2"
L96="--transpose --k 5 --h 32 --w 32 --cin 96 --cout 96 --n 1000 --reps 2"
L192="--k 3 --h 16 --w 16 --cin 192 --cout 192 --n 1000 --reps 2"
echo "== D10 base"; python tools/bench_layer.py $D10 2>&1 | head -24
echo "== D10 noload noepi"; APVAE_PATCH_DEBUG=14 python tools/bench_layer.py $D10 2>&1 | head -16
echo "== D10 noload nomma"; APVAE_PATCH_DEBUG=16 python tools/bench_layer.py $D10 2>&1 | head -16
echo "== L96x5"; python tools/bench_layer.py $L96 2>&1 | head -20
echo "== L192"; python tools/bench_layer.py $L192 2>&1 | head -20
