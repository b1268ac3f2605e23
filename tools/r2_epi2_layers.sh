#!/bin/bash
# one vs two epilogue warpgroups, layer by layer (bench_layer: 5 timed launches each, 2368 samples)
run() { for E in -1 1; do python tools/bench_layer.py "$@" --n 2368 --reps 5 --tuning patch_epi2=$E 2>&1 | tail -1 | sed "s/^/epi2=$E  /"; done; }
echo "== G level_8/conv_1"; run --transpose --k 3 --h 32 --w 32 --cin 96 --cout 96 --res even
echo "== G level_8/shortcut"; run --transpose --k 1 --h 16 --w 16 --cin 192 --cout 96 --outs raw
echo "== G level_3/shortcut"; run --transpose --k 1 --h 8 --w 8 --cin 192 --cout 192 --outs raw
echo "== G level_0/conv_0"; run --transpose --k 3 --h 8 --w 8 --cin 2 --cout 192 --outs act
echo "== D 16->16"; run --k 3 --h 32 --w 32 --cin 16 --cout 16 --outs act
echo "== D 16->32"; run --k 3 --h 32 --w 32 --cin 16 --cout 32
echo "== D 32->32"; run --k 3 --h 32 --w 32 --cin 32 --cout 32 --outs act
echo "== D 32->64"; run --k 3 --h 32 --w 32 --cin 32 --cout 64
echo "== D 64->64"; run --k 3 --h 32 --w 32 --cin 64 --cout 64 --outs act
echo "== D 64->128 s2"; run --k 3 --stride 2 --h 32 --w 32 --cin 64 --cout 128
echo "== B 64->64 @16"; run --transpose --k 3 --h 16 --w 16 --cin 64 --cout 64 --res same
echo "== B 32->32 @32"; run --transpose --k 3 --h 32 --w 32 --cin 32 --cout 32 --res same
echo "== B 32->32 5x5 @64"; run --transpose --k 5 --h 64 --w 64 --cin 32 --cout 32 --res same
