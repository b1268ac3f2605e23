#!/bin/bash
set -e
L192="--k 3 --h 16 --w 16 --cin 192 --cout 192 --n 1024 --reps 3"
L96="--transpose --k 5 --h 32 --w 32 --cin 96 --cout 96 --n 512 --reps 3"
export APVAE_PATCH_CLUSTER=1
APVAE_PATCH=0 python tools/bench_layer.py $L192
python tools/bench_layer.py $L192
APVAE_PATCH_MACC=2 python tools/bench_layer.py $L96
NCU="ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s 3 -c 1"
APVAE_PATCH=0 $NCU -o gpurun_out/prof_tap192 python tools/bench_layer.py $L192
$NCU -o gpurun_out/prof_patch192 python tools/bench_layer.py $L192
APVAE_PATCH_MACC=2 $NCU -o gpurun_out/prof_patch96 python tools/bench_layer.py $L96
