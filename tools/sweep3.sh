#!/bin/bash
L192="--k 3 --h 16 --w 16 --cin 192 --cout 192 --n 1024 --outs none"
L96="--transpose --k 5 --h 32 --w 32 --cin 96 --cout 96 --n 512 --outs none"
B="python tools/bench_layer.py"
export APVAE_PATCH_CLUSTER=1
for dbg in 0 1 2 3; do echo "== debug=$dbg"; APVAE_PATCH_DEBUG=$dbg $B $L192; APVAE_PATCH_DEBUG=$dbg APVAE_PATCH_MACC=2 $B $L96; APVAE_PATCH_DEBUG=$dbg APVAE_PATCH_MACC=1 $B $L96; done
echo "== stages"
for bs in 2 4 8 12; do APVAE_PATCH_BSTAGES=$bs APVAE_PATCH_MACC=2 $B $L96; done
for as in 1 2 4; do APVAE_PATCH_ASTAGES=$as APVAE_PATCH_MACC=2 $B $L96; done
for bs in 2 3 5; do APVAE_PATCH_BSTAGES=$bs $B $L192; done
APVAE_PATCH=0 $B $L192
APVAE_PATCH=0 $B $L96
