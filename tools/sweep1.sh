#!/bin/bash
# kernel-variant sweep on two representative layers (results -> gpurun_out/sweep1.log)
L192="--k 3 --h 16 --w 16 --cin 192 --cout 192 --n 256"
L96="--transpose --k 5 --h 32 --w 32 --cin 96 --cout 96 --n 256"
B="python tools/bench_layer.py"
for outs in raw,act none; do
  echo "== outs=$outs"
  APVAE_PATCH=0 $B $L192 --outs $outs
  for c in 1 2 4; do APVAE_PATCH_CLUSTER=$c $B $L192 --outs $outs; done
  APVAE_PATCH=0 $B $L96 --outs $outs
  for m in 1 2; do for c in 1 2 4; do APVAE_PATCH_MACC=$m APVAE_PATCH_CLUSTER=$c $B $L96 --outs $outs; done; done
done
echo "== other shapes"
$B --k 3 --h 32 --w 32 --cin 64 --cout 64 --n 256
$B --k 3 --h 32 --w 32 --cin 16 --cout 16 --n 256
$B --k 3 --h 16 --w 16 --cin 128 --cout 128 --n 256
$B --k 3 --h 16 --w 16 --cin 256 --cout 256 --n 256
APVAE_PATCH=0 $B --k 3 --h 16 --w 16 --cin 256 --cout 256 --n 256
$B --k 1 --h 16 --w 16 --cin 192 --cout 192 --n 256
