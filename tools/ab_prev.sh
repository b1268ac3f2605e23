#!/bin/bash
P=adversarial-prior-vae_b200
cp $P/libapvae_b200.so /tmp/new.so
CASES=patch192,patch96x5,patch96x3_even,nar_D30_a4,nar_D20_a4,nar_D10_a4
for r in 1 2; do
  cp $P/libapvae_prev.so $P/libapvae_b200.so; echo "== prev run $r"; python tools/layer_suite.py --only $CASES 2>&1 | awk '{print $1, $(NF-3)}'
  cp /tmp/new.so $P/libapvae_b200.so; echo "== new run $r"; python tools/layer_suite.py --only $CASES 2>&1 | awk '{print $1, $(NF-3)}'
  echo "== new epi1 run $r"; APVAE_PATCH_EPI=1 python tools/layer_suite.py --only $CASES 2>&1 | awk '{print $1, $(NF-3)}'
done
