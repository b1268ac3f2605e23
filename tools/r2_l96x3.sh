#!/bin/bash
# What bounds G_theta/level_8/conv_1 (96->96 3x3 transposed @32x32 + half-resolution residual; 41.5 % tensor-pipe activity)?
# Probe runs (results are garbage, timing only): remove one actor at a time; then clock-stamp traces of CTA 0.
L="--transpose --k 3 --h 32 --w 32 --cin 96 --cout 96 --n 2368 --reps 5"
for R in even none; do
  echo "== res=$R base";        python tools/bench_layer.py $L --res $R 2>&1 | tail -1
  echo "== res=$R no loads";    python tools/bench_layer.py $L --res $R --tuning debug=1 2>&1 | tail -1
  echo "== res=$R no epilogue work"; python tools/bench_layer.py $L --res $R --tuning debug=4 2>&1 | tail -1
  echo "== res=$R no stores";   python tools/bench_layer.py $L --res $R --tuning debug=5 2>&1 | tail -1
  echo "== res=$R no MMAs";     python tools/bench_layer.py $L --res $R --tuning debug=6 2>&1 | tail -1
done
echo "== raw only"; python tools/bench_layer.py $L --res even --outs raw 2>&1 | tail -1
echo "== macc 1";   python tools/bench_layer.py $L --res even --tuning patch_macc=1 2>&1 | tail -1
echo "== single CTA"; python tools/bench_layer.py $L --res even --tuning patch_2cta=-1 2>&1 | tail -1
echo "== 5x5 reference shape"; python tools/bench_layer.py --transpose --k 5 --h 32 --w 32 --cin 96 --cout 96 --n 2368 --reps 5 --res same 2>&1 | tail -1
echo "== trace"; python tools/bench_layer.py $L --res even --tuning trace=1 2>&1 | head -30
