#!/bin/bash
# Round-2 multi-GPU measurements on N GPUs of one box (tools/gpu_retry.sh <t> 'bash tools/scale.sh N' --gpus N):
#   weak scaling  (BASELINE configs[1] per GPU), strong scaling (configs[2]: 10 000 images through Evaluator.run, ragged
#   shards), and with SWEEP=1 the z_dim sweep of configs[4] (NLL + MALA throughput on arch B CIFAR shapes).
N=${1:-1}
O=gpurun_out
mkdir -p $O
RUN="python"
[ "$N" -gt 1 ] && RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541"
timeout 600 $RUN bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --no-op-times --no-mala > $O/r2_bench_${N}gpu.log 2> $O/r2_bench_${N}gpu.err
echo "weak rc=$?"; tail -1 $O/r2_bench_${N}gpu.log | cut -c1-200
timeout 900 $RUN bench.py --gpus $N --images 10000 --warmup 3 --no-cpu-baseline --no-op-times --no-mala > $O/r2_strong_${N}gpu.log 2> $O/r2_strong_${N}gpu.err
echo "strong rc=$?"; tail -1 $O/r2_strong_${N}gpu.log | cut -c1-200
if [ -n "$SWEEP" ]; then bash tools/sweep_zdim.sh $N $O/r2_sweep_zdim_${N}gpu.jsonl; fi
nvidia-smi --query-gpu=index,name,clocks.max.sm --format=csv,noheader | head -8
