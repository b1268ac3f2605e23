#!/bin/bash
# BASELINE configs[4]: z_dim sweep on CIFAR-shape data (arch B = dmm_for_sampling.py; arch A needs d % 64 == 0):
# NLL evaluation (K = 1000) and energy-gradient (MALA) sampling throughput.  N = number of GPUs (default 1).
N=${1:-1}
OUT=${2:-gpurun_out/sweep_zdim_${N}gpu.log}
: > $OUT
RUN="python"
[ "$N" -gt 1 ] && RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
for d in 32 64 128 256 512; do
  timeout 300 $RUN bench.py --gpus $N --arch B --dataset cifar10 --z-dim $d --steps 3 --warmup 3 --no-cpu-baseline --no-op-times 2>/dev/null | grep '^{' >> $OUT
  timeout 300 $RUN tools/bench_mala.py --arch B --dataset cifar10 --z-dim $d --chains 1024 --steps 20 2>/dev/null | grep '^{' >> $OUT
done
python - $OUT <<'PY'
import json, sys
for l in open(sys.argv[1]):
    j = json.loads(l)
    print('%-90s %12.1f %s  (n_gpus %s)' % (j['metric'][:90], j['value'], j['unit'], j.get('n_gpus')))
PY
