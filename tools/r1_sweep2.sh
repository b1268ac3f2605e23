#!/bin/bash
# z_dim sweep extremes on 2 GPUs (BASELINE configs[4] is quoted at 8; the path shards with no data-path collective)
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
OUT=gpurun_out/r1_sweep_zdim_2gpu.jsonl
: > $OUT
for d in 32 512; do
  timeout 200 $RUN bench.py --gpus 2 --arch B --dataset cifar10 --z-dim $d --steps 3 --warmup 3 --no-cpu-baseline --no-op-times 2>/dev/null | grep '^{' >> $OUT
  timeout 200 $RUN tools/bench_mala.py --arch B --dataset cifar10 --z-dim $d --chains 1024 --steps 20 2>/dev/null | grep '^{' >> $OUT
done
python - $OUT <<'PY'
import json, sys
for l in open(sys.argv[1]):
    j = json.loads(l)
    print('%-90s %12.1f %s  (n_gpus %s)' % (j['metric'][:90], j['value'], j['unit'], j.get('n_gpus')))
PY
