#!/bin/bash
timeout 600 python -m pytest tests/test_model_gpu.py -x -q -m gpu -k "networks or is_nll or precise" 2>&1 | tail -2
timeout 300 python bench.py --steps 4 --no-cpu-baseline > gpurun_out/bench_heads3.log 2>&1
tail -1 gpurun_out/bench_heads3.log | python -c "
import json,sys
j=json.loads(sys.stdin.read()); print(j['value']); [print(o) for o in j['ops'] if 'hbm_gbs' in o]"
