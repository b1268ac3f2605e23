#!/bin/bash
# Round-1 final measurements on one B200 (bench defaults): GPU tests, bench line, reference arm, --set full capture of the
# dominant op (G_theta/level_9/conv_1 = launch 781 of `bench.py --steps 1 --warmup 3`), sampler benches, z_dim sweep.
O=gpurun_out
timeout 1500 python -m pytest tests -q -m gpu > $O/final_gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -2 $O/final_gpu_tests.log
timeout 900 python bench.py > $O/r1_bench_default.log 2>&1; echo "bench rc=$?"; tail -1 $O/r1_bench_default.log | cut -c1-260
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/r1_bench_reference.log 2>&1; echo "ref rc=$?"; tail -1 $O/r1_bench_reference.log | cut -c1-200
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-op-times"
timeout 600 ncu --set full --clock-control none --import-source on -s 781 -c 1 -f -o /tmp/r1_dom $CMD > $O/prof_ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i /tmp/r1_dom.ncu-rep --page raw --csv > $O/r1_dom_conv1_raw.csv 2>/dev/null
ncu -i /tmp/r1_dom.ncu-rep --page details --csv > $O/r1_dom_conv1_details.csv 2>/dev/null
ncu -i /tmp/r1_dom.ncu-rep --page source --csv > $O/r1_dom_conv1_source.csv 2>/dev/null
ls -la /tmp/r1_dom.ncu-rep; [ $(stat -c %s /tmp/r1_dom.ncu-rep) -lt 30000000 ] && cp /tmp/r1_dom.ncu-rep $O/r1_dom_conv1.ncu-rep
timeout 600 python tools/bench_mala.py > $O/mala_bf16.log 2>&1; tail -1 $O/mala_bf16.log | cut -c1-300
timeout 600 python tools/bench_mala.py --precise > $O/mala_precise.log 2>&1; tail -1 $O/mala_precise.log | cut -c1-300
bash tools/sweep_zdim.sh 1 $O/r1_sweep_zdim_1gpu.jsonl
