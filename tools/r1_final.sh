#!/bin/bash
# Round-1 final measurements on one B200 (bench defaults): GPU tests, bench line, reference arm, ncu launch list of the
# bench command, sampler benches, z_dim sweep.  (tools/r1_profile.sh makes the per-launch metric window and the --set full
# capture of the dominant op.)
O=gpurun_out
timeout 1500 python -m pytest tests -q -m gpu > $O/final_gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -2 $O/final_gpu_tests.log
timeout 900 python bench.py > $O/r1_bench_default.log 2>&1; echo "bench rc=$?"; tail -1 $O/r1_bench_default.log | cut -c1-260
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/r1_bench_reference.log 2>&1; echo "ref rc=$?"; tail -1 $O/r1_bench_reference.log | cut -c1-200
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-op-times"
timeout 600 $CMD > $O/prof_plain.log 2>&1; tail -1 $O/prof_plain.log | cut -c60-130
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file $O/r1_launch_list.csv $CMD > $O/prof_ncu_list.log 2>&1
echo "launch list rc=$? lines=$(wc -l < $O/r1_launch_list.csv)"
timeout 600 python tools/bench_mala.py --op-times > $O/mala_bf16.log 2>&1; tail -1 $O/mala_bf16.log | cut -c1-300
timeout 600 python tools/bench_mala.py --precise > $O/mala_precise.log 2>&1; tail -1 $O/mala_precise.log | cut -c1-300
bash tools/sweep_zdim.sh 1 $O/r1_sweep_zdim_1gpu.jsonl
