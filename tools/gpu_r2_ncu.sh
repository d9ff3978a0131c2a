#!/bin/bash
# Round-2 profiling evidence: launch list of the bench command (after it exits 0 without ncu) and a fresh
# `ncu --set full` of the lookup kernel (regenerates profiles/corr_traffic.json).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
FLAGS="--steps 2 --warmup 3 --no-sequence --no-cpu-baseline --no-reference-ext --no-sustained"
timeout 600 python bench.py $FLAGS > gpurun_out/r2_ncu_bench_plain.json 2> gpurun_out/r2_ncu_bench_plain.err; echo "plain rc=$?"
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r02_launches_bench.csv python bench.py $FLAGS > gpurun_out/r2_ncu_bench.log 2>&1; echo "ncu list rc=$?"
timeout 300 python tools/time_corr.py --iters 5 > gpurun_out/r2_ncu_time_corr.jsonl 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:corr_fwd_mma -c 2 -f -o gpurun_out/r02_corr_fwd_full python tools/time_corr.py --iters 1 > gpurun_out/r2_ncu_corr_full.log 2>&1; echo "ncu full rc=$?"
timeout 900 python -m pytest tests/test_refdpvo_gpu.py -x -q -k "b200_dpvo_class" > gpurun_out/r2_ncu_pytest_seq.log 2>&1; echo "pytest seq rc=$?"; tail -n 3 gpurun_out/r2_ncu_pytest_seq.log
timeout 600 python tools/refdpvo.py --backend dropin --pipeline b200 --mode fps --frames 1000 --out gpurun_out/r2_ncu_b200_fps.json > gpurun_out/r2_ncu_b200_fps.log 2>&1; cut -c1-200 gpurun_out/r2_ncu_b200_fps.json
ls -la gpurun_out/r02_corr_fwd_full.ncu-rep gpurun_out/r02_launches_bench.csv
