#!/bin/bash
# config 5 at size (1,000 / 1,500 keyframes; the reference extension beside it at 1,000), launch list of one global BA
# and ncu --set full of its new kernels (tcgen05 bulk update, panel, Schur product)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python tools/time_global_ba.py --frames 1000 --ref 2>gpurun_out/gba1000.err | tee gpurun_out/r02_global_ba_1000.json | cut -c1-700
timeout 600 python tools/time_global_ba.py --frames 1500 2>gpurun_out/gba1500.err | tee gpurun_out/r02_global_ba_1500.json | cut -c1-500
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r02_ncu_launches_global_ba.csv \
  python tools/time_global_ba.py --frames 1000 --iters 1 > gpurun_out/ncu_gba_launches.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"chol_syrk128_umma_kernel|chol_panel_kernel|eff_schur_kernel" --launch-skip 20 --launch-count 5 \
  -f -o gpurun_out/r02_gba_kernels python tools/time_global_ba.py --frames 1000 --iters 1 > gpurun_out/ncu_gba_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out/*.ncu-rep gpurun_out/r02_ncu_launches_global_ba.csv
