#!/bin/bash
# ncu --set full of the blocked Cholesky's kernels with WARM caches (--cache-control none: the panel kernel's one-warp
# sections are instruction-fetch bound when every replay starts cold, which is not how they run in the factorisation)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python tools/time_chol.py 1000 2>&1 | cut -c1-1200
timeout 600 ncu --set full --clock-control none --cache-control none --import-source on -k regex:"chol_syrk128_umma_kernel|chol_panel_kernel" --launch-skip 40 --launch-count 4 \
  -f -o gpurun_out/r02_chol_warm python tools/time_chol.py 5994 > gpurun_out/ncu_chol.log 2>&1
echo "ncu chol rc=$?"; tail -n 3 gpurun_out/ncu_chol.log | cut -c1-300
ls -la gpurun_out/*.ncu-rep
