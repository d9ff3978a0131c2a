#!/bin/bash
# ncu --set full of the global BA's Schur kernel (late in the 1,000-frame sequence) and of the blocked Cholesky's kernels
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:eff_schur_kernel --launch-skip 620 --launch-count 1 \
  -f -o gpurun_out/r02_eff_schur python tools/trace_global_ba.py --sequence 1000 > gpurun_out/ncu_eff_schur.log 2>&1
echo "ncu eff_schur rc=$?"; tail -n 3 gpurun_out/ncu_eff_schur.log | cut -c1-300
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"chol_syrk128_kernel|chol_panel_kernel" --launch-skip 8 --launch-count 4 \
  -f -o gpurun_out/r02_chol python tools/time_chol.py 5994 > gpurun_out/ncu_chol.log 2>&1
echo "ncu chol rc=$?"; tail -n 3 gpurun_out/ncu_chol.log | cut -c1-300
ls -la gpurun_out/*.ncu-rep
