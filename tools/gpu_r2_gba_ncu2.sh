#!/bin/bash
# ncu --set full of the global BA's assembly and Schur kernels on the 1,000-keyframe problem
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"ba_assemble_kernel|eff_schur_kernel" --launch-skip 2 --launch-count 2 \
  -f -o gpurun_out/r02_gba_asm_schur python tools/time_global_ba.py --frames 1000 --iters 1 > gpurun_out/ncu_gba_asm.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/r02_gba_asm_schur.ncu-rep
