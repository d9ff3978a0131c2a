#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_ba_gpu.py tests/test_lietorch_gpu.py tests/test_viewer_gpu.py -x -q > gpurun_out/r2c6_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 15 gpurun_out/r2c6_pytest.log
timeout 600 python tools/time_global_ba.py --frames 1000 --ref > gpurun_out/r2c6_global_ba_1000.json 2> gpurun_out/r2c6_global_ba.err; cat gpurun_out/r2c6_global_ba_1000.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:corr_bwd_fast -c 2 -o gpurun_out/r02_corr_bwd_fast python tools/time_corr_bwd.py > gpurun_out/r2c6_ncu_bwd.log 2>&1; echo "ncu bwd rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"ba_assemble_kernel|eff_schur_kernel" -c 4 -o gpurun_out/r02_ba_global python tools/time_global_ba.py --frames 1000 --iters 1 > gpurun_out/r2c6_ncu_ba.log 2>&1; echo "ncu ba rc=$?"
ls -la gpurun_out/*.ncu-rep
