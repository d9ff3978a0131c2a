mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:linear384 -c 12 -o gpurun_out/prof_update_gemm -f python tools/time_update.py --profile > gpurun_out/ncu_update_gemm.log 2>&1; echo rc=$?; tail -2 gpurun_out/ncu_update_gemm.log
