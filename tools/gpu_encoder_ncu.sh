mkdir -p gpurun_out
timeout 300 python tools/run_encoder_once.py 2>&1 | tail -1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_mma_kernel --launch-skip 20 -c 10 -o gpurun_out/enc_conv_full -f python tools/run_encoder_once.py > gpurun_out/enc_ncu_full.log 2>&1; echo rc=$?
ls -la gpurun_out/enc_conv_full.ncu-rep
