mkdir -p gpurun_out
timeout 300 python tools/run_encoder_once.py 2>&1 | tail -2
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/enc_launches.csv python tools/run_encoder_once.py > gpurun_out/enc_ncu.log 2>&1; echo rc=$?
