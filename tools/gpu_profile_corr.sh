# ncu --set full of the correlation lookup + launch list of the default bench command (stream launches).
mkdir -p gpurun_out
timeout 300 python tools/time_corr.py --iters 5 > gpurun_out/time_corr_p.log 2>&1; echo rc=$?; tail -1 gpurun_out/time_corr_p.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:corr_fwd -c 2 -f -o gpurun_out/corr_full_r01b python tools/time_corr.py --iters 1 > gpurun_out/ncu_corr_full.log 2>&1; echo ncu rc=$?
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-update-iteration --no-graph > gpurun_out/b_ng.log 2>&1; echo rc=$?
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_bench_r01b.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-update-iteration --no-graph > gpurun_out/ncu_final.log 2>&1; echo rc=$?
ls -la gpurun_out/*.ncu-rep
