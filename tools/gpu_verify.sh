# Full check on one B200: GPU test suite, smoke(), both bench arms, ncu launch list of the full frame.
# gpurun --timeout 3000 -- "bash tools/gpu_verify.sh"
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/bench_o.json 2>gpurun_out/bench_o.err; echo rc=$?; tail -3 gpurun_out/bench_o.err
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_o_ref.json 2>gpurun_out/bench_o_ref.err; echo rc=$?; cat gpurun_out/bench_o_ref.json | cut -c1-600
timeout 300 python tools/run_fullframe_once.py 3 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_fullframe.csv python tools/run_fullframe_once.py 3 > gpurun_out/ncu_ff.log 2>&1; echo ncu rc=$?
