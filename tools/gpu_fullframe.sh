# hotpath.FullFrame tests + the bench line with its extras.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fullframe_gpu.py -x -q 2>&1 | tail -30
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_n.json 2>gpurun_out/bench_n.err; echo rc=$?; python -c "
import json; d=json.load(open('gpurun_out/bench_n.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['kernels_per_step']); print(d['update_iteration']); print(d['image_frame']); print(d['full_frame'])"; tail -3 gpurun_out/bench_n.err
