mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_encoder_gpu.py -m gpu -x -q 2>&1 | tail -6
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_img.json 2>gpurun_out/bench_img.err; echo rc=$?
python -c "
import json; d=json.load(open('gpurun_out/bench_img.json')); print(d['value'], d['e2e']['value'], d['image_frame'], d['update_iteration']['ms_per_step'])"; tail -3 gpurun_out/bench_img.err
