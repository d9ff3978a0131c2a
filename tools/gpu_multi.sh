# 2 GPUs: training all-reduce timing, bench under torchrun, full-frame tests, 1-GPU bench.  gpurun --gpus 2 -- "bash tools/gpu_multi.sh"
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 tools/time_allreduce.py 2>gpurun_out/allreduce.err | tee gpurun_out/allreduce_2gpu.json; tail -2 gpurun_out/allreduce.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus 2 --steps 50 --warmup 5 > gpurun_out/bench_2gpu.json 2>gpurun_out/bench_2gpu.err; echo rc=$?; wc -l gpurun_out/bench_2gpu.json; cut -c1-160 gpurun_out/bench_2gpu.json
timeout 900 python -m pytest tests/test_fullframe_gpu.py -x -q 2>&1 | tail -3
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_q.json 2>gpurun_out/bench_q.err; echo rc=$?; python -c "
import json; d=json.load(open('gpurun_out/bench_q.json')); print(d['value'], d['e2e']['value']); print(d['full_frame'])"; tail -3 gpurun_out/bench_q.err
