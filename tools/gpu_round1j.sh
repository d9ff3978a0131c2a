mkdir -p gpurun_out
for v in "" "DPVO_BENCH_STREAM=1" "DPVO_BENCH_NO_OVERLAP=1"; do
echo "--- $v"; env $v timeout 600 python bench.py --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['ms_per_step_stream_launch'])"
done
