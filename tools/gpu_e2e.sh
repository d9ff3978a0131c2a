for v in "" "DPVO_BENCH_STREAM=1" "DPVO_BENCH_NO_OVERLAP=1"; do
echo "--- $v"; env $v timeout 600 python bench.py --no-cpu-baseline --no-update-iteration 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('value',round(d['value']), 'e2e', round(d['e2e']['value']), d['launch_mode'])"
done
timeout 600 python tools/time_e2e.py 2>&1 | tail -2
