#!/bin/bash
# full GPU suite + smoke + the default bench line (what the driver runs at round end)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/ba_accuracy.jsonl gpurun_out/refdpvo_parity.jsonl gpurun_out/reftrain_parity.jsonl
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/r2_full_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 8 gpurun_out/r2_full_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_full_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/r2_full_smoke.log
timeout 1500 python bench.py > gpurun_out/r2_full_bench.json 2> gpurun_out/r2_full_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_full_bench.json'))
print({k:d.get(k) for k in ("value","ms_per_step","steps","gpu_launches")}); print("e2e", d["e2e"]["value"]); print("sustained", d["sustained"]["value"], d["sustained"]["clocks"])
print("roof", {k:d["roofline"].get(k) for k in ("kernel","frac","dram_frac","ms_per_launch")})
print("refext", {k:d["reference_cuda_ext"].get(k) for k in ("ms_per_update_reference","ms_per_update_ours","ratio")})
s=d.get("sequence",{}); print("seq", {k:v for k,v in s.items() if k not in ("arms","what")}); print("cpu", d.get("cpu_baseline",{}).get("value"))
PY
