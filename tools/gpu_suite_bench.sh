#!/bin/bash
# One gpurun call: the GPU suite, smoke, and the bench of the configurations given as arguments (default: 4).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -8 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log
for c in ${@:-4}; do
  timeout 900 python bench.py --config $c --no-cpu > gpurun_out/bench_c$c.json 2> gpurun_out/bench_c$c.err; echo "bench c$c rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_c$c.json").read().strip().splitlines()[-1])
    print("c$c", {k:d[k] for k in ("value","ms_per_step","ms_per_eval","kernel_ms","gpu_launches")}, "roof", d["roofline"]["bound"], round(d["roofline"]["frac"],4), "e2e", d["e2e"]["value"], d["e2e"].get("matches_device_path"))
except Exception as e: print("c$c parse failed", e)
PY
done
