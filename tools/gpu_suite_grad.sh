#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -4 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py --config 4 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_c4.json").read().strip().splitlines()[-1])
print({k:d[k] for k in ("value","ms_per_eval")}, "roof", round(d["roofline"]["frac"],4), "e2e", d["e2e"]["value"], "cpu", d["cpu_baseline"]["value"])
PY
bash tools/gpu_grad_scaling.sh
