#!/bin/bash
# One gpurun call: GPU parity suite, smoke, bench of every BASELINE configuration (own arm) + the reference arm.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
nproc >> gpurun_out/gpu.txt; lscpu | grep "Model name" >> gpurun_out/gpu.txt
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
for c in 4 2 3 5; do
  timeout 900 python bench.py --config $c > gpurun_out/bench_c$c.json 2> gpurun_out/bench_c$c.err; echo "bench c$c rc=$?" >> gpurun_out/bench_c$c.err
done
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
tail -5 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log
for c in 4 2 3 5; do tail -2 gpurun_out/bench_c$c.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_c$c.json").read().strip().splitlines()[-1])
    print("c$c", {k:d[k] for k in ("value","ms_per_step","ms_per_eval","kernel_ms","gpu_launches")}, "roof", d["roofline"]["bound"], round(d["roofline"]["frac"],4), "e2e", d["e2e"]["value"], d["e2e"]["matches_device_path"], "cpu", d.get("cpu_baseline",{}).get("value"))
except Exception as e: print("c$c parse failed", e)
PY
done
