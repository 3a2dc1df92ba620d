#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python tools/bench_grad_scaling.py > gpurun_out/grad_scaling.log 2> gpurun_out/grad_scaling.err; echo "rc=$?"
python - <<'PY'
import json
for l in open("gpurun_out/grad_scaling.log"):
    if l.startswith("{"):
        d=json.loads(l); print(d["topology"], d["P"], round(d["forward_ms"],3), round(d["reverse_ms"],3))
    else: print(l.strip()[:150])
PY
tail -3 gpurun_out/grad_scaling.err
