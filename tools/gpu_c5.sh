#!/bin/bash
mkdir -p gpurun_out
python bench.py --config 5 --no-cpu > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_c5.json").read().strip().splitlines()[-1])
print({k:d[k] for k in ("value","ms_per_step","ms_per_eval")}, d["roofline"]["frac"], d["config"]["kernel"])
PY
python tools/bench_configs.py c5 2>&1 | tail -3
HGF_NO_TMA=1 python tools/bench_configs.py c5 2>&1 | tail -2
