#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -k "components or wide or config5" > gpurun_out/pytest_n2.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_n2.log
tail -30 gpurun_out/pytest_n2.log
python bench.py --config 5 --no-cpu > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "rc=$?"; tail -3 gpurun_out/bench_c5.err
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_c5.json").read().strip().splitlines()[-1])
print({k:d[k] for k in ("value","ms_per_step","ms_per_eval")}, d["roofline"]["frac"], d["config"]["kernel"])
PY
