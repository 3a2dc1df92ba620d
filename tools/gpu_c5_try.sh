#!/bin/bash
# One gpurun call: config 5 bench, then the GPU suite and smoke.
mkdir -p gpurun_out
timeout 600 python bench.py --config 5 --no-cpu > gpurun_out/c5_try.json 2> gpurun_out/c5_try.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/c5_try.json").read().strip().splitlines()[-1])
print("c5", d["config"].get("kernel"), "ms_per_eval", d["ms_per_eval"], "value", d["value"], "roof", d["roofline"]["frac"])
PY
bash tools/gpu_quick.sh
