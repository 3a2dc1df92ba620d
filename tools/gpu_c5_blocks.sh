#!/bin/bash
# One gpurun call: config 5 with the component kernel under other register caps (tools/bench_c5_blocks.py), and the stock bench.
mkdir -p gpurun_out
timeout 600 python bench.py --config 5 --no-cpu > gpurun_out/c5_stock.json 2> gpurun_out/c5_stock.err
for b in "$@"; do timeout 600 python tools/bench_c5_blocks.py $b > gpurun_out/c5_blocks_$b.json 2> gpurun_out/c5_blocks_$b.err; done
python - "$@" <<'PY'
import json, sys
for tag in ["stock"] + [f"blocks_{b}" for b in sys.argv[1:]]:
    try:
        d = json.loads(open(f"gpurun_out/c5_{tag}.json").read().strip().splitlines()[-1])
        print(tag, d["config"].get("kernel"), "ms_per_eval", d["ms_per_eval"], "value", d["value"])
    except Exception as e:
        print(tag, "failed", e)
PY
