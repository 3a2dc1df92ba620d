#!/bin/bash
# One gpurun call: the headline bench with its cpu_baseline leg, and the reference arm (the instance-vectorised CPU arm)
# for configs 4, 3, 2 and 5 on the GPU box's host cores.
mkdir -p gpurun_out
lscpu | grep -E "Model name|^CPU\(s\)|Flags" | cut -c1-400 > gpurun_out/cpu.txt; nproc >> gpurun_out/cpu.txt
timeout 600 python bench.py --config 4 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench rc=$?"
for c in 4 3 2 5; do
  timeout 600 python bench.py --impl reference --config $c --steps 5 --warmup 2 > gpurun_out/bench_ref_c$c.json 2> gpurun_out/bench_ref_c$c.err; echo "ref c$c rc=$?"
done
cat gpurun_out/cpu.txt | cut -c1-200
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_c4.json").read().strip().splitlines()[-1])
print("c4 value %.3e e2e %.3e cpu %s"%(d["value"], d["e2e"]["value"], json.dumps(d["cpu_baseline"])[:700]))
for c in (4,3,2,5):
    try:
        r=json.loads(open(f"gpurun_out/bench_ref_c{c}.json").read().strip().splitlines()[-1])
        print(c, "ref value %.3e"%r["value"], r["config"]["same_config"], r["cpu_baseline"]["cores"], "%.3e %.3e"%(r["cpu_baseline"]["value_scalar_float32"], r["cpu_baseline"]["value_scalar_float64"]), r["ms_per_step"])
    except Exception as e: print(c, "failed", e)
PY
tail -3 gpurun_out/bench_ref_c4.err
