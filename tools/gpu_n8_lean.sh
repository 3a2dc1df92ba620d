#!/bin/bash
# Multi-GPU call, lean: bench config 4 (in-kernel all-reduce), 3, 5, 2 at N GPUs, default steps.
N=${1:-8}
mkdir -p gpurun_out
run() { name=$1; shift; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29520 + RANDOM % 100)) bench.py --gpus $N "$@" > gpurun_out/bench_${name}_n$N.json 2> gpurun_out/bench_${name}_n$N.err; echo "rc=$?" >> gpurun_out/bench_${name}_n$N.err; }
run c4_fused --no-cpu
run c3 --config 3 --no-cpu
run c5 --config 5 --no-cpu
run c2 --config 2 --no-cpu
for f in c4_fused c3 c5 c2; do tail -1 gpurun_out/bench_${f}_n$N.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_${f}_n$N.json").read().strip().splitlines()[-1])
    print("$f", {k:d[k] for k in ("value","ms_per_step","ms_per_eval","kernel_ms","gpu_launches")}, "e2e", d["e2e"]["value"], d["clocks"])
except Exception as e: print("$f parse failed", e)
PY
done
