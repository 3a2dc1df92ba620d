#!/bin/bash
# Config 4 at N GPUs with the shipped library: in-kernel all-reduce, then ncclAllReduce.
N=${1:-8}
mkdir -p gpurun_out
run() { name=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29520 + RANDOM % 100)) bench.py --gpus $N --steps 20 --warmup 5 "$@" > gpurun_out/bench_${name}_n$N.json 2> gpurun_out/bench_${name}_n$N.err; echo "rc=$?" >> gpurun_out/bench_${name}_n$N.err; }
run c4_fused
HGF_BENCH_COLLECTIVE=nccl run c4_nccl
for f in c4_fused c4_nccl; do tail -1 gpurun_out/bench_${f}_n$N.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_${f}_n$N.json").read().strip().splitlines()[-1])
    print("$f", {k:d[k] for k in ("value","ms_per_step","ms_per_eval","kernel_ms","gpu_launches")}, [round(x,3) for x in d["per_rank_ms_per_step"]["all"]], "e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e"].get("matches_device_path"), d["config"].get("collective_max_rel_err_vs_nccl"))
except Exception as e: print("$f parse failed", e)
PY
done
