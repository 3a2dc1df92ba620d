#!/bin/bash
# Final multi-GPU pass (8 GPUs): the tests that need >= 2 GPUs, config 4 (fused all-reduce, then NCCL), config 5, and the
# reference arm launched the way the driver launches it (torchrun: rank 0 alone works and prints).
N=${1:-8}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -k "two_gpus or fused" > gpurun_out/pytest_multi.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_multi.log
tail -4 gpurun_out/pytest_multi.log
run() { name=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29520 + RANDOM % 100)) bench.py --gpus $N --steps 20 --warmup 5 "$@" > gpurun_out/bench_${name}_n$N.json 2> gpurun_out/bench_${name}_n$N.err; echo "rc=$?" >> gpurun_out/bench_${name}_n$N.err; }
run c4_fused
HGF_BENCH_COLLECTIVE=nccl run c4_nccl
run c5 --config 5
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29677 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err; echo "ref rc=$?"
for f in c4_fused c4_nccl c5; do tail -1 gpurun_out/bench_${f}_n$N.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_${f}_n$N.json").read().strip().splitlines()[-1])
    print("$f", {k:d[k] for k in ("value","ms_per_step","ms_per_eval","kernel_ms","gpu_launches")}, [round(x,3) for x in d["per_rank_ms_per_step"]["all"]], "e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e"].get("matches_device_path"), d["config"].get("collective_max_rel_err_vs_nccl"), d["clocks"])
except Exception as e: print("$f parse failed", e)
PY
done
wc -l gpurun_out/bench_ref_n$N.json; cut -c1-400 gpurun_out/bench_ref_n$N.json
