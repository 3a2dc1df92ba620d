#!/bin/bash
# Multi-GPU call (gpurun --gpus N): the fused all-reduce check, the N-GPU tests, bench at N GPUs (fused and NCCL).
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/fused_reduce_check.py > gpurun_out/fused_check_n$N.log 2>&1; echo "rc=$?" >> gpurun_out/fused_check_n$N.log
timeout 600 python -m pytest tests -m gpu -q -k "fused or two_gpus" > gpurun_out/pytest_multi_n$N.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_multi_n$N.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_c4_n${N}_fused.json 2> gpurun_out/bench_c4_n${N}_fused.err; echo "rc=$?" >> gpurun_out/bench_c4_n${N}_fused.err
HGF_BENCH_COLLECTIVE=nccl timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_c4_n${N}_nccl.json 2> gpurun_out/bench_c4_n${N}_nccl.err; echo "rc=$?" >> gpurun_out/bench_c4_n${N}_nccl.err
tail -3 gpurun_out/fused_check_n$N.log; tail -3 gpurun_out/pytest_multi_n$N.log
for f in fused nccl; do tail -1 gpurun_out/bench_c4_n${N}_$f.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_c4_n${N}_$f.json").read().strip().splitlines()[-1])
    print("$f", {k:d[k] for k in ("value","ms_per_step","ms_per_eval","kernel_ms","gpu_launches")}, d["per_rank_ms_per_step"], "e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e"]["matches_device_path"], d["config"]["collective_max_rel_err_vs_nccl"])
except Exception as e: print("$f parse failed", e)
PY
done
