#!/bin/bash
# One gpurun call, ONE ncu invocation: --set full capture (with source) of config $1's kernel ($2 = launches to skip, default 2).
mkdir -p gpurun_out
C=${1:-5}; S=${2:-2}
B="python bench.py --steps 2 --warmup 1 --no-cpu --evals-per-step 1 --config $C"
$B > gpurun_out/ncu_plain_c$C.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:scan_kernel -s $S -c 1 -f -o gpurun_out/c${C}_full $B > gpurun_out/ncu_full_c$C.log 2>&1
echo "rc=$?"; ls -la gpurun_out | tail -4
