#!/bin/bash
# One gpurun call, ONE ncu invocation: --set full capture (with source) of the config-4 logp+grad kernel.
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 1 --no-cpu --evals-per-step 1 --config 4"
$B > gpurun_out/ncu_plain_c4.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:scan_kernel -s 2 -c 1 -f -o gpurun_out/c4_logp_grad $B > gpurun_out/ncu_full_c4.log 2>&1
ls -la gpurun_out | tail -5
