#!/bin/bash
# One short gpurun call: shared-reciprocal bit check, then the GPU suite and smoke.
mkdir -p gpurun_out
timeout 600 python tools/shared_rcp_check.py > gpurun_out/shared_rcp_check.log 2>&1; echo "rc=$?" >> gpurun_out/shared_rcp_check.log
tail -3 gpurun_out/shared_rcp_check.log
bash tools/gpu_quick.sh
