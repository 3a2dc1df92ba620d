#!/bin/bash
mkdir -p gpurun_out
timeout 300 ./tools/ubench/rcp_scaling > gpurun_out/rcp_scaling.log 2>&1; echo "rcp rc=$?"; cat gpurun_out/rcp_scaling.log
