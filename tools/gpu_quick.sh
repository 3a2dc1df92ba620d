#!/bin/bash
# One short gpurun call: the GPU suite (optionally a -k filter as $1) and smoke.
mkdir -p gpurun_out
if [ -n "$1" ]; then timeout 1500 python -m pytest tests -m gpu -q -k "$1" > gpurun_out/pytest_gpu.log 2>&1; else timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; fi
echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -30 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log
