#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -k "reverse or gradient" > gpurun_out/pytest_rev.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_rev.log
tail -25 gpurun_out/pytest_rev.log
timeout 900 python tools/bench_grad_scaling.py > gpurun_out/grad_scaling.log 2> gpurun_out/grad_scaling.err; echo "rc=$?"
cat gpurun_out/grad_scaling.log | cut -c1-330; tail -5 gpurun_out/grad_scaling.err
