#!/bin/bash
# One gpurun call: the reverse-sweep / gradient GPU tests (16-node network included), then the gradient-scaling bench.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x -k "reverse or gradient or no_alloc" > gpurun_out/pytest_rev.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_rev.log
tail -25 gpurun_out/pytest_rev.log
timeout 1200 python tools/bench_grad_scaling.py > gpurun_out/grad_scaling.log 2> gpurun_out/grad_scaling.err; echo "rc=$?"
grep -E "^#|forward filter alone|ghgf16" gpurun_out/grad_scaling.log | cut -c1-330; tail -5 gpurun_out/grad_scaling.err
