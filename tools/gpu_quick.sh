#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python tools/bench_specialise.py > gpurun_out/specialise.log 2>&1
timeout 300 python bench.py --no-cpu > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err
timeout 300 python tools/bench_ef.py > gpurun_out/bench_ef.log 2>&1
