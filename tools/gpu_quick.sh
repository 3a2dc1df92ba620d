#!/bin/bash
# quick GPU loop: parity tests, config benchmarks with and without TMA staging, headline bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python tools/bench_configs.py > gpurun_out/configs_tma.log 2>&1
HGF_NO_TMA=1 timeout 300 python tools/bench_configs.py > gpurun_out/configs_notma.log 2>&1
timeout 300 python bench.py --no-cpu > gpurun_out/bench_tma.json 2> gpurun_out/bench_tma.err
HGF_NO_TMA=1 timeout 300 python bench.py --no-cpu > gpurun_out/bench_notma.json 2> gpurun_out/bench_notma.err
