#!/bin/bash
# quick GPU loop: parity tests (+ TMA tests first), config benchmarks with and without TMA write-back
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k tma > gpurun_out/pytest_tma.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_tma.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python tools/bench_configs.py c5 > gpurun_out/configs_c5_tma.log 2>&1
HGF_NO_TMA=1 timeout 300 python tools/bench_configs.py c5 > gpurun_out/configs_c5_notma.log 2>&1
