#!/bin/bash
# the whole GPU suite and the parity report under the validation build (IEEE division, libdevice, no FMA contraction)
mkdir -p gpurun_out
HGF_MATH=accurate timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_accurate.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_accurate.log
HGF_MATH=accurate timeout 600 python tools/parity_report.py > gpurun_out/parity_accurate.json 2> gpurun_out/parity_accurate.err
timeout 600 python tools/parity_report.py > gpurun_out/parity_sfu.json 2> gpurun_out/parity_sfu.err
HGF_MATH=accurate timeout 300 python bench.py --no-cpu > gpurun_out/bench_accurate.json 2> gpurun_out/bench_accurate.err
HGF_MATH=accurate timeout 300 python tools/bench_configs.py c3 > gpurun_out/configs_accurate.log 2>&1
