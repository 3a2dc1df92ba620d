#!/bin/bash
# One gpurun call: GPU parity tests, smoke, bench, parity report, ncu launch list and full captures of the top kernels.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
nproc >> gpurun_out/gpu.txt; lscpu | grep "Model name" >> gpurun_out/gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 600 python tools/bench_configs.py > gpurun_out/configs.log 2>&1
timeout 600 python tools/parity_report.py > gpurun_out/parity_sfu.json 2> gpurun_out/parity_sfu.err
HGF_MATH=accurate timeout 600 python tools/parity_report.py > gpurun_out/parity_accurate.json 2> gpurun_out/parity_accurate.err
if [ "$1" != "noncu" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:scan_kernel -s 3 -c 1 -o gpurun_out/c4_logp_grad -f python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/ncu_full_c4.log 2>&1
REPS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:scan_kernel -s 1 -c 1 -o gpurun_out/c3_fwd -f python tools/bench_configs.py c3 > gpurun_out/ncu_full_c3.log 2>&1
REPS=1 C5_N=32768 timeout 900 ncu --set full --clock-control none --import-source on -k regex:scan_kernel -s 4 -c 1 -o gpurun_out/c5_traj -f python tools/bench_configs.py c5 > gpurun_out/ncu_full_c5.log 2>&1
fi
ls -la gpurun_out
