#!/bin/bash
# One gpurun call: GPU suite, smoke, config-4 bench, then ONE ncu pass (pipe / issue / DRAM counters of the config-4 kernel).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -4 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py --config 4 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_c4.json").read().strip().splitlines()[-1])
print({k:d[k] for k in ("value","ms_per_eval")}, "roof", round(d["roofline"]["frac"],4), "e2e", d["e2e"]["value"], "cpu", d["cpu_baseline"]["value"])
PY
M=gpu__time_duration.sum,sm__inst_executed_pipe_xu.sum,sm__pipe_xu_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fma.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size,launch__block_size,dram__bytes_read.sum,dram__bytes_write.sum,sm__cycles_elapsed.avg,sm__cycles_active.avg
B="python bench.py --steps 2 --warmup 1 --no-cpu --evals-per-step 1 --config 4"
$B > gpurun_out/ncu_plain_c4.log 2>&1 && ncu --metrics $M --clock-control none -k regex:scan_kernel -s 2 -c 1 --csv --log-file gpurun_out/counters_c4.csv $B > gpurun_out/ncu_counters_c4.log 2>&1
tail -20 gpurun_out/counters_c4.csv | cut -c1-300
