#!/bin/bash
# One gpurun call, ONE ncu invocation: pipe / issue / DRAM counters of the config-3 forward kernel.
mkdir -p gpurun_out
M=gpu__time_duration.sum,sm__inst_executed_pipe_xu.sum,sm__pipe_xu_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fma.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size,launch__block_size,dram__bytes_read.sum,dram__bytes_write.sum,sm__cycles_elapsed.avg,sm__cycles_active.avg
B="python bench.py --steps 2 --warmup 1 --no-cpu --evals-per-step 1 --config 3"
$B > gpurun_out/ncu_plain_c3.log 2>&1 && ncu --metrics $M --clock-control none -k regex:scan_kernel -s 2 -c 1 --csv --log-file gpurun_out/counters_c3.csv $B > gpurun_out/ncu_counters_c3.log 2>&1
grep -E "inst_executed|issue_active|time_duration|cycles_active.avg" gpurun_out/counters_c3.csv | awk -F'","' '{print $(NF-2), $NF}' | tr -d '"'
