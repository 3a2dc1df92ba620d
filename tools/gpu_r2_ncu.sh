#!/bin/bash
# One gpurun call (one GPU): the ncu evidence of the round -- per-launch times of the stock bench, the pipe / DRAM counters
# north_star asks for (configs 3, 4, 5) and `--set full` captures of the config-4 and config-5 kernels with source.
# Every profiled command line first exits 0 without ncu (B200_PROFILING.md).
mkdir -p gpurun_out
M=gpu__time_duration.sum,sm__inst_executed_pipe_xu.sum,sm__pipe_xu_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fma.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size,launch__block_size,dram__bytes_read.sum,dram__bytes_write.sum,sm__cycles_elapsed.avg,sm__cycles_active.avg
B="python bench.py --steps 2 --warmup 1 --no-cpu --evals-per-step 1"
$B --config 4 > gpurun_out/ncu_plain_c4.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c4.csv $B --config 4 > gpurun_out/ncu_launches_c4.log 2>&1
for c in 4 3; do
  $B --config $c > gpurun_out/ncu_plain_c$c.log 2>&1 && ncu --metrics $M --clock-control none -k regex:scan_kernel -s 2 -c 1 --csv --log-file gpurun_out/counters_c$c.csv $B --config $c > gpurun_out/ncu_counters_c$c.log 2>&1
done
export HGF_BENCH_C5_STEPS=250
$B --config 5 > gpurun_out/ncu_plain_c5.log 2>&1 && ncu --metrics $M --clock-control none -k regex:scan_kernel -s 4 -c 2 --csv --log-file gpurun_out/counters_c5.csv $B --config 5 > gpurun_out/ncu_counters_c5.log 2>&1
$B --config 5 > gpurun_out/ncu_plain_c5.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:scan_kernel -s 4 -c 1 -f -o gpurun_out/c5_traj $B --config 5 > gpurun_out/ncu_full_c5.log 2>&1
unset HGF_BENCH_C5_STEPS
$B --config 4 > gpurun_out/ncu_plain_c4.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:scan_kernel -s 2 -c 1 -f -o gpurun_out/c4_logp_grad $B --config 4 > gpurun_out/ncu_full_c4.log 2>&1
ls -la gpurun_out | tail -20
