#!/bin/bash
# One gpurun call: the whole GPU suite, the no-allocation entry point's test alone in a fresh process, smoke, the
# forward-mode / reverse-sweep crossover and the headline bench.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_noalloc_alone.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python tools/bench_grad_scaling.py > gpurun_out/grad_scaling.log 2> gpurun_out/grad_scaling.err; echo "grad scaling rc=$?"
timeout 600 python bench.py --config 4 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench rc=$?"
tail -6 gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_noalloc_alone.log; tail -2 gpurun_out/smoke.log
cut -c1-400 gpurun_out/grad_scaling.log; tail -3 gpurun_out/grad_scaling.err
cut -c1-600 gpurun_out/bench_c4.json
