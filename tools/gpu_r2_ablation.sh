#!/bin/bash
# One gpurun call: the math-policy ablation (tools/parity_ablation.py); per-instance results dumped for analysis.
mkdir -p gpurun_out
HGF_ABLATION_DUMP=gpurun_out/ablation_dump.npz timeout 900 python tools/parity_ablation.py "$@" > gpurun_out/parity_ablation.jsonl 2> gpurun_out/parity_ablation.err; echo "ablation rc=$?" >> gpurun_out/parity_ablation.err
tail -3 gpurun_out/parity_ablation.err; wc -l gpurun_out/parity_ablation.jsonl; ls -la gpurun_out
