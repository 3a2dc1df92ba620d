#!/bin/bash
# Refresh of the math-policy ablation with the shipped kernels (packed tangents, ConstAttributes): the round-1 policy and its
# one-knob variants, the shipped policy and its one-knob variants, the validation build.  Objects are pre-built on the CPU box.
mkdir -p gpurun_out
export HGF_ABLATION_BOX=32768
export HGF_ABLATION_EXTRA='{"resp0": ["-DHGF_RESP_MODE=0"], "resp0_div2": ["-DHGF_RESP_MODE=0", "-DHGF_DIV_MODE=2"], "resp0_exp2": ["-DHGF_RESP_MODE=0", "-DHGF_EXP_MODE=2"], "resp0_log1": ["-DHGF_RESP_MODE=0", "-DHGF_LOG_MODE=1"], "resp0_nofmad": ["-DHGF_RESP_MODE=0", "-fmad=false"], "resp0_pow2": ["-DHGF_RESP_MODE=0", "-DHGF_POW_MODE=2"]}'
timeout 1500 python tools/parity_ablation.py sfu resp0 resp0_div2 resp0_exp2 resp0_log1 resp0_nofmad resp0_pow2 div1 exp2 log1 nofmad div1_exp2_log1_nofmad accurate > gpurun_out/parity_ablation.jsonl 2> gpurun_out/parity_ablation.err; echo "rc=$?"
python - <<'PY'
import json
for l in open("gpurun_out/parity_ablation.jsonl"):
    if not l.startswith("{"): continue
    d=json.loads(l)
    s=d["smoke"]; c=d["config4"]
    print(f'{d["variant"]:26s} box32768: grad-kernel >1e-4 {s["logp"]["n_gt_1e-4"]:4d}  fwd-kernel {s["logp_fwd_kernel"]["n_gt_1e-4"]:4d}  max {s["logp"]["max"]:.1e} | config4: >1e-4 {c["logp"]["n_gt_1e-4"]:4d}  logp+grad {c["logp_grad_ms"]["min"]:.3f} ms  fwd {c["forward_ms"]["min"]:.3f} ms')
PY
tail -3 gpurun_out/parity_ablation.err
