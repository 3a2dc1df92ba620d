"""Parity evidence for profiles/: distribution of the relative error of the CUDA path (through the C ABI)
against the float32 CPU oracle, split by how rounding-sensitive each instance is (float32 oracle vs float64
oracle of the SAME algorithm).  Run once per math build: `HGF_MATH=accurate python tools/parity_report.py`.
"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
import nets
from oracle import network as onet
from pyhgf_b200 import load_data, response
from pyhgf_b200.network import Network


def q(x):
    x = np.asarray(x)
    return {"n": int(x.size), "max": float(x.max()), "p99": float(np.quantile(x, 0.99)), "p90": float(np.quantile(x, 0.9)),
            "p50": float(np.median(x)), "frac<=1e-4": float((x <= 1e-4).mean()), "frac<=1e-5": float((x <= 1e-5).mean())}


def rel(a, b, floor=0.0):
    return np.abs(a - b) / (np.abs(b) + floor)


out = {"math": os.environ.get("HGF_MATH", "sfu")}
u, y = load_data("binary")
rng = np.random.default_rng(0)
n = 4096
om2, om3, temp = rng.uniform(-6, -3, n), rng.uniform(-6, -3, n), np.exp(rng.uniform(-0.3, 0.8, n))
p = np.stack([om2, om3, temp]).astype(np.float32).astype(np.float64)
keys = [(1, "tonic_volatility"), (2, "tonic_volatility"), ("response_parameter", 0)]
res = nets.binary3(Network).batch(u, dict(zip(keys, p)), response_function=response.binary_softmax_inverse_temperature,
                                  response_function_inputs=y, gradient=True, record=[(1, "mean"), (2, "mean"), (1, "precision"), (2, "precision")])
got = res.surprise.double().cpu().numpy()
o32 = nets.binary3(onet.Network).create_belief_propagation_fn()
o64 = nets.binary3(onet.Network, dtype="float64").create_belief_propagation_fn()
slots = [o32.slot(1, "tonic_volatility"), o32.slot(2, "tonic_volatility"), o32.slot(-1, "response_parameter", 0)]
s32 = o32.batch_surprise(u, slots, p, "binary_softmax_inverse_temperature", resp_inputs=y)
s64 = o64.batch_surprise(u, slots, p, "binary_softmax_inverse_temperature", resp_inputs=y)
_, g32 = o32.batch_grad(u, slots, p, "binary_softmax_inverse_temperature", resp_inputs=y, dtype=32)
_, g64 = o64.batch_grad(u, slots, p, "binary_softmax_inverse_temperature", resp_inputs=y)
sens = rel(s32, s64)
well = sens <= 1e-5
b = {"instances": n, "well_conditioned(|f32/f64-1|<=1e-5 in the oracle)": int(well.sum()),
     "oracle_f32_vs_f64_all": q(sens), "cuda_vs_oracle_f32_well": q(rel(got, s32)[well]), "cuda_vs_oracle_f32_all": q(rel(got, s32)),
     "cuda_vs_oracle_f64_well": q(rel(got, s64)[well])}
for i, k in enumerate(keys):
    gg = res.gradient[k].double().cpu().numpy()
    sc = np.abs(g64[i]).max() * 1e-3
    b[f"grad{i}_cuda_vs_dual_f32_well"] = q(rel(gg, g32[i], sc)[well])
    b[f"grad{i}_cuda_vs_dual_f64_well"] = q(rel(gg, g64[i], sc)[well])
    b[f"grad{i}_dual_f32_vs_f64_well"] = q(rel(g32[i], g64[i], sc)[well])
out["binary3_logp_grad_box[-6,-3]"] = b

# trajectories of the single reference network (per-trial means / precisions)
for name, build, data in [("binary3", nets.binary3, u), ("continuous3", nets.continuous3, load_data("continuous"))]:
    g = build(Network).input_data(data)
    o = build(onet.Network).input_data(data)
    tr = {}
    for node in range(1, len(g.edges)):
        for k in nets.BELIEFS:
            a = np.asarray(g.node_trajectories[node][k], dtype=np.float64)
            c = np.asarray(o.node_trajectories[node][k], dtype=np.float64)
            tr[f"{node}.{k}"] = float(np.nanmax(rel(a, c, 1e-7)))
    out[f"{name}_trajectory_max_rel"] = tr

x = load_data("continuous")
n = 4096
om = [np.clip(rng.normal(m, 1.0, n), m - 3, m + 3) for m in (-13.0, -2.0, -2.0)]
pc = np.stack(om).astype(np.float32).astype(np.float64)
keys = [(i, "tonic_volatility") for i in (1, 2, 3)]
res = nets.continuous3(Network).batch(x, dict(zip(keys, pc)), response_function=response.first_level_gaussian_surprise)
got = res.surprise.double().cpu().numpy()
o32 = nets.continuous3(onet.Network).create_belief_propagation_fn()
o64 = nets.continuous3(onet.Network, dtype="float64").create_belief_propagation_fn()
slots = [o32.slot(i, "tonic_volatility") for i in (1, 2, 3)]
s32 = o32.batch_surprise(x, slots, pc, "first_level_gaussian_surprise")
s64 = o64.batch_surprise(x, slots, pc, "first_level_gaussian_surprise")
fin = np.isfinite(s32) & np.isfinite(s64) & np.isfinite(got)
sens = rel(s32, s64, 1.0)
well = fin & (sens <= 1e-5)
out["continuous3_logp_config2_sample"] = {
    "instances": n, "finite_pattern_equal": bool(np.array_equal(np.isfinite(got), np.isfinite(s32))), "well_conditioned": int(well.sum()),
    "oracle_f32_vs_f64_all": q(sens[fin]), "cuda_vs_oracle_f32_well": q(rel(got, s32, 1.0)[well]), "cuda_vs_oracle_f32_all": q(rel(got, s32, 1.0)[fin])}
print(json.dumps(out, indent=1))
