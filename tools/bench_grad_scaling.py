"""logp + gradient cost against the number of differentiated parameters P: forward-mode tangents in the filter's registers
(ceil(P / 4) passes) versus the generated checkpointed reverse sweep (csrc/hgf_reverse.cuh), on the BASELINE topologies.
65 536 instances; binary3: 320 trials (config 4's shape), continuous3: 614 observations (config 2's); ghgf16 (config 5's network): 32 768 x 300.

    python tools/bench_grad_scaling.py > profiles/r2_grad_scaling.log
"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from pyhgf_b200 import load_data, response
from pyhgf_b200.network import Network

dev, N = "cuda", 65536
rng = np.random.default_rng(0)


def col(lo, hi, log=False):
    v = rng.uniform(lo, hi, N)
    return torch.as_tensor(np.exp(v) if log else v, dtype=torch.float32, device=dev)


def time_launch(res, reps=7):
    for _ in range(4):
        res.relaunch()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); res.relaunch(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


cases = {}
u, y = load_data("binary")
cases["binary3"] = dict(
    net=lambda: bench.net_binary3(Network), series=u, resp=response.binary_softmax_inverse_temperature, resp_inputs=y,
    params=[((1, "tonic_volatility"), col(-6, -3)), ((2, "tonic_volatility"), col(-6, -3)), (("response_parameter", 0), col(-0.3, 0.8, True)),
            ((1, "mean"), col(-0.5, 0.5)), ((2, "mean"), col(-0.5, 0.5)), ((1, "precision"), col(-0.5, 0.5, True)), ((2, "precision"), col(-0.5, 0.5, True)),
            ((1, "tonic_drift"), col(-0.01, 0.01)), ((2, "tonic_drift"), col(-0.01, 0.01)), ((1, "autoconnection_strength"), col(0.9, 1.0)),
            ((2, "autoconnection_strength"), col(0.9, 1.0)), ((2, "volatility_coupling_children", 0), col(0.8, 1.2)),
            ((1, "volatility_coupling_parents", 0), col(0.8, 1.2)), ((1, "value_coupling_children", 0), col(0.9, 1.1)),
            ((0, "value_coupling_parents", 0), col(0.9, 1.1)), (("precision_clipping_value",), col(-14, -13, True))])
x = load_data("continuous")
cases["continuous3"] = dict(
    net=lambda: bench.net_continuous3(Network), series=x, resp=response.first_level_gaussian_surprise, resp_inputs=None,
    params=[((1, "tonic_volatility"), col(-14, -12)), ((2, "tonic_volatility"), col(-3, -1)), ((3, "tonic_volatility"), col(-3, -1)),
            ((2, "mean"), col(0.5, 1.5)), ((3, "mean"), col(0.5, 1.5)), ((2, "precision"), col(2, 3, True)), ((3, "precision"), col(2, 3, True)),
            ((1, "tonic_drift"), col(-1e-3, 1e-3)), ((2, "tonic_drift"), col(-0.01, 0.01)), ((3, "tonic_drift"), col(-0.01, 0.01)),
            ((2, "autoconnection_strength"), col(0.9, 1.0)), ((3, "autoconnection_strength"), col(0.9, 1.0)),
            ((2, "volatility_coupling_children", 0), col(0.8, 1.2)), ((1, "volatility_coupling_parents", 0), col(0.8, 1.2)),
            ((3, "volatility_coupling_children", 0), col(0.8, 1.2)), ((2, "volatility_coupling_parents", 0), col(0.8, 1.2))])

# BASELINE config 5's 16-node network (volatile-state nodes): 32 768 instances x 300 steps; forward mode runs through the opcode
# interpreter (no register-resident gradient kernels are built for it), the reverse sweep is generated (Rev_ghgf16)
N16 = 32768
cases["ghgf16"] = dict(
    net=lambda: bench.net_ghgf16(Network), series=bench.ar1_series(300, 4, seed=1), resp=response.total_gaussian_surprise, resp_inputs=None, n=N16,
    params=[((i, "tonic_volatility"), col(-5, -3)[:N16]) for i in range(4, 12)] + [((i, "tonic_volatility_vol"), col(-5, -3)[:N16]) for i in range(12, 16)] +
           [((12, "mean"), col(-0.3, 0.3)[:N16]), ((13, "mean"), col(-0.3, 0.3)[:N16]), ((14, "mean_vol"), col(-0.3, 0.3)[:N16]), ((15, "mean_vol"), col(-0.3, 0.3)[:N16])])

print(f"# {torch.cuda.get_device_name(0)}; N = {N} instances (ghgf16: {N16}); ms per logp+grad launch (median of 7 after 4 warm-up launches, inputs resident)")
for name, c in cases.items():
    net = c["net"]()
    kw = dict(response_function=c["resp"])
    if c["resp_inputs"] is not None:
        kw["response_function_inputs"] = c["resp_inputs"]
    fwd0 = time_launch(net.batch(c["series"], dict(c["params"][:1]), **kw))
    print(f"{name}: forward filter alone (logp, no gradient) {fwd0:.3f} ms")
    for P in ((4, 8, 16) if name == "ghgf16" else (1, 2, 3, 4, 6, 8, 9, 12, 16)):
        p = dict(c["params"][:P])
        row = {"topology": name, "P": P, "forward_only_ms": fwd0}
        for mode in ("forward", "reverse"):
            res = net.batch(c["series"], p, gradient=mode, **kw)
            row[mode + "_ms"] = time_launch(res)
            row[mode + "_launches"] = None
            if mode == "forward":
                g_f = {k: v.clone() for k, v in res.gradient.items()}
            else:
                worst = 0.0
                for k, v in res.gradient.items():
                    sc = g_f[k].abs().median().clamp_min(1e-30)
                    worst = max(worst, float(((v - g_f[k]).abs() / (g_f[k].abs() + sc)).median()))
                row["median_rel_diff_reverse_vs_forward"] = worst
        row["reverse_over_forward"] = row["reverse_ms"] / row["forward_ms"]
        print(json.dumps(row), flush=True)
