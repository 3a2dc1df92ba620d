"""Scratch diagnostic (GPU box): float32 / float64 CUDA gradients vs the Dual<double> oracle."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import network as onet
from pyhgf_b200 import load_data, response
from pyhgf_b200.network import Network

u, y = load_data("binary")
rng = np.random.default_rng(0)
n = 4096
om2, om3 = rng.uniform(-6, -2, n), rng.uniform(-6, -2, n)
temp = np.exp(rng.uniform(-0.3, 0.8, n))
mk = lambda cls, **kw: cls(**kw).add_nodes(kind="binary-state").add_nodes(value_children=0).add_nodes(volatility_children=1)
o = mk(onet.Network).create_belief_propagation_fn()
slots = [o.slot(1, "tonic_volatility"), o.slot(2, "tonic_volatility"), o.slot(-1, "response_parameter", 0)]
p = np.stack([om2, om3, temp]).astype(np.float32).astype(np.float64)
keys = [(1, "tonic_volatility"), (2, "tonic_volatility"), ("response_parameter", 0)]
for dt in ("float32", "float64"):
    ws, wg = o.batch_grad(u, slots, p, "binary_softmax_inverse_temperature", resp_inputs=y, dtype=int(dt[-2:]))
    g = mk(Network, dtype=dt)
    res = g.batch(u, dict(zip(keys, p)), response_function=response.binary_softmax_inverse_temperature, response_function_inputs=y, gradient=True)
    s = res.surprise.double().cpu().numpy()
    print(dt, "surprise max rel", np.abs(s / ws - 1).max())
    for i, k in enumerate(keys):
        gg = res.gradient[k].double().cpu().numpy()
        err = np.abs(gg - wg[i])
        rel = err / (np.abs(wg[i]) + 1e-30)
        j = np.argmax(err)
        print(dt, k, "scale", np.abs(wg[i]).max(), "max abs err", err.max(), "at", p[:, j], "want", wg[i][j], "got", gg[j],
              "median rel", np.median(rel), "p99 rel", np.quantile(rel, 0.99), "frac>1e-3", (rel > 1e-3).mean())
