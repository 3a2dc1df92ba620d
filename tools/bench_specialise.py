"""Interpreter vs on-demand specialisation for a topology that has no built-in kernel: a CONNECTED 16-node
generalised HGF (as BASELINE config 5's, but volatile node 14 is the value parent of both 12 and 13 and node 15
sits above 14).  `python tools/bench_specialise.py build` only compiles (CPU box: warms the on-disk cache)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import nets
from pyhgf_b200.network import Network


def connected16(cls, **kw):
    net = cls(**kw)
    net.add_nodes(n_nodes=4, precision=100.0)
    for i in range(4):
        net.add_nodes(value_children=i, tonic_volatility=-4.0)
    for i in range(4):
        net.add_nodes(volatility_children=4 + i, tonic_volatility=-4.0)
    net.add_nodes(kind="volatile-state", value_children=[4, 5], autoconnection_strength=1.0)
    net.add_nodes(kind="volatile-state", value_children=[6, 7], autoconnection_strength=1.0)
    net.add_nodes(kind="volatile-state", value_children=[12, 13], autoconnection_strength=1.0)
    net.add_nodes(kind="volatile-state", value_children=14, autoconnection_strength=1.0)
    return net


if len(sys.argv) > 1 and sys.argv[1] == "build":
    print(connected16(Network).specialise(verbose=True))
    for b, g in ((nets.multi_child, 2), (nets.volatile2, 1), (nets.nonlinear, 2)):
        print(b.__name__, b(Network).specialise(gradient_parameters=g, verbose=True))
    sys.exit(0)

import torch
T, n = 2000, 32768
x = nets.ar1_series(T)
rng = np.random.default_rng(1)
params = {(i, "tonic_volatility"): torch.as_tensor(rng.normal(-4, 1.0, n), dtype=torch.float32, device="cuda") for i in range(4, 12)}
record = [(i, k) for i in range(16) for k in nets.BELIEFS]
for mode in ("interpreter", "specialised"):
    net = connected16(Network)
    if mode == "specialised":
        net.specialise()
    for rec in ([], record):
        res = net.batch(x, params, record=rec)
        torch.cuda.synchronize()
        ts = []
        for _ in range(3):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); res.relaunch(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
        ms = min(ts)
        print(f"connected16 N={n} T={T} records={len(rec)}: kernel={res.kernel_name} {ms:.2f} ms -> {n*T*16/ms/1e6:.1f} G node-updates/s", flush=True)
        del res
        torch.cuda.empty_cache()
