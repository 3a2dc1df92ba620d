"""Scratch (GPU box): time forward / logp+grad launches of the 3-level binary HGF at several batch sizes."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pyhgf_b200 import load_data, response
from pyhgf_b200.network import Network

u, y = load_data("binary")
reps = int(os.environ.get("REPS", "5"))
sizes = [int(s) for s in os.environ.get("SIZES", "65536,262144,1048576").split(",")]
net = Network().add_nodes(kind="binary-state").add_nodes(value_children=0).add_nodes(volatility_children=1)
keys = [(1, "tonic_volatility"), (2, "tonic_volatility"), ("response_parameter", 0)]
for n in sizes:
    rng = np.random.default_rng(0)
    p = [torch.as_tensor(rng.uniform(-6, -3, n), dtype=torch.float32, device="cuda"),
         torch.as_tensor(rng.uniform(-6, -3, n), dtype=torch.float32, device="cuda"),
         torch.as_tensor(np.exp(rng.uniform(-0.3, 0.8, n)), dtype=torch.float32, device="cuda")]
    for grad in (False, True):
        res = net.batch(u, dict(zip(keys, p)), response_function=response.binary_softmax_inverse_temperature,
                        response_function_inputs=y, gradient=grad)
        torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); res.relaunch(); b.record(); torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        ms = min(ts)
        print(f"n={n:8d} grad={int(grad)} kernel={res.kernel_name} best {ms:.3f} ms  -> {n*320/ms/1e6:.2f} G instance-steps/s, {n*320*3/ms/1e6:.1f} G node-updates/s", flush=True)
