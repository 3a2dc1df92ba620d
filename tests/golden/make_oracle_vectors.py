"""Write tests/golden/oracle_vectors.npz: float64 results of the CPU oracle (oracle/hgf_oracle.cpp, pinned to the
reference's own goldens by tests/test_oracle_golden.py) on the reference's data series, for the BASELINE topologies.
The reference itself cannot be imported in the authoring container (no jax), so these vectors come from the pinned
restatement; they are committed so that the GPU parity tests also hold against fixed numbers.

    python tests/golden/make_oracle_vectors.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
sys.path.insert(0, os.path.join(HERE, ".."))
import nets  # noqa: E402
from oracle import conditioning  # noqa: E402
from oracle import network as onet  # noqa: E402

d = np.load(os.path.join(HERE, "..", "..", "pyhgf_b200", "data", "datasets.npz"))
u, y, x = d["binary_u"], d["binary_y"], d["continuous"]
out = {}

# trajectories of the single reference networks (float64), every belief leaf
for name, build, data in (("binary3", nets.binary3, u), ("continuous3", nets.continuous3, x), ("ghgf16", nets.ghgf16, nets.ar1_series(200))):
    net = build(onet.Network, dtype="float64").input_data(data)
    for node in range(len(net.edges)):
        for k in nets.BELIEFS:
            out[f"traj/{name}/{node}/{k}"] = np.asarray(net.node_trajectories[node][k], dtype=np.float64)

# logp + gradient of the 3-level binary HGF with the inverse-temperature response (BASELINE config 4's evaluation)
rng = np.random.default_rng(20260101)
n = 64
p = np.stack([rng.uniform(-6, -3.5, n), rng.uniform(-6, -3.5, n), np.exp(rng.uniform(-0.3, 0.8, n))])
p = p.astype(np.float32).astype(np.float64)
o = nets.binary3(onet.Network, dtype="float64").create_belief_propagation_fn()
slots = [o.slot(1, "tonic_volatility"), o.slot(2, "tonic_volatility"), o.slot(-1, "response_parameter", 0)]
s, g = o.batch_grad(u, slots, p, "binary_softmax_inverse_temperature", resp_inputs=y)
o32 = nets.binary3(onet.Network).create_belief_propagation_fn()
s32 = o32.batch_surprise(u, slots, p, "binary_softmax_inverse_temperature", resp_inputs=y)
out["config4/params"], out["config4/surprise"], out["config4/gradient"], out["config4/surprise_f32"] = p, s, g, s32


# the float32 oracle's gradients (Dual<float>) and, per instance, how far its own float32 results move when the filter
# parameters move by 1e-6 relative (oracle/conditioning.py): the bound the float32 launches are held to
def _f32(q):
    s_, g_ = o32.batch_grad(u, slots, q, "binary_softmax_inverse_temperature", resp_inputs=y, dtype=32)
    return np.concatenate([s_[None], g_])


g32 = _f32(p)[1:]
floor = np.concatenate([[0.0], 1e-4 * np.abs(g32).max(axis=1)])[:, None]
base, sens = conditioning.sensitivity(_f32, p, rows=[0, 1], n_draws=16, floor=floor)
assert np.array_equal(base[0], s32)
out["config4/gradient_f32"], out["config4/sensitivity_f32"] = g32, sens

# config 2: first-level Gaussian surprise of the 4-node continuous HGF
pc = np.stack([np.clip(rng.normal(m, 1.0, n), m - 2, m + 2) for m in (-13.0, -2.0, -2.0)]).astype(np.float32).astype(np.float64)
oc = nets.continuous3(onet.Network, dtype="float64").create_belief_propagation_fn()
slots_c = [oc.slot(i, "tonic_volatility") for i in (1, 2, 3)]
out["config2/params"] = pc
out["config2/surprise"] = oc.batch_surprise(x, slots_c, pc, "first_level_gaussian_surprise")
oc32 = nets.continuous3(onet.Network).create_belief_propagation_fn()
out["config2/surprise_f32"], out["config2/sensitivity_f32"] = conditioning.sensitivity(
    lambda q: oc32.batch_surprise(x, slots_c, q, "first_level_gaussian_surprise"), pc, n_draws=16, floor=100.0)

path = os.path.join(HERE, "oracle_vectors.npz")
np.savez_compressed(path, **out)
print("wrote", path, len(out), "arrays", os.path.getsize(path), "bytes")
