"""Diagnosis aid: trajectories + pointwise surprise of chosen smoke-box instances under chosen math-policy variants
(tools/parity_ablation.py), saved for comparison with the oracle off-line.
    python tools/diag_instance.py OUT.npz IDX[,IDX...] VARIANT [VARIANT...]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np
import parity_ablation as pa
from pyhgf_b200 import load_data, response
from pyhgf_b200 import specialise as sp

out, idx, variants = sys.argv[1], [int(i) for i in sys.argv[2].split(",")], sys.argv[3:]
u, y = load_data("binary")
rng = np.random.default_rng(0)
n = 4096
p = np.stack([rng.uniform(-6, -3, n), rng.uniform(-6, -3, n), np.exp(rng.uniform(-0.3, 0.8, n))]).astype(np.float32)[:, idx]
rec = [(0, "expected_mean"), (1, "mean"), (1, "expected_mean"), (1, "precision"), (1, "expected_precision"), (2, "mean"), (2, "expected_mean"), (2, "precision"), (2, "expected_precision")]  # slot order: the static kernels record these
dump = {"params": p, "idx": np.array(idx)}
for v in variants:
    b3, _ = pa.networks()
    sp.specialise(b3, float32=[(0, 0), (2, 1)], extra_flags=pa.VARIANTS[v])
    r = b3.batch(u, dict(zip(pa.KEYS, p)), response_function=response.binary_softmax_inverse_temperature, response_function_inputs=y,
                 record=rec, pointwise=True)
    dump[f"{v}/logp"] = r.surprise.cpu().numpy()
    dump[f"{v}/pointwise"] = r.pointwise.cpu().numpy()
    for k in rec:
        dump[f"{v}/{k[0]}/{k[1]}"] = r.trajectories[k].cpu().numpy()
    print(v, r.kernel_name, dump[f"{v}/logp"], r.trajectories[(2, "mean")][:, -1].cpu().numpy())
np.savez_compressed(out, **dump)
