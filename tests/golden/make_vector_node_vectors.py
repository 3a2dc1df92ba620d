"""Write tests/golden/vector_nodes.npz: results of the pinned CPU restatements of the categorical and exponential-family
state nodes (oracle/categorical.py, oracle/ef_state.py; pinned by tests/test_oracle_golden.py to tests/test_categorical.py
and tests/test_nodes/test_exponential_family.py of the reference), committed so that the GPU parity tests also hold
against fixed numbers.  The reference cannot be imported in the authoring container (no jax).

    python tests/golden/make_vector_node_vectors.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import categorical as cat  # noqa: E402
from oracle import ef_state as ef  # noqa: E402

out = {}
np.random.seed(123)
x3 = np.array([np.random.multinomial(n=1, pvals=[0.1, 0.2, 0.7]) for _ in range(10)], dtype=float)  # tests/test_categorical.py:12-16
x5 = np.random.default_rng(7).multinomial(1, [0.05, 0.15, 0.2, 0.25, 0.35], size=80).astype(float)
for name, x, n in (("reference3", x3, 3), ("five", x5, 5)):
    out[f"categorical/{name}/x"] = x
    for dtype in ("float32", "float64"):
        _, c = cat.categorical_filter(x, n, None, dtype=dtype)
        for key in ("alpha", "mean", "kl_divergence", "surprise"):
            out[f"categorical/{name}/{dtype}/{key}"] = c[key]

spiral = ef.spiral_data()
out["ef/spiral"] = spiral
out["ef/fixed/float32/xis"] = ef.ef_filter(spiral, 3.0, 2, "multivariate-normal")
out["ef/fixed/float64/xis"] = ef.ef_filter(spiral, 3.0, 2, "multivariate-normal", dtype=np.float64)
out["ef/fixed_univariate/float64/xis"] = ef.ef_filter(spiral[:, 0], 7.5, 1, "normal", dtype=np.float64)
dyn = ef.ef_dynamic_filter(spiral[:300], 2, "multivariate-normal", "hgf-2", dtype=np.float64)  # default attributes
out["ef/hgf2/float64/xis"], out["ef/hgf2/float64/nus"] = dyn["xis"], dyn["nus"]

path = os.path.join(HERE, "vector_nodes.npz")
np.savez_compressed(path, **out)
print("wrote", path, len(out), "arrays", os.path.getsize(path), "bytes")
