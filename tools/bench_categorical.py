"""Categorical state node, batched: the implied binary HGFs (interpreter, then an on-demand register-resident kernel)
plus hgf_categorical_update; and that kernel alone against the measured HBM copy peak."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pyhgf_b200 import Network, categorical  # noqa: E402


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    rng = np.random.default_rng(0)
    n, N, T = 4, 32768, 320
    x = rng.multinomial(1, [0.1, 0.2, 0.3, 0.4], size=T).astype(np.float32)
    om = torch.as_tensor(rng.uniform(-5.0, -2.0, N), dtype=torch.float32, device="cuda")
    net = Network().add_nodes(kind="categorical-state", node_parameters={"n_categories": n})
    params = {(1 + n + k, "tonic_volatility"): om for k in range(n)}
    record = [(0, "kl_divergence"), (0, "surprise")]
    run = lambda: net.batch(x, params, record=record)  # noqa: E731
    for label in ("interpreter", "specialised"):
        if label == "specialised":
            if sys.argv[1:] == ["nobuild"]:
                break
            net.specialise()
        ms = timed(run)
        r = run()
        print(json.dumps({"path": label, "kernel": r.kernel_name, "n_categories": n, "n_instances": N, "n_steps": T, "ms": round(ms, 3),
                          "node_updates_per_s": (2 * n + 1) * N * T / ms * 1e3,
                          "kl_finite_fraction": float(torch.isfinite(r.trajectories[(0, "kl_divergence")]).double().mean()),
                          "kl_mean_finite": float(torch.nanmean(r.trajectories[(0, "kl_divergence")]))}), flush=True)
    # the categorical step alone
    r = net.batch(x, params, record=[(1 + k, "expected_mean") for k in range(n)] + [(1 + n + k, "mean") for k in range(n)])
    em = torch.stack([r.trajectories[(1 + k, "expected_mean")].t() for k in range(n)]).contiguous()  # the filter's own output
    pm = torch.stack([r.trajectories[(1 + n + k, "mean")].t() for k in range(n)]).contiguous()
    ms = timed(lambda: net.batch(x, params, record=[(1 + k, "expected_mean") for k in range(n)] + [(1 + n + k, "mean") for k in range(n)]))
    print(json.dumps({"path": "implied filter alone (8 recorded leaves)", "ms": round(ms, 3)}), flush=True)
    obs = torch.as_tensor(x, device="cuda")[None].contiguous()
    for vec in (False, True):
        ms = timed(lambda: categorical.update(bits=32, device=torch.device("cuda"), observations=obs, expected_mean=em, parent_mean=pm,
                                              alpha0=np.ones(n), input_group_size=0, want_vectors=vec))
        moved = (2 * n + 2 + (2 * n if vec else 0)) * T * N * 4
        print(json.dumps({"path": "hgf_categorical_update", "alpha_and_mean_written": vec, "ms": round(ms, 4), "GBps": moved / ms / 1e6,
                          "steps_x_instances_per_s": T * N / ms * 1e3}), flush=True)


if __name__ == "__main__":
    main()
