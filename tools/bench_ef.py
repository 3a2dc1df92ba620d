"""Throughput of hgf_ef_filter (ef-state nodes, fixed learning rate): instances x steps per second, and the HBM
write rate of the trajectory case against the measured copy peak (MEASURED_PEAKS.json when present)."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pyhgf_b200 import Network  # noqa: E402


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
    T = 1000
    for d, N, record in ((2, 65536, True), (2, 1 << 20, False), (1, 1 << 20, False), (8, 1 << 18, False)):
        K = d + d * (d + 1) // 2
        x = torch.as_tensor(rng.normal(size=(T, d)), dtype=torch.float32, device="cuda")
        nus = torch.as_tensor(rng.uniform(1, 10, N), dtype=torch.float32, device="cuda")
        net = Network().add_nodes(kind="ef-state", dimension=d, distribution="normal" if d == 1 else "multivariate-normal")
        ms = timed(lambda: net.batch(x, {(0, "nus"): nus}, record=[(0, "xis")] if record else [], last_attributes=True))
        out = {"dimension": d, "n_instances": N, "n_steps": T, "trajectory": record, "ms": round(ms, 4),
               "instance_steps_per_s": N * T / ms * 1e3}
        if record:
            out["trajectory_write_GBps"] = K * T * N * 4 / ms / 1e6
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
