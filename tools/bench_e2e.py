"""e2e leg of bench.py (hgf_logp_grad_host, pinned host buffers) under different chunk-pipeline settings."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch

    from pyhgf_b200.network import Network

    torch.cuda.set_device(0)
    u, y, params = bench.synth_config4(0)
    net = Network().add_nodes(kind="binary-state").add_nodes(value_children=0).add_nodes(volatility_children=1)
    net.create_belief_propagation_fn()
    for chunks in sys.argv[1:] or ["4", "0", "2", "3", "6", "8", "4"]:
        os.environ["HGF_HOST_CHUNKS"] = chunks
        best = []
        for _ in range(3):
            dt, h2d, d2h, _ = bench.e2e_leg(net, net.plan, u, y, params, 50, 10, 0)
            best.append(1e3 * dt / 50)
        print(f"HGF_HOST_CHUNKS={chunks}: {min(best):.4f} ms/step (runs: {', '.join(f'{b:.4f}' for b in best)})", flush=True)


if __name__ == "__main__":
    main()
