"""Config 5 with the 8-node component's kernel compiled for another occupancy bound.

    python tools/bench_c5_blocks.py <blocks per SM | registers per thread (> 32)> [--build-only]

Builds an on-demand specialisation of the component (the plain recording filter) with -DHGF_MID_MIN_BLOCKS=<blocks>
(hgf_scan.cuh::scan_min_blocks: the register cap is 65 536 / (64 x blocks)), registers it -- it then serves the
component's plans instead of the built-in kernel -- and runs the stock bench of config 5.  One value per process.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    blocks = int(sys.argv[1])  # > 32: a register cap (-DHGF_SCAN_MAXNREG) instead
    build_only = "--build-only" in sys.argv
    import bench
    from pyhgf_b200 import specialise as sp
    from pyhgf_b200.network import Network

    net = bench.net_ghgf16(Network)
    net.create_belief_propagation_fn()
    comp, _ = net._component_network(net.connected_components()[0])
    comp.create_belief_propagation_fn()
    name = sp.specialise(comp, float32=[(0, 0)], extra_flags=[f"-DHGF_SCAN_MAXNREG={blocks}" if blocks > 32 else f"-DHGF_MID_MIN_BLOCKS={blocks}"], load=not build_only)
    print(f"blocks={blocks} kernel={name}", file=sys.stderr, flush=True)
    if build_only:
        return
    sys.argv = ["bench.py", "--config", "5", "--no-cpu"]
    bench.main()


if __name__ == "__main__":
    main()
