"""Per-config throughput of the BASELINE configurations on one GPU (scratch / DESIGN.md numbers).

    python tools/bench_configs.py [c2] [c3] [c5]
"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
import nets
from pyhgf_b200 import load_data, response
from pyhgf_b200.network import Network

which = sys.argv[1:] or ["c2", "c3", "c5"]
reps = int(os.environ.get("REPS", "5"))
dev = "cuda"


def timeit(res, reps=reps):
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); res.relaunch(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return min(ts)


if "c2" in which:
    x = load_data("continuous")
    n = 65536
    rng = np.random.default_rng(0)
    om = [torch.as_tensor(np.clip(rng.normal(m, 1.0, n), m - 3, m + 3), dtype=torch.float32, device=dev) for m in (-13.0, -2.0, -2.0)]
    net = nets.continuous3(Network)
    res = net.batch(x, {(1, "tonic_volatility"): om[0], (2, "tonic_volatility"): om[1], (3, "tonic_volatility"): om[2]},
                    response_function=response.first_level_gaussian_surprise)
    ms = timeit(res)
    print(f"C2 continuous3 N={n} T=614 fwd logp: kernel={res.kernel_name} {ms:.3f} ms -> {n*614*4/ms/1e6:.1f} G node-updates/s", flush=True)
    res = net.batch(x, {(1, "tonic_volatility"): om[0], (2, "tonic_volatility"): om[1], (3, "tonic_volatility"): om[2]},
                    response_function=response.first_level_gaussian_surprise, gradient=True)
    ms = timeit(res)
    print(f"C2 continuous3 N={n} T=614 logp+grad(3): kernel={res.kernel_name} {ms:.3f} ms -> {n*614*4/ms/1e6:.1f} G node-updates/s", flush=True)

if "c3" in which:
    u, _ = load_data("binary")
    side = 1024
    om2 = torch.linspace(-8, 0, side, device=dev).repeat_interleave(side)
    om3 = torch.linspace(-8, 0, side, device=dev).repeat(side)
    net = nets.binary3(Network)
    res = net.batch(u, {(1, "tonic_volatility"): om2, (2, "tonic_volatility"): om3}, response_function=response.first_level_binary_surprise)
    ms = timeit(res)
    n = side * side
    print(f"C3 binary3 grid N={n} T=320 fwd logp: kernel={res.kernel_name} {ms:.3f} ms -> {n*320*3/ms/1e6:.1f} G node-updates/s", flush=True)

if "c5" in which:
    T = int(os.environ.get("C5_T", "2000"))
    n = int(os.environ.get("C5_N", "32768"))
    x = nets.ar1_series(T)
    net = nets.ghgf16(Network)
    rng = np.random.default_rng(1)
    params = {(i, "tonic_volatility"): torch.as_tensor(rng.normal(-4, 1.0, n), dtype=torch.float32, device=dev) for i in range(4, 12)}
    for i in range(12, 16):
        params[(i, "tonic_volatility_vol")] = torch.as_tensor(rng.normal(-4, 1.0, n), dtype=torch.float32, device=dev)
    record = [(i, k) for i in range(16) for k in nets.BELIEFS + (("mean_vol", "expected_mean_vol", "precision_vol", "expected_precision_vol") if i >= 12 else ())]
    for rec in ([], record):
        if os.environ.get("HGF_FORCE_INTERPRETER") == "1":
            net = nets.ghgf16(Network)
        res = net.batch(x, params, record=rec)
        ms = timeit(res, reps=3)
        gb = len(rec) * T * n * 4 / 1e9
        print(f"C5 ghgf16 N={n} T={T} records={len(rec)} ({gb:.1f} GB): kernel={res.kernel_name} {ms:.2f} ms -> {n*T*16/ms/1e6:.1f} G node-updates/s, {gb/ms*1e3:.0f} GB/s", flush=True)
        del res
        torch.cuda.empty_cache()
