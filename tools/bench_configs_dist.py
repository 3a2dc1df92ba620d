"""BASELINE configs 2, 3 and 5 under torchrun (one rank per GPU, no data-path collective):
  C3: the 1 024 x 1 024 (omega_2, omega_3) grid, rows sharded over the ranks (strong scaling);
  C2: 65 536 tonic-volatility samples sharded over the ranks (strong scaling);
  C5: 32 768 instances of the 16-node gHGF per rank with all 80 belief leaves recorded (262 144 on 8 GPUs).
Device time per launch = max over ranks (CUDA events, barrier on both sides); rank 0 prints one JSON line per config.

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/bench_configs_dist.py
"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch, torch.distributed as dist
import nets
from pyhgf_b200 import load_data, response
from pyhgf_b200.distributed import shard_bounds
from pyhgf_b200.network import Network

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    os.environ.pop("NCCL_DEBUG", None)  # the image sets NCCL_DEBUG=VERSION: keep the banner off stdout
    dist.init_process_group("nccl", device_id=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(res, reps=5):
    for _ in range(2):
        flush.zero_(); res.relaunch()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ts = []
    for _ in range(reps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); res.relaunch(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    t = torch.tensor([float(np.mean(ts))], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def report(name, units, ms, **kw):
    if rank == 0:
        print(json.dumps({"config": name, "n_gpus": world, "node_updates_per_s": units / (ms * 1e-3), "ms": ms, **kw}), flush=True)


# ---- C3
u, _ = load_data("binary")
side = 1024
lo, hi = shard_bounds(side, world, rank)
om2 = torch.linspace(-8, 0, side, device=dev)[lo:hi].repeat_interleave(side)
om3 = torch.linspace(-8, 0, side, device=dev).repeat(hi - lo)
res = nets.binary3(Network).batch(u, {(1, "tonic_volatility"): om2, (2, "tonic_volatility"): om3}, response_function=response.first_level_binary_surprise)
report("C3 binary3 grid 1M x 320, fwd logp (strong scaling)", side * side * 320 * 3, timed(res), kernel=res.kernel_name)
del res
# ---- C2
x = load_data("continuous")
n = 65536
rng = np.random.default_rng(0)
lo, hi = shard_bounds(n, world, rank)
om = [torch.as_tensor(np.clip(rng.normal(m, 1.0, n), m - 3, m + 3)[lo:hi], dtype=torch.float32, device=dev) for m in (-13.0, -2.0, -2.0)]
res = nets.continuous3(Network).batch(x, {(i + 1, "tonic_volatility"): om[i] for i in range(3)}, response_function=response.first_level_gaussian_surprise)
report("C2 continuous3 65 536 x 614, fwd logp (strong scaling)", n * 614 * 4, timed(res), kernel=res.kernel_name)
del res
# ---- C5
T, n = 2000, 32768
xs = nets.ar1_series(T)
rng = np.random.default_rng(1 + rank)
params = {(i, "tonic_volatility"): torch.as_tensor(rng.normal(-4, 1.0, n), dtype=torch.float32, device=dev) for i in range(4, 12)}
for i in range(12, 16):
    params[(i, "tonic_volatility_vol")] = torch.as_tensor(rng.normal(-4, 1.0, n), dtype=torch.float32, device=dev)
record = [(i, k) for i in range(16) for k in nets.BELIEFS + (("mean_vol", "expected_mean_vol", "precision_vol", "expected_precision_vol") if i >= 12 else ())]
res = nets.ghgf16(Network).batch(xs, params, record=record)
ms = timed(res, reps=3)
report("C5 ghgf16 32 768 per GPU x 2000, 80 leaves recorded (weak scaling)", world * n * T * 16, ms, kernel=res.kernel_name,
       trajectory_GB_per_gpu=len(record) * T * n * 4 / 1e9, hbm_write_GBps_per_gpu=len(record) * T * n * 4 / 1e9 / (ms * 1e-3))
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
