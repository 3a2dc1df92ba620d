"""2+ GPU check (torchrun): the in-kernel all-reduce over peer memory (FusedReduce) against NCCL, for several
consecutive launches (epochs), with and without the multicast mapping; then both timed on BASELINE config 4's shape.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/fused_reduce_check.py
"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from pyhgf_b200 import response
from pyhgf_b200.network import Network
from pyhgf_b200.distributed import FusedReduce

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
out = {"world": world}


def make(n_part, n_chain, T, seed):
    rng = np.random.default_rng(seed)
    u = (rng.uniform(size=(n_part, T)) > 0.5).astype(np.float32)
    y = (rng.uniform(size=(n_part, T)) > 0.5).astype(np.float32)
    n = n_part * n_chain
    p = torch.as_tensor(np.stack([rng.uniform(-6, -3, n), rng.uniform(-6, -3, n), np.exp(rng.uniform(-0.3, 0.8, n))]),
                        dtype=torch.float32, device=dev)
    return u, y, p


def launch(u, y, p, n_chain, fused=None):
    n = p.shape[1]
    net = Network().add_nodes(kind="binary-state").add_nodes(value_children=0).add_nodes(volatility_children=1)
    keys = [(1, "tonic_volatility"), (2, "tonic_volatility"), ("response_parameter", 0)]
    group = (torch.arange(n, device=dev, dtype=torch.int32) % n_chain).contiguous()
    return net.batch(torch.from_numpy(u).to(dev), {k: p[i] for i, k in enumerate(keys)},
                     response_function=response.binary_softmax_inverse_temperature, response_function_inputs=torch.from_numpy(y).to(dev),
                     input_group_size=n_chain, response_group_size=n_chain, gradient=True, group=group, n_groups=n_chain,
                     out_scale=-1.0, fused_reduce=fused)


# ---- correctness over epochs, ragged sizes
worst = 0.0
for multicast in (False, True):
    n_chain = 48
    fr = FusedReduce(n_chain, 32, dev, multicast=multicast)
    out[f"multicast_ptr_{int(multicast)}"] = bool(fr.multicast_ptr)
    u, y, p = make(37 + rank, n_chain, 200, 100 + rank)
    ref = launch(u, y, p, n_chain)
    res = launch(u, y, p, n_chain, fused=fr)
    for epoch in range(6):
        if epoch:
            p.add_(0.01 * (epoch + rank))  # parameters change in place between launches
            ref.relaunch(); res.relaunch()
        want = ref.reduced.clone()
        dist.all_reduce(want)
        torch.cuda.synchronize()
        got = res.reduced.clone()
        err = float(((got - want).abs() / want.abs()).max())
        worst = max(worst, err)
        assert err < 1e-5, (multicast, epoch, err, got[:4], want[:4])
    del fr
out["max_rel_err_vs_nccl"] = worst

# ---- timing on config 4's shape (per GPU: 1024 participants x 64 chains x 320 trials)
u, y, p = make(1024, 64, 320, 7 + rank)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for name in ("nccl", "fused", "fused_multicast"):
    fr = None if name == "nccl" else FusedReduce(64, 32, dev, multicast=name.endswith("multicast"))
    res = launch(u, y, p, 64, fused=fr)
    def step():
        res.relaunch()
        if fr is None:
            dist.all_reduce(res.reduced)
    for _ in range(5):
        flush.zero_(); step()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(20)]
    for a, b in ev:
        flush.zero_(); a.record(); step(); b.record()
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in ev) / len(ev)
    t = torch.tensor([ms], device=dev, dtype=torch.float64); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    out[f"ms_per_step_{name}"] = float(t.item())
    dist.barrier()
if rank == 0:
    print(json.dumps(out))
dist.barrier()
dist.destroy_process_group()
