"""The N>1 host logic on CPU: world_size-2 gloo processes shard participants and all-reduce the per-chain
sums.  The per-rank evaluation is a stand-in with the evaluator's interface (the CUDA op needs a GPU); what
is under test is the partition, the slicing of parameters and which gradients get reduced."""

import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def test_shard_bounds_partition_everything():
    from pyhgf_b200.distributed import shard_bounds, shard_groups

    for n in (0, 1, 7, 1024, 1027):
        for w in (1, 2, 3, 8):
            spans = [shard_bounds(n, w, r) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    assert shard_groups(10, 64, 2, 1) == (5, 10, 320, 640)
    with pytest.raises(ValueError):
        shard_bounds(4, 2, 2)


class FakeEvaluator:
    """logp_{p,c} = -(a_p * theta_{p,c} - t)^2 summed over trials' mean; analytic gradients."""

    def __init__(self, input_data, response_function, response_function_inputs, n_chains):
        self.a = torch.as_tensor(np.asarray(input_data)).double().mean(1)  # (n_local,)
        self.c = n_chains

    def _terms(self, theta, t):
        theta = torch.as_tensor(theta).double().reshape(len(self.a), self.c)
        r = self.a[:, None] * theta - torch.as_tensor(t).double()
        return theta, r

    def __call__(self, theta, t):
        _, r = self._terms(theta, t)
        return (-(r**2)).sum(0)

    def value_and_grad(self, theta, t):
        _, r = self._terms(theta, t)
        return (-(r**2)).sum(0), {"theta": -2 * r * self.a[:, None], "t": (2 * r).sum(0) if np.ndim(t) else (2 * r).sum()}


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from pyhgf_b200.distributed import ShardedLogp

    rng = np.random.default_rng(0)
    n, c, T = 7, 3, 5
    data = rng.normal(size=(n, T))
    theta = rng.normal(size=(n, c))
    t = rng.normal(size=(c,))
    op = ShardedLogp(data, None, None, n_chains=c, evaluator_factory=FakeEvaluator)
    total = op(theta=theta, t=t)
    total2, grads = op.value_and_grad(theta=theta, t=t)
    full = FakeEvaluator(data, None, None, c)
    want, want_g = full.value_and_grad(theta, t)
    ok = torch.allclose(total, want) and torch.allclose(total2, want)
    ok = ok and torch.allclose(grads["t"], want_g["t"])  # shared parameter: reduced
    ok = ok and torch.allclose(grads["theta"], want_g["theta"][op.local_slice])  # per participant: sharded
    ok = ok and (op.local_slice == slice(0, 4) if rank == 0 else op.local_slice == slice(4, 7))
    out[rank] = bool(ok)
    dist.destroy_process_group()


def test_sharded_logp_two_ranks_gloo():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    assert out[0] and out[1]
