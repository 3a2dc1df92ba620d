"""Instance-sharded evaluation across the GPUs of one box.

Network instances are independent (no exchange inside the time loop), so the
batch shards contiguously over ranks -- by participant for grouped inputs, so
that a participant's series and all its chains live on one GPU -- with no
data-path collective.  The only exchange is the sum of the per-group (per-chain)
log-probabilities and of gradients of *shared* parameters: one small
all-reduce over NCCL / NVLink (``torch.distributed``), enqueued on the stream
of the kernel that produced the sums.  The reference has no multi-device path
("Parallelism strategies"): this layer is new.
"""

from __future__ import annotations

from typing import Optional, Sequence

import torch
import torch.distributed as dist


def shard_bounds(n_items: int, world_size: int, rank: int) -> tuple:
    """Contiguous, balanced partition: the first ``n_items % world_size`` ranks get one extra item."""
    if world_size < 1 or not 0 <= rank < world_size:
        raise ValueError("invalid rank / world_size")
    base, extra = divmod(n_items, world_size)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_groups(n_groups: int, group_size: int, world_size: int, rank: int) -> tuple:
    """Shard whole groups (participants); returns (first group, last group, first instance, last instance)."""
    g0, g1 = shard_bounds(n_groups, world_size, rank)
    return g0, g1, g0 * group_size, g1 * group_size


def all_reduce_sums(tensors: Sequence[torch.Tensor], group: Optional[dist.ProcessGroup] = None) -> None:
    """Sum small per-chain / shared-parameter tensors over ranks, in place (one flat all-reduce)."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    tensors = [t for t in tensors if t is not None]
    if not tensors:
        return
    if len(tensors) == 1:
        dist.all_reduce(tensors[0], op=dist.ReduceOp.SUM, group=group)
        return
    flat = torch.cat([t.reshape(-1) for t in tensors])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    off = 0
    for t in tensors:
        t.copy_(flat[off:off + t.numel()].view_as(t))
        off += t.numel()


class FusedReduce:
    """Workspace of the IN-KERNEL all-reduce of the per-group (per-chain) sums (``hgf_call_t::peer_workspace``).

    Each rank allocates ``hgf_fused_reduce_bytes`` of symmetric memory (CUDA VMM, mapped into every peer by
    ``torch.distributed._symmetric_memory``); the kernel's last block adds the rank's sums into every peer's copy
    over NVLink (one ``multimem.red`` per value through the NVSwitch multicast mapping when the driver offers one),
    publishes "deposited" to every peer and the kernel ends: it never waits for the other ranks, and there is no NCCL
    call on the path.  The consumer waits: ``result()`` enqueues a one-block kernel on the current stream that
    returns once every rank's sums of that launch have arrived, and hands back the ``[n_groups]`` view.  Pass the
    object to ``Network.batch(..., fused_reduce=...)``; every rank must launch the same number of times (the epoch
    counter is part of the protocol), and ranks may drift apart by up to ``RING // 2 - 1`` launches.
    """

    RING = 16  # HGF_FUSED_RING

    def __init__(self, n_groups: int, bits: int = 32, device=None, group: Optional[dist.ProcessGroup] = None,
                 multicast: bool = True):
        import ctypes as C

        import torch.distributed._symmetric_memory as symm_mem

        from . import _lib

        L = _lib.lib()
        group = dist.group.WORLD if group is None else group
        self.n_groups, self.bits = int(n_groups), int(bits)
        self.world_size, self.rank = dist.get_world_size(group), dist.get_rank(group)
        if self.world_size > 8:
            raise ValueError("the fused reduction spans the GPUs of one box (<= 8)")
        nbytes = int(L.hgf_fused_reduce_bytes(self.n_groups, self.bits))
        device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        try:
            self.buffer = symm_mem.empty(nbytes, dtype=torch.uint8, device=device)
        except RuntimeError:  # older torch: the group must be enabled first
            symm_mem.enable_symm_mem_for_group(group.group_name)
            self.buffer = symm_mem.empty(nbytes, dtype=torch.uint8, device=device)
        self.buffer.zero_()
        self.handle = symm_mem.rendezvous(self.buffer, group.group_name)
        self.peer_ptrs = (C.c_void_p * self.world_size)(*[int(p) for p in self.handle.buffer_ptrs])
        mc = 0
        if multicast:
            try:
                mc = int(self.handle.multicast_ptr or 0)
            except (RuntimeError, AttributeError):
                mc = 0
        self.multicast_ptr = mc
        self.epoch = 0
        self._L = L
        torch.cuda.synchronize(device)
        dist.barrier(group)  # every rank's workspace is zeroed before anyone's first launch

    def attach(self, call) -> int:
        """Fill the fused-reduction fields of an ``hgf_call_t`` for the next launch; returns its epoch."""
        import ctypes as C

        e = self.epoch
        self.epoch += 1
        call.n_groups, call.world_size, call.rank, call.epoch = self.n_groups, self.world_size, self.rank, e & 0xFFFFFFFF
        call.peer_workspace = C.cast(self.peer_ptrs, C.POINTER(C.c_void_p))
        call.multicast_workspace = self.multicast_ptr or None
        call.reduced = None
        return e

    def result(self, epoch: Optional[int] = None, wait: bool = True, stream=None) -> torch.Tensor:
        """The all-reduced ``[n_groups]`` sums of launch ``epoch`` (default: the last one).  ``wait`` enqueues, on
        ``stream`` (default: the current one), the wait for every rank's deposit of that launch; work enqueued on
        that stream afterwards sees the complete sums.  The view stays valid for the next ``RING // 2 - 1``
        launches on this workspace (the slot is then zeroed for reuse): clone it to keep it longer."""
        e = self.epoch - 1 if epoch is None else epoch
        if e < 0:
            raise RuntimeError("no launch has used this workspace yet")
        if e < self.epoch - (self.RING // 2 - 1):
            raise RuntimeError(f"the sums of launch {e} have been recycled ({self.epoch} launches so far, {self.RING // 2 - 1} kept)")
        if wait:
            from . import _lib

            dev = self.buffer.device
            with torch.cuda.device(dev):
                s = torch.cuda.current_stream(dev) if stream is None else stream
                _lib.check(self._L.hgf_fused_reduce_wait(self.buffer.data_ptr(), self.n_groups, self.bits, self.world_size,
                                                         e & 0xFFFFFFFF, s.cuda_stream))
        off = int(self._L.hgf_fused_reduce_result_offset(self.n_groups, self.bits, e & 0xFFFFFFFF))
        w = self.bits // 8
        return self.buffer[off:off + self.n_groups * w].view(torch.float32 if self.bits == 32 else torch.float64)


class ShardedLogp:
    """``HGFDistribution`` over the ranks of a process group.

    Every rank constructs it with the FULL ``input_data`` / ``response_function_inputs`` ``(N, T)`` and
    calls it with FULL parameter arrays ``(N,)`` / ``(N, n_chains)``; each rank evaluates its contiguous
    block of participants on its own GPU and the per-chain sums are all-reduced.  Per-participant
    gradients stay sharded (``local_slice`` tells which rows a rank owns).
    """

    def __init__(self, input_data, response_function, response_function_inputs=None, *, n_chains: int = 1, group=None,
                 evaluator_factory=None, **kwargs):
        self.group = group
        initialised = dist.is_available() and dist.is_initialized()
        self.world = dist.get_world_size(group) if initialised else 1
        self.rank = dist.get_rank(group) if initialised else 0
        n = len(input_data)
        self.lo, self.hi = shard_bounds(n, self.world, self.rank)
        self.n_participants, self.n_chains = n, n_chains
        if evaluator_factory is None:
            from .distribution import HGFDistribution as evaluator_factory
        resp = None if response_function_inputs is None else response_function_inputs[self.lo:self.hi]
        self.local = evaluator_factory(input_data=input_data[self.lo:self.hi], response_function=response_function,
                                       response_function_inputs=resp, n_chains=n_chains, **kwargs)

    @property
    def local_slice(self) -> slice:
        return slice(self.lo, self.hi)

    def _local_params(self, params: dict) -> dict:
        out = {}
        for k, v in params.items():
            shape = tuple(getattr(v, "shape", ()))
            out[k] = v[self.lo:self.hi] if len(shape) >= 1 and shape[0] == self.n_participants else v
        return out

    def __call__(self, **params):
        total = self.local(**self._local_params(params))
        total = total.clone() if torch.is_tensor(total) else torch.as_tensor(total)
        all_reduce_sums([total], self.group)
        return total

    def value_and_grad(self, **params):
        """(all-reduced sum of logp [per chain], {name: gradient}); gradients of per-participant parameters
        are this rank's rows, gradients of shared (scalar / per-chain) parameters are all-reduced."""
        total, grads = self.local.value_and_grad(**self._local_params(params))
        total = total.clone()
        shared = []
        for k, v in params.items():
            shape = tuple(getattr(v, "shape", ()))
            if not (len(shape) >= 1 and shape[0] == self.n_participants):
                grads[k] = grads[k].clone()
                shared.append(grads[k])
        all_reduce_sums([total] + shared, self.group)
        return total, grads
