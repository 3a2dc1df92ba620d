"""Benchmark of the batched belief-propagation path (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Headline workload = BASELINE config 4 (the configuration the target is quoted on): 3-level binary HGF,
logp + gradient wrt (omega_2, omega_3, inverse temperature), 1 024 participants x 64 chains x 320 trials
PER GPU (weak scaling: participants are sharded, no data-path collective; the 64 per-chain log-probability
sums are all-reduced with NCCL).  A "step" is one fused logp+grad evaluation of the whole batch.
One JSON line on stdout (rank 0); see the contract in the task description.
"""

from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_PARTICIPANTS, N_CHAINS, N_TRIALS, N_NODES = 1024, 64, 320, 3
# algorithmic work of one 3-level binary step (forward), + inverse-temperature response
MUFU_PER_STEP = 55 + 5
MUFU_PEAK_PER_CLK_PER_SM, N_SM = 16, 148
# dram__bytes_read.sum + dram__bytes_write.sum of one launch of this kernel on this workload, from the
# `ncu --set full` capture summarised in profiles/r1_ncu_c4_logp_grad.txt (3.70 MB read, 0 B written: the
# 1.05 MB of results stay in L2 until evicted).  A constant from a profile, not measured in this run.
NCU_DRAM_BYTES_PER_LAUNCH = 3703808


def synth_config4(rank: int, n_participants: int = N_PARTICIPANTS, n_chains: int = N_CHAINS, n_trials: int = N_TRIALS):
    """Synthetic participants: block-wise Bernoulli contingencies switching every 40
    trials, decisions drawn from block-consistent noisy choices, chain parameters = truth + N(0, 0.5^2)."""
    rng = np.random.default_rng(1000 + rank)
    blocks = rng.integers(0, 2, size=(n_participants, (n_trials + 39) // 40))
    p = np.where(np.repeat(blocks, 40, axis=1)[:, :n_trials] == 1, 0.8, 0.2)
    u = (rng.uniform(size=p.shape) < p).astype(np.float32)
    y = (rng.uniform(size=p.shape) < 0.3 + 0.4 * p).astype(np.float32)
    om2_true = rng.normal(-4.0, 1.0, size=(n_participants, 1))
    om3_true = rng.normal(-4.0, 1.0, size=(n_participants, 1))
    logt_true = rng.normal(0.5, 0.5, size=(n_participants, 1))
    noise = rng.normal(0.0, 0.5, size=(3, n_participants, n_chains))
    om2 = np.clip(om2_true + noise[0], -8.0, -1.0)
    om3 = np.clip(om3_true + noise[1], -8.0, -1.0)
    temp = np.exp(np.clip(logt_true + noise[2], -1.5, 2.0))
    params = np.stack([om2.reshape(-1), om3.reshape(-1), temp.reshape(-1)]).astype(np.float32)
    return u, y, params


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md) through NVML
    (nvidia_ml_py, one sample every ~2 ms from a thread: the timed region of a default run lasts ~16 ms, too short
    for `nvidia-smi -lms`); falls back to one `nvidia-smi` query if NVML cannot be loaded."""

    REASONS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}

    def __init__(self, index: int):
        self.index, self.sm, self.mask, self.max_mhz, self.thread, self.stop = index, [], 0, None, None, threading.Event()
        self.nvml = None
        try:
            import pynvml

            pynvml.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(visible.split(",")[index]) if visible and visible.split(",")[index].strip().isdigit() else index
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception:  # noqa: BLE001
            self.nvml = None

    def _sample(self):
        n = self.nvml
        try:
            self.sm.append(float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
            self.mask |= int(n.nvmlDeviceGetCurrentClocksEventReasons(self.handle))
        except Exception:  # noqa: BLE001
            pass

    def _run(self):
        while not self.stop.is_set():
            self._sample()
            self.stop.wait(0.002)

    def __enter__(self):
        if self.nvml is not None:
            self.thread = threading.Thread(target=self._run, daemon=True)
            self.thread.start()
        return self

    def __exit__(self, *exc):
        if self.thread is not None:
            self._sample()  # at least one sample taken while the last kernels drain
            self.stop.set()
            self.thread.join(timeout=2)

    def summary(self):
        if self.nvml is None:
            try:
                q = "clocks.sm,clocks.max.sm"
                r = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                   capture_output=True, text=True, timeout=20).stdout.split(",")
                return {"sm_mhz": float(r[0]), "sm_max_mhz": float(r[1]), "reasons": [], "samples": 1, "source": "nvidia-smi after the timed region"}
            except Exception:  # noqa: BLE001
                return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        reasons = sorted(k for k, bit in self.REASONS.items() if self.mask & bit)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_mhz, "reasons": reasons,
                "samples": len(self.sm), "source": "NVML, every ~2 ms during the timed region"}


def oracle_logp_grad(u, y, params, n_chains, n_threads=0):
    """The CPU leg (cpu_baseline / --impl reference): the oracle port of the same evaluation.
    Forward-mode Dual<double> logp+grad with OpenMP over instances (oracle/hgf_oracle.cpp)."""
    from oracle import network as onet

    o = (onet.Network().add_nodes(kind="binary-state").add_nodes(value_children=0).add_nodes(volatility_children=1)
         .create_belief_propagation_fn())
    slots = [o.slot(1, "tonic_volatility"), o.slot(2, "tonic_volatility"), o.slot(-1, "response_parameter", 0)]
    t0 = time.perf_counter()
    s, g = o.batch_grad(u.astype(np.float64), slots, params.astype(np.float64), "binary_softmax_inverse_temperature",
                        resp_inputs=y.astype(np.float64), group_size=n_chains, n_threads=n_threads)
    return time.perf_counter() - t0, s, g


def cpu_baseline_leg(budget_s: float = 12.0):
    """Bounded sample of the headline workload on the host cores (oracle port, all threads)."""
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    n_part = 8
    u, y, params = synth_config4(0, n_participants=n_part)
    dt, _, _ = oracle_logp_grad(u, y, params, N_CHAINS, n_threads=cores)  # probe (also warms the library)
    per_part = max(dt / n_part, 1e-6)
    n_part = int(min(N_PARTICIPANTS, max(8, budget_s / per_part)))
    u, y, params = synth_config4(0, n_participants=n_part)
    dt, _, _ = oracle_logp_grad(u, y, params, N_CHAINS, n_threads=cores)
    n_inst = n_part * N_CHAINS
    return {"value": n_inst * N_TRIALS * N_NODES / dt, "unit": "node-updates/s", "cores": cores, "kind": "port",
            "sample": f"{n_part} participants x {N_CHAINS} chains x {N_TRIALS} trials logp+grad (Dual<double>, OpenMP), {dt:.2f} s",
            "logp_grad_evals_per_s": n_inst / dt}


def run_reference(args):
    """--impl reference: the reference's algorithm on the host cores.  pyhgf's own implementation is
    Python/JAX (+ Rust); neither jax/jaxlib nor cargo exist in this image (no wheel in /opt/wheelhouse),
    so the timed arm is the oracle port (oracle/hgf_oracle.cpp) with all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # torchrun exports OMP_NUM_THREADS=1 to its workers; this arm is the CPU implementation with ALL host threads
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    n_part = 8
    u, y, params = synth_config4(0, n_participants=n_part)
    dt, _, _ = oracle_logp_grad(u, y, params, N_CHAINS, n_threads=cores)
    budget = 60.0 / max(1, args.steps + args.warmup)  # whole run within a few minutes
    n_part = int(min(N_PARTICIPANTS, max(8, min(budget, 8.0) / max(dt / n_part, 1e-6))))
    u, y, params = synth_config4(0, n_participants=n_part)
    for _ in range(args.warmup):
        oracle_logp_grad(u, y, params, N_CHAINS, n_threads=cores)
    times = [oracle_logp_grad(u, y, params, N_CHAINS, n_threads=cores)[0] for _ in range(args.steps)]
    total = float(sum(times))
    n_inst = n_part * N_CHAINS
    value = n_inst * N_TRIALS * N_NODES * args.steps / total
    sample = f"{n_part} participants x {N_CHAINS} chains x {N_TRIALS} trials per step"
    line = {
        "impl": "reference", "metric": "node_updates_per_s", "value": value, "unit": "node-updates/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "config4: 3-level binary HGF logp+grad, participants x 64 chains x 320 trials (bounded sample)",
                   "sample": sample},
        "cpu_baseline": {"value": value, "unit": "node-updates/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "node-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "logp_grad_evals_per_s": n_inst * args.steps / total, "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def e2e_leg(net, plan, u, y, params, steps, warmup, device_index):
    """The same evaluation through the C ABI with HOST (pinned) buffers: hgf_logp_grad_host copies the
    inputs in, runs, copies logp / gradient / per-chain sums out and synchronises -- all inside the timing."""
    import torch

    from pyhgf_b200 import _lib

    L = _lib.lib()
    n = params.shape[1]
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()  # noqa: E731
    h_u, h_y, h_p = pin(u), pin(y), pin(params)
    h_attr = pin(net._flatten(plan))
    h_group = pin((np.arange(n) % N_CHAINS).astype(np.int32))
    h_logp = torch.empty(n, dtype=torch.float32).pin_memory()
    h_grad = torch.empty((3, n), dtype=torch.float32).pin_memory()
    h_red = torch.zeros(N_CHAINS, dtype=torch.float32).pin_memory()
    slots = (C.c_int32 * 3)(plan.slot(1, "tonic_volatility"), plan.slot(2, "tonic_volatility"), plan.slot(-1, "response_parameter", 0))
    call = _lib.Call()
    call.dtype, call.n_instances, call.n_steps = 32, n, u.shape[1]
    call.input_group_size = call.response_group_size = N_CHAINS
    call.input_data, call.response_inputs = h_u.data_ptr(), h_y.data_ptr()
    call.attributes = h_attr.data_ptr()
    call.n_params, call.param_slots, call.params = 3, C.cast(slots, C.POINTER(C.c_int32)), h_p.data_ptr()
    call.response_fn = _lib.RESPONSES["binary_softmax_inverse_temperature"]
    call.surprise_sum, call.gradient, call.reduced, call.group = h_logp.data_ptr(), h_grad.data_ptr(), h_red.data_ptr(), h_group.data_ptr()
    call.out_scale = -1.0
    call.n_groups = N_CHAINS
    h2d = sum(t.numel() * t.element_size() for t in (h_u, h_y, h_p, h_attr, h_group, h_red))
    d2h = sum(t.numel() * t.element_size() for t in (h_logp, h_grad, h_red))
    for _ in range(warmup):
        h_red.zero_()
        _lib.check(L.hgf_logp_grad_host(plan.handle, C.byref(call), device_index))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        h_red.zero_()
        _lib.check(L.hgf_logp_grad_host(plan.handle, C.byref(call), device_index))
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return dt, h2d, d2h, h_logp.numpy().copy()


def run_gpu(args):
    import torch
    import torch.distributed as dist

    from pyhgf_b200 import _lib, response
    from pyhgf_b200.network import Network

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # stdout carries ONE JSON line: NCCL prints its version banner there at any NCCL_DEBUG level >= VERSION, so its
        # log goes to stderr (and is off unless HGF_NCCL_DEBUG asks for a level)
        os.environ["NCCL_DEBUG_FILE"] = "/dev/stderr"
        if os.environ.get("HGF_NCCL_DEBUG"):
            os.environ["NCCL_DEBUG"] = os.environ["HGF_NCCL_DEBUG"]
        else:
            os.environ.pop("NCCL_DEBUG", None)
        dist.init_process_group("nccl", device_id=dev)
    _lib.lib()

    u, y, params = synth_config4(rank)
    n = params.shape[1]
    net = Network().add_nodes(kind="binary-state").add_nodes(value_children=0).add_nodes(volatility_children=1)
    net.create_belief_propagation_fn()
    d_u, d_y = torch.from_numpy(u).to(dev), torch.from_numpy(y).to(dev)
    d_p = torch.from_numpy(params).to(dev)
    group = (torch.arange(n, device=dev, dtype=torch.int32) % N_CHAINS).contiguous()
    keys = [(1, "tonic_volatility"), (2, "tonic_volatility"), ("response_parameter", 0)]
    # N > 1: the 64 per-chain sums are all-reduced INSIDE the kernel over NVLink peer memory (FusedReduce:
    # multimem.red through the NVSwitch multicast mapping); HGF_BENCH_COLLECTIVE=nccl (or no symmetric memory on
    # the box) falls back to one ncclAllReduce after the kernel
    fused, collective = None, "none"
    if world > 1:
        collective = "ncclAllReduce of 64 per-chain sums after the kernel"
        if os.environ.get("HGF_BENCH_COLLECTIVE", "fused") == "fused":
            try:
                from pyhgf_b200.distributed import FusedReduce

                fused = FusedReduce(N_CHAINS, 32, dev)
                collective = ("in-kernel all-reduce of 64 per-chain sums over NVLink peer memory (" +
                              ("multimem.red, NVSwitch multicast" if fused.multicast_ptr else "per-peer red") + "), no NCCL call")
            except Exception as exc:  # noqa: BLE001 - symmetric memory unavailable: keep NCCL
                collective += f" (fused path unavailable: {type(exc).__name__})"

    def evaluate(fr):
        return net.batch(d_u, {k: d_p[i] for i, k in enumerate(keys)}, response_function=response.binary_softmax_inverse_temperature,
                         response_function_inputs=d_y, input_group_size=N_CHAINS, response_group_size=N_CHAINS, gradient=True,
                         group=group, n_groups=N_CHAINS, out_scale=-1.0, fused_reduce=fr)

    res = evaluate(fused)
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def step():
        res.relaunch()
        if world > 1 and fused is None:
            dist.all_reduce(res.reduced)  # the 64 per-chain sums: the path's only exchange

    for _ in range(max(args.warmup, 0)):
        flush.zero_()
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    mids = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    launches0 = _lib.launch_count()
    with ClockSampler(local) as clocks:
        wall0 = time.perf_counter()
        for i in range(args.steps):
            flush.zero_()  # L2 flush between timed iterations (outside the event pair)
            starts[i].record()
            res.relaunch()
            mids[i].record()
            if world > 1 and fused is None:
                dist.all_reduce(res.reduced)
            ends[i].record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        wall = time.perf_counter() - wall0
    launches = _lib.launch_count() - launches0
    step_ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))
    kern_ms = sum(s.elapsed_time(m) for s, m in zip(starts, mids)) / args.steps
    t = torch.tensor([step_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    step_ms = float(t.item())

    collective_err = None
    if world > 1:  # the reduced sums against a plain NCCL all-reduce of an un-fused launch (outside the timing)
        got = res.reduced.clone()
        ref = evaluate(None)
        want = ref.reduced.clone()
        dist.all_reduce(want)
        torch.cuda.synchronize()
        collective_err = float(((got - want).abs() / want.abs().clamp_min(1e-30)).max())

    e2e_dt, h2d, d2h, e2e_logp = e2e_leg(net, net.plan, u, y, params, args.steps, max(args.warmup, 3), local)
    same = np.allclose(e2e_logp, res.surprise.cpu().numpy(), rtol=1e-6, equal_nan=True)
    t2 = torch.tensor([e2e_dt], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t2, op=dist.ReduceOp.MAX)
    e2e_dt = float(t2.item())

    if rank == 0:
        units = world * n * N_TRIALS * N_NODES  # node-updates per step, all ranks
        value = units * args.steps / (step_ms * 1e-3)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        # algorithmic bytes of one launch: params in, logp + gradient out, each participant's series once
        alg_bytes = n * (3 * 4 + 4 * 4) + u.nbytes + y.nbytes + n * 4
        achieved = alg_bytes / (kern_ms * 1e-3) / 1e9
        ck = clocks.summary()
        sm_max = ck["sm_max_mhz"] or float(peaks.get("sm_max_mhz", 1965.0))
        mufu_peak = N_SM * MUFU_PEAK_PER_CLK_PER_SM * sm_max * 1e6
        mufu = n * N_TRIALS * MUFU_PER_STEP / (kern_ms * 1e-3)
        line = {
            "metric": "node_updates_per_s", "value": value, "unit": "node-updates/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": step_ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "config4: 3-level binary HGF logp+grad (omega_2, omega_3, inverse temperature), "
                                   f"{N_PARTICIPANTS} participants x {N_CHAINS} chains x {N_TRIALS} trials per GPU",
                       "instances_per_gpu": n, "trials": N_TRIALS, "nodes": N_NODES, "kernel": res.kernel_name,
                       "l2": "flushed between timed iterations (256 MiB write)", "collective": collective, "collective_max_rel_err_vs_nccl": collective_err},
            "logp_grad_evals_per_s": world * n * args.steps / (step_ms * 1e-3),
            "wall_ms_per_step": 1e3 * wall / args.steps,
            "kernel_ms": kern_ms,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                         "traffic": NCU_DRAM_BYTES_PER_LAUNCH, "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback",
                         "note": "the kernel is FP32/SFU-bound, not HBM-bound: see roofline_compute"},
            "roofline_compute": {"bound": "mufu", "achieved": mufu, "peak": mufu_peak, "unit": "op/s", "frac": mufu / mufu_peak,
                                 "note": f"algorithmic forward count {MUFU_PER_STEP} MUFU-class ops per instance-step (SURVEY 8d); "
                                         "gradient work not counted; peak = 148 SM x 16/clk x max SM clock"},
            "e2e": {"value": units * args.steps / e2e_dt, "unit": "node-updates/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": 1e3 * e2e_dt / args.steps, "api": "hgf_logp_grad_host (C ABI, pinned host buffers)",
                    "matches_device_path": bool(same)},
            "gpu_launches": int(launches),
            "clocks": ck,
        }
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline_leg()
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
