"""Benchmark of the batched belief-propagation path (BASELINE.json's metric: node-updates/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config {2,3,4,5}] [--impl reference]

--config picks the BASELINE.json configuration (default 4, the one the target is quoted on):
  2  4-node continuous HGF on the USD-CHF series (614 obs), forward logp, 65 536 samples SHARDED over the GPUs (strong)
  3  3-level binary HGF, forward logp over the 1 024 x 1 024 (omega_2, omega_3) grid x 320 trials, grid rows SHARDED (strong)
  4  3-level binary HGF, logp + gradient wrt (omega_2, omega_3, inverse temperature), 1 024 participants x 64 chains x
     320 trials PER GPU (weak); the 64 per-chain sums are all-reduced over NVLink by the kernel itself
  5  16-node generalised HGF, 32 768 instances PER GPU x 2 000 steps, all 80 belief leaves recorded (21 GB per GPU; weak)
A "step" is `evals_per_step` consecutive evaluations of the whole batch on resident inputs (a sampler's iteration is
several logp+grad evaluations; it also makes the timed region of the stock K = 20 run longer than 0.2 s); the L2 is
flushed between steps.  One JSON line on stdout (rank 0); see the contract in the task description.
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
N_SM, MUFU_PER_CLK_PER_SM = 148, 16
# algorithmic SFU-class operations (exp, log, division) of ONE instance-step, counted from the reference's
# formulas.  3-level binary: 19 exp + 8 log + 28 div = 55; the inverse-temperature response adds 2 log2 + 2 exp2 + 1 div;
# the forward-mode gradient of (omega_2, omega_3, temperature) needs 15 more that no algebra removes (one reciprocal per
# logarithm's derivative (8), per logaddexp weight (3) and per pow derivative (2), exp + reciprocal of the Lambert-W JVP).
# 4-node continuous: 37 exp + 16 log + 50 div.
MUFU = {"binary3_fwd": 55, "binary3_logp_grad": 55 + 5 + 15, "continuous3_fwd": 103}


# ----------------------------------------------------------------------------------------------------------------------
# synthetic inputs
def synth_config4(rank: int, n_participants: int = N_PARTICIPANTS, n_chains: int = N_CHAINS, n_trials: int = N_TRIALS):
    """Synthetic participants (C4): block-wise Bernoulli contingencies switching every 40 trials, decisions drawn from
    block-consistent noisy choices, chain parameters = truth + N(0, 0.5^2)."""
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


def ar1_series(n_steps, n_series=4, phi=0.98, sigma=0.1, seed=0):
    rng = np.random.default_rng(seed)
    x = np.zeros((n_steps, n_series))
    e = rng.normal(0.0, sigma, size=(n_steps, n_series))
    for t in range(1, n_steps):
        x[t] = phi * x[t - 1] + e[t]
    return x.astype(np.float32)


def shard(n_items, world, rank):
    base, extra = divmod(n_items, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


# networks, built through the Network surface of whichever implementation `cls` is (the CUDA front-end or the oracle)
def net_binary3(cls):
    return cls().add_nodes(kind="binary-state").add_nodes(value_children=0).add_nodes(volatility_children=1)


def net_continuous3(cls, x0=1.04):
    return (cls().add_nodes(precision=1e4)
            .add_nodes(value_children=([0], [1.0]), node_parameters={"mean": x0, "precision": 1e4, "tonic_volatility": -13.0, "tonic_drift": 0.0})
            .add_nodes(volatility_children=([1], [1.0]), node_parameters={"mean": 1.0, "precision": 1e1, "tonic_volatility": -2.0, "tonic_drift": 0.0})
            .add_nodes(volatility_children=([2], [1.0]), node_parameters={"mean": 1.0, "precision": 1e1, "tonic_volatility": -2.0, "tonic_drift": 0.0}))


def net_ghgf16(cls):
    net = cls()
    net.add_nodes(n_nodes=4, precision=100.0)
    for i in range(4):
        net.add_nodes(value_children=i, tonic_volatility=-4.0)
    for i in range(4):
        net.add_nodes(volatility_children=4 + i, tonic_volatility=-4.0)
    net.add_nodes(kind="volatile-state", value_children=[4, 5], autoconnection_strength=1.0)
    net.add_nodes(kind="volatile-state", value_children=[6, 7], autoconnection_strength=1.0)
    net.add_nodes(kind="volatile-state", value_children=12, autoconnection_strength=1.0)
    net.add_nodes(kind="volatile-state", value_children=13, autoconnection_strength=1.0)
    return net


BELIEFS = ("mean", "expected_mean", "precision", "expected_precision")
VOL_BELIEFS = ("mean_vol", "expected_mean_vol", "precision_vol", "expected_precision_vol")


class Workload:
    """One BASELINE configuration: its network, this rank's shard of the synthetic batch (host arrays), the keys of the
    per-instance parameters, and what a launch computes."""

    def __init__(self, config: int, rank: int, world: int, scale: float = 1.0):
        from pyhgf_b200 import load_data

        self.config, self.rank, self.world = config, rank, world
        self.gradient, self.record, self.response, self.resp_inputs, self.group_size = False, [], None, None, 0
        self.n_chains = 0
        if config == 4:
            self.build, self.n_nodes, self.scaling, self.evals_per_step = net_binary3, 3, "weak", 16
            n_part = max(1, int(round(N_PARTICIPANTS * scale)))
            self.series, self.resp_inputs, self.params = synth_config4(rank, n_participants=n_part)
            self.keys = [(1, "tonic_volatility"), (2, "tonic_volatility"), ("response_parameter", 0)]
            self.gradient, self.response, self.group_size, self.n_chains = True, "binary_softmax_inverse_temperature", N_CHAINS, N_CHAINS
            self.n_global = world * self.params.shape[1]
            self.workload = (f"config4: 3-level binary HGF logp+grad (omega_2, omega_3, inverse temperature), {n_part} participants x "
                             f"{N_CHAINS} chains x {N_TRIALS} trials per GPU")
            self.mufu_per_step, self.kernel_kind = MUFU["binary3_logp_grad"], "logp+grad"
        elif config == 3:
            self.build, self.n_nodes, self.scaling, self.evals_per_step = net_binary3, 3, "strong", 4
            side = max(8, int(round(1024 * scale ** 0.5)))
            lo, hi = shard(side, world, rank)
            grid = np.linspace(-8, 0, side, dtype=np.float32)
            self.series = load_data("binary")[0].astype(np.float32)
            self.params = np.stack([np.repeat(grid[lo:hi], side), np.tile(grid, hi - lo)])
            self.keys = [(1, "tonic_volatility"), (2, "tonic_volatility")]
            self.response = "first_level_binary_surprise"
            self.n_global = side * side
            self.workload = f"config3: 3-level binary HGF forward logp over the {side} x {side} (omega_2, omega_3) grid x 320 trials, grid rows sharded"
            self.mufu_per_step, self.kernel_kind = MUFU["binary3_fwd"], "forward logp"
        elif config == 2:
            self.build, self.n_nodes, self.scaling, self.evals_per_step = net_continuous3, 4, "strong", 16
            n = max(64, int(round(65536 * scale)))
            rng = np.random.default_rng(0)
            cols = [np.clip(rng.normal(m, 1.0, n), m - 3, m + 3).astype(np.float32) for m in (-13.0, -2.0, -2.0)]
            lo, hi = shard(n, world, rank)
            self.series = load_data("continuous").astype(np.float32)
            self.params = np.stack([c[lo:hi] for c in cols])
            self.keys = [(1, "tonic_volatility"), (2, "tonic_volatility"), (3, "tonic_volatility")]
            self.response = "first_level_gaussian_surprise"
            self.n_global = n
            self.workload = f"config2: 4-node continuous HGF forward logp, {n} tonic-volatility samples x 614 observations (USD-CHF), samples sharded"
            self.mufu_per_step, self.kernel_kind = MUFU["continuous3_fwd"], "forward logp"
        elif config == 5:
            self.build, self.n_nodes, self.scaling, self.evals_per_step = net_ghgf16, 16, "weak", 2
            n, T = max(64, int(round(32768 * scale))), int(os.environ.get("HGF_BENCH_C5_STEPS", "2000"))  # (the override: short ncu captures)
            rng = np.random.default_rng(1 + rank)
            self.series = ar1_series(T)
            self.keys = [(i, "tonic_volatility") for i in range(4, 12)] + [(i, "tonic_volatility_vol") for i in range(12, 16)]
            self.params = rng.normal(-4.0, 1.0, size=(len(self.keys), n)).astype(np.float32)
            self.record = [(i, k) for i in range(16) for k in BELIEFS + (VOL_BELIEFS if i >= 12 else ())]
            self.n_global = world * n
            self.workload = f"config5: 16-node generalised HGF, {n} instances per GPU x {T} steps, all {len(self.record)} belief leaves recorded"
            self.mufu_per_step, self.kernel_kind = None, "filter + trajectory write-back"
        else:
            raise SystemExit("--config must be 2, 3, 4 or 5")
        self.n_local = self.params.shape[1]
        self.n_trials = self.series.shape[-2] if config == 5 else self.series.shape[-1]
        self.units_per_eval = self.n_global * self.n_trials * self.n_nodes  # node-updates of one evaluation, all ranks

    # bytes one launch must move: parameters in, results out, each series once; + the trajectories
    def algorithmic_bytes(self):
        P = len(self.keys)
        b = self.n_local * 4 * (P + 1 + (P if self.gradient else 0)) + self.series.nbytes
        if self.resp_inputs is not None:
            b += self.resp_inputs.nbytes
        if self.n_chains:
            b += self.n_local * 4
        return b + len(self.record) * self.n_trials * self.n_local * 4

    def launch(self, dev, fused=None):
        import torch

        from pyhgf_b200 import response
        from pyhgf_b200.network import Network

        net = self.build(Network)
        net.create_belief_propagation_fn()
        d_p = torch.from_numpy(self.params).to(dev)
        kw = {}
        if self.response:
            kw["response_function"] = getattr(response, self.response)
        if self.resp_inputs is not None:
            kw["response_function_inputs"] = torch.from_numpy(self.resp_inputs).to(dev)
        if self.group_size:
            kw.update(input_group_size=self.group_size, response_group_size=self.group_size)
        if self.n_chains:
            kw.update(group=(torch.arange(self.n_local, device=dev, dtype=torch.int32) % self.n_chains).contiguous(), n_groups=self.n_chains,
                      out_scale=-1.0, fused_reduce=fused)
        res = net.batch(torch.from_numpy(self.series).to(dev), {k: d_p[i] for i, k in enumerate(self.keys)}, gradient=self.gradient,
                        record=self.record, **kw)
        return net, res


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md) through NVML (nvidia_ml_py,
    one sample every ~5 ms from a thread); falls back to one `nvidia-smi` query if NVML cannot be loaded."""

    REASONS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}

    def __init__(self, index: int, period: float = 0.005):
        self.index, self.sm, self.mask, self.max_mhz, self.thread, self.stop, self.period = index, [], 0, None, None, threading.Event(), period
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
            self.stop.wait(self.period)

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
                "samples": len(self.sm), "source": "NVML, every ~5 ms during the timed region"}


# ----------------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the same evaluation (the reference itself -- Python/JAX + a Rust crate -- cannot run on
# this image: no jax / jaxlib wheel, no cargo)
def host_cores():
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


CPU_VARIANTS = {
    "simd": "float32, instance-vectorised (AVX-512 / AVX2 vectors of instances, libmvec), OpenMP over vectors; logp+grad by the generated reverse sweep "
            "(oracle/hgf_simd.cpp) -- what jit(vmap(grad)) on XLA-CPU does",
    "f32": "float32 scalar per instance (Dual<float> forward-mode tangents for logp+grad), OpenMP over instances (oracle/hgf_oracle.cpp)",
    "f64": "float64 scalar per instance (Dual<double>), OpenMP over instances",
}


def best_cpu_variant(config: int) -> str:
    """The fastest CPU arm that covers the configuration: the instance-vectorised one covers the networks of all four."""
    return "simd"


def oracle_eval(w: Workload, variant: str, n_threads: int):
    """One evaluation of workload `w`'s LOCAL shard by a CPU restatement of the reference's algorithm; returns seconds.
    variant: see CPU_VARIANTS ("simd" is the arm the headline ratio is taken against where it exists)."""
    from oracle import network as onet

    o = w.build(onet.Network if variant != "f64" else (lambda: onet.Network(dtype="float64"))).create_belief_propagation_fn()
    slots = []
    for k in w.keys:
        slots.append(o.slot(-1, "response_parameter", k[1]) if k[0] == "response_parameter" else o.slot(k[0], k[1]))
    if variant == "simd":
        from oracle import simd

        simd.lib()  # (built where it runs: not part of the timed region)
        t0 = time.perf_counter()
        simd.batch(o, w.series, slots, w.params, w.response, resp_inputs=w.resp_inputs, group_size=w.group_size, gradient=w.gradient,
                   n_threads=n_threads)
        return time.perf_counter() - t0
    series = w.series.astype(np.float64)
    params = w.params.astype(np.float64)
    ri = None if w.resp_inputs is None else w.resp_inputs.astype(np.float64)
    t0 = time.perf_counter()
    if w.gradient:
        o.batch_grad(series, slots, params, w.response, resp_inputs=ri, group_size=w.group_size, n_threads=n_threads, dtype=32 if variant == "f32" else 64)
    else:
        # config 5 has no response function: the filter alone (no trajectory is recorded by the oracle's batch path)
        o.batch_surprise(series, slots, params, w.response or "first_level_gaussian_surprise", resp_inputs=ri, group_size=w.group_size, n_threads=n_threads)
    return time.perf_counter() - t0


def cpu_sample(config: int, budget_s: float, variant: str = "f32"):
    """A bounded sample of the configuration on the host cores: scale the batch so that one evaluation takes ~budget_s."""
    cores = host_cores()
    probe = Workload(config, 0, 1, scale=1.0 / (16 if variant == "simd" else 128))  # (large enough to amortise the thread start-up)
    dt = oracle_eval(probe, variant, cores)  # also warms the library
    dt = oracle_eval(probe, variant, cores)
    per_unit = dt / max(probe.units_per_eval, 1)
    full = Workload(config, 0, 1)
    scale = min(1.0, budget_s / max(per_unit * full.units_per_eval, 1e-9))
    w = full if scale >= 1.0 else Workload(config, 0, 1, scale=scale)
    return w, cores


def cpu_baseline_leg(config: int, budget_s: float = 12.0):
    variant = best_cpu_variant(config)
    w, cores = cpu_sample(config, budget_s if variant != "simd" else 3.0, variant)
    oracle_eval(w, variant, cores)
    dt = min(oracle_eval(w, variant, cores) for _ in range(2 if variant == "simd" else 1))
    out = {"value": w.units_per_eval / dt, "unit": "node-updates/s", "cores": cores, "kind": "port", "variant": CPU_VARIANTS[variant],
           "sample": f"{w.workload} -- one evaluation, {dt:.2f} s", "cpu": cpu_model()}
    if variant == "simd":
        ws, _ = cpu_sample(config, 4.0, "f32")
        out["scalar_value"] = ws.units_per_eval / oracle_eval(ws, "f32", cores)
        out["scalar_variant"] = CPU_VARIANTS["f32"]
    return out


def run_reference(args):
    """--impl reference: the reference's algorithm on the host cores.  pyhgf's own implementation is Python/JAX (+ Rust);
    neither jax/jaxlib nor cargo exist in this image, so the timed arm is a C++ restatement of its algorithm under oracle/, on
    all host threads: the instance-vectorised float32 arm (oracle/hgf_simd.cpp -- float32 SIMD across the batch and reverse-mode
    gradients, as XLA-CPU would run jit(vmap(grad))) where it covers the configuration, else the scalar port; the scalar
    float32 / float64 figures are reported beside it."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    variant = best_cpu_variant(args.config)
    budget = min(8.0, 90.0 / max(1, args.steps + args.warmup))  # whole run within a few minutes
    w, cores = cpu_sample(args.config, budget, variant)
    for _ in range(args.warmup):
        oracle_eval(w, variant, cores)
    times = [oracle_eval(w, variant, cores) for _ in range(args.steps)]
    total = float(sum(times))
    value = w.units_per_eval * args.steps / total
    ws, _ = cpu_sample(args.config, 4.0, "f32")
    t32 = oracle_eval(ws, "f32", cores)
    t64 = oracle_eval(ws, "f64", cores)
    full = Workload(args.config, 0, 1)
    same = w.units_per_eval == full.units_per_eval
    sample = f"{w.workload} -- one evaluation per step" + ("" if same else " (bounded sample of the configuration)")
    line = {
        "impl": "reference", "metric": "node_updates_per_s", "value": value, "unit": "node-updates/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True,
        "scaling": w.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": w.workload, "sample": sample, "same_config": bool(same)},
        "cpu_baseline": {"value": value, "unit": "node-updates/s", "cores": cores, "kind": "port", "variant": CPU_VARIANTS[variant],
                         "sample": sample, "cpu": cpu_model(), "value_scalar_float32": ws.units_per_eval / t32,
                         "value_scalar_float64": ws.units_per_eval / t64,
                         "note": "none of these is the reference itself (JAX / Rust are not installable here): all time C++ restatements of its algorithm"},
        "e2e": {"value": value, "unit": "node-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------------------------------
def e2e_leg(w: Workload, steps, warmup, device_index, fused):
    """The same step through the C ABI with HOST (pinned) buffers: every evaluation hgf_logp_grad_host / hgf_filter_host
    copies the inputs in, runs, copies logp / gradient / trajectories / per-chain sums out and synchronises -- all inside the
    timing.  On several GPUs the call also carries the exchange: the per-chain sums it returns are the ALL-REDUCED ones."""
    import torch

    from pyhgf_b200 import _lib

    from pyhgf_b200.network import Network

    L = _lib.lib()
    net = w.build(Network)
    net.create_belief_propagation_fn()
    plan = net._get_plan(w.series.shape[1] if w.series.ndim == 2 and w.config == 5 else 1)
    n, T, P = w.n_local, w.n_trials, len(w.keys)
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()  # noqa: E731
    h_u, h_p, h_attr = pin(w.series), pin(w.params), pin(net._flatten(plan))
    ins, outs = [h_u, h_p, h_attr], []
    slots = (C.c_int32 * P)(*[net._resolve_slot(plan, k) for k in w.keys])
    call = _lib.Call()
    call.dtype, call.n_instances, call.n_steps = 32, n, T
    call.input_group_size = call.response_group_size = w.group_size
    call.input_data, call.attributes = h_u.data_ptr(), h_attr.data_ptr()
    call.n_params, call.param_slots, call.params = P, C.cast(slots, C.POINTER(C.c_int32)), h_p.data_ptr()
    call.response_fn = _lib.RESPONSES[w.response]
    if w.resp_inputs is not None:
        h_y = pin(w.resp_inputs)
        ins.append(h_y)
        call.response_inputs = h_y.data_ptr()
    call.out_scale = -1.0 if w.n_chains else 1.0
    h_logp = None
    if w.response:
        h_logp = torch.empty(n, dtype=torch.float32).pin_memory()
        outs.append(h_logp)
        call.surprise_sum = h_logp.data_ptr()
    if w.gradient:
        h_grad = torch.empty((P, n), dtype=torch.float32).pin_memory()
        outs.append(h_grad)
        call.gradient = h_grad.data_ptr()
    rec = None
    if w.record:
        rec = (C.c_int32 * len(w.record))(*[net._resolve_slot(plan, k) for k in w.record])
        h_traj = torch.empty((len(w.record), T, n), dtype=torch.float32).pin_memory()
        outs.append(h_traj)
        call.n_records, call.record_slots, call.trajectories = len(w.record), C.cast(rec, C.POINTER(C.c_int32)), h_traj.data_ptr()
    h_red = None
    if w.n_chains:
        h_group = pin((np.arange(n) % w.n_chains).astype(np.int32))
        h_red = torch.zeros(w.n_chains, dtype=torch.float32).pin_memory()
        ins += [h_group] + ([] if fused is not None else [h_red])
        outs.append(h_red)
        call.group, call.n_groups = h_group.data_ptr(), w.n_chains
    fn = L.hgf_logp_grad_host if w.gradient else L.hgf_filter_host
    h2d = sum(t.numel() * t.element_size() for t in ins)
    d2h = sum(t.numel() * t.element_size() for t in outs)

    def evaluate():
        if h_red is not None:
            if fused is not None:
                fused.attach(call)  # world / rank / epoch / peer workspaces; the host `reduced` then receives the all-reduced sums
            else:
                h_red.zero_()
            call.reduced = h_red.data_ptr()
        _lib.check(fn(plan.handle, C.byref(call), device_index))

    R = w.evals_per_step
    for _ in range(warmup):
        evaluate()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps * R):
        evaluate()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return dt, h2d * R, d2h * R, (None if h_logp is None else h_logp.numpy().copy()), (None if h_red is None else h_red.numpy().copy())


def measured_traffic(config: int):
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of this configuration's kernel, from the committed
    `ncu --set full` capture (profiles/r2_ncu_counters.json, written by tools/ncu_counters.py); None if there is none."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "r2_ncu_counters.json")))
        e = d[f"config{config}"]
        return float(e["dram__bytes_read.sum"]) + float(e["dram__bytes_write.sum"]), "profiles/r2_ncu_counters.json (one ncu capture of this launch; not measured in this run)"
    except (OSError, KeyError, ValueError, TypeError):
        return None, None


def run_gpu(args):
    import torch
    import torch.distributed as dist

    from pyhgf_b200 import _lib

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # stdout carries ONE JSON line: NCCL's log (whatever level the caller set) goes to stderr
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    _lib.lib()

    w = Workload(args.config, rank, world)
    R = args.evals_per_step or w.evals_per_step
    w.evals_per_step = R
    # N > 1, config 4: the 64 per-chain sums are all-reduced INSIDE the kernel over NVLink peer memory (FusedReduce:
    # multimem.red through the NVSwitch multicast mapping; the kernel deposits and exits, the consumer waits);
    # HGF_BENCH_COLLECTIVE=nccl (or no symmetric memory on the box) falls back to one ncclAllReduce after each kernel.
    # Configs 2, 3, 5 have no exchange at all (independent instances, results stay sharded).
    fused, collective = None, "none"
    if world > 1 and w.n_chains:
        collective = "ncclAllReduce of 64 per-chain sums after each kernel"
        if os.environ.get("HGF_BENCH_COLLECTIVE", "fused") == "fused":
            try:
                from pyhgf_b200.distributed import FusedReduce

                fused = FusedReduce(w.n_chains, 32, dev)
                collective = ("in-kernel all-reduce of 64 per-chain sums over NVLink peer memory (" +
                              ("multimem.red, NVSwitch multicast" if fused.multicast_ptr else "per-peer red") +
                              "): each launch deposits and exits, the step's end waits for every rank's deposits; no NCCL call")
            except Exception as exc:  # noqa: BLE001 - symmetric memory unavailable: keep NCCL
                collective += f" (fused path unavailable: {type(exc).__name__})"

    net, res = w.launch(dev, fused)
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def step():
        for _ in range(R):
            res.relaunch(consume=False)
            if world > 1 and w.n_chains and fused is None:
                dist.all_reduce(res.reduced)  # the 64 per-chain sums: the path's only exchange
        if fused is not None:
            res.reduced = fused.result()  # every rank's deposits of the step's last launch (hence of all of them) have arrived

    for _ in range(max(args.warmup, 0)):
        flush.zero_()
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    launches0 = _lib.launch_count()
    with ClockSampler(local) as clocks:
        wall0 = time.perf_counter()
        for i in range(args.steps):
            flush.zero_()  # L2 flush between timed steps (outside the event pair)
            starts[i].record()
            step()
            ends[i].record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        wall = time.perf_counter() - wall0
    launches = _lib.launch_count() - launches0
    per_step = [s.elapsed_time(e) for s, e in zip(starts, ends)]
    own_ms = float(sum(per_step))
    t = torch.tensor([own_ms], device=dev, dtype=torch.float64)
    all_ms = [t.clone() for _ in range(world)]
    if world > 1:
        dist.all_gather(all_ms, t)
    rank_ms = [float(x.item()) / args.steps for x in all_ms] if world > 1 else [own_ms / args.steps]
    step_ms = max(rank_ms) * args.steps  # timed on the device, MAX over ranks

    collective_err = None
    if world > 1 and w.n_chains:  # the reduced sums against a plain NCCL all-reduce of an un-fused launch (outside the timing)
        got = res.reduced.clone()
        _, ref = w.launch(dev, None)
        want = ref.reduced.clone()
        dist.all_reduce(want)
        torch.cuda.synchronize()
        collective_err = float(((got - want).abs() / want.abs().clamp_min(1e-30)).max())

    # config 5 returns 21 GB per evaluation: its end-to-end leg runs an eighth of the instances (2.6 GB of pinned host
    # memory per rank instead of 21) for at most 3 steps; the value is that batch's own throughput
    w_e2e, e2e_steps = (w, args.steps) if args.config != 5 else (Workload(5, rank, world, scale=1.0 / 8), max(1, min(args.steps, 3)))
    w_e2e.evals_per_step = R
    e2e_dt, h2d, d2h, e2e_logp, e2e_red = e2e_leg(w_e2e, e2e_steps, min(max(args.warmup, 1), 3), local, fused)
    same = None if (e2e_logp is None or w_e2e is not w) else bool(np.allclose(e2e_logp, res.surprise.cpu().numpy(), rtol=1e-6, equal_nan=True))
    if e2e_red is not None and world > 1 and fused is not None:
        torch.cuda.synchronize()
        same = bool(same and np.allclose(e2e_red, res.reduced.cpu().numpy(), rtol=1e-5))  # the host call returned the ALL-REDUCED sums
    t2 = torch.tensor([e2e_dt / e2e_steps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t2, op=dist.ReduceOp.MAX)
    e2e_step_s = float(t2.item())

    if rank == 0:
        units = w.units_per_eval * R  # node-updates of one step, all ranks
        value = units * args.steps / (step_ms * 1e-3)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        ck = clocks.summary()
        sm_max = ck["sm_max_mhz"] or float(peaks.get("sm_max_mhz", 1965.0))
        kern_ms = own_ms / (args.steps * R)  # rank 0's average launch (configs 2-4: one kernel per evaluation)
        alg_bytes = w.algorithmic_bytes()
        hbm = alg_bytes / (kern_ms * 1e-3) / 1e9
        traffic, traffic_src = measured_traffic(args.config)
        roof_hbm = {"bound": "hbm", "achieved": hbm, "peak": hbm_peak, "unit": "GB/s", "frac": hbm / hbm_peak, "traffic": traffic,
                    "traffic_source": traffic_src, "algorithmic_bytes_per_launch": alg_bytes,
                    "peak_source": "MEASURED_PEAKS.json (copy bandwidth)" if peaks else "fallback 6.65 TB/s"}
        if w.mufu_per_step:
            mufu_peak = N_SM * MUFU_PER_CLK_PER_SM * sm_max * 1e6
            mufu = w.n_local * w.n_trials * w.mufu_per_step / (kern_ms * 1e-3)
            roofline = {"bound": "mufu", "achieved": mufu / 1e9, "peak": mufu_peak / 1e9, "unit": "Gop/s", "frac": mufu / mufu_peak, "traffic": traffic,
                        "traffic_source": traffic_src,
                        "note": f"{w.kernel_kind}: {w.mufu_per_step} algorithmic SFU-class operations (exp / log / division, gradient work included) per "
                                "instance-step (SURVEY.md 8d, DESIGN.md 4) x instances x trials / launch time; peak = 148 SM x 16 MUFU lanes/clk x max SM "
                                "clock.  The launch moves a few MB: its HBM figure is in roofline_hbm"}
            extra = {"roofline_hbm": roof_hbm}
        else:
            roofline, extra = roof_hbm, {}
        line = {
            "metric": "node_updates_per_s", "value": value, "unit": "node-updates/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": step_ms / args.steps, "higher_is_better": True, "scaling": w.scaling,
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": w.workload, "baseline_config": args.config, "evals_per_step": R, "instances_per_gpu": w.n_local, "trials": w.n_trials,
                       "nodes": w.n_nodes, "kernel": res.kernel_name, "l2": "flushed between timed steps (256 MiB write)",
                       "collective": collective, "collective_max_rel_err_vs_nccl": collective_err},
            "ms_per_eval": step_ms / (args.steps * R),
            "evals_per_s": w.n_global * R * args.steps / (step_ms * 1e-3),
            "wall_ms_per_step": 1e3 * wall / args.steps,
            "kernel_ms": kern_ms,
            "per_rank_ms_per_step": {"min": min(rank_ms), "median": float(np.median(rank_ms)), "max": max(rank_ms), "all": rank_ms},
            "per_step_ms_rank0": {"min": min(per_step), "median": float(np.median(per_step)), "max": max(per_step)},
            "roofline": roofline, **extra,
            "e2e": {"value": w_e2e.units_per_eval * R / e2e_step_s, "unit": "node-updates/s", "instances_per_gpu": w_e2e.n_local, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": 1e3 * e2e_step_s, "steps": e2e_steps,
                    "api": ("hgf_logp_grad_host" if w.gradient else "hgf_filter_host") + " (C ABI, pinned host buffers" +
                           ("; the per-chain sums it returns are the ALL-REDUCED ones: every call includes the exchange)" if fused is not None
                            else "; no cross-GPU exchange in this leg: the HGF_BENCH_COLLECTIVE=nccl arm reduces on the device path only)" if world > 1 and w.n_chains
                            else ")"), "matches_device_path": same},
            "gpu_launches": int(launches),
            "clocks": ck,
        }
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline_leg(args.config)
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
    ap.add_argument("--config", type=int, default=4, choices=[2, 3, 4, 5], help="BASELINE.json configuration (default 4: the headline)")
    ap.add_argument("--evals-per-step", type=int, default=0, help="evaluations of the batch per timed step (0: the configuration's default)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
