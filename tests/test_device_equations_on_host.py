"""The DEVICE update equations (pyhgf_b200/csrc/hgf_math.cuh, hgf_ops.cuh, hgf_scan.cuh are __host__ __device__)
compiled for the host by tools/hostsim and run against the oracle -- the CUDA source itself is tied to the oracle on a
machine without a GPU:

  * with IEEE / libm primitives the kernel's time loop reproduces the float32 oracle's summed surprise BIT FOR BIT
    (same equations, same operation order), and its forward-mode tangents give the oracle's gradients;
  * with the SFU primitives emulated inside their documented error bounds and the SHIPPED math policy, every instance
    stays inside the parity bound the GPU tests assert (tests/parity.py); with the response's last-bit arithmetic
    switched off (round 1's policy) it does not.
"""

import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools", "hostsim"))

import nets  # noqa: E402
from oracle import conditioning  # noqa: E402
from oracle import network as onet  # noqa: E402
from parity import assert_conditioned  # noqa: E402


@pytest.fixture(scope="module")
def box(datasets):
    import run as hostsim

    rng = np.random.default_rng(7)
    n = 1024
    p = np.stack([rng.uniform(-6, -3, n), rng.uniform(-6, -3, n), np.exp(rng.uniform(-0.3, 0.8, n))])
    return hostsim, hostsim.Box("box", datasets["u"], datasets["y"], p, 0)


def test_ieee_build_of_the_device_equations_is_the_oracle(box):
    hostsim, b = box
    logp, grad = b.run(hostsim.build(["@"]))
    assert np.array_equal(logp.astype(np.float32), b.want.astype(np.float32))  # bit for bit
    for i in range(3):  # the kernel's closed-form JVPs against the oracle's operation-by-operation Dual<float>
        scale = np.abs(b.want_g[i]).max()
        rel = np.abs(grad[i] - b.want_g[i]) / (np.abs(b.want_g[i]) + 1e-4 * scale)
        assert np.median(rel) < 1e-6 and np.quantile(rel, 0.99) < 1e-3, (i, float(np.median(rel)), float(rel.max()))


def test_shipped_policy_with_emulated_sfu_stays_inside_the_parity_bound(box):
    hostsim, b = box
    o32 = nets.binary3(onet.Network).create_belief_propagation_fn()
    slots = [o32.slot(1, "tonic_volatility"), o32.slot(2, "tonic_volatility"), o32.slot(-1, "response_parameter", 0)]
    want, sens = conditioning.sensitivity(
        lambda q: o32.batch_surprise(b.u.astype(np.float64), slots, q, "binary_softmax_inverse_temperature", resp_inputs=b.y.astype(np.float64)),
        b.p.astype(np.float64), rows=[0, 1], n_draws=16)
    shipped, _ = b.run(hostsim.build([]))  # the defaults of hgf_math.cuh
    stats = assert_conditioned(shipped, want, sens, 1e-4, what="shipped policy, emulated SFU", max_relaxed=0.02)
    round1, _ = b.run(hostsim.build(["-DHGF_RESP_MODE=0"]))
    err = np.abs(round1 / want - 1)
    assert (err > conditioning.tolerance(sens, 1e-4, 4.0)).sum() > stats["n_over_rtol"] + 2  # what the policy fixes


def test_generated_reverse_sweep_is_the_oracles_gradient(box, datasets):
    """csrc/hgf_reverse.cuh + generated/reverse_*.h (pyhgf_b200/revgen.py) compiled for the host, float64: the summed
    surprise is the oracle's bit for bit and all eight gradients -- volatilities, inverse temperature, initial beliefs,
    tonic drift, autoconnection -- are the Dual<double> oracle's to 1e-10; same for the 4-node continuous HGF with a
    volatility coupling among the parameters."""
    import ctypes as C

    from pyhgf_b200.network import Network

    hostsim, _ = box
    lib = hostsim.build(["@"])
    fp = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    rng = np.random.default_rng(3)
    n = 48
    u, y, x = (np.ascontiguousarray(datasets[k], dtype=np.float64) for k in ("u", "y", "x"))
    x = np.ascontiguousarray(x[:300])
    cases = [
        (nets.binary3, "sim_binary3_reverse_f64", u, y, "binary_softmax_inverse_temperature",
         [(1, "tonic_volatility"), (2, "tonic_volatility"), ("response_parameter", 0), (1, "mean"), (2, "precision"), (1, "tonic_drift"),
          (2, "autoconnection_strength"), (1, "precision")],
         [rng.uniform(-6, -3, n), rng.uniform(-6, -3, n), np.exp(rng.uniform(-0.3, 0.8, n)), rng.uniform(-0.5, 0.5, n), np.exp(rng.uniform(-0.5, 0.5, n)),
          rng.uniform(-0.01, 0.01, n), rng.uniform(0.9, 1.0, n), np.exp(rng.uniform(-0.5, 0.5, n))]),
        (nets.continuous3, "sim_continuous3_reverse_f64", x, None, "first_level_gaussian_surprise",
         [(1, "tonic_volatility"), (2, "tonic_volatility"), (3, "tonic_volatility"), (2, "volatility_coupling_children", 0), (2, "tonic_drift"), (3, "mean")],
         [rng.uniform(-12, -9, n), rng.uniform(-3, -1, n), rng.uniform(-3, -1, n), rng.uniform(0.8, 1.2, n), rng.uniform(-0.01, 0.01, n), rng.uniform(0.5, 1.5, n)]),
    ]
    for build, entry, series, resp, rname, keys, vals in cases:
        net = build(Network, dtype="float64")
        net.create_belief_propagation_fn()
        plan = net._get_plan(1)
        attrs = np.ascontiguousarray(net._flatten(plan), dtype=np.float64)
        slots = np.array([net._resolve_slot(plan, k) for k in keys], dtype=np.int32)
        p = np.ascontiguousarray(np.stack(vals))
        logp, grad = np.zeros(n), np.zeros((len(keys), n))
        if resp is not None:
            getattr(lib, entry)(n, len(series), 0, fp(series), fp(resp), fp(attrs), len(keys), fp(slots), fp(p), 5, fp(logp), fp(grad))
        else:
            getattr(lib, entry)(n, len(series), fp(series), fp(attrs), len(keys), fp(slots), fp(p), fp(logp), fp(grad))
        o = build(onet.Network, dtype="float64").create_belief_propagation_fn()
        oslots = [o.slot(-1, "response_parameter", k[1]) if k[0] == "response_parameter" else o.slot(*k) for k in keys]
        want_s, want_g = o.batch_grad(series, oslots, p, rname, resp_inputs=resp)
        assert np.array_equal(logp, want_s), build.__name__
        for i, k in enumerate(keys):
            assert np.allclose(grad[i], want_g[i], rtol=1e-9, atol=1e-10 * np.abs(want_g[i]).max()), (build.__name__, k)


def test_generated_reverse_sweep_of_the_16_node_network(box):
    """BASELINE config 5's network (continuous + volatile-state nodes, two components): the generated reverse sweep compiled
    for the host, float64, against the Dual<double> oracle -- all 12 volatilities and six other attributes (an initial belief, a
    volatility-level precision, an autoconnection, a value and a volatility coupling, a drift), under both Gaussian response
    functions (total_gaussian_surprise sums every continuous node)."""
    import ctypes as C

    from pyhgf_b200.network import Network

    hostsim, _ = box
    lib = hostsim.build(["@", "-DHOSTSIM_GHGF16"])
    fp = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    rng = np.random.default_rng(5)
    n, T = 24, 150
    x = np.ascontiguousarray(nets.ar1_series(T, 4, seed=3), dtype=np.float64)
    net = nets.ghgf16(Network, dtype="float64")
    net.create_belief_propagation_fn()
    plan = net._get_plan(4)
    attrs = np.ascontiguousarray(net._flatten(plan), dtype=np.float64)
    o = nets.ghgf16(onet.Network, dtype="float64").create_belief_propagation_fn()
    draw = {"tonic_volatility": lambda: rng.uniform(-5, -3, n), "tonic_volatility_vol": lambda: rng.uniform(-5, -3, n),
            "mean": lambda: rng.uniform(-0.3, 0.3, n), "precision_vol": lambda: np.exp(rng.uniform(-0.3, 0.3, n)),
            "autoconnection_strength": lambda: rng.uniform(0.9, 1.0, n), "tonic_drift": lambda: rng.uniform(-0.01, 0.01, n)}
    all_keys = ([(i, "tonic_volatility") for i in range(4, 12)] + [(i, "tonic_volatility_vol") for i in range(12, 16)] +
                [(12, "mean"), (14, "precision_vol"), (13, "autoconnection_strength"), (5, "value_coupling_parents", 0),
                 (9, "volatility_coupling_children", 0), (6, "tonic_drift")])
    for fn, rname in ((2, "total_gaussian_surprise"), (1, "first_level_gaussian_surprise")):
        for lo in range(0, len(all_keys), 6):  # (the oracle carries at most 8 tangents)
            keys = all_keys[lo:lo + 6]
            p = np.ascontiguousarray(np.stack([draw.get(k[1], lambda: rng.uniform(0.8, 1.2, n))() for k in keys]))
            slots = np.array([net._resolve_slot(plan, k) for k in keys], dtype=np.int32)
            logp, grad = np.zeros(n), np.zeros((len(keys), n))
            lib.sim_ghgf16_reverse_f64(n, T, fp(x), fp(attrs), len(keys), fp(slots), fp(p), fn, fp(logp), fp(grad))
            want_s, want_g = o.batch_grad(x, [o.slot(*k) for k in keys], p, rname)
            assert np.array_equal(logp, want_s), rname
            for i, k in enumerate(keys):
                assert np.allclose(grad[i], want_g[i], rtol=1e-9, atol=1e-10 * max(np.abs(want_g[i]).max(), 1e-300)), (rname, k)


@pytest.mark.parametrize("case", ["continuous3/eHGF", "continuous3/standard", "volatile2/eHGF", "volatile2/standard", "volatile2/unbounded",
                                  "multi_child/eHGF", "multi_child/standard", "multi_child/unbounded",
                                  "multi_child/unbounded+mean-field", "multi_child/eHGF+mean-field", "volatile2/unbounded+mean-field",
                                  "continuous3/standard+mean-field",
                                  "coupling/tanh-sin", "coupling/sigmoid-gelu", "coupling/relu-leaky_relu", "coupling/gelu-identity", "coupling/sin-tanh",
                                  "coupling/volatile-tanh", "coupling/volatile-gelu"])
def test_generated_reverse_sweep_covers_every_update_scheme(case):
    """The sweep of a topology under each of the reference's three volatility-update schemes (unbounded / eHGF / standard;
    get_update_sequence.py:176-222), their mean-field variants, and every non-linear value-coupling function, generated and
    compiled for the host on the spot (tools/hostsim/revcheck.py): a volatile-state node, a continuous HGF, a user topology with
    two inputs, couplings among the parameters and irregular time steps, and chains with tanh / sin / sigmoid / gelu / relu /
    leaky-relu / identity couplings under continuous and volatile-state parents.  logp is the float64 oracle's bit for bit, every
    gradient the Dual<double> oracle's to 1e-10."""
    import revcheck

    cases = revcheck.CASES or revcheck._cases()
    equal, worst = revcheck.compare(**cases[case])
    assert equal
    assert worst < 1e-10, worst
