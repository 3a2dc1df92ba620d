"""Parity of the CUDA path (through the C ABI) with the CPU oracle and with the reference's goldens.

Tolerances (BASELINE.json north_star): float32, <= 1e-5 relative on per-trial means and precisions,
<= 1e-4 relative on summed surprise / logp and on gradients, asserted on EVERY instance of every batch: where the
oracle's own float32 result moves by more than the tolerance under a 1e-6 relative change of its parameters
(oracle/conditioning.py) the instance's bound is a small multiple of that sensitivity instead (tests/parity.py).
float64 launches are held to 1e-9.  Everything here needs a B200: run with ``-m gpu``.
"""

import ctypes as C
import json
import os

import numpy as np
import pytest

import nets
from oracle import conditioning
from oracle import network as onet
from parity import AMPLIFICATION_GRAD, assert_conditioned

pytestmark = pytest.mark.gpu

RTOL_TRAJ = 1e-5
RTOL_SUM = 1e-4


@pytest.fixture(scope="module")
def G():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from pyhgf_b200 import _lib

    _lib.lib()  # the CUDA library must be present: there is no fallback to skip to
    import pyhgf_b200.network as g

    return g


@pytest.fixture(scope="module")
def R():
    from pyhgf_b200 import response

    return response


def assert_traj_close(gnet, onet_, nodes, rtol, atol=1e-7, keys=nets.BELIEFS, sens=None):
    """Every recorded belief at every trial; `sens` (traj_sensitivity): the oracle's own reproducibility of each value."""
    for node in nodes:
        for k in keys:
            a = np.asarray(gnet.node_trajectories[node][k], dtype=np.float64)
            b = np.asarray(onet_.node_trajectories[node][k], dtype=np.float64)
            assert a.shape == b.shape
            assert np.array_equal(np.isnan(a), np.isnan(b)), (node, k)
            tol = rtol if sens is None else conditioning.tolerance(sens[(node, k)], rtol, 4.0)
            with np.errstate(invalid="ignore"):
                ok = (np.abs(a - b) <= atol + tol * np.abs(b)) | (np.isnan(a) & np.isnan(b)) | (a == b)
            assert ok.all(), (node, k, float(np.nanmax(np.abs(a - b) / (np.abs(b) + atol))))


def traj_sensitivity(run_oracle, nodes, keys=nets.BELIEFS, eps=1e-6, n_draws=4, seed=5):
    """`run_oracle(perturb)` builds the oracle network, calls `perturb(net)` on it before the data are fed, runs it and
    returns it.  Per recorded value: the largest relative change over a few runs whose tonic volatilities moved by
    `eps` relative -- how reproducible the float32 oracle's own trajectory is (oracle/conditioning.py)."""
    base = run_oracle(lambda net: None)
    rng = np.random.default_rng(seed)
    sens = {(n, k): np.zeros(np.shape(base.node_trajectories[n][k])) for n in nodes for k in keys}

    def perturb(net):
        for node, attrs in net.attributes.items():
            if isinstance(node, int) and node >= 0 and "tonic_volatility" in attrs:
                attrs["tonic_volatility"] = float(attrs["tonic_volatility"]) * (1.0 + eps * rng.uniform(-1, 1))

    for _ in range(n_draws):
        other = run_oracle(perturb)
        for key in sens:
            b = np.asarray(base.node_trajectories[key[0]][key[1]], dtype=np.float64)
            c = np.asarray(other.node_trajectories[key[0]][key[1]], dtype=np.float64)
            with np.errstate(invalid="ignore", divide="ignore"):
                d = np.abs(c - b) / (np.abs(b) + 1e-7)
            sens[key] = np.maximum(sens[key], np.nan_to_num(d, nan=0.0, posinf=0.0))
    return sens


# ------------------------------------------------------------------ reference goldens, straight on the GPU
def test_golden_binary_one_step(G):  # tests/test_binary.py:30-70
    net = (G.Network().add_nodes(kind="binary-state").add_nodes(value_children=0, mean=1.0, tonic_volatility=1.0)
           .add_nodes(volatility_children=1, mean=1.0, tonic_volatility=1.0).input_data(np.ones(1)))
    a = net.last_attributes
    for k, v in zip(["mean", "expected_mean", "expected_precision"], [1.0, 0.7310586, 0.19661193]):
        assert np.isclose(a[0][k], v, rtol=1e-5)
    for k, v in zip(["mean", "expected_mean", "precision", "expected_precision"], [2.2378633, 1.0, 0.21726260, 0.02065066]):
        assert np.isclose(a[1][k], v, rtol=1e-5)
    for k, v in zip(["mean", "expected_mean", "precision", "expected_precision"], [0.37627372, 1.0, 0.39340553, 0.26894143]):
        assert np.isclose(a[2][k], v, rtol=1e-5)


def test_golden_binary_scan(G, datasets):  # tests/test_binary.py:73-94
    net = nets.binary3(G.Network).input_data(datasets["u"])
    for k, v in zip(nets.BELIEFS, [-2.3191204, -2.171629, 0.6937843, 0.6019279]):
        assert np.isclose(net.node_trajectories[1][k][-1], v, rtol=1e-5)
    for k, v in zip(nets.BELIEFS, [2.478936, 2.4795845, 1.7116449, 1.6612335]):
        assert np.isclose(net.node_trajectories[2][k][-1], v, rtol=1e-5)


def test_golden_continuous_scan(G, datasets):  # tests/test_continuous.py:101-131
    net = nets.continuous2(G.Network).input_data(datasets["x"])
    for node, vals in {0: [0.8241, 0.7945459, 1.0, 0.7623690], 1: [0.7978380, 0.7945459, 6.8440757, 6.0817065],
                       2: [-2.2633543, -2.2633660, 1.4760352, 1.4371926]}.items():
        for k, v in zip(nets.BELIEFS, vals):
            assert np.isclose(net.node_trajectories[node][k][-1], v, rtol=1e-5), (node, k)


def test_golden_surprises(G, R, datasets):  # tests/test_model.py:113-268, tests/test_responses.py:10-50, README.md:115
    x, u, y = datasets["x"], datasets["u"], datasets["y"]
    two = (G.Network().add_nodes(precision=1e4)
           .add_nodes(value_children=([0], [1.0]), node_parameters={"mean": x[0], "precision": 1e4, "tonic_volatility": -3.0, "tonic_drift": 0.0})
           .add_nodes(volatility_children=([1], [1.0]), node_parameters={"mean": 0.0, "precision": 1e1, "tonic_volatility": -3.0, "tonic_drift": 0.0})
           .input_data(x))
    assert np.isclose(two.surprise(R.first_level_gaussian_surprise).sum(), -1924.7515, rtol=1e-5)
    three = nets.continuous3(G.Network).input_data(x)
    assert np.isclose(three.surprise(R.first_level_gaussian_surprise).sum(), -2034.3989, rtol=1e-5)
    assert np.isclose(R.total_gaussian_surprise(three).sum(), -2535.604, rtol=1e-5)
    b2 = (G.Network().add_nodes(kind="binary-state", node_parameters={"mean": 0.0, "precision": 0.0})
          .add_nodes(value_children=([0], [1.0]), node_parameters={"mean": 0.5, "precision": 1e4, "tonic_volatility": -6.0, "tonic_drift": 0.0})
          .input_data(u))
    assert np.isclose(b2.surprise(R.first_level_binary_surprise).sum(), 215.58821, rtol=1e-5)
    soft = (G.Network().add_nodes(kind="binary-state", node_parameters={"mean": 0.5, "precision": 0.0})
            .add_nodes(value_children=([0], [1.0]), node_parameters={"mean": 0.0, "precision": 1.0, "tonic_volatility": -6.0, "tonic_drift": 0.0})
            .input_data(u))
    assert np.isclose(soft.surprise(R.binary_softmax, response_function_inputs=y).sum(), 195.81573, rtol=1e-5)
    s = soft.surprise(R.binary_softmax_inverse_temperature, response_function_inputs=y, response_function_parameters=2.0)
    assert np.isclose(s.sum(), 188.77818, rtol=1e-5)
    readme = nets.binary2(G.Network).input_data(u)
    s = readme.surprise(R.binary_softmax_inverse_temperature, response_function_inputs=y, response_function_parameters=4.0)
    assert np.isclose(s.sum(), 138.8992462158203, rtol=1e-5)


def test_saturating_binary_is_finite(G, R, datasets):  # tests/test_model.py:237-268: rounding-sensitive at the 1 % level
    three = (G.Network().add_nodes(kind="binary-state", node_parameters={"mean": 0.0, "precision": 0.0})
             .add_nodes(value_children=([0], [1.0]), node_parameters={"mean": 0.5, "precision": 1e4, "tonic_volatility": -6.0, "tonic_drift": 0.0})
             .add_nodes(volatility_children=([1], [1.0]), node_parameters={"mean": 0.0, "precision": 1e1, "tonic_volatility": -2.0, "tonic_drift": 0.0})
             .input_data(datasets["u"]))
    s = three.surprise(R.first_level_binary_surprise).sum()
    assert np.isfinite(s) and 0.9 * 1242.3856 < s < 1.1 * 1256.755


# ------------------------------------------------------------------ full trajectories vs the oracle
CASES = [
    ("binary3", nets.binary3, "u", {}),
    ("binary2", nets.binary2, "u", {}),
    ("continuous2", nets.continuous2, "x", {}),
    ("continuous3", nets.continuous3, "x", {}),
    ("volatile2", nets.volatile2, "x", {}),
    ("tanh", nets.tanh_coupling, "x", {}),
    ("multi_child", nets.multi_child, "ar2", {}),
    ("ghgf16", nets.ghgf16, "ar4", {}),
]
for vu in ("standard", "eHGF"):
    for mf in (False, True):
        CASES.append((f"continuous2-{vu}-mf{int(mf)}", nets.continuous2, "x", dict(volatility_updates=vu, mean_field_updates=mf)))
        CASES.append((f"volatile2-{vu}-mf{int(mf)}", nets.volatile2, "x", dict(volatility_updates=vu, mean_field_updates=mf)))
CASES.append(("continuous2-unbounded-mf1", nets.continuous2, "x", dict(mean_field_updates=True)))
CASES.append(("binary3-ehgf", nets.binary3, "u", dict(volatility_updates="eHGF")))


def _series(name, datasets):
    if name in ("u", "x"):
        return datasets[name]
    return nets.ar1_series(400, n_series=int(name[2:]))


@pytest.mark.parametrize("dtype", ["float32", "float64"])
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_trajectories_match_oracle(G, datasets, case, dtype):
    _, build, series, kw = case
    data = _series(series, datasets)
    g = build(G.Network, dtype=dtype, **kw).input_data(data)
    o = build(onet.Network, dtype=dtype, **kw).input_data(data)
    rtol = RTOL_TRAJ if dtype == "float32" else 1e-9
    if dtype == "float32" and case[0] == "tanh":
        rtol = 3e-5  # CUDA tanhf vs glibc tanhf (<= 2 ulp each) compounded over 614 recursive steps
    keys = nets.BELIEFS
    assert_traj_close(g, o, range(g.n_nodes), rtol, atol=1e-6 if dtype == "float32" else 1e-12, keys=keys)
    for node, e in enumerate(g.edges):
        if e.node_type == 6:
            assert_traj_close(g, o, [node], rtol, atol=1e-6 if dtype == "float32" else 1e-12,
                              keys=("mean_vol", "expected_mean_vol", "precision_vol", "expected_precision_vol"))
    # the reference's record_trajectories=False contract (tests/test_continuous.py:134-165)
    g2 = build(G.Network, dtype=dtype, **kw).input_data(data, record_trajectories=False)
    assert g2.node_trajectories is None
    for node in range(g.n_nodes):
        for k in ("mean", "precision"):
            # (jnp.isclose in the reference's test: the two calls may run different kernel variants --
            # register-resident vs. interpreter -- whose mul+add contraction differs in the last bits)
            assert np.isclose(g2.last_attributes[node][k], g.last_attributes[node][k], rtol=RTOL_TRAJ if dtype == "float32" else 1e-12)
            assert g.last_attributes[node][k] == g.node_trajectories[node][k][-1]


def test_time_steps_and_missing_observations(G, datasets):
    rng = np.random.default_rng(3)
    x = datasets["x"][:200]
    ts = rng.uniform(0.5, 2.0, size=200)
    obs = (rng.uniform(size=200) > 0.2).astype(int)
    for vu in ("unbounded", "eHGF", "standard"):
        g = nets.continuous2(G.Network, volatility_updates=vu).input_data(x, time_steps=ts, observed=obs)
        o = nets.continuous2(onet.Network, volatility_updates=vu).input_data(x, time_steps=ts, observed=obs)
        assert_traj_close(g, o, range(3), RTOL_TRAJ, atol=1e-6)
    u = datasets["u"]
    obs = (rng.uniform(size=320) > 0.3).astype(int)
    g = nets.binary3(G.Network).input_data(u, observed=obs)
    o = nets.binary3(onet.Network).input_data(u, observed=obs)
    assert_traj_close(g, o, range(3), RTOL_TRAJ, atol=1e-6)


def test_nan_propagates_like_the_reference(G, R):
    # omega so large that exp overflows: the reference signals in-band (NaN beliefs -> +inf surprise)
    data = np.array([0.0, 1.0] * 20)
    g = nets.binary3(G.Network, om2=80.0, om3=80.0).input_data(data)
    o = nets.binary3(onet.Network, om2=80.0, om3=80.0).input_data(data)
    for node in range(3):
        for k in nets.BELIEFS:
            a, b = g.node_trajectories[node][k], o.node_trajectories[node][k]
            assert np.array_equal(np.isnan(a), np.isnan(b)), (node, k)
            assert np.array_equal(np.isinf(a), np.isinf(b)), (node, k)
    sg = g.surprise(R.first_level_binary_surprise)
    so = o.surprise(onet.first_level_binary_surprise)
    assert np.array_equal(np.isinf(sg), np.isinf(so))


def test_total_gaussian_surprise_is_inf_at_every_trial_once_node1_was_nan(G, R, datasets):
    """pyhgf/response.py:86-89: `jnp.where(jnp.any(jnp.isnan(node_trajectories[1]["mean"])), jnp.inf, surprise)` acts on
    the whole per-trial vector, not only on its sum -- single network and batched pointwise output alike."""
    x = datasets["x"][:60].copy()
    bad = x.copy()
    bad[30] = np.nan  # a NaN observation: node 1's posterior mean is NaN from trial 30 on, finite before
    o = nets.continuous3(onet.Network).input_data(bad)
    want = o.surprise(onet.total_gaussian_surprise)
    m1 = np.asarray(o.node_trajectories[1]["mean"])
    assert np.isfinite(m1[:30]).all() and np.isnan(m1[30:]).all() and np.isinf(want).all()
    g = nets.continuous3(G.Network).input_data(bad)
    assert np.isinf(g.surprise(R.total_gaussian_surprise)).all()
    res = nets.continuous3(G.Network).batch(np.stack([x, bad, x]), {(2, "tonic_volatility"): np.array([-2.0, -2.0, -3.0])}, input_group_size=1,
                                            response_function=R.total_gaussian_surprise, pointwise=True)
    pw = res.pointwise.cpu().numpy()
    assert np.isfinite(pw[0]).all() and np.isfinite(pw[2]).all() and np.isinf(pw[1]).all()
    assert np.isinf(res.surprise.cpu().numpy()[1]) and np.isfinite(res.surprise.cpu().numpy()[[0, 2]]).all()
    clean = nets.continuous3(onet.Network, om2=-2.0).input_data(x).surprise(onet.total_gaussian_surprise)
    assert np.allclose(pw[0], clean, rtol=1e-4, atol=1e-4)


# ------------------------------------------------------------------ batched surprise and gradients
def _grid(n, lo=-7.0, hi=-1.0, seed=0):
    rng = np.random.default_rng(seed)
    return rng.uniform(lo, hi, size=n)


@pytest.mark.parametrize("n", [1, 31, 128, 1000])
def test_batched_binary_logp_matches_oracle(G, R, datasets, n):
    u, y = datasets["u"], datasets["y"]
    om2, om3, temp = _grid(n, -7, -2.5, 1), _grid(n, -7, -2.5, 2), np.exp(_grid(n, -0.5, 1.0, 3))
    g = nets.binary3(G.Network)
    res = g.batch(u, {(1, "tonic_volatility"): om2, (2, "tonic_volatility"): om3, ("response_parameter", 0): temp},
                  response_function=R.binary_softmax_inverse_temperature, response_function_inputs=y, pointwise=True,
                  last_attributes=True, record=[(1, "mean"), (2, "precision")])
    o = nets.binary3(onet.Network).create_belief_propagation_fn()
    slots = [o.slot(1, "tonic_volatility"), o.slot(2, "tonic_volatility"), o.slot(-1, "response_parameter", 0)]

    def oracle(p):  # summed surprise and the two recorded beliefs at the last trial, [3][n]
        s_, last = o.batch_surprise(u, slots, p, "binary_softmax_inverse_temperature", resp_inputs=y, return_last=True)
        return np.stack([s_, last[:, o.slot(1, "mean")], last[:, o.slot(2, "precision")]])

    p = np.stack([om2, om3, temp]).astype(np.float32).astype(np.float64)
    (want, m1, p2), sens = conditioning.sensitivity(oracle, p, rows=[0, 1], floor=1e-6)
    got = res.surprise.double().cpu().numpy()
    assert tuple(res.pointwise.shape) == (n, 320)
    # the box reaches omega = -2.5, where the second-level belief starts to run away: every instance is
    # held to the stated tolerance or, where the oracle itself is not reproducible to it, to its own sensitivity
    assert_conditioned(got, want, sens[0], RTOL_SUM, what="surprise")
    assert_conditioned(res.pointwise.sum(1).double().cpu().numpy(), want, sens[0], RTOL_SUM, what="pointwise")
    for got_t, want_v, sn in [(res.last_attributes[(1, "mean")], m1, sens[1]), (res.trajectories[(1, "mean")][:, -1], m1, sens[1]),
                              (res.trajectories[(2, "precision")][:, -1], p2, sens[2])]:
        assert_conditioned(got_t.double().cpu().numpy(), want_v, sn, RTOL_TRAJ, floor=1e-6, what="last beliefs")


def test_batched_continuous_logp_matches_oracle(G, R, datasets):
    x = datasets["x"]
    n = 777
    rng = np.random.default_rng(0)
    om = [np.clip(rng.normal(m, 1.0, n), m - 3, m + 3) for m in (-13.0, -2.0, -2.0)]
    g = nets.continuous3(G.Network)
    res = g.batch(x, {(1, "tonic_volatility"): om[0], (2, "tonic_volatility"): om[1], (3, "tonic_volatility"): om[2]},
                  response_function=R.first_level_gaussian_surprise)
    o = nets.continuous3(onet.Network).create_belief_propagation_fn()
    slots = [o.slot(i, "tonic_volatility") for i in (1, 2, 3)]
    want = o.batch_surprise(x, slots, np.stack(om), "first_level_gaussian_surprise")
    got = res.surprise.double().cpu().numpy()
    finite = np.isfinite(want)
    assert np.array_equal(finite, np.isfinite(got))
    assert np.allclose(got[finite], want[finite], rtol=RTOL_SUM, atol=1e-2)  # sums cross zero: ~2e3 magnitude terms


@pytest.mark.parametrize("dtype,rtol", [("float64", 1e-7), ("float32", RTOL_SUM)])
def test_gradient_matches_dual_oracle(G, R, datasets, dtype, rtol):
    u, y = datasets["u"], datasets["y"]
    n = 200
    # well-conditioned box (the reference's own tests and tutorials sit at omega in [-6, -3])
    om2, om3, temp = _grid(n, -6.5, -3.5, 1), _grid(n, -6.5, -3.5, 2), np.exp(_grid(n, -0.3, 0.8, 3))
    mean1 = _grid(n, -0.5, 0.5, 4)
    g = nets.binary3(G.Network, dtype=dtype)
    keys = [(1, "tonic_volatility"), (2, "tonic_volatility"), ("response_parameter", 0), (1, "mean"), (2, "precision")]
    vals = [om2, om3, temp, mean1, np.exp(_grid(n, -0.5, 0.5, 5))]
    res = g.batch(u, dict(zip(keys, vals)), response_function=R.binary_softmax_inverse_temperature,
                  response_function_inputs=y, gradient=True)
    o = nets.binary3(onet.Network, dtype=dtype).create_belief_propagation_fn()
    slots = [o.slot(1, "tonic_volatility"), o.slot(2, "tonic_volatility"), o.slot(-1, "response_parameter", 0),
             o.slot(1, "mean"), o.slot(2, "precision")]
    p = np.stack(vals)
    if dtype == "float32":
        p = p.astype(np.float32).astype(np.float64)
    # like-for-like precision: the float32 kernel is checked against the Dual<float> oracle (north_star: <= 1e-4
    # relative on summed logp and gradients), the float64 kernel against Dual<double>

    def oracle(q):
        s_, g_ = o.batch_grad(u, slots, q, "binary_softmax_inverse_temperature", resp_inputs=y, dtype=32 if dtype == "float32" else 64)
        return np.concatenate([s_[None], g_])

    if dtype == "float64":
        want = oracle(p)
        want_s, want_g = want[0], want[1:]
        assert np.allclose(res.surprise.double().cpu().numpy(), want_s, rtol=rtol)
        for i, k in enumerate(keys):
            got = res.gradient[k].double().cpu().numpy()
            scale = np.abs(want_g[i]).max()
            assert np.allclose(got, want_g[i], rtol=rtol, atol=rtol * scale), (k, np.abs(got - want_g[i]).max(), scale)
        return
    # float32: every instance, logp and every gradient, to the stated 1e-4 (a gradient that crosses zero is measured
    # against |g| + 1e-4 max|g|) or to the oracle's own reproducibility where that is worse
    scales = np.abs(oracle(p)[1:]).max(axis=1)
    floor = np.concatenate([[0.0], rtol * scales])[:, None]
    want, sens = conditioning.sensitivity(oracle, p, rows=[0, 1, 3, 4], n_draws=16, floor=floor)
    want_s, want_g = want[0], want[1:]
    assert_conditioned(res.surprise.double().cpu().numpy(), want_s, sens[0], rtol, what="logp", max_relaxed=0.02)
    for i, k in enumerate(keys):
        got = res.gradient[k].double().cpu().numpy()
        assert_conditioned(got, want_g[i], sens[1 + i], rtol, amplification=AMPLIFICATION_GRAD, floor=rtol * scales[i], what=k)
        # the gradient summed over instances (what a sampler's all-reduce carries) is held to the stated 1e-4 as well
        assert np.isclose(got.sum(), want_g[i].sum(), rtol=3 * rtol, atol=rtol * np.abs(want_g[i]).sum() / n ** 0.5), (k, got.sum(), want_g[i].sum())
    # and the float32 gradient is a faithful gradient: close to the float64 one
    _, g64 = o.batch_grad(u, slots, p, "binary_softmax_inverse_temperature", resp_inputs=y)
    for i, k in enumerate(keys):
        got = res.gradient[k].double().cpu().numpy()
        rel = np.abs(got - g64[i]) / (np.abs(g64[i]) + 1e-3 * np.abs(g64[i]).max())
        assert np.median(rel) < 1e-4, (k, np.median(rel))


@pytest.mark.parametrize("dtype", ["float64", "float32"])
def test_reverse_sweep_matches_dual_oracle(G, R, datasets, dtype):
    """The generated checkpointed reverse sweep (csrc/hgf_reverse.cuh, pyhgf_b200/revgen.py; the reference: jax.grad over the
    scan, pyhgf/model/network.py:612-624) against the forward-mode Dual oracle: 8 parameters of the 3-level binary HGF --
    volatilities, inverse temperature, initial beliefs, drift, autoconnection -- and 6 of the 4-node continuous HGF
    including a volatility coupling.  float64 to 1e-7; float32 per instance to the stated 1e-4 / the oracle's sensitivity."""
    u, y, x = datasets["u"], datasets["y"], datasets["x"][:300]
    n = 96
    cases = [
        (nets.binary3, u, y, "binary_softmax_inverse_temperature", R.binary_softmax_inverse_temperature,
         [(1, "tonic_volatility"), (2, "tonic_volatility"), ("response_parameter", 0), (1, "mean"), (2, "precision"), (1, "tonic_drift"),
          (2, "autoconnection_strength"), (1, "precision")],
         [_grid(n, -6, -3.5, 1), _grid(n, -6, -3.5, 2), np.exp(_grid(n, -0.3, 0.8, 3)), _grid(n, -0.5, 0.5, 4), np.exp(_grid(n, -0.5, 0.5, 5)),
          _grid(n, -0.01, 0.01, 6), _grid(n, 0.9, 1.0, 7), np.exp(_grid(n, -0.5, 0.5, 8))]),
        (nets.continuous3, x, None, "first_level_gaussian_surprise", R.first_level_gaussian_surprise,
         [(1, "tonic_volatility"), (2, "tonic_volatility"), (3, "tonic_volatility"), (2, "volatility_coupling_children", 0),
          (1, "volatility_coupling_parents", 0), (2, "tonic_drift")],
         [_grid(n, -12, -9, 1), _grid(n, -3, -1, 2), _grid(n, -3, -1, 3), _grid(n, 0.8, 1.2, 4), _grid(n, 0.8, 1.2, 5), _grid(n, -0.01, 0.01, 6)]),
    ]
    for build, data, resp_in, rname, rfn, keys, vals in cases:
        p = np.stack(vals)
        if dtype == "float32":
            p = p.astype(np.float32).astype(np.float64)
        kw = {} if resp_in is None else {"response_function_inputs": resp_in}
        res = build(G.Network, dtype=dtype).batch(data, dict(zip(keys, p)), response_function=rfn, gradient="reverse", **kw)
        fwd = build(G.Network, dtype=dtype).batch(data, dict(zip(keys, p)), response_function=rfn, gradient="forward", **kw)
        o = build(onet.Network, dtype=dtype).create_belief_propagation_fn()
        slots = [o.slot(-1, "response_parameter", k[1]) if k[0] == "response_parameter" else o.slot(*k) for k in keys]

        def oracle(q):
            s_, g_ = o.batch_grad(data, slots, q, rname, resp_inputs=resp_in, dtype=32 if dtype == "float32" else 64)
            return np.concatenate([s_[None], g_])

        if dtype == "float64":
            want = oracle(p)
            assert np.allclose(res.surprise.cpu().numpy(), want[0], rtol=1e-9)
            for i, k in enumerate(keys):
                got = res.gradient[k].cpu().numpy()
                assert np.allclose(got, want[1 + i], rtol=1e-7, atol=1e-7 * np.abs(want[1 + i]).max()), (build.__name__, k)
                assert np.allclose(got, fwd.gradient[k].cpu().numpy(), rtol=1e-7, atol=1e-7 * np.abs(want[1 + i]).max())
            continue
        scales = np.abs(oracle(p)[1:]).max(axis=1)
        floor = np.concatenate([[0.0 if resp_in is not None else 100.0], RTOL_SUM * scales])[:, None]
        filt = [i for i, k in enumerate(keys) if k[0] != "response_parameter"]
        want, sens = conditioning.sensitivity(oracle, p, rows=filt, n_draws=16, floor=floor)
        assert_conditioned(res.surprise.double().cpu().numpy(), want[0], sens[0], RTOL_SUM, floor=float(floor[0, 0]), what="logp (reverse)")
        for i, k in enumerate(keys):
            assert_conditioned(res.gradient[k].double().cpu().numpy(), want[1 + i], sens[1 + i], RTOL_SUM, amplification=AMPLIFICATION_GRAD,
                               floor=RTOL_SUM * scales[i], what=(build.__name__, k))


def test_reverse_sweep_dispatch(G, R, datasets):
    """gradient=True takes the reverse sweep as soon as forward mode would need a second pass (here 13 parameters: one launch
    instead of four), keeps forward mode for 4 filter parameters + 1 response parameter, and the two kernels agree; asking
    for the sweep where it cannot run is an error, not a silent fallback."""
    from pyhgf_b200 import _lib

    u, y = datasets["u"], datasets["y"]
    n = 64
    keys = [(1, "tonic_volatility"), (2, "tonic_volatility"), ("response_parameter", 0), (1, "mean"), (2, "mean"), (1, "precision"), (2, "precision"),
            (1, "tonic_drift"), (2, "tonic_drift"), (2, "autoconnection_strength"), (1, "autoconnection_strength"),
            (2, "volatility_coupling_children", 0), (1, "volatility_coupling_parents", 0)]
    vals = [_grid(n, -6, -3.5, 1), _grid(n, -6, -3.5, 2), np.exp(_grid(n, -0.3, 0.8, 3)), _grid(n, -0.5, 0.5, 4), _grid(n, -0.5, 0.5, 5),
            np.exp(_grid(n, -0.5, 0.5, 6)), np.exp(_grid(n, -0.5, 0.5, 7)), _grid(n, -0.01, 0.01, 8), _grid(n, -0.01, 0.01, 9), _grid(n, 0.9, 1.0, 10),
            _grid(n, 0.9, 1.0, 11), _grid(n, 0.8, 1.2, 12), _grid(n, 0.8, 1.2, 13)]
    kw = dict(response_function=R.binary_softmax_inverse_temperature, response_function_inputs=y)
    net = nets.binary3(G.Network, dtype="float64")
    before = _lib.launch_count()
    auto = net.batch(u, dict(zip(keys, vals)), gradient=True, **kw)
    assert _lib.launch_count() - before == 1          # the reverse sweep: one kernel
    before = _lib.launch_count()
    fwd = net.batch(u, dict(zip(keys, vals)), gradient="forward", **kw)
    assert _lib.launch_count() - before == 4          # 13 parameters, 4 tangents per float64 pass: four forward-mode passes
    for k in keys:
        a, b = auto.gradient[k].cpu().numpy(), fwd.gradient[k].cpu().numpy()
        assert np.allclose(a, b, rtol=1e-7, atol=1e-7 * np.abs(b).max()), k
    net32 = nets.binary3(G.Network, dtype="float32")
    for n_keys, launches in ((5, 1), (6, 1)):  # 4 filter parameters + the temperature: ONE forward-mode pass; 5 + 1: the sweep
        sub = dict(zip(keys[:n_keys], vals[:n_keys]))
        before = _lib.launch_count()
        few = net32.batch(u, sub, gradient=True, **kw)
        assert _lib.launch_count() - before == launches
        assert few._call.gradient_mode == 0
        before = _lib.launch_count()
        ref = net32.batch(u, sub, gradient="forward", **kw)
        assert _lib.launch_count() - before == (1 if n_keys == 5 else 2)
        for k in keys[:n_keys]:
            a, b = few.gradient[k].cpu().numpy(), ref.gradient[k].cpu().numpy()
            if n_keys == 5:
                assert np.array_equal(a, b), k  # the same kernel
            else:
                assert np.allclose(a, b, rtol=1e-3, atol=1e-3 * np.abs(b).max()), k
    with pytest.raises(_lib.BackendError, match="reverse"):
        net.batch(u, dict(zip(keys[:3], vals[:3])), gradient="reverse", record=[(1, "mean")], **kw)
    with pytest.raises(_lib.BackendError, match="no generated reverse sweep"):
        nets.volatile2(G.Network).batch(datasets["x"][:50], {(1, "tonic_volatility_vol"): np.full(4, -3.0)},
                                        response_function=R.first_level_gaussian_surprise, gradient="reverse")


def test_reverse_sweep_16_node_network(G, R):
    """BASELINE config 5's network (volatile-state nodes included): d logp / d (all 12 volatilities + 4 more attributes) by
    the generated reverse sweep in ONE launch (gradient=True picks it: forward mode would need 4 passes through the
    interpreter), against the Dual<double> oracle in float64 and against the float64 launch in float32."""
    from pyhgf_b200 import _lib

    n, T = 96, 200
    x = nets.ar1_series(T, 4, seed=3)
    keys = ([(i, "tonic_volatility") for i in range(4, 12)] + [(i, "tonic_volatility_vol") for i in range(12, 16)] +
            [(12, "mean"), (13, "autoconnection_strength"), (5, "value_coupling_parents", 0), (9, "volatility_coupling_children", 0)])
    vals = ([_grid(n, -5, -3, 1 + i) for i in range(12)] + [_grid(n, -0.3, 0.3, 20), _grid(n, 0.9, 1.0, 21), _grid(n, 0.8, 1.2, 22), _grid(n, 0.8, 1.2, 23)])
    for rfn, rname in ((R.total_gaussian_surprise, "total_gaussian_surprise"), (R.first_level_gaussian_surprise, "first_level_gaussian_surprise")):
        net = nets.ghgf16(G.Network, dtype="float64")
        before = _lib.launch_count()
        res = net.batch(x, dict(zip(keys, vals)), response_function=rfn, gradient=True)
        assert res.kernel_name == "static:ghgf16"
        assert _lib.launch_count() - before == 1
        o = nets.ghgf16(onet.Network, dtype="float64").create_belief_propagation_fn()
        p = np.stack(vals)
        for lo in range(0, len(keys), 8):  # (the oracle carries at most 8 tangents)
            sl = slice(lo, lo + 8)
            want_s, want_g = _oracle_grad_subset(o, x, keys, p, sl, rname)
            assert np.allclose(res.surprise.cpu().numpy(), want_s, rtol=1e-9)
            for i, k in enumerate(keys[sl]):
                got = res.gradient[k].cpu().numpy()
                assert np.allclose(got, want_g[i], rtol=1e-7, atol=1e-7 * max(np.abs(want_g[i]).max(), 1e-300)), (rname, k)
        # a FEW parameters: forward mode would be one pass, but this topology has no register-resident gradient kernel (it would
        # be the interpreter's): gradient=True takes the sweep, bit for bit the explicit "reverse" result
        few = dict(zip(keys[:3], vals[:3]))
        a = net.batch(x, few, response_function=rfn, gradient=True)
        b = net.batch(x, few, response_function=rfn, gradient="reverse")
        f = net.batch(x, few, response_function=rfn, gradient="forward")
        for k in few:
            assert np.array_equal(a.gradient[k].cpu().numpy(), b.gradient[k].cpu().numpy()), k
            assert np.allclose(a.gradient[k].cpu().numpy(), f.gradient[k].cpu().numpy(), rtol=1e-7, atol=1e-7 * float(f.gradient[k].abs().max())), k
        r32 = nets.ghgf16(G.Network, dtype="float32").batch(x, dict(zip(keys, vals)), response_function=rfn, gradient="reverse")
        s64 = res.surprise.cpu().numpy()
        assert np.median(np.abs(r32.surprise.double().cpu().numpy() - s64) / np.abs(s64)) < 1e-5
        for k in keys:
            g64 = res.gradient[k].cpu().numpy()
            err = np.abs(r32.gradient[k].double().cpu().numpy() - g64) / np.maximum(np.abs(g64), 1e-3 * max(np.abs(g64).max(), 1e-300))
            assert np.median(err) < 1e-3, (rname, k, float(np.median(err)))


def _oracle_grad_subset(o, x, keys, p, sl, rname):
    """Dual oracle gradient wrt the parameters keys[sl] with EVERY parameter of `keys` set per instance: the oracle carries at
    most 8 tangents, so the parameters outside the subset are written into the network's attributes, one instance at a time."""
    n = p.shape[1]
    all_slots = [o.slot(*k) for k in keys]
    sub = list(range(*sl.indices(len(keys))))
    rest = [i for i in range(len(keys)) if i not in sub]
    want_s, want_g = np.zeros(n), np.zeros((len(sub), n))
    for j in range(n):
        saved = {}
        for i in rest:
            node, key = keys[i][0], keys[i][1]
            a = o.attributes[node]
            if len(keys[i]) == 3:
                saved[i] = np.array(a[key], dtype=np.float64).copy()
                arr = np.array(a[key], dtype=np.float64)
                arr[keys[i][2]] = p[i, j]
                a[key] = arr
            else:
                saved[i] = a[key]
                a[key] = float(p[i, j])
        s_, g_ = o.batch_grad(x, [all_slots[i] for i in sub], p[sub][:, j:j + 1], rname)
        want_s[j], want_g[:, j] = s_[0], g_[:, 0]
        for i in rest:
            o.attributes[keys[i][0]][keys[i][1]] = saved[i]
    return want_s, want_g


def test_gradient_continuous_and_coupling(G, R, datasets):
    x = datasets["x"][:300]
    n = 64
    g = nets.continuous3(G.Network, dtype="float64")
    keys = [(1, "tonic_volatility"), (2, "tonic_volatility"), (3, "tonic_volatility"), (2, "volatility_coupling_children", 0),
            (1, "volatility_coupling_parents", 0), (2, "tonic_drift")]
    vals = [_grid(n, -12, -9, 1), _grid(n, -3, -1, 2), _grid(n, -3, -1, 3), _grid(n, 0.8, 1.2, 4), _grid(n, 0.8, 1.2, 5),
            _grid(n, -0.01, 0.01, 6)]
    res = g.batch(x, dict(zip(keys, vals)), response_function=R.first_level_gaussian_surprise, gradient=True)
    o = nets.continuous3(onet.Network, dtype="float64").create_belief_propagation_fn()
    slots = [o.slot(1, "tonic_volatility"), o.slot(2, "tonic_volatility"), o.slot(3, "tonic_volatility"),
             o.slot(2, "volatility_coupling_children", 0), o.slot(1, "volatility_coupling_parents", 0), o.slot(2, "tonic_drift")]
    want_s, want_g = o.batch_grad(x, slots, np.stack(vals), "first_level_gaussian_surprise")
    assert np.allclose(res.surprise.cpu().numpy(), want_s, rtol=1e-8)
    for i, k in enumerate(keys):
        got = res.gradient[k].cpu().numpy()
        assert np.allclose(got, want_g[i], rtol=1e-6, atol=1e-7 * np.abs(want_g[i]).max()), k


def test_grouped_series_and_chain_reduction(G, R):
    """Config 4's layout: participants x chains, per-participant series, per-chain sums reduced on the device."""
    rng = np.random.default_rng(5)
    n_part, n_chain, T = 6, 5, 120
    u = (rng.uniform(size=(n_part, T)) > 0.5).astype(float)
    y = (rng.uniform(size=(n_part, T)) > 0.5).astype(float)
    n = n_part * n_chain
    om2, om3, temp = _grid(n, -5, -2, 1), _grid(n, -5, -2, 2), np.exp(_grid(n, -0.3, 0.8, 3))
    import torch

    group = torch.arange(n, dtype=torch.int32) % n_chain
    g = nets.binary3(G.Network)
    res = g.batch(u, {(1, "tonic_volatility"): om2, (2, "tonic_volatility"): om3, ("response_parameter", 0): temp},
                  response_function=R.binary_softmax_inverse_temperature, response_function_inputs=y,
                  input_group_size=n_chain, response_group_size=n_chain, group=group, n_groups=n_chain, out_scale=-1.0,
                  gradient=True)
    o = nets.binary3(onet.Network).create_belief_propagation_fn()
    slots = [o.slot(1, "tonic_volatility"), o.slot(2, "tonic_volatility"), o.slot(-1, "response_parameter", 0)]
    want = o.batch_surprise(u, slots, np.stack([om2, om3, temp]), "binary_softmax_inverse_temperature", resp_inputs=y,
                            group_size=n_chain)
    logp = res.surprise.double().cpu().numpy()  # out_scale = -1
    assert np.allclose(logp, -want, rtol=RTOL_SUM)
    per_chain = -want.reshape(n_part, n_chain).sum(0)
    assert np.allclose(res.reduced.double().cpu().numpy(), per_chain, rtol=RTOL_SUM)


def test_distribution_op(G, R, datasets):
    from pyhgf_b200.distribution import HGFDistribution, HGFPointwise

    u, y = datasets["u"], datasets["y"]
    op = HGFDistribution(n_levels=2, model_type="binary", input_data=u[None, :], response_function=R.binary_softmax,
                         response_function_inputs=y[None, :])
    # tests/test_responses.py:10-30 -- omega_2 = -6 gives 195.81573 with the default initial beliefs
    assert np.isclose(float(op(tonic_volatility_2=-6.0)), -195.81573, rtol=1e-5)
    n = 9
    uu = np.repeat(u[None, :], n, 0)
    yy = np.repeat(y[None, :], n, 0)
    om2 = np.linspace(-6, -2, n)
    op3 = HGFDistribution(n_levels=3, model_type="binary", input_data=uu, response_function=R.binary_softmax_inverse_temperature,
                          response_function_inputs=yy)
    total, grads = op3.value_and_grad(tonic_volatility_2=om2, response_function_parameters=np.full(n, 2.0))
    o = nets.binary3(onet.Network).create_belief_propagation_fn()
    slots = [o.slot(1, "tonic_volatility"), o.slot(-1, "response_parameter", 0)]
    want_s, want_g = o.batch_grad(u, slots, np.stack([om2, np.full(n, 2.0)]), "binary_softmax_inverse_temperature", resp_inputs=y)
    assert np.isclose(float(total), -want_s.sum(), rtol=RTOL_SUM)
    assert np.allclose(grads["tonic_volatility_2"].double().cpu().numpy(), -want_g[0], rtol=1e-3, atol=1e-4 * np.abs(want_g[0]).max())
    pw = HGFPointwise(n_levels=3, model_type="binary", input_data=uu, response_function=R.binary_softmax_inverse_temperature,
                      response_function_inputs=yy)(tonic_volatility_2=om2, response_function_parameters=np.full(n, 2.0))
    assert tuple(pw.shape) == (n, 320)
    assert np.allclose(pw.sum(1).double().cpu().numpy(), -want_s, rtol=RTOL_SUM)


# ------------------------------------------------------------------ the C ABI with host buffers
def test_c_abi_host_entry_point(G, datasets):
    from pyhgf_b200 import _lib

    u = datasets["u"].astype(np.float32)
    net = nets.binary3(G.Network)
    net.create_belief_propagation_fn()
    plan = net.plan
    n = 300
    om2 = _grid(n, -6, -1, 1).astype(np.float32)
    attr = net._flatten(plan)
    slots = (C.c_int32 * 1)(plan.slot(1, "tonic_volatility"))
    out = np.zeros(n, np.float32)
    call = _lib.Call()
    call.dtype, call.n_instances, call.n_steps = 32, n, 320
    call.input_data = u.ctypes.data
    call.attributes = attr.ctypes.data
    call.n_params, call.param_slots, call.params = 1, C.cast(slots, C.POINTER(C.c_int32)), om2.ctypes.data
    call.response_fn = _lib.RESPONSES["first_level_binary_surprise"]
    call.surprise_sum = out.ctypes.data
    call.out_scale = 1.0
    before = _lib.launch_count()
    _lib.check(_lib.lib().hgf_filter_host(plan.handle, C.byref(call), 0))
    assert _lib.launch_count() == before + 1
    o = nets.binary3(onet.Network).create_belief_propagation_fn()
    want = o.batch_surprise(u, [o.slot(1, "tonic_volatility")], om2[None, :], "first_level_binary_surprise")
    assert np.allclose(out, want, rtol=RTOL_SUM)
    # invalid arguments are errors, not crashes
    call.n_params = 99
    assert _lib.lib().hgf_filter_host(plan.handle, C.byref(call), 0) != 0
    assert b"n_params" in _lib.lib().hgf_last_error()


def test_empty_batch_and_empty_series(G, R, datasets):
    g = nets.binary3(G.Network)
    res = g.batch(datasets["u"], {(1, "tonic_volatility"): np.zeros(0)}, response_function=R.first_level_binary_surprise,
                  n_instances=0)
    assert res.surprise.numel() == 0
    res = g.batch(np.zeros(0), {(1, "tonic_volatility"): np.array([-3.0, -4.0])}, response_function=R.first_level_binary_surprise)
    assert np.array_equal(res.surprise.cpu().numpy(), np.zeros(2, np.float32))


# ------------------------------------------------------------------ full-size properties (BASELINE configs)
def test_config3_grid_properties(G, R, datasets):
    """1M-point omega grid x 320 trials: spot rows against the oracle, determinism, permutation equivariance."""
    import torch

    u = datasets["u"]
    side = 1024
    om2 = np.repeat(np.linspace(-8, 0, side), side)
    om3 = np.tile(np.linspace(-8, 0, side), side)
    g = nets.binary3(G.Network)
    p = {(1, "tonic_volatility"): torch.as_tensor(om2, dtype=torch.float32, device="cuda"),
         (2, "tonic_volatility"): torch.as_tensor(om3, dtype=torch.float32, device="cuda")}
    a = g.batch(u, p, response_function=R.first_level_binary_surprise).surprise
    b = g.batch(u, p, response_function=R.first_level_binary_surprise).surprise
    assert torch.equal(a, b) or torch.equal(torch.nan_to_num(a), torch.nan_to_num(b))
    perm = torch.randperm(side * side, device="cuda")
    c = g.batch(u, {k: v[perm] for k, v in p.items()}, response_function=R.first_level_binary_surprise).surprise
    assert torch.equal(torch.nan_to_num(a[perm], nan=-1.0), torch.nan_to_num(c, nan=-1.0))
    idx = np.random.default_rng(0).choice(side * side, 512, replace=False)
    o = nets.binary3(onet.Network).create_belief_propagation_fn()
    oslots = [o.slot(1, "tonic_volatility"), o.slot(2, "tonic_volatility")]
    ps = np.stack([om2[idx], om3[idx]]).astype(np.float32).astype(np.float64)
    want, sens = conditioning.sensitivity(lambda q: o.batch_surprise(u, oslots, q, "first_level_binary_surprise"), ps)
    got = a[torch.as_tensor(idx, device="cuda")].double().cpu().numpy()
    finite = np.isfinite(want)
    assert np.array_equal(finite, np.isfinite(got))
    # the grid reaches omega = 0: beyond omega = -2.5 the beliefs saturate and float32 itself is not reproducible
    # (reference tests/test_model.py:237-268 documents that regime); everything below it is held to 1e-4 per instance
    calm = finite & (om2[idx] <= -2.5) & (om3[idx] <= -2.5)
    assert calm.sum() > 100
    assert_conditioned(got[calm], want[calm], sens[calm], RTOL_SUM, what="config3 sample", max_relaxed=0.05)
    rel = np.abs(got[finite] - want[finite]) / np.abs(want[finite])
    assert np.median(rel) < 1e-5


def test_config5_trajectories_layout(G, datasets):
    """16-node gHGF with recorded trajectories: every recorded leaf of a batch instance equals the single run."""
    x = nets.ar1_series(300)
    g = nets.ghgf16(G.Network)
    n = 260
    rng = np.random.default_rng(1)
    params = {(i, "tonic_volatility"): rng.normal(-4, 0.5, n) for i in range(4, 12)}
    record = [(i, k) for i in range(16) for k in nets.BELIEFS]
    res = g.batch(x, params, record=record)
    for n_i in (0, 129, 259):
        single = nets.ghgf16(onet.Network)
        for i in range(4, 12):
            single.attributes[i]["tonic_volatility"] = float(np.float32(params[(i, "tonic_volatility")][n_i]))
        single.input_data(x)
        for i in (4, 9, 12, 15):
            for k in nets.BELIEFS:
                got = res.trajectories[(i, k)][n_i].double().cpu().numpy()
                assert np.allclose(got, single.node_trajectories[i][k], rtol=RTOL_TRAJ, atol=1e-6), (n_i, i, k)


@pytest.mark.parametrize("net_name,n", [("ghgf16", 260), ("ghgf16", 1000), ("binary3", 64), ("multi_child", 333), ("ghgf16_all", 200)])
@pytest.mark.parametrize("dtype", ["float32", "float64"])
def test_tma_trajectory_writeback_is_bit_identical(G, datasets, net_name, n, dtype, monkeypatch):
    """Trajectories written through shared-memory staging + one TMA tensor store per block-step equal, bit for
    bit, those written with plain coalesced stores (HGF_NO_TMA=1) -- including the ragged last block (n not a
    multiple of 64, clipped by the tensor map) and batches whose row stride is not 16-byte aligned (n = 333:
    the library must fall back to plain stores by itself)."""
    import torch

    rng = np.random.default_rng(7)
    if net_name == "binary3":
        data, build, nodes = datasets["u"], nets.binary3, (1, 2)
    elif net_name == "multi_child":  # interpreter path, two input columns
        data, build, nodes = nets.ar1_series(200, n_series=2), nets.multi_child, (2, 3, 4, 5)
    else:
        data, build, nodes = nets.ar1_series(150), nets.ghgf16, tuple(range(4, 12))
    params = {(i, "tonic_volatility"): rng.normal(-4, 0.5, n) for i in nodes}
    record = [(i, k) for i in nodes for k in nets.BELIEFS]
    if net_name == "ghgf16_all":  # all 80 leaves: 80 KB of staging tiles in float64 (> 48 KB: opt-in dynamic shared memory)
        record = [(i, k) for i in range(16) for k in nets.BELIEFS + (("mean_vol", "expected_mean_vol", "precision_vol", "expected_precision_vol") if i >= 12 else ())]
    out = {}
    for flag in ("0", "1"):
        monkeypatch.setenv("HGF_NO_TMA", flag)
        res = build(G.Network, dtype=dtype).batch(data, params, record=record)
        torch.cuda.synchronize()
        out[flag] = {k: res.trajectories[k].clone() for k in record}
    for k in record:
        a, b = out["0"][k], out["1"][k]
        assert a.shape == (n, data.shape[0])
        assert torch.equal(torch.nan_to_num(a, nan=123.0), torch.nan_to_num(b, nan=123.0)), k


def test_fused_all_reduce_two_gpus():
    """The in-kernel all-reduce over NVLink peer memory against NCCL (tools/fused_reduce_check.py under torchrun):
    needs two GPUs, so it is skipped on the single-GPU box."""
    import os
    import subprocess
    import sys

    import torch

    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29533", os.path.join(root, "tools", "fused_reduce_check.py")], capture_output=True, text=True,
                       timeout=600, cwd=root)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert '"max_rel_err_vs_nccl"' in r.stdout


def test_to_pandas_matches_the_notebook_goldens_and_the_oracle(G, datasets):
    """`Network.to_pandas()` on the CUDA trajectories: 1.1-Binary_HGF.ipynb cell 27, 1.3-Continuous_HGF.ipynb cell 16."""
    g = nets.binary3(G.Network).input_data(datasets["u"])
    o = nets.binary3(onet.Network).input_data(datasets["u"])
    dg, do = g.to_pandas(), o.to_pandas()
    assert list(dg.columns) == list(do.columns)
    for col, gold in (("x_0_surprise", 206.60437), ("x_1_surprise", 291.26907), ("x_2_surprise", 238.23419)):
        assert np.isclose(dg[col].sum(), gold, rtol=RTOL_SUM), (col, dg[col].sum())
        assert np.allclose(dg[col], do[col], rtol=1e-4, atol=1e-5)
    x = datasets["x"]
    build = lambda cls: (cls().add_nodes(precision=1e4).add_nodes(precision=1e4, mean=x[0], tonic_volatility=-13.0, value_children=0)  # noqa: E731
                         .add_nodes(precision=1e1, tonic_volatility=-2.0, volatility_children=1).input_data(x))
    dg = build(G.Network).to_pandas()
    for col, gold in (("x_0_surprise", -2071.4644), ("x_1_surprise", -2446.2734), ("x_2_surprise", 665.03186)):
        assert np.isclose(dg[col].sum(), gold, rtol=RTOL_SUM), (col, dg[col].sum())
    assert np.isclose(dg.total_surprise.sum(), -2071.4644 - 2446.2734 + 665.03186, rtol=RTOL_SUM)


# ------------------------------------------------------------------ BASELINE configs at full size: size-independent properties
def _config4(n_part=1024, n_chain=64, T=320, seed=0):
    rng = np.random.default_rng(seed)
    blocks = rng.integers(0, 2, size=(n_part, (T + 39) // 40))
    p = np.where(np.repeat(blocks, 40, axis=1)[:, :T] == 1, 0.8, 0.2)
    u = (rng.uniform(size=p.shape) < p).astype(np.float32)
    y = (rng.uniform(size=p.shape) < 0.3 + 0.4 * p).astype(np.float32)
    n = n_part * n_chain
    params = np.stack([rng.uniform(-6, -3, n), rng.uniform(-6, -3, n), np.exp(rng.uniform(-0.3, 0.8, n))]).astype(np.float32)
    return u, y, params


def test_config4_full_size_properties(G, R):
    """1 024 participants x 64 chains x 320 trials, logp + gradient in one launch (BASELINE config 4):
    (i) the per-chain sums reduced on the device equal the sums of the per-instance logp; (ii) an instance's logp and
    gradient do not depend on the batch around it (a 2-participant sub-batch reproduces them bit for bit);
    (iii) the gradient wrt the inverse temperature is the derivative of logp (central difference in float64 on a
    sample); (iv) a random sample matches the oracle."""
    import torch

    n_part, n_chain = 1024, 64
    u, y, p = _config4(n_part, n_chain)
    n = n_part * n_chain
    keys = [(1, "tonic_volatility"), (2, "tonic_volatility"), ("response_parameter", 0)]
    group = torch.arange(n, device="cuda", dtype=torch.int32) % n_chain

    def run(u_, y_, p_, dtype="float32", **kw):
        return nets.binary3(G.Network, dtype=dtype).batch(u_, {k: p_[i] for i, k in enumerate(keys)},
                                                          response_function=R.binary_softmax_inverse_temperature, response_function_inputs=y_,
                                                          input_group_size=n_chain, response_group_size=n_chain, gradient=True, out_scale=-1.0, **kw)

    res = run(u, y, p, group=group, n_groups=n_chain)
    logp = res.surprise  # out_scale = -1: log-probabilities
    assert logp.shape == (n,) and torch.isfinite(logp).all()
    want = logp.double().view(n_part, n_chain).sum(0)
    assert torch.allclose(res.reduced.double(), want, rtol=2e-5)                      # (i) float32 atomics vs float64 sum
    lo = 517
    sub = run(u[lo:lo + 2], y[lo:lo + 2], p[:, lo * n_chain:(lo + 2) * n_chain])
    assert torch.equal(sub.surprise, logp[lo * n_chain:(lo + 2) * n_chain])           # (ii)
    for k in keys:
        assert torch.equal(sub.gradient[k], res.gradient[k][lo * n_chain:(lo + 2) * n_chain])
    # (iii) every gradient is the derivative of logp: central differences in float64, all three parameters
    sel = slice(lo * n_chain, lo * n_chain + 32)
    p64 = p[:, sel].astype(np.float64)
    u1, y1 = u[lo:lo + 1], y[lo:lo + 1]
    f = lambda q: run(u1, y1, q, dtype="float64").surprise.cpu().numpy()  # noqa: E731
    r64 = run(u1, y1, p64, dtype="float64")
    for j, k in enumerate(keys):
        h = 1e-6
        up, dn = p64.copy(), p64.copy()
        up[j] += h
        dn[j] -= h
        fd = (f(up) - f(dn)) / (2 * h)
        g64 = r64.gradient[k].cpu().numpy()
        assert np.allclose(g64, fd, rtol=2e-5, atol=2e-5 * np.abs(fd).max()), (k, np.abs(g64 - fd).max())
        g32 = res.gradient[k][sel].double().cpu().numpy()
        assert np.median(np.abs(g32 - g64) / (np.abs(g64) + 1e-3 * np.abs(g64).max())) < 1e-4, k
    # (iv) a random sample against the oracle, every sampled instance
    idx = np.sort(np.random.default_rng(3).choice(n, 256, replace=False))
    o = nets.binary3(onet.Network).create_belief_propagation_fn()
    slots = [o.slot(1, "tonic_volatility"), o.slot(2, "tonic_volatility"), o.slot(-1, "response_parameter", 0)]
    got = -logp[torch.as_tensor(idx, device="cuda")].double().cpu().numpy()

    def oracle(q):  # instance i of the sample reads participant idx[i] // n_chain's series
        return np.array([o.batch_surprise(u[i // n_chain], slots, q[:, [c]], "binary_softmax_inverse_temperature", resp_inputs=y[i // n_chain])[0]
                         for c, i in enumerate(idx)])

    want, sens = conditioning.sensitivity(oracle, p[:, idx].astype(np.float64), rows=[0, 1])
    # config 4's chains spread over omega in [-8, -1]: the saturating part of that box is not reproducible in
    # float32 (the bound follows the oracle's own sensitivity there); no instance is dropped
    assert_conditioned(got, want, sens, RTOL_SUM, what="config4 sample")


def test_config2_full_size_properties(G, R, datasets):
    """65 536 tonic-volatility samples x 614 observations, 4-node continuous HGF (BASELINE config 2): deterministic,
    batch-independent, and a random sample matches the oracle."""
    import torch

    x = datasets["x"]
    n = 65536
    rng = np.random.default_rng(0)
    om = [np.clip(rng.normal(m, 1.0, n), m - 3, m + 3).astype(np.float32) for m in (-13.0, -2.0, -2.0)]
    keys = [(i, "tonic_volatility") for i in (1, 2, 3)]
    run = lambda cols: nets.continuous3(G.Network).batch(x, dict(zip(keys, cols)), response_function=R.first_level_gaussian_surprise).surprise  # noqa: E731
    a, b = run(om), run(om)
    assert torch.equal(torch.nan_to_num(a, nan=1.0, posinf=2.0), torch.nan_to_num(b, nan=1.0, posinf=2.0))
    sl = slice(40000, 40100)
    c = run([v[sl] for v in om])
    assert torch.equal(torch.nan_to_num(a[sl], nan=1.0, posinf=2.0), torch.nan_to_num(c, nan=1.0, posinf=2.0))
    idx = rng.choice(n, 300, replace=False)
    o = nets.continuous3(onet.Network).create_belief_propagation_fn()
    slots = [o.slot(i, "tonic_volatility") for i in (1, 2, 3)]
    want = o.batch_surprise(x, slots, np.stack([v[idx] for v in om]).astype(np.float64), "first_level_gaussian_surprise")
    got = a[torch.as_tensor(idx, device="cuda")].double().cpu().numpy()
    fin = np.isfinite(want)
    assert np.array_equal(fin, np.isfinite(got))
    assert np.allclose(got[fin], want[fin], rtol=RTOL_SUM, atol=1e-2)


def test_config5_full_size_properties(G):
    """16-node gHGF, 32 768 instances (one GPU's share of config 5) x 400 steps with all 80 belief leaves recorded
    through the TMA write-back: instances with equal parameters give equal trajectories wherever they sit in the batch,
    `expected_mean` of an autoconnected node at step t equals its `mean` at step t-1, and precisions stay positive."""
    import torch

    T, n = 400, 32768
    x = nets.ar1_series(T)
    rng = np.random.default_rng(1)
    base = {i: rng.normal(-4, 1.0, n).astype(np.float32) for i in range(4, 12)}
    for v in base.values():
        v[n - 37] = v[5]  # a twin of instance 5 in the ragged tail of another block
    params = {(i, "tonic_volatility"): base[i] for i in range(4, 12)}
    record = [(i, k) for i in range(16) for k in nets.BELIEFS + (("mean_vol", "expected_mean_vol", "precision_vol", "expected_precision_vol") if i >= 12 else ())]
    assert len(record) == 80
    res = nets.ghgf16(G.Network).batch(x, params, record=record)
    for k in record:
        tr = res.trajectories[k]
        assert tr.shape == (n, T)
        assert torch.equal(torch.nan_to_num(tr[5], nan=7.0), torch.nan_to_num(tr[n - 37], nan=7.0)), k
    for node in (4, 5, 6, 7):  # value parents: lambda = 1, no drift, no value parents of their own ... except 12/13's coupling
        pr = res.trajectories[(node, "precision")]
        assert (pr[torch.isfinite(pr)] > 0).all()
    for node in (8, 9, 10, 11):  # volatility parents: mu_hat(t) = mu(t-1)
        m, mh = res.trajectories[(node, "mean")], res.trajectories[(node, "expected_mean")]
        ok = torch.isfinite(m[:, :-1]) & torch.isfinite(mh[:, 1:])
        assert torch.equal(m[:, :-1][ok], mh[:, 1:][ok])


@pytest.mark.parametrize("build", [nets.multi_child, nets.volatile2, nets.nonlinear], ids=lambda b: b.__name__)
def test_on_demand_specialisation_matches_the_interpreter(G, R, build, monkeypatch):
    """A user topology specialised on demand (register-resident kernels, pyhgf_b200/specialise.py) gives the
    interpreter's trajectories, logp and gradients."""
    import torch

    n_in = len(build(G.Network).input_idxs)
    x = nets.ar1_series(200, n_series=n_in)
    x = x if n_in > 1 else x[:, 0]
    n = 300
    rng = np.random.default_rng(3)
    cont = [i for i, e in enumerate(build(G.Network).edges) if e.node_type == 2 and (e.value_children or e.volatility_children)]
    params = {(i, "tonic_volatility"): rng.normal(-4, 0.3, n) for i in cont[:2]} or {(1, "tonic_volatility_vol"): rng.normal(-4, 0.3, n)}
    record = [(i, k) for i in range(len(build(G.Network).edges)) for k in nets.BELIEFS]
    out = {}
    for mode in ("interpreter", "specialised"):
        net = build(G.Network)
        if mode == "specialised":
            name = net.specialise(gradient_parameters=len(params))
            assert net.plan.kernel_name == name
        else:
            monkeypatch.setenv("HGF_FORCE_INTERPRETER", "1")
            net.create_belief_propagation_fn()
            assert net.plan.kernel_name == "interpreter"
        res = net.batch(x, params, record=record, response_function=R.total_gaussian_surprise, gradient=True)
        assert res.kernel_name == net.plan.kernel_name
        torch.cuda.synchronize()
        out[mode] = res
        monkeypatch.delenv("HGF_FORCE_INTERPRETER", raising=False)
    a, b = out["interpreter"], out["specialised"]
    assert torch.allclose(a.surprise, b.surprise, rtol=1e-5, equal_nan=True)
    for k in params:
        scale = float(a.gradient[k].abs().max())
        assert torch.allclose(a.gradient[k], b.gradient[k], rtol=1e-3, atol=1e-4 * scale, equal_nan=True), k
    for k in record:
        assert torch.allclose(a.trajectories[k], b.trajectories[k], rtol=RTOL_TRAJ, atol=1e-6, equal_nan=True), k


@pytest.mark.parametrize("scheme", ["unbounded", "eHGF"])
def test_on_demand_specialisation_with_reverse_sweep(G, R, scheme):
    """A USER topology (two inputs under a shared value parent, itself under a value parent and a two-level volatility
    chain) specialised with `reverse=True`: the sweep generated and compiled on demand gives the Dual<double> oracle's
    gradient wrt 7 parameters in one launch, under the default scheme and under the eHGF."""
    from pyhgf_b200 import _lib, revgen

    x = nets.ar1_series(200, n_series=2)
    n = 64
    keys = [(2, "tonic_volatility"), (3, "tonic_volatility"), (4, "tonic_volatility"), (5, "tonic_volatility"), (2, "value_coupling_children", 1),
            (4, "volatility_coupling_children", 0), (3, "mean")]
    vals = [_grid(n, -4, -2, 1), _grid(n, -6, -4, 2), _grid(n, -4, -2, 3), _grid(n, -5, -3, 4), _grid(n, 0.5, 0.9, 5), _grid(n, 0.8, 1.2, 6),
            _grid(n, -0.3, 0.3, 7)]
    net = nets.multi_child(G.Network, dtype="float64", volatility_updates=scheme)
    name = net.specialise(float64=True, reverse=True)
    assert name.startswith("specialised:") and net.plan.kernel_name == name
    before = _lib.launch_count()
    res = net.batch(x, dict(zip(keys, vals)), response_function=R.total_gaussian_surprise, gradient=True)
    assert res.kernel_name == name and _lib.launch_count() - before == 1
    o = nets.multi_child(onet.Network, dtype="float64", volatility_updates=scheme).create_belief_propagation_fn()
    want_s, want_g = o.batch_grad(x, [o.slot(*k) for k in keys], np.stack(vals), "total_gaussian_surprise")
    got_s = res.surprise.cpu().numpy()
    ok = np.isfinite(want_s)  # (under the eHGF a few of these filters blow up: +inf surprise on both sides, no gradient to compare)
    assert ok.sum() > n // 2 and np.array_equal(np.isfinite(got_s), ok)
    assert np.allclose(got_s[ok], want_s[ok], rtol=1e-9)
    for i, k in enumerate(keys):
        assert np.allclose(res.gradient[k].cpu().numpy()[ok], want_g[i][ok], rtol=1e-7, atol=1e-7 * np.abs(want_g[i][ok]).max()), k
    assert revgen.Unsupported is not None  # (node kinds outside the scan -- ef-state, categorical -- raise it; none is built here)


def test_shared_reciprocals_against_a_build_without_them():
    """hgf_math.cuh::Recip -- one MUFU.RCP per denominator of the unbounded update serving a / x, c / x and a / (2 x) -- against
    a build of the same kernels WITHOUT it (-DHGF_NO_SHARED_RCP): the gradient kernels return the same bits, logp and
    gradients; the plain filter differs by the placement of one rounding (nvcc fuses the other product of a two-product
    sum) and stays within 1e-5 on 99 % of the instances; two builds with the same flags are identical
    (tools/shared_rcp_check.py, in its own process: a registered specialisation would serve every later plan of the
    topology in this one; tools/ubench/rcp_scaling.cu is the primitive's own check)."""
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "tools", "shared_rcp_check.py")], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert not any(line["control_same_flags_differing"].values()), line
    assert not any(v for w, v in line["values_differing_in_any_bit"].items() if w != "filter logp"), line


def test_generative_rollouts(G, datasets):
    """`Network.sample` (pyhgf/utils/sample.py:18-116): in-kernel draws from the input nodes' predictions.
    The random stream is Philox, not JAX's threefry, so the check is (i) the draws have the predicted distribution,
    (ii) a rollout is exactly the filter run on its own samples (fed to the oracle as external observations),
    (iii) same key -> same rollouts, another key -> different ones, (iv) rollouts start from `last_attributes`."""
    import torch

    n, T = 8192, 40
    g = nets.binary3(G.Network)
    pred = g.sample(time_steps=np.ones(T), n_predictions=n, rng_key=7)
    u = pred[0]["mean"]
    assert u.shape == (n, T) and bool(((u == 0) | (u == 1)).all())
    mu0 = pred[0]["expected_mean"][:, 0]
    assert torch.allclose(mu0, torch.full_like(mu0, 0.5))                      # default prior: sigmoid(0)
    assert abs(float(u[:, 0].mean()) - 0.5) < 4 * 0.5 / n ** 0.5               # (i) Bernoulli(0.5)
    # calibration over all steps: E[u - mu_hat] = 0, Var = E[mu_hat (1 - mu_hat)]
    resid = (u - pred[0]["expected_mean"]).double()
    assert abs(float(resid.mean())) < 4 * 0.5 / (n * T) ** 0.5
    for i in (0, 4097, n - 1):                                                  # (ii)
        o = nets.binary3(onet.Network).input_data(u[i].cpu().numpy().astype(float))
        for node in (1, 2):
            for k in nets.BELIEFS:
                got = pred[node][k][i].double().cpu().numpy()
                assert np.allclose(got, o.node_trajectories[node][k], rtol=RTOL_TRAJ, atol=1e-6), (i, node, k)
    again = g.sample(time_steps=np.ones(T), n_predictions=n, rng_key=7)        # (iii)
    other = g.sample(time_steps=np.ones(T), n_predictions=n, rng_key=8)
    assert torch.equal(again[0]["mean"], u) and not torch.equal(other[0]["mean"], u)

    # continuous input: standardised residuals of the draws are N(0, 1)
    c = nets.continuous2(G.Network).input_data(datasets["x"][:100])            # (iv) start from the filtered state
    pred = c.sample(time_steps=np.ones(20), n_predictions=n, rng_key=1)
    x, m, p = (pred[0][k].double() for k in ("mean", "expected_mean", "expected_precision"))
    zed = (x - m) * p.sqrt()
    assert abs(float(zed.mean())) < 4 / (n * 20) ** 0.5
    assert abs(float(zed.var()) - 1.0) < 0.05
    assert abs(float((zed.abs() < 1.96).double().mean()) - 0.95) < 0.01
    want = float(c.last_attributes[1]["mean"])                                  # lambda = 1, no drift: mu_hat(first step) = last mean
    assert torch.allclose(pred[1]["expected_mean"][:, 0].double(), torch.full((n,), want, dtype=torch.float64, device="cuda"), rtol=1e-6)


@pytest.mark.parametrize("n_part,n_chain,shared", [(800, 50, False), (1, 40000, True), (1024, 64, False)])
def test_host_entry_point_chunk_pipeline(G, n_part, n_chain, shared, monkeypatch):
    """hgf_logp_grad_host on a large batch: the chunked copy/compute pipeline (HGF_HOST_CHUNKS slices of whole series groups, one stream
    each, ScanArgs::ld row strides) returns exactly what the one-shot path (HGF_HOST_CHUNKS=0) returns, for grouped
    series whose group size does not divide the block size, and for one shared series."""
    from pyhgf_b200 import _lib

    rng = np.random.default_rng(2)
    T = 96
    n = n_part * n_chain
    u = (rng.uniform(size=(n_part, T)) > 0.5).astype(np.float32)
    y = (rng.uniform(size=(n_part, T)) > 0.5).astype(np.float32)
    params = np.stack([rng.uniform(-6, -3, n), rng.uniform(-6, -3, n), np.exp(rng.uniform(-0.3, 0.8, n))]).astype(np.float32)
    group = (np.arange(n) % 7).astype(np.int32)
    net = nets.binary3(G.Network)
    net.create_belief_propagation_fn()
    plan = net.plan
    attr = net._flatten(plan)
    slots = (C.c_int32 * 3)(plan.slot(1, "tonic_volatility"), plan.slot(2, "tonic_volatility"), plan.slot(-1, "response_parameter", 0))
    out = {}
    for flag in ("4", "0", "3"):
        monkeypatch.setenv("HGF_HOST_CHUNKS", flag)
        logp, grad, red = np.zeros(n, np.float32), np.zeros((3, n), np.float32), np.zeros(7, np.float32)
        call = _lib.Call()
        call.dtype, call.n_instances, call.n_steps = 32, n, T
        call.input_group_size = call.response_group_size = 0 if shared else n_chain
        call.input_data, call.response_inputs, call.attributes = u.ctypes.data, y.ctypes.data, attr.ctypes.data
        call.n_params, call.param_slots, call.params = 3, C.cast(slots, C.POINTER(C.c_int32)), params.ctypes.data
        call.response_fn = _lib.RESPONSES["binary_softmax_inverse_temperature"]
        call.surprise_sum, call.gradient, call.reduced, call.group = logp.ctypes.data, grad.ctypes.data, red.ctypes.data, group.ctypes.data
        call.out_scale = -1.0
        before = _lib.launch_count()
        _lib.check(_lib.lib().hgf_logp_grad_host(plan.handle, C.byref(call), 0))
        out[flag] = (logp, grad, red, _lib.launch_count() - before)
    assert out["4"][3] == 4 and out["0"][3] == 1 and out["3"][3] == 3  # slices = launches; 0 is the one-shot path
    for flag in ("4", "3"):
        assert np.array_equal(out[flag][0], out["0"][0]) and np.array_equal(out[flag][1], out["0"][1])
        assert np.allclose(out[flag][2], out["0"][2], rtol=1e-5)       # atomics: order differs
    assert np.isfinite(out["4"][0]).all()


def test_committed_golden_vectors(G, R, datasets):
    """The CUDA path against tests/golden/oracle_vectors.npz (float64 oracle results, committed): float64 launches to
    1e-9, float32 launches to the stated tolerances on the instances that are well-conditioned (float32 and float64
    oracle agree to 1e-5); no oracle code runs in this test."""
    import os

    v = np.load(os.path.join(os.path.dirname(__file__), "golden", "oracle_vectors.npz"))
    for name, build, data in (("binary3", nets.binary3, datasets["u"]), ("continuous3", nets.continuous3, datasets["x"]),
                              ("ghgf16", nets.ghgf16, nets.ar1_series(200))):
        for dtype, rtol, atol in (("float64", 1e-9, 1e-12), ("float32", RTOL_TRAJ, 1e-6)):
            g = build(G.Network, dtype=dtype).input_data(data)
            for node in range(len(g.edges)):
                for k in nets.BELIEFS:
                    got = np.asarray(g.node_trajectories[node][k], dtype=np.float64)
                    assert np.allclose(got, v[f"traj/{name}/{node}/{k}"], rtol=rtol, atol=atol, equal_nan=True), (name, dtype, node, k)
    keys = [(1, "tonic_volatility"), (2, "tonic_volatility"), ("response_parameter", 0)]
    p = v["config4/params"]
    for dtype in ("float64", "float32"):
        res = nets.binary3(G.Network, dtype=dtype).batch(datasets["u"], dict(zip(keys, p)), response_function=R.binary_softmax_inverse_temperature,
                                                         response_function_inputs=datasets["y"], gradient=True)
        s = res.surprise.double().cpu().numpy()
        if dtype == "float64":
            assert np.allclose(s, v["config4/surprise"], rtol=1e-9)
            for i, k in enumerate(keys):
                assert np.allclose(res.gradient[k].cpu().numpy(), v["config4/gradient"][i], rtol=1e-7, atol=1e-9)
        else:
            # float32 against the committed float32 oracle results (Dual<float> gradients), every instance, with the
            # committed per-instance sensitivities of that oracle (tests/golden/make_oracle_vectors.py)
            assert_conditioned(s, v["config4/surprise_f32"], v["config4/sensitivity_f32"][0], RTOL_SUM, what="config4 golden")
            for i, k in enumerate(keys):
                gg, want = res.gradient[k].double().cpu().numpy(), v["config4/gradient_f32"][i]
                assert_conditioned(gg, want, v["config4/sensitivity_f32"][1 + i], RTOL_SUM, amplification=AMPLIFICATION_GRAD,
                                   floor=RTOL_SUM * np.abs(want).max(), what=k)
    keys = [(i, "tonic_volatility") for i in (1, 2, 3)]
    for dtype in ("float64", "float32"):
        res = nets.continuous3(G.Network, dtype=dtype).batch(datasets["x"], dict(zip(keys, v["config2/params"])),
                                                             response_function=R.first_level_gaussian_surprise)
        s, want = res.surprise.double().cpu().numpy(), v["config2/surprise"]
        fin = np.isfinite(want)
        assert np.array_equal(fin, np.isfinite(s))
        if dtype == "float64":
            assert np.allclose(s[fin], want[fin], rtol=1e-8, atol=1e-7)
        else:
            # the summed Gaussian surprise crosses zero (terms of magnitude ~2e3): relative to |s| + 100
            assert_conditioned(s, v["config2/surprise_f32"], v["config2/sensitivity_f32"], RTOL_SUM, floor=100.0, what="config2 golden")


def test_plain_c_caller_of_the_abi(G, datasets, tmp_path):
    """examples/c_abi_binary3.c: a C program (gcc, no Python / torch in the process) builds the 3-level binary HGF's
    topology by hand, calls hgf_plan_create / hgf_filter_host and gets the oracle's surprise."""
    import os
    import subprocess

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib_dir = os.path.join(root, "pyhgf_b200", "lib")
    exe = str(tmp_path / "c_abi_binary3")
    subprocess.run(["gcc", "-O2", "-I", os.path.join(root, "include"), os.path.join(root, "examples", "c_abi_binary3.c"), "-L", lib_dir,
                    "-lpyhgf_cuda", f"-Wl,-rpath,{lib_dir}", "-o", exe], check=True)
    np.savetxt(tmp_path / "u.txt", datasets["u"], fmt="%d")
    n = 257
    r = subprocess.run([exe, str(tmp_path / "u.txt"), str(n), str(tmp_path / "out.txt")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "kernel: static:binary3" in r.stdout and "launches: 1" in r.stdout
    out = np.loadtxt(tmp_path / "out.txt")
    assert out.shape == (n, 3)
    o = nets.binary3(onet.Network).create_belief_propagation_fn()
    want = o.batch_surprise(datasets["u"], [o.slot(1, "tonic_volatility"), o.slot(2, "tonic_volatility")], out[:, :2].T.copy(),
                            "first_level_binary_surprise")
    _, sens = conditioning.sensitivity(lambda q: o.batch_surprise(datasets["u"], [o.slot(1, "tonic_volatility"), o.slot(2, "tonic_volatility")], q,
                                                                  "first_level_binary_surprise"), out[:, :2].T.copy())
    assert_conditioned(out[:, 2], want, sens, RTOL_SUM, what="C caller")
    # the defaults written out by hand in the C file are the front-end's: omega = -4 reproduces the reference golden
    assert np.isfinite(out[:, 2]).all()


def test_auto_specialise_opt_in(G, R, monkeypatch):
    """HGF_AUTO_SPECIALISE=1: the first batched call of a topology without a built-in kernel builds and uses its
    register-resident specialisation instead of the interpreter."""
    x = nets.ar1_series(50, n_series=1)[:, 0]
    p = {(2, "tonic_volatility"): np.full(70, -3.0)}
    a = nets.tanh_coupling(G.Network).batch(x, p, response_function=R.first_level_gaussian_surprise)
    monkeypatch.setenv("HGF_AUTO_SPECIALISE", "1")
    b = nets.tanh_coupling(G.Network).batch(x, p, response_function=R.first_level_gaussian_surprise)
    assert b.kernel_name.startswith("specialised:")
    assert np.allclose(a.surprise.cpu().numpy(), b.surprise.cpu().numpy(), rtol=1e-5, equal_nan=True)


def test_disconnected_network_is_split_over_components(G, monkeypatch):
    """BASELINE config 5's 16-node network has two connected components: `Network.batch` runs them as two concurrent
    8-node launches (register-resident `static:ghgf8`) and returns the trajectories of the single 16-node launch."""
    import torch

    net = nets.ghgf16(G.Network)
    assert net.connected_components() == [[0, 1, 4, 5, 8, 9, 12, 14], [2, 3, 6, 7, 10, 11, 13, 15]]
    x = nets.ar1_series(300)
    n = 700
    rng = np.random.default_rng(4)
    params = {(i, "tonic_volatility"): rng.normal(-4, 0.5, n) for i in range(4, 12)}
    params.update({(i, "tonic_volatility_vol"): rng.normal(-4, 0.5, n) for i in range(12, 16)})
    record = [(i, k) for i in range(16) for k in nets.BELIEFS + (("mean_vol", "precision_vol") if i >= 12 else ())]
    obs = (rng.uniform(size=x.shape) > 0.05).astype(np.uint8)
    out = {}
    for name, split, fuse in (("fused", "1", "1"), ("streams", "1", "0"), ("single", "0", "1")):
        monkeypatch.setenv("HGF_SPLIT_COMPONENTS", split)
        monkeypatch.setenv("HGF_FUSE_COMPONENTS", fuse)
        res = nets.ghgf16(G.Network).batch(x, params, record=record, observed=obs, time_steps=np.full(300, 0.5))
        res.relaunch()
        torch.cuda.synchronize()
        out[name] = res
    # the two halves are isomorphic: ONE launch of the 8-node kernel over 2 x 700 threads; with fusing off, one launch per
    # component on its own stream; with splitting off, the 16-node kernel
    assert out["fused"].kernel_name == "components(2) in one launch: static:ghgf8"
    assert out["streams"].kernel_name == "components(2): static:ghgf8 | static:ghgf8" and out["single"].kernel_name == "static:ghgf16"
    for k in record:
        a, b, c = out["fused"].trajectories[k], out["single"].trajectories[k], out["streams"].trajectories[k]
        assert a.shape == b.shape == c.shape == (n, 300)
        assert torch.equal(torch.nan_to_num(a), torch.nan_to_num(c)), k   # the same kernel, launched once or per component
        # the 8-node and the 16-node kernel are two float32 evaluations in different instruction order (missing observations,
        # dt = 0.5, 300 steps): held to 1e-4 here; each is held to 1e-5 against the oracle in test_config5_trajectories_layout
        assert torch.allclose(a, b, rtol=1e-4, atol=1e-5, equal_nan=True), k
    # a connected network, a call with a response function or a gradient are not split
    assert nets.binary3(G.Network).connected_components() == [[0, 1, 2]]


def test_wide_x10_network_runs_as_one_launch_of_the_component_kernel(G, datasets):
    """The reference's "wide (x10)" benchmark network (docs/paper.ipynb cell 46: ten 3-level binary HGFs side by side, 30
    nodes, each fed its own input column): the ten components of every instance run as 10 x N threads of the
    register-resident 3-level kernel in one launch, and every component's trajectories are the single 3-level network's."""
    import torch

    u = datasets["u"]
    wide = G.Network()
    for i in range(10):
        wide.add_nodes(kind="binary-state").add_nodes(value_children=3 * i).add_nodes(volatility_children=3 * i + 1)
    n = 40
    rng = np.random.default_rng(9)
    series = np.stack([np.roll(u, 7 * i) for i in range(10)], axis=1)  # a different column per component
    params = {(3 * i + 1, "tonic_volatility"): rng.uniform(-6, -3.5, n) for i in range(10)}
    params.update({(3 * i + 2, "tonic_volatility"): rng.uniform(-6, -3.5, n) for i in range(10)})
    record = [(j, k) for j in range(30) for k in ("mean", "expected_mean")]
    res = wide.batch(series, params, record=record)
    assert res.kernel_name == "components(10) in one launch: static:binary3"
    for i in (0, 4, 9):
        one = nets.binary3(G.Network).batch(series[:, i], {(1, "tonic_volatility"): params[(3 * i + 1, "tonic_volatility")],
                                                            (2, "tonic_volatility"): params[(3 * i + 2, "tonic_volatility")]},
                                            record=[(j, k) for j in range(3) for k in ("mean", "expected_mean")])
        for j in range(3):
            for k in ("mean", "expected_mean"):
                assert torch.equal(torch.nan_to_num(res.trajectories[(3 * i + j, k)]), torch.nan_to_num(one.trajectories[(j, k)])), (i, j, k)
    o = nets.binary3(onet.Network, om2=float(np.float32(params[(13, "tonic_volatility")][5])), om3=float(np.float32(params[(14, "tonic_volatility")][5])))
    o.input_data(series[:, 4])
    got = res.trajectories[(13, "mean")][5].double().cpu().numpy()
    assert np.allclose(got, o.node_trajectories[1]["mean"], rtol=RTOL_TRAJ, atol=1e-6)


# ---------------------------------------------------------------- ef-state nodes (hgf_ef_filter)
def test_ef_state_reference_golden(G):
    """tests/test_nodes/test_exponential_family.py:14-66 through the drop-in surface: same calls, same golden."""
    from oracle import ef_state as ef

    x = ef.spiral_data()
    uni = G.Network().add_nodes(kind="ef-state", learning="generalised-filtering").input_data(input_data=x[:, 0])
    assert np.isfinite(uni.node_trajectories[0]["xis"]).all() and uni.node_trajectories[0]["xis"].shape == (1000, 2)
    net = G.Network().add_nodes(kind="ef-state", nus=3.0, learning="generalised-filtering",
                                distribution="multivariate-normal", dimension=2).input_data(input_data=x)
    golden = np.array([3.4652710e01, -1.0609777e00, 1.2103647e03, -3.6398651e01, 3.3951855e00], dtype=np.float32)
    got = net.node_trajectories[0]["xis"]
    assert got.dtype == np.float32 and np.isclose(got[-1], golden).all()
    want = ef.ef_filter(x, 3.0, 2, "multivariate-normal")
    assert np.allclose(got, want, rtol=RTOL_TRAJ, atol=1e-6)
    assert np.allclose(net.last_attributes[0]["xis"], want[-1], rtol=RTOL_TRAJ, atol=1e-6)
    assert np.allclose(net.node_trajectories[0]["observation_ss"][7], ef.sufficient_statistics(x[7], 2), rtol=1e-6)
    assert np.allclose(net.node_trajectories[0]["mean"], x.astype(np.float32))


@pytest.mark.parametrize("bits", [32, 64])
@pytest.mark.parametrize("d", [1, 2, 3, 5, 8])
def test_ef_state_batch_matches_oracle(G, d, bits):
    """Instances that differ in nus and in their series, with unobserved steps, every built dimension."""
    import torch
    from oracle import ef_state as ef

    rng = np.random.default_rng(40 + d)
    dtype, name = (np.float32, "float32") if bits == 32 else (np.float64, "float64")
    Gs, g, T = 3, 67, 41  # 201 instances: not a multiple of the block size
    x = rng.normal(0.5, 2.0, size=(Gs, T, d))
    mask = rng.random((Gs, T)) > 0.2
    nus = rng.uniform(0.5, 20.0, Gs * g)
    dist = "normal" if d == 1 else "multivariate-normal"
    net = G.Network(dtype=name).add_nodes(kind="ef-state", dimension=d, distribution=dist)
    r = net.batch(x, {(0, "nus"): nus}, observed=mask, input_group_size=g, record=[(0, "xis")], last_attributes=True)
    traj = r.trajectories[(0, "xis")].cpu().numpy()
    last = r.last_attributes[(0, "xis")].cpu().numpy()
    K = d + d * (d + 1) // 2
    assert traj.shape == (Gs * g, T, K) and last.shape == (Gs * g, K)
    rtol = RTOL_TRAJ if bits == 32 else 1e-12
    for n in (0, 1, 66, 67, 130, 200):
        want = ef.ef_filter(x[n // g], dtype(nus[n]), d, dist, observed=mask[n // g], dtype=dtype)
        assert np.allclose(traj[n], want, rtol=rtol, atol=rtol), (n, np.abs(traj[n] - want).max())
        assert np.array_equal(last[n], traj[n, -1])
    # a shared series, no mask, no trajectory: only the final statistics come back
    r2 = net.batch(x[0], {(0, "nus"): nus[:50]}, last_attributes=True)
    assert r2.trajectories == {}
    want = ef.ef_filter(x[0], dtype(nus[7]), d, dist, dtype=dtype)
    assert np.allclose(r2.last_attributes[(0, "xis")][7].cpu().numpy(), want[-1], rtol=rtol, atol=rtol)
    assert torch.cuda.is_available()


def test_ef_state_two_input_nodes_and_errors(G):
    from oracle import ef_state as ef
    from pyhgf_b200 import _lib

    rng = np.random.default_rng(5)
    x = rng.normal(size=(30, 3))
    net = (G.Network().add_nodes(kind="ef-state", nus=2.0)
           .add_nodes(kind="ef-state", nus=9.0, dimension=2, distribution="multivariate-normal"))
    before = _lib.launch_count()
    net.input_data(x)
    assert _lib.launch_count() - before == 2  # one launch per node
    assert np.allclose(net.node_trajectories[0]["xis"], ef.ef_filter(x[:, 0], 2.0, 1, "normal"), rtol=RTOL_TRAJ, atol=1e-6)
    assert np.allclose(net.node_trajectories[1]["xis"], ef.ef_filter(x[:, 1:], 9.0, 2, "multivariate-normal"), rtol=RTOL_TRAJ, atol=1e-6)
    with pytest.raises(ValueError, match="column"):
        net.input_data(x[:, :2])
    c = _lib.EfCall()
    c.dtype, c.n_instances, c.n_steps, c.dimension = 32, 4, 3, 9
    assert _lib.lib().hgf_ef_filter(C.byref(c), None) != 0 and b"dimension" in _lib.lib().hgf_last_error()
    c.dimension = 2
    assert _lib.lib().hgf_ef_filter(C.byref(c), None) != 0 and b"null input" in _lib.lib().hgf_last_error()
    c.n_instances = 0
    assert _lib.lib().hgf_ef_filter(C.byref(c), None) == 0  # empty batch: nothing to do


# ---------------------------------------------------------------- input_custom_sequence (switching plans)
def test_custom_sequence_reference_case(G, datasets):
    """tests/test_model.py:270-317: two copies of the network's own sequence, random branch indexes -- the result is
    the plain filter's."""
    u = datasets["u"]
    net = nets.binary3(G.Network).create_belief_propagation_fn()
    seq = net.update_sequence
    idx = np.random.default_rng(3).binomial(1, 0.5, len(u))
    net.scan_fn = None
    net.input_custom_sequence(update_branches=(seq, seq), branches_idx=idx, input_data=u)
    plain = nets.binary3(G.Network).input_data(u)
    for node in range(3):
        for k in ("mean", "precision", "expected_mean", "expected_precision"):
            assert np.allclose(net.node_trajectories[node][k], plain.node_trajectories[node][k], rtol=RTOL_TRAJ, atol=1e-7, equal_nan=True)


@pytest.mark.parametrize("dtype", ["float32", "float64"])
@pytest.mark.parametrize("build", [nets.binary3, nets.continuous3], ids=["binary3", "continuous3"])
def test_custom_sequence_matches_oracle(G, datasets, build, dtype):
    """Branches that drop the top level's update, or every update: trajectories vs the oracle's switching scan."""
    rng = np.random.default_rng(11)
    u = datasets["u"][:200] if build is nets.binary3 else datasets["x"][:200]
    net = build(G.Network, dtype=dtype).create_belief_propagation_fn()
    o = build(onet.Network, dtype=dtype).create_belief_propagation_fn()
    top = len(net.edges) - 1
    gseq, oseq = net.update_sequence, o.update_sequence
    gb = (gseq, (gseq.prediction_steps, tuple(s for s in gseq.update_steps if s[0] != top)), (gseq.prediction_steps, ()))
    ob = (oseq, (oseq[0], tuple(s for s in oseq[1] if s[0] != top)), (oseq[0], ()))
    idx = rng.integers(-1, 4, len(u))  # out-of-range values are clamped, as lax.switch does
    ts = rng.uniform(0.5, 2.0, len(u))
    net.input_custom_sequence(gb, idx, u, time_steps=ts)
    o.input_custom_sequence(ob, idx, u, time_steps=ts)
    sens = None
    if dtype == "float32":  # random time steps and frozen levels: some trials amplify rounding beyond 1e-5 in the oracle itself

        def run_oracle(perturb):
            o2 = build(onet.Network, dtype=dtype).create_belief_propagation_fn()
            perturb(o2)
            return o2.input_custom_sequence(ob, idx, u, time_steps=ts)

        sens = traj_sensitivity(run_oracle, range(len(net.edges)))
    assert_traj_close(net, o, range(len(net.edges)), RTOL_TRAJ if dtype == "float32" else 1e-9, sens=sens)
    assert np.isclose(net.last_attributes[top]["mean"], o.last_attributes[top]["mean"], rtol=1e-5)
    # the top level really was frozen on the steps that skipped it
    frozen = np.flatnonzero(np.clip(idx, 0, 2)[1:] == 2) + 1
    assert np.array_equal(net.node_trajectories[top]["mean"][frozen], net.node_trajectories[top]["mean"][frozen - 1])


def test_custom_sequence_abi_errors(G):
    from pyhgf_b200 import _lib

    net = nets.binary3(G.Network).create_belief_propagation_fn()
    seq = net.update_sequence
    with pytest.raises(ValueError, match="one entry per observation"):
        net.input_custom_sequence((seq, seq), [0, 1], np.ones(5))
    with pytest.raises(_lib.BackendError, match="single update sequence"):
        net._launch(values=net._format_series(np.ones(4), None, net.device, False)[0], observed=None, time_steps=None, input_group_size=0,
                    n_instances=1, params={}, response_fn=0, response_inputs=None, response_group_size=0, response_parameters=None,
                    record_slots=[], want_last=True, want_sum=False, want_pointwise=False, want_grad=False, group=None, n_groups=0,
                    branch_idx=[0, 0, 0, 0])


# ---------------------------------------------------------------- categorical-state nodes (hgf_categorical_update)
def test_categorical_reference_golden(G):
    """tests/test_categorical.py:9-88 through the drop-in surface: same calls, same goldens."""
    np.random.seed(123)
    x = np.array([np.random.multinomial(n=1, pvals=[0.1, 0.2, 0.7]) for _ in range(10)], dtype=float)
    net = G.Network().add_nodes(kind="categorical-state", node_parameters={"n_categories": 3, "binary_parameters": {"tonic_volatility_2": -2.0}})
    net.input_data(input_data=x)
    net.to_pandas()
    assert np.isclose(net.node_trajectories[0]["kl_divergence"].sum(), 1.2642910)
    assert np.isclose(net.node_trajectories[0]["surprise"].sum(), 12.418026)
    assert net.node_trajectories[0]["alpha"].shape == (10, 3) and net.node_trajectories[4]["mean"].shape == (10,)
    assert np.array_equal(net.node_trajectories[2]["mean"], x[:, 1].astype(np.float32))
    assert np.isclose(net.last_attributes[0]["kl_divergence"], net.node_trajectories[0]["kl_divergence"][-1])


@pytest.mark.parametrize("dtype", ["float32", "float64"])
@pytest.mark.parametrize("n", [2, 3, 5, 9])
def test_categorical_matches_oracle(G, n, dtype):
    from oracle import categorical as cat

    rng = np.random.default_rng(100 + n)
    p = rng.dirichlet(np.ones(n))
    x = rng.multinomial(1, p, size=60).astype(float)
    net = G.Network(dtype=dtype).add_nodes(kind="categorical-state", node_parameters={"n_categories": n}).input_data(x)
    onet_, c = cat.categorical_filter(x, n, None, dtype=dtype)
    rt = RTOL_TRAJ if dtype == "float32" else 1e-9
    for k in range(n):  # the implied binary HGFs: binary k is node 1 + k here, node k in the oracle's implied network
        for key in ("expected_mean", "mean"):
            assert np.allclose(net.node_trajectories[1 + n + k][key], onet_.node_trajectories[n + k][key], rtol=rt, atol=1e-7)
        assert np.allclose(net.node_trajectories[1 + k]["expected_mean"], onet_.node_trajectories[k]["expected_mean"], rtol=rt, atol=1e-7)
    got = net.node_trajectories[0]
    # nu divides by (sigmoid(mu) - mu_hat), a difference of nearby numbers: alpha carries the amplified rounding of both runs
    amp = 1e-3 if dtype == "float32" else 1e-7
    assert np.allclose(got["alpha"], c["alpha"], rtol=amp), np.abs(got["alpha"] / c["alpha"] - 1).max()
    assert np.allclose(got["mean"], c["mean"], rtol=rt)
    assert np.allclose(got["surprise"], c["surprise"], rtol=RTOL_SUM if dtype == "float32" else 1e-9)
    assert np.allclose(got["kl_divergence"], c["kl_divergence"], rtol=10 * amp, atol=amp)
    assert np.isclose(got["kl_divergence"].sum(), c["kl_divergence"].sum(), rtol=10 * amp)


def test_categorical_batch(G):
    """Instances that differ in the tonic volatility of the implied continuous nodes; per-instance result = single runs."""
    rng = np.random.default_rng(8)
    n, N, T = 4, 96, 50
    x = rng.multinomial(1, [0.1, 0.2, 0.3, 0.4], size=T).astype(float)
    om = rng.uniform(-5.0, -2.0, N)
    net = G.Network().add_nodes(kind="categorical-state", node_parameters={"n_categories": n})
    params = {(1 + n + k, "tonic_volatility"): om for k in range(n)}
    r = net.batch(x, params, record=[(0, "kl_divergence"), (0, "surprise"), (0, "alpha"), (1 + n, "mean")])
    kl, s, alpha = (r.trajectories[(0, k)].cpu().numpy() for k in ("kl_divergence", "surprise", "alpha"))
    assert kl.shape == (N, T) and alpha.shape == (N, T, n) and r.trajectories[(1 + n, "mean")].shape == (N, T)
    for i in (0, 17, 95):
        one = G.Network().add_nodes(kind="categorical-state", node_parameters={"n_categories": n})
        for k in range(n):
            one.attributes[1 + n + k]["tonic_volatility"] = float(om[i])
        one.input_data(x)
        assert np.allclose(kl[i], one.node_trajectories[0]["kl_divergence"], rtol=1e-4, atol=1e-6)
        assert np.allclose(s[i], one.node_trajectories[0]["surprise"], rtol=1e-5)
        assert np.allclose(alpha[i], one.node_trajectories[0]["alpha"], rtol=1e-4)


@pytest.mark.parametrize("learning,dtype", [("hgf-1", "float32"), ("hgf-2", "float32"), ("hgf-3", "float32"), ("hgf-2", "float64")])
def test_ef_state_hgf_learning_matches_oracle(G, learning, dtype):
    """learning='hgf-*' (tests/test_nodes/test_exponential_family.py:69-95 builds the hgf-2 network with these edits but
    asserts nothing about it): the per-step learning rate implied by the statistic-tracking HGFs, against the oracle."""
    from oracle import ef_state as ef

    x = ef.spiral_data()[:300]
    npd = np.float32 if dtype == "float32" else np.float64
    stride = {"hgf-1": 2, "hgf-2": 3, "hgf-3": 4}[learning]
    net = G.Network(dtype=dtype).add_nodes(kind="ef-state", learning=learning, distribution="multivariate-normal", dimension=2)

    def edit(attributes, shift):  # the reference test's edits: B nodes' volatility, A nodes' precision, the covariance trackers' means
        for k in range(5):
            attributes[1 + stride * k + 1 - shift]["tonic_volatility"] = -2.0
            attributes[1 + stride * k - shift]["precision"] = 0.01
        for k in (2, 3, 4):
            attributes[1 + stride * k + stride - 1 - shift]["mean"] = 10.0

    edit(net.attributes, 0)
    mask = np.arange(len(x)) % 11 != 5
    net.input_data(x, observed=mask)
    want = ef.ef_dynamic_filter(x, 2, "multivariate-normal", learning, edit=lambda o: edit(o.attributes, 1), observed=mask, dtype=npd)
    got = net.node_trajectories[0]
    assert got["xis"].shape == (300, 5) and got["nus"].shape == (300,) and np.isfinite(got["xis"]).all()
    # nu divides differences of nearby numbers: its rounding is amplified, and xis inherits it through 1 / (1 + nu)
    rt = 2e-3 if dtype == "float32" else 1e-8
    # on an unobserved step the parents keep their predictions, so the expectation differential is 0 and nu is not finite
    assert not np.isfinite(got["nus"][~mask]).any() and not np.isfinite(want["nus"][~mask]).any()
    assert np.allclose(got["nus"][mask], want["nus"][mask], rtol=rt, atol=rt), np.abs(got["nus"][mask] - want["nus"][mask]).max()
    assert np.allclose(got["xis"], want["xis"], rtol=rt, atol=rt)
    assert np.array_equal(got["xis"][5], got["xis"][4])  # unobserved step: blank update
    a_ref = want["network"].node_trajectories[0]["expected_mean"]
    assert np.allclose(net.node_trajectories[1]["expected_mean"], a_ref, rtol=RTOL_TRAJ if dtype == "float32" else 1e-9, atol=1e-6)


def test_vector_nodes_against_committed_vectors(G):
    """Categorical and ef-state nodes against tests/golden/vector_nodes.npz; no oracle code runs in this test."""
    import os

    v = np.load(os.path.join(os.path.dirname(__file__), "golden", "vector_nodes.npz"))
    for name, n in (("reference3", 3), ("five", 5)):
        for dtype, amp in (("float32", 1e-3), ("float64", 1e-7)):
            net = G.Network(dtype=dtype).add_nodes(kind="categorical-state", node_parameters={"n_categories": n})
            got = net.input_data(v[f"categorical/{name}/x"]).node_trajectories[0]
            assert np.allclose(got["alpha"], v[f"categorical/{name}/{dtype}/alpha"], rtol=amp)
            assert np.allclose(got["surprise"], v[f"categorical/{name}/{dtype}/surprise"], rtol=RTOL_SUM if dtype == "float32" else 1e-9)
            assert np.allclose(got["kl_divergence"], v[f"categorical/{name}/{dtype}/kl_divergence"], rtol=10 * amp, atol=amp)
    x = v["ef/spiral"]
    for dtype, rt in (("float32", RTOL_TRAJ), ("float64", 1e-12)):
        net = G.Network(dtype=dtype).add_nodes(kind="ef-state", distribution="multivariate-normal", dimension=2).input_data(x)
        assert np.allclose(net.node_trajectories[0]["xis"], v[f"ef/fixed/{dtype}/xis"], rtol=rt, atol=rt)
    uni = G.Network(dtype="float64").add_nodes(kind="ef-state", nus=7.5).input_data(x[:, 0])
    assert np.allclose(uni.node_trajectories[0]["xis"], v["ef/fixed_univariate/float64/xis"], rtol=1e-12)
    dyn = G.Network(dtype="float64").add_nodes(kind="ef-state", learning="hgf-2", distribution="multivariate-normal", dimension=2)
    got = dyn.input_data(x[:300]).node_trajectories[0]
    assert np.allclose(got["nus"], v["ef/hgf2/float64/nus"], rtol=1e-8) and np.allclose(got["xis"], v["ef/hgf2/float64/xis"], rtol=1e-8)


def test_vector_nodes_batch_masks_and_dynamic_ef(G):
    """Network.batch with an observation mask (categorical) and with a dynamic ef-state node: per-instance results equal
    the single-series runs with the same attributes."""
    from oracle import ef_state as ef

    rng = np.random.default_rng(21)
    n, N, T = 3, 70, 40
    x = rng.multinomial(1, [0.2, 0.3, 0.5], size=T).astype(float)
    mask = (rng.random(T) > 0.25).astype(int)
    om = rng.uniform(-5.0, -2.5, N)
    net = G.Network().add_nodes(kind="categorical-state", node_parameters={"n_categories": n})
    r = net.batch(x, {(1 + n + k, "tonic_volatility"): om for k in range(n)}, observed=mask, record=[(0, "surprise"), (0, "alpha")])
    for i in (0, 33, 69):
        one = G.Network().add_nodes(kind="categorical-state", node_parameters={"n_categories": n})
        for k in range(n):
            one.attributes[1 + n + k]["tonic_volatility"] = float(om[i])
        one.input_data(x, observed=mask)
        assert np.allclose(r.trajectories[(0, "surprise")][i].cpu().numpy(), one.node_trajectories[0]["surprise"], rtol=1e-5)
        assert np.allclose(r.trajectories[(0, "alpha")][i].cpu().numpy(), one.node_trajectories[0]["alpha"], rtol=1e-4, equal_nan=True)

    xs = ef.spiral_data()[:120]
    om_b = rng.uniform(-4.0, -1.5, 48)
    dyn = G.Network(dtype="float64").add_nodes(kind="ef-state", learning="hgf-2", distribution="multivariate-normal", dimension=2)
    b_nodes = [2, 5, 8, 11, 14]
    rb = dyn.batch(xs, {(b, "tonic_volatility"): om_b for b in b_nodes}, record=[(0, "xis"), (0, "nus"), (1, "expected_mean")])
    assert rb.trajectories[(0, "xis")].shape == (48, 120, 5) and rb.trajectories[(0, "nus")].shape == (48, 120)
    for i in (0, 20, 47):
        one = G.Network(dtype="float64").add_nodes(kind="ef-state", learning="hgf-2", distribution="multivariate-normal", dimension=2)
        for b in b_nodes:
            one.attributes[b]["tonic_volatility"] = float(om_b[i])
        one.input_data(xs)
        assert np.allclose(rb.trajectories[(0, "xis")][i].cpu().numpy(), one.node_trajectories[0]["xis"], rtol=1e-9)
        assert np.allclose(rb.trajectories[(0, "nus")][i].cpu().numpy(), one.node_trajectories[0]["nus"], rtol=1e-9)
        assert np.allclose(rb.trajectories[(1, "expected_mean")][i].cpu().numpy(), one.node_trajectories[1]["expected_mean"], rtol=1e-10)
