"""Pin the CPU oracle against the reference's own golden vectors.

Each test names the reference test it restates.  Values are the float32 JAX
goldens unless noted; the tolerance is the reference's (`jnp.isclose` default:
rtol 1e-5, atol 1e-8) except where the oracle's libm `exp/log` and XLA-CPU's
differ by an ulp over hundreds of recursive steps (then rtol 2e-6..1e-5 is
still met; see the per-test comment).
"""

import numpy as np
import pytest

from oracle.network import (
    Network,
    binary_softmax,
    binary_softmax_inverse_temperature,
    binary_surprise,
    first_level_binary_surprise,
    first_level_gaussian_surprise,
    gaussian_surprise,
    lambert_w0,
    total_gaussian_surprise,
)

RTOL = 1e-5


def close(a, b, rtol=RTOL, atol=1e-8):
    return np.isclose(a, b, rtol=rtol, atol=atol)


# --------------------------------------------------------------------------- tests/test_math.py:15-43
def test_binary_surprise_kats():
    assert close(binary_surprise(1.0, 0.8), 0.22314355)
    assert close(binary_surprise(0.0, 0.8), 1.60943791)
    assert close(binary_surprise(0.0, 0.0), 1.0132795e-06)
    assert close(binary_surprise(0.0, 1.0), 13.802319)
    assert close(binary_surprise(1.0, 0.2), 1.609438)  # tests/test_binary.py:21-27


def test_gaussian_surprise_kat():
    assert close(gaussian_surprise(1.0, 0.0, 1.0), 1.4189385)


def test_lambert_w0_identity():
    for z in [0.0, 1e-3, 0.5, 1.0, 10.0, 1e4, 1e10]:
        w = lambert_w0(z, 64)
        assert abs(w * np.exp(w) - z) <= 1e-9 * max(1.0, z)


# --------------------------------------------------------------------------- tests/test_binary.py:30-70
def test_update_binary_input_parents():
    net = (
        Network()
        .add_nodes(kind="binary-state")
        .add_nodes(value_children=0, mean=1.0, tonic_volatility=1.0)
        .add_nodes(volatility_children=1, mean=1.0, tonic_volatility=1.0)
        .input_data(np.ones(1))
    )
    a = net.last_attributes
    for k, v in zip(["mean", "expected_mean", "expected_precision"], [1.0, 0.7310586, 0.19661193]):
        assert close(a[0][k], v)
    for k, v in zip(["mean", "expected_mean", "precision", "expected_precision"], [2.2378633, 1.0, 0.21726260, 0.02065066]):
        assert close(a[1][k], v)
    for k, v in zip(["mean", "expected_mean", "precision", "expected_precision"], [0.37627372, 1.0, 0.39340553, 0.26894143]):
        assert close(a[2][k], v)


# --------------------------------------------------------------------------- tests/test_binary.py:73-94
def test_binary_scan_loop(datasets):
    net = (
        Network()
        .add_nodes(kind="binary-state")
        .add_nodes(value_children=0, mean=0.0, tonic_volatility=-4.0)
        .add_nodes(volatility_children=1, mean=0.0, tonic_volatility=-4.0)
        .input_data(datasets["u"])
    )
    # 320 recursive float32 steps: libm-vs-XLA exp/log ulps leave ~1e-6 relative drift
    for k, v in zip(["mean", "expected_mean", "precision", "expected_precision"],
                    [-2.3191204, -2.171629, 0.6937843, 0.6019279]):
        assert close(net.node_trajectories[1][k][-1], v, rtol=5e-6)
    for k, v in zip(["mean", "expected_mean", "precision", "expected_precision"],
                    [2.478936, 2.4795845, 1.7116449, 1.6612335]):
        assert close(net.node_trajectories[2][k][-1], v, rtol=5e-6)


# --------------------------------------------------------------------------- tests/test_binary.py:97-124
@pytest.mark.parametrize("pattern", ["blocks", "alternating"])
def test_volatile_examples_no_nan(pattern):
    data = np.array([1.0] * 50 + [0.0] * 50) if pattern == "blocks" else np.array([0.0, 1.0] * 50)
    net = (
        Network(volatility_updates="unbounded")
        .add_nodes(kind="binary-state")
        .add_nodes(value_children=[0], tonic_volatility=5.0)
        .input_data(data)
    )
    assert not np.any(np.isnan(net.node_trajectories[0]["expected_mean"]))


# --------------------------------------------------------------------------- tests/test_continuous.py:9-56
def test_one_node_hgf():
    net = Network().add_nodes().add_nodes(value_children=0).input_data(np.array([0.2]))
    for k, v in zip(["precision", "expected_precision", "mean", "expected_mean"], [1.0, 1.0, 0.2, 0.0]):
        assert close(net.node_trajectories[0][k][0], v)
    for k, v in zip(["precision", "expected_precision", "mean", "expected_mean"], [1.9820137, 0.98201376, 0.10090748, 0.0]):
        assert close(net.node_trajectories[1][k][0], v)


def test_two_nodes_hgf():
    net = Network().add_nodes().add_nodes(value_children=0).add_nodes(volatility_children=0).input_data(np.array([0.2]))
    for k, v in zip(["precision", "expected_precision", "mean", "expected_mean"], [1.0, 0.2715764, 0.2, 0.0]):
        assert close(net.node_trajectories[0][k][0], v)
    for k, v in zip(["precision", "expected_precision", "mean", "expected_mean"], [1.2535901, 0.98201376, 0.0433278, 0.0]):
        assert close(net.node_trajectories[1][k][0], v)
    for k, v in zip(["precision", "expected_precision", "mean", "expected_mean"], [1.1065601, 0.98201376, -0.13515851, 0.0]):
        assert close(net.node_trajectories[2][k][0], v)


# --------------------------------------------------------------------------- tests/test_continuous.py:59-98
def test_nonlinear_coupling_fn():
    net = (
        Network()
        .add_nodes(n_nodes=2)
        .add_nodes(value_children=0, coupling_fn=("identity",))
        .add_nodes(value_children=1, coupling_fn=("sin",))
    )
    net.input_data(np.array([[0.2, 0.4], [0.2, 0.4]]))
    for k, v in zip(["precision", "expected_precision", "mean", "expected_mean"], [2.9125834, 1.9125834, 0.13492969, 0.10090748]):
        assert close(net.node_trajectories[2][k][1], v)
    for k, v in zip(["precision", "expected_precision", "mean", "expected_mean"], [2.9124038, 1.9125834, 0.2689424, 0.20181496]):
        assert close(net.node_trajectories[3][k][1], v)


# --------------------------------------------------------------------------- tests/test_continuous.py:101-131
def test_continuous_scan_loop(datasets):
    net = Network().add_nodes().add_nodes(value_children=0).add_nodes(volatility_children=0).input_data(datasets["x"])
    # 614 recursive float32 steps
    for k, v in zip(["precision", "expected_precision", "mean", "expected_mean"], [1.0, 0.7623690, 0.8241, 0.7945459]):
        assert close(net.node_trajectories[0][k][-1], v, rtol=5e-6)
    for k, v in zip(["precision", "expected_precision", "mean", "expected_mean"], [6.8440757, 6.0817065, 0.7978380, 0.7945459]):
        assert close(net.node_trajectories[1][k][-1], v, rtol=5e-6)
    for k, v in zip(["precision", "expected_precision", "mean", "expected_mean"], [1.4760352, 1.4371926, -2.2633543, -2.2633660]):
        assert close(net.node_trajectories[2][k][-1], v, rtol=5e-6)


# --------------------------------------------------------------------------- tests/test_continuous.py:134-165
def test_record_trajectories(datasets):
    a = Network().add_nodes().add_nodes(value_children=0).input_data(datasets["x"], record_trajectories=True)
    b = Network().add_nodes().add_nodes(value_children=0).input_data(datasets["x"], record_trajectories=False)
    assert len(a.node_trajectories[0]["mean"]) == 614 and b.node_trajectories is None
    for i in (0, 1):
        for k in ("mean", "precision"):
            assert a.last_attributes[i][k] == b.last_attributes[i][k]


# --------------------------------------------------------------------------- tests/test_model.py:113-197
def test_continuous_hgf_surprise(datasets):
    x = datasets["x"]
    two = (
        Network()
        .add_nodes(precision=1e4)
        .add_nodes(value_children=([0], [1.0]),
                   node_parameters={"mean": x[0], "precision": 1e4, "tonic_volatility": -3.0, "tonic_drift": 0.0})
        .add_nodes(volatility_children=([1], [1.0]),
                   node_parameters={"mean": 0.0, "precision": 1e1, "tonic_volatility": -3.0, "tonic_drift": 0.0})
        .create_belief_propagation_fn()
    )
    two.input_data(x)
    assert close(two.surprise(first_level_gaussian_surprise).sum(), -1924.7515)
    assert len(two.node_trajectories[1]["mean"]) == 614
    three = (
        Network()
        .add_nodes(precision=1e4)
        .add_nodes(value_children=([0], [1.0]),
                   node_parameters={"mean": 1.04, "precision": 1e4, "tonic_volatility": -13.0, "tonic_drift": 0.0})
        .add_nodes(volatility_children=([1], [1.0]),
                   node_parameters={"mean": 1.0, "precision": 1e1, "tonic_volatility": -2.0, "tonic_drift": 0.0})
        .add_nodes(volatility_children=([2], [1.0]),
                   node_parameters={"mean": 1.0, "precision": 1e1, "tonic_volatility": -2.0, "tonic_drift": 0.0})
        .create_belief_propagation_fn()
    )
    three.input_data(x)
    assert close(three.surprise(first_level_gaussian_surprise).sum(), -2034.3989)
    assert close(total_gaussian_surprise(three).sum(), -2535.604)


# --------------------------------------------------------------------------- tests/test_model.py:200-268
def test_binary_hgf_surprise(datasets):
    u = datasets["u"]
    two = (
        Network()
        .add_nodes(kind="binary-state", node_parameters={"mean": 0.0, "precision": 0.0})
        .add_nodes(kind="continuous-state", value_children=([0], [1.0]),
                   node_parameters={"mean": 0.5, "precision": 1e4, "tonic_volatility": -6.0, "tonic_drift": 0.0})
        .create_belief_propagation_fn()
    )
    two.input_data(u)
    assert close(two.surprise(first_level_binary_surprise).sum(), 215.58821)
    three = (
        Network()
        .add_nodes(kind="binary-state", node_parameters={"mean": 0.0, "precision": 0.0})
        .add_nodes(kind="continuous-state", value_children=([0], [1.0]),
                   node_parameters={"mean": 0.5, "precision": 1e4, "tonic_volatility": -6.0, "tonic_drift": 0.0})
        .add_nodes(volatility_children=([1], [1.0]),
                   node_parameters={"mean": 0.0, "precision": 1e1, "tonic_volatility": -2.0, "tonic_drift": 0.0})
        .create_belief_propagation_fn()
    )
    three.input_data(u)
    # saturating regime (clip at 1e-6, pi=1e4): rounding-sensitive at the 1 % level in
    # float64 but the float32 op order reproduces the golden.
    assert close(three.surprise(first_level_binary_surprise).sum(), 1242.3856)


# --------------------------------------------------------------------------- tests/test_responses.py:10-50 and README.md:115
def test_binary_responses(datasets):
    u, y = datasets["u"], datasets["y"]
    net = (
        Network()
        .add_nodes(kind="binary-state", node_parameters={"mean": 0.5, "precision": 0.0})
        .add_nodes(kind="continuous-state", value_children=([0], [1.0]),
                   node_parameters={"mean": 0.0, "precision": 1.0, "tonic_volatility": -6.0, "tonic_drift": 0.0})
        .create_belief_propagation_fn()
        .input_data(u)
    )
    assert close(net.surprise(binary_softmax, response_function_inputs=y).sum(), 195.81573)
    s = net.surprise(binary_softmax_inverse_temperature, response_function_inputs=y, response_function_parameters=2.0)
    assert close(s.sum(), 188.77818)
    readme = Network().add_nodes(kind="binary-state").add_nodes(kind="continuous-state", value_children=0).input_data(u)
    s = readme.surprise(binary_softmax_inverse_temperature, response_function_inputs=y, response_function_parameters=4.0)
    assert close(s.sum(), 138.8992462158203)


# --------------------------------------------------------------------------- tests/test_updates/posterior/test_posterior_update_continuous.py:13-99
def test_continuous_posterior_updates():
    net = (
        Network().add_nodes().add_nodes(value_children=0).add_nodes(volatility_children=1).add_nodes(volatility_children=2)
    )
    attributes, _, _ = net.get_network()
    from copy import deepcopy

    a = deepcopy(attributes)
    a[0]["temp"]["value_prediction_error"] = 1.0357
    a[0]["mean"] = 1.0357
    assert close(net.apply_op("POST_CONT", 1, a)[1]["mean"], 0.51785)

    def vol_case(extra_time=False):
        a = deepcopy(attributes)
        if extra_time:
            a[-1]["time_step"] = 1.0
            a[1]["temp"]["current_variance"] = 1.0
        a[1]["temp"]["effective_precision"] = 0.01798621006309986
        a[1]["temp"]["value_prediction_error"] = 0.5225493907928467
        a[1]["temp"]["volatility_prediction_error"] = -0.23639076948165894
        a[1]["expected_precision"] = 0.9820137619972229
        a[1]["mean"] = 0.5225493907928467
        a[1]["precision"] = 1.9820137023925781
        return a

    out = net.apply_op("POST_CONT", 2, vol_case())
    assert close(out[2]["mean"], -0.0021212, rtol=1e-4) and close(out[2]["precision"], 1.0022112)
    # (the reference's eHGF 'value update' check at :61-64 reads node 1 after updating node 2: a no-op, skipped)
    out = net.apply_op("POST_CONT_EHGF", 2, vol_case(True))
    assert close(out[2]["mean"], -0.00212589, rtol=1e-5, atol=1e-8) and close(out[2]["precision"], 1.0022061)
    a = deepcopy(attributes)
    a[-1]["time_step"] = 1.0
    a[1]["temp"]["current_variance"] = 1.0
    a[1]["expected_precision"] = 0.9820137619972229
    a[1]["mean"] = 0.5225493907928467
    a[1]["precision"] = 1.9820137023925781
    out = net.apply_op("POST_CONT_UNBOUNDED", 2, a)
    assert close(out[2]["mean"], 0.01330645) and close(out[2]["precision"], 1.0161277)


# --------------------------------------------------------------------------- tests/test_utils.py:70-92, tests/test_model.py:26-41
def test_update_sequence_lengths():
    n1 = Network().add_nodes(kind="binary-state").add_nodes(value_children=0).create_belief_propagation_fn()
    assert len(n1.update_sequence[0]) == 2 and len(n1.update_sequence[1]) == 2
    n2 = (
        Network(volatility_updates="standard").add_nodes().add_nodes(value_children=0).add_nodes(volatility_children=1)
        .create_belief_propagation_fn()
    )
    assert len(n2.update_sequence[0]) == 3 and len(n2.update_sequence[1]) == 4
    custom = (
        Network()
        .add_nodes(kind="continuous-state").add_nodes(kind="binary-state")
        .add_nodes(value_children=0).add_nodes(value_children=1)
        .add_nodes(value_children=[2, 3]).add_nodes(value_children=4)
        .add_nodes(volatility_children=[2, 3]).add_nodes(volatility_children=2)
        .add_nodes(volatility_children=7).add_nodes(value_parents=8).add_nodes(volatility_parents=9)
    )
    custom.input_data(np.ones((10, 2)), observed=np.ones((10, 2)))
    assert custom.input_idxs == (0, 1, 10)  # node 10 is auto-detected but unfed
    assert len(custom.update_sequence[0]) == 11


# --------------------------------------------------------------------------- tests/test_binary.rs:20-252 (float64, eHGF, tol 1e-10)
def test_rust_f64_ehgf_binary_goldens():
    two = (
        Network(volatility_updates="eHGF", dtype="float64")
        .add_nodes(kind="binary-state")
        .add_nodes(value_children=0, mean=1.0, tonic_volatility=1.0)
        .input_data(np.array([1.0]))
    )
    t = two.node_trajectories
    for k, v in [("expected_mean", 0.731058578630005), ("precision", 0.196611933241482)]:
        assert abs(t[0][k][0] - v) < 1e-10
    assert abs(t[0]["temp"]["value_prediction_error"][0] - 1.367879441171442) < 1e-10
    for k, v in [("mean", 1.577681201748482), ("expected_mean", 1.0), ("precision", 0.465553354611477),
                 ("expected_precision", 0.268941421369995)]:
        assert abs(t[1][k][0] - v) < 1e-10
    three = (
        Network(volatility_updates="eHGF", dtype="float64")
        .add_nodes(kind="binary-state")
        .add_nodes(value_children=0, mean=1.0, tonic_volatility=1.0)
        .add_nodes(volatility_children=1, mean=1.0, tonic_volatility=1.0)
        .input_data(np.array([1.0, 0.0]))
    )
    t = three.node_trajectories
    s0 = {1: dict(mean=2.237863416338325, expected_mean=1.0, precision=0.217262597650345, expected_precision=0.020650664408863),
          2: dict(mean=-0.590073315695407, expected_mean=1.0, precision=0.537517033851118, expected_precision=0.268941421369995)}
    s1 = {0: dict(mean=0.0, expected_mean=0.903598504654528, precision=0.087108247040629, expected_precision=0.087108247040629),
          1: dict(mean=-4.287959232795414, expected_mean=2.237863416338325, precision=0.138465072257898, expected_precision=0.051356825217268),
          2: dict(mean=2.133602312876886, expected_mean=-0.590073315695407, precision=0.789268236136739, expected_precision=0.218403176385631)}
    for node, d in s0.items():
        for k, v in d.items():
            assert abs(t[node][k][0] - v) < 1e-10, (node, k)
    for node, d in s1.items():
        for k, v in d.items():
            assert abs(t[node][k][1] - v) < 1e-10, (node, k)


# --------------------------------------------------------------------------- tests/test_continuous.rs:16-133 (float64, eHGF, tol 1e-5)
def test_rust_f64_ehgf_continuous_goldens():
    net = (
        Network(volatility_updates="eHGF", dtype="float64")
        .add_nodes().add_nodes(value_children=0).add_nodes(volatility_children=0)
        .input_data(np.array([0.2]))
    )
    t = net.node_trajectories
    for node, d in {0: dict(precision=1.0, expected_precision=0.27157641, mean=0.2, expected_mean=0.0),
                    1: dict(precision=1.25359020, expected_precision=0.98201376, mean=0.04332778, expected_mean=0.0),
                    2: dict(precision=1.09553182, expected_precision=0.98201376, mean=-0.16509254, expected_mean=0.0)}.items():
        for k, v in d.items():
            assert abs(t[node][k][0] - v) < 1e-5, (node, k)


# --------------------------------------------------------------------------- tests/test_nodes/test_volatile.py:112-139
@pytest.mark.parametrize("vu", ["standard", "eHGF", "unbounded"])
@pytest.mark.parametrize("mf", [False, True])
def test_volatile_node_matches_explicit(datasets, vu, mf):
    x = datasets["x"]
    vol = (
        Network(volatility_updates=vu, mean_field_updates=mf)
        .add_nodes().add_nodes(kind="volatile-state", value_children=0, autoconnection_strength=1.0).input_data(x)
    )
    exp = (
        Network(volatility_updates=vu, mean_field_updates=mf)
        .add_nodes().add_nodes(value_children=0, tonic_volatility=0.0).add_nodes(volatility_children=1).input_data(x)
    )
    for node in (0, 1):
        for k in ["mean", "expected_mean", "precision", "expected_precision"]:
            assert np.allclose(vol.node_trajectories[node][k], exp.node_trajectories[node][k], rtol=1e-5), (node, k)
    for vk, ek in {"mean_vol": "mean", "expected_mean_vol": "expected_mean", "precision_vol": "precision",
                   "expected_precision_vol": "expected_precision"}.items():
        assert np.allclose(vol.node_trajectories[1][vk], exp.node_trajectories[2][ek], rtol=1e-4, atol=1e-6), vk


# --------------------------------------------------------------------------- gradient oracle self-check (unpinned by the reference)
def _three_level(dtype="float64"):
    return (
        Network(dtype=dtype)
        .add_nodes(kind="binary-state")
        .add_nodes(value_children=0)
        .add_nodes(volatility_children=1)
        .create_belief_propagation_fn()
    )


def test_dual_gradient_matches_finite_differences(datasets):
    u, y = datasets["u"], datasets["y"]
    net = _three_level()
    slots = [net.slot(1, "tonic_volatility"), net.slot(2, "tonic_volatility"), net.slot(-1, "response_parameter", 0),
             net.slot(1, "mean"), net.slot(2, "precision")]
    base = np.array([[-3.0], [-2.5], [1.7], [0.1], [1.3]])
    s, g = net.batch_grad(u, slots, base, "binary_softmax_inverse_temperature", resp_inputs=y)
    for q in range(len(slots)):
        h = 1e-6
        up, dn = base.copy(), base.copy()
        up[q] += h
        dn[q] -= h
        net.dtype = 64
        fu = net.batch_surprise(u, slots, up, "binary_softmax_inverse_temperature", resp_inputs=y)
        fd = net.batch_surprise(u, slots, dn, "binary_softmax_inverse_temperature", resp_inputs=y)
        fdiff = (fu[0] - fd[0]) / (2 * h)
        assert np.isclose(g[q, 0], fdiff, rtol=2e-5, atol=1e-7), (q, g[q, 0], fdiff)


def test_batch_matches_single(datasets):
    u = datasets["u"]
    net = _three_level("float32")
    om2 = np.linspace(-6, -1, 7)
    om3 = np.linspace(-5, -1, 7)
    params = np.stack([om2, om3])
    s = net.batch_surprise(u, [net.slot(1, "tonic_volatility"), net.slot(2, "tonic_volatility")], params,
                           "first_level_binary_surprise")
    for n in range(7):
        single = (
            Network().add_nodes(kind="binary-state").add_nodes(value_children=0, tonic_volatility=om2[n])
            .add_nodes(volatility_children=1, tonic_volatility=om3[n]).input_data(u)
        )
        ref = single.surprise(first_level_binary_surprise)
        # batch sums in float32 inside the scan; the trajectory path sums float32 terms in float64
        assert np.isclose(s[n], ref.sum(), rtol=2e-6)


# --------------------------------------------------------------------------- docs/source/notebooks goldens (to_pandas)
def test_to_pandas_notebook_goldens(datasets):
    """1.1-Binary_HGF.ipynb cell 27 and 1.3-Continuous_HGF.ipynb cell 16: per-node surprise sums of
    `Network.to_pandas()` (pyhgf/utils/to_pandas.py:88-112)."""
    net = (Network().add_nodes(kind="binary-state").add_nodes(value_children=0).add_nodes(volatility_children=1)
           .input_data(datasets["u"]))
    df = net.to_pandas()
    assert close(df.x_0_surprise.sum(), 206.60437)
    assert close(df.x_1_surprise.sum(), 291.26907)
    assert close(df.x_2_surprise.sum(), 238.23419)
    assert close(df.total_surprise.sum(), 206.60437 + 291.26907 + 238.23419)
    x = datasets["x"]
    net = (Network().add_nodes(precision=1e4)
           .add_nodes(precision=1e4, mean=x[0], tonic_volatility=-13.0, value_children=0)
           .add_nodes(precision=1e1, tonic_volatility=-2.0, volatility_children=1).input_data(x))
    df = net.to_pandas()
    assert close(df.x_0_surprise.sum(), -2071.4644)
    assert close(df.x_1_surprise.sum(), -2446.2734)
    assert close(df.x_2_surprise.sum(), 665.03186)
    assert list(df.columns[:2]) == ["time_steps", "time"] and "x_1_expected_precision" in df.columns


# --------------------------------------------------------------------------- committed fixtures (tests/golden/)
def test_oracle_reproduces_the_committed_vectors():
    """tests/golden/oracle_vectors.npz (written by tests/golden/make_oracle_vectors.py) is what the oracle gives today."""
    import os

    import nets

    v = np.load(os.path.join(os.path.dirname(__file__), "golden", "oracle_vectors.npz"))
    d = np.load(os.path.join(os.path.dirname(__file__), "..", "pyhgf_b200", "data", "datasets.npz"))
    net = nets.binary3(Network, dtype="float64").input_data(d["binary_u"])
    for node in (1, 2):
        for k in nets.BELIEFS:
            assert np.allclose(net.node_trajectories[node][k], v[f"traj/binary3/{node}/{k}"], rtol=1e-12, atol=0, equal_nan=True)
    o = nets.binary3(Network, dtype="float64").create_belief_propagation_fn()
    slots = [o.slot(1, "tonic_volatility"), o.slot(2, "tonic_volatility"), o.slot(-1, "response_parameter", 0)]
    s, g = o.batch_grad(d["binary_u"], slots, v["config4/params"], "binary_softmax_inverse_temperature", resp_inputs=d["binary_y"])
    assert np.allclose(s, v["config4/surprise"], rtol=1e-12) and np.allclose(g, v["config4/gradient"], rtol=1e-9)


def test_exponential_family_fixed_learning_rate_golden():
    """tests/test_nodes/test_exponential_family.py:14-66: the univariate run has no NaN, the bivariate spiral ends at
    the xis the reference asserts."""
    from oracle import ef_state as ef

    x = ef.spiral_data()
    tr = ef.ef_filter(x, 3.0, 2, "multivariate-normal")
    golden = np.array([3.4652710e01, -1.0609777e00, 1.2103647e03, -3.6398651e01, 3.3951855e00], dtype=np.float32)
    assert np.isclose(tr[-1], golden).all()
    uni = ef.ef_filter(x[:, 0], 3.0, 1, "normal")
    assert uni.shape == (1000, 2) and np.isfinite(uni).all()
    # the update is an exponential moving average of t(x) with weight 1/(1+nu): closed form after T steps
    w = 1.0 / 4.0
    t = np.stack([x[:, 0], x[:, 0] ** 2], 1)
    k = np.arange(len(x))[::-1]
    closed = (w * (1 - w) ** k)[:, None] * t
    expected = closed.sum(0) + (1 - w) ** len(x) * np.array([0.0, 1.0])
    assert np.allclose(ef.ef_filter(x[:, 0], 3.0, 1, "normal", dtype=np.float64)[-1], expected, rtol=1e-12)
    mask = (np.arange(len(x)) % 3 != 0)
    held = ef.ef_filter(x, 3.0, 2, "multivariate-normal", observed=mask)
    assert np.array_equal(held[3], held[2]) and not np.array_equal(held[4], held[3])


def test_categorical_state_node_golden():
    """tests/test_categorical.py:9-88: summed KL divergence and surprise of the 3-category example.  The reference builds
    the implied binary HGFs from its default parameters whatever ``binary_parameters`` says (add_nodes.py:532-537)."""
    from oracle import categorical as cat

    np.random.seed(123)
    x = np.array([np.random.multinomial(n=1, pvals=[0.1, 0.2, 0.7]) for _ in range(10)], dtype=float)
    _, c = cat.categorical_filter(x, 3, None, dtype="float32")
    assert np.isclose(c["kl_divergence"].sum(), 1.2642910) and np.isclose(c["surprise"].sum(), 12.418026)
    _, c64 = cat.categorical_filter(x, 3, None, dtype="float64")
    assert np.isclose(c64["kl_divergence"].sum(), 1.2642910) and np.isclose(c64["surprise"].sum(), 12.418026)
    assert np.allclose(c["alpha"], c64["alpha"], rtol=1e-4) and (c64["alpha"] > 0).all()
    _, other = cat.categorical_filter(x, 3, {"tonic_volatility_2": -2.0}, dtype="float32")
    assert not np.isclose(other["surprise"].sum(), 12.418026)  # the override would change the answer


def test_vector_node_restatements_reproduce_the_committed_vectors():
    """tests/golden/vector_nodes.npz (tests/golden/make_vector_node_vectors.py) against the restatements as they are now."""
    import os

    from oracle import categorical as cat
    from oracle import ef_state as ef

    v = np.load(os.path.join(os.path.dirname(__file__), "golden", "vector_nodes.npz"))
    assert np.isclose(v["categorical/reference3/float32/kl_divergence"].sum(), 1.2642910)
    assert np.isclose(v["categorical/reference3/float32/surprise"].sum(), 12.418026)
    assert np.isclose(v["ef/fixed/float32/xis"][-1], [3.4652710e01, -1.0609777e00, 1.2103647e03, -3.6398651e01, 3.3951855e00]).all()
    _, c = cat.categorical_filter(v["categorical/five/x"], 5, None, dtype="float64")
    assert np.allclose(c["kl_divergence"], v["categorical/five/float64/kl_divergence"], rtol=1e-12)
    assert np.allclose(ef.ef_filter(v["ef/spiral"], 3.0, 2, "multivariate-normal", dtype=np.float64), v["ef/fixed/float64/xis"], rtol=1e-13)
    d = ef.ef_dynamic_filter(v["ef/spiral"][:300], 2, "multivariate-normal", "hgf-2", dtype=np.float64)
    assert np.allclose(d["nus"], v["ef/hgf2/float64/nus"], rtol=1e-10, equal_nan=True)
