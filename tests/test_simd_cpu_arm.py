"""The instance-vectorised float32 CPU arm of bench.py (oracle/hgf_simd.cpp: SIMD across instances, libmvec, the generated
reverse sweep) computes what the scalar oracle computes.  It is a TIMED BASELINE, never a parity checker, so the bar is
statistical: libmvec's exp / log / pow differ from libm's in the last ulps, which ill-conditioned instances amplify
(oracle/conditioning.py); the bulk of every batch must agree to float32 round-off."""

import numpy as np
import pytest

import nets
from oracle import network as onet

simd = pytest.importorskip("oracle.simd")

RF = "binary_softmax_inverse_temperature"


def _rel(got, want, floor=0.0):
    want = np.asarray(want, dtype=np.float64)
    return np.abs(np.asarray(got, dtype=np.float64) - want) / np.maximum(np.abs(want), floor + 1e-30)


def _bulk(err, what):
    err = err[np.isfinite(err)]
    assert np.median(err) <= 2e-6, (what, float(np.median(err)))
    assert np.quantile(err, 0.9) <= 1e-4, (what, float(np.quantile(err, 0.9)))


@pytest.fixture(scope="module")
def binary_case(datasets):
    rng = np.random.default_rng(7)
    n = 203  # not a multiple of the vector width
    p = np.stack([rng.uniform(-6, -3, n), rng.uniform(-6, -3, n), np.exp(rng.uniform(-0.3, 0.8, n))]).astype(np.float32)
    o = nets.binary3(onet.Network).create_belief_propagation_fn()
    slots = [o.slot(1, "tonic_volatility"), o.slot(2, "tonic_volatility"), o.slot(-1, "response_parameter", 0)]
    return o, slots, p, datasets["u"], datasets["y"]


def test_vector_width_is_a_simd_width():
    assert simd.width() in (8, 16)


def test_logp_and_reverse_sweep_gradient_match_the_scalar_oracle(binary_case):
    o, slots, p, u, y = binary_case
    want_s, want_g = o.batch_grad(u, slots, p.astype(np.float64), RF, resp_inputs=y, dtype=32)
    got_s, got_g = simd.batch(o, u, slots, p, RF, resp_inputs=y, gradient=True)
    assert got_s.shape == want_s.shape and got_g.shape == want_g.shape
    _bulk(_rel(got_s, want_s), "logp")
    for i in range(3):
        _bulk(_rel(got_g[i], want_g[i], floor=1e-3 * np.abs(want_g[i]).max()), f"gradient {i}")
    # the forward-only call is the same forward sweep
    assert np.array_equal(simd.batch(o, u, slots, p, RF, resp_inputs=y), got_s)
    # one thread or many: the same numbers
    one_s, one_g = simd.batch(o, u, slots, p, RF, resp_inputs=y, gradient=True, n_threads=1)
    assert np.array_equal(one_s, got_s) and np.array_equal(one_g, got_g)


def test_series_grouped_inside_a_vector(binary_case, datasets):
    """Instances of one SIMD vector that read different series (group size 5 does not divide the width)."""
    o, slots, p, u, y = binary_case
    rng = np.random.default_rng(3)
    n_groups = 41
    us = (rng.uniform(size=(n_groups, len(u))) < 0.5).astype(np.float32)
    ys = (rng.uniform(size=(n_groups, len(u))) < 0.5).astype(np.float32)
    want = o.batch_surprise(us.astype(np.float64), slots, p.astype(np.float64), RF, resp_inputs=ys.astype(np.float64), group_size=5)
    got = simd.batch(o, us, slots, p, RF, resp_inputs=ys, group_size=5)
    _bulk(_rel(got, want), "grouped logp")


def test_continuous_hgf_forward_and_gradient(datasets):
    x = datasets["x"][:300]
    rng = np.random.default_rng(2)
    n = 100
    p = np.stack([np.clip(rng.normal(m, 1.0, n), m - 3, m + 3) for m in (-13.0, -2.0, -2.0)]).astype(np.float32)
    o = nets.continuous3(onet.Network).create_belief_propagation_fn()
    slots = [o.slot(1, "tonic_volatility"), o.slot(2, "tonic_volatility"), o.slot(3, "tonic_volatility")]
    want_s, want_g = o.batch_grad(x, slots, p.astype(np.float64), "first_level_gaussian_surprise", dtype=32)
    got_s, got_g = simd.batch(o, x, slots, p, "first_level_gaussian_surprise", gradient=True)
    assert np.array_equal(np.isnan(got_s), np.isnan(want_s))
    _bulk(_rel(got_s, want_s), "continuous logp")
    for i in range(3):
        _bulk(_rel(got_g[i], want_g[i], floor=1e-3 * np.nanmax(np.abs(want_g[i]))), f"continuous gradient {i}")
    # irregular sampling
    ts = rng.uniform(0.5, 2.0, len(x))
    want = o.batch_surprise(x, slots, p.astype(np.float64), "first_level_gaussian_surprise", time_steps=ts)
    got = simd.batch(o, x, slots, p, "first_level_gaussian_surprise", time_steps=ts)
    _bulk(_rel(got, want), "continuous logp, time steps")


def test_16_node_network_with_volatile_nodes():
    """BASELINE config 5's network: the filter alone (what its CPU arm times), logp and the reverse-sweep gradient."""
    x = nets.ar1_series(200, 4, seed=3)
    o = nets.ghgf16(onet.Network).create_belief_propagation_fn()
    keys = [(i, "tonic_volatility") for i in range(4, 10)] + [(12, "tonic_volatility_vol")]
    rng = np.random.default_rng(1)
    n = 70
    p = np.stack([rng.uniform(-5, -3, n) for _ in keys]).astype(np.float32)
    slots = [o.slot(*k) for k in keys]
    want_s, want_g = o.batch_grad(x, slots, p.astype(np.float64), "total_gaussian_surprise", dtype=32)
    got_s, got_g = simd.batch(o, x, slots, p, "total_gaussian_surprise", gradient=True)
    _bulk(_rel(got_s, want_s), "16-node logp")
    for i in range(len(keys)):
        _bulk(_rel(got_g[i], want_g[i], floor=1e-3 * np.abs(want_g[i]).max()), f"16-node gradient {keys[i]}")
    assert np.array_equal(simd.batch(o, x, slots, p, None), np.zeros(n, dtype=np.float32))  # no response function: the filter alone


def test_other_topologies_are_refused():
    o = nets.volatile2(onet.Network).create_belief_propagation_fn()
    with pytest.raises(NotImplementedError):
        simd.batch(o, np.zeros(10), [o.slot(1, "tonic_volatility")], np.full((1, 4), -3.0), "first_level_gaussian_surprise")
