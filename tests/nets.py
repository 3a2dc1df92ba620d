"""Network builders shared by the parity tests: each takes the Network CLASS (the CPU oracle's or the CUDA
front-end's -- the reference's ``NETWORK_CLASSES`` pattern, tests/test_nodes/test_binary.py:9) and builds
the same graph with it."""

import numpy as np


def binary3(cls, om2=-4.0, om3=-4.0, **kw):
    return (cls(**kw).add_nodes(kind="binary-state").add_nodes(value_children=0, tonic_volatility=om2)
            .add_nodes(volatility_children=1, tonic_volatility=om3))


def binary2(cls, om2=-4.0, **kw):
    return cls(**kw).add_nodes(kind="binary-state").add_nodes(value_children=0, tonic_volatility=om2)


def continuous2(cls, **kw):
    """tests/test_continuous.py:101-131: input, value parent, volatility parent of the input."""
    return cls(**kw).add_nodes().add_nodes(value_children=0).add_nodes(volatility_children=0)


def continuous3(cls, x0=1.04, om1=-13.0, om2=-2.0, om3=-2.0, **kw):
    """tests/test_model.py:150-180 / notebook 1.3 cell 23: BASELINE config 2's network."""
    return (
        cls(**kw)
        .add_nodes(precision=1e4)
        .add_nodes(value_children=([0], [1.0]), node_parameters={"mean": x0, "precision": 1e4, "tonic_volatility": om1, "tonic_drift": 0.0})
        .add_nodes(volatility_children=([1], [1.0]), node_parameters={"mean": 1.0, "precision": 1e1, "tonic_volatility": om2, "tonic_drift": 0.0})
        .add_nodes(volatility_children=([2], [1.0]), node_parameters={"mean": 1.0, "precision": 1e1, "tonic_volatility": om3, "tonic_drift": 0.0})
    )


def nonlinear(cls, **kw):
    """tests/test_continuous.py:59-98"""
    return (cls(**kw).add_nodes(n_nodes=2).add_nodes(value_children=0, coupling_fn=("identity",))
            .add_nodes(value_children=1, coupling_fn=("sin",)))


def tanh_coupling(cls, **kw):
    """tests/test_nodes/test_continuous.py:58-76 style: tanh value coupling + volatility parent."""
    return (cls(**kw).add_nodes().add_nodes(value_children=0, coupling_fn=("tanh",)).add_nodes(volatility_children=1))


def volatile2(cls, **kw):
    """tests/test_nodes/test_volatile.py:112-139"""
    return cls(**kw).add_nodes().add_nodes(kind="volatile-state", value_children=0, autoconnection_strength=1.0)


def multi_child(cls, **kw):
    """Two inputs sharing a value parent that has a volatility parent, plus a second-level value parent
    (interior Gaussian child: Schur / RTS terms)."""
    return (
        cls(**kw)
        .add_nodes(n_nodes=2, precision=50.0)
        .add_nodes(value_children=([0, 1], [1.0, 0.7]), tonic_volatility=-3.0)
        .add_nodes(value_children=2, tonic_volatility=-5.0)
        .add_nodes(volatility_children=2, tonic_volatility=-3.0)
        .add_nodes(volatility_children=4, tonic_volatility=-4.0)
    )


def ghgf16(cls, **kw):
    """BASELINE config 5: 16-node generalised HGF -- continuous inputs 0-3, value parents
    4-7, volatility parents 8-11 of 4-7, volatile-state value parents 12,13 of (4,5),(6,7), volatile-state
    value parents 14,15 of 12,13."""
    net = cls(**kw)
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


def ar1_series(n_steps, n_series=4, phi=0.98, sigma=0.1, seed=0):
    rng = np.random.default_rng(seed)
    x = np.zeros((n_steps, n_series))
    e = rng.normal(0.0, sigma, size=(n_steps, n_series))
    for t in range(1, n_steps):
        x[t] = phi * x[t - 1] + e[t]
    return x


BELIEFS = ("mean", "expected_mean", "precision", "expected_precision")
