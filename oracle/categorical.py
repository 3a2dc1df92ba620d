"""TEST INFRASTRUCTURE ONLY -- CPU restatement of pyhgf's categorical state node.

Follows ``add_categorical_state`` (pyhgf/model/add_nodes.py:473-545), ``fill_categorical_state_node``
(pyhgf/utils/fill_categorical_state_node.py:13-81: one binary HGF per category -- binary node, continuous value parent --
and ONE volatility parent shared by the continuous nodes), ``categorical_state_prediction_error``
(pyhgf/updates/prediction_error/categorical.py:12-47: the observation vector is handed to the binary nodes) and
``categorical_state_update`` (pyhgf/updates/posterior/categorical.py:12-91: after the binary HGFs have updated, the
implied Dirichlet concentration, its KL divergence from the previous one and the summed binary surprise), with
``dirichlet_kullback_leibler`` (pyhgf/math.py:307-341) and ``binary_surprise`` (pyhgf/math.py).

The implied binary HGFs run through the C++ oracle (oracle/network.py); the categorical step is NumPy in the requested
dtype.  Parity status: PINNED by tests/test_categorical.py:9-88 (Σ kl_divergence = 1.2642910, Σ surprise = 12.418026),
asserted in tests/test_oracle_golden.py.  Only tests/ may import this module.
"""
import numpy as np
from scipy.special import digamma, gamma

from oracle import network as onet


def binary_parameters(n_categories: int, overrides=None) -> dict:  # add_nodes.py:496-506
    p = {"n_categories": n_categories, "precision_1": 1.0, "precision_2": 1.0, "precision_3": 1.0,
         "mean_1": 1.0 / n_categories, "mean_2": -np.log(n_categories - 1), "mean_3": 0.0,
         "tonic_volatility_2": -4.0, "tonic_volatility_3": -4.0}
    p.update(overrides or {})
    return p


def implied_network(n_categories: int, overrides=None, dtype="float32", **kw):
    """The binary HGFs under the categorical node, WITHOUT that node: binary 0..n-1 (the inputs), continuous n..2n-1,
    shared volatility parent 2n."""
    p = binary_parameters(n_categories, overrides)
    n = n_categories
    net = onet.Network(dtype=dtype, **kw)
    net.add_nodes(kind="binary-state", n_nodes=n, node_parameters={"mean": p["mean_1"], "precision": p["precision_1"]})
    for i in range(n):
        net.add_nodes(value_children=i, node_parameters={"mean": p["mean_2"], "precision": p["precision_2"],
                                                         "tonic_volatility": p["tonic_volatility_2"]})
    net.add_nodes(volatility_children=[n + i for i in range(n)],
                  node_parameters={"mean": p["mean_3"], "precision": p["precision_3"], "tonic_volatility": p["tonic_volatility_3"]})
    return net


def dirichlet_kullback_leibler(alpha_1, alpha_2, dtype):  # math.py:337-341
    a1, a2 = alpha_1.astype(dtype), alpha_2.astype(dtype)
    with np.errstate(all="ignore"):
        return dtype(
            np.log(dtype(gamma(a1.sum(dtype=dtype)) / gamma(a2.sum(dtype=dtype))))
            + np.sum(np.log((gamma(a2) / gamma(a1)).astype(dtype)), dtype=dtype)
            + np.sum((a1 - a2) * (digamma(a1).astype(dtype) - dtype(digamma(a1.sum(dtype=dtype)))), dtype=dtype)
        )


def categorical_update(observations, expected_mean, parent_mean, alpha0, dtype=np.float32):
    """posterior/categorical.py:44-89 for every step.  observations, expected_mean, parent_mean: (T, n).
    -> dict(alpha (T, n), mean (T, n), kl_divergence (T,), surprise (T,))."""
    x = np.asarray(observations, dtype=dtype)
    em = np.asarray(expected_mean, dtype=dtype)
    pm = np.asarray(parent_mean, dtype=dtype)
    T, n = x.shape
    out = {"alpha": np.empty((T, n), dtype), "mean": np.empty((T, n), dtype), "kl_divergence": np.empty(T, dtype),
           "surprise": np.empty(T, dtype)}
    alpha_prev = np.asarray(alpha0, dtype=dtype)
    with np.errstate(all="ignore"):
        for t in range(T):
            updated = (dtype(1) / (dtype(1) + np.exp(-pm[t]))).astype(dtype)
            pe = x[t] - em[t]
            delta_xi = updated - em[t]
            nu = pe / delta_xi - dtype(1)
            alpha = (nu * em[t] + dtype(1)).astype(dtype)
            out["kl_divergence"][t] = dirichlet_kullback_leibler(alpha_prev, alpha, dtype)
            emc = np.clip(em[t], dtype(1e-6), dtype(1.0 - 1e-6))  # binary_surprise(clipping=True), pyhgf/math.py:248-249
            out["surprise"][t] = np.sum(np.where(x[t] != 0, -np.log(emc), -np.log(dtype(1) - emc)), dtype=dtype)
            out["alpha"][t] = alpha
            out["mean"][t] = updated
            alpha_prev = alpha
    return out


def categorical_filter(input_data, n_categories, overrides=None, dtype="float32", **kw):
    """``Network().add_nodes(kind="categorical-state", ...).input_data(input_data)``: (implied network, categorical dict)."""
    net = implied_network(n_categories, overrides, dtype=dtype, **kw)
    net.input_data(np.asarray(input_data, dtype=np.float64))
    n = n_categories
    em = np.stack([net.node_trajectories[i]["expected_mean"] for i in range(n)], axis=1)
    pm = np.stack([net.node_trajectories[n + i]["mean"] for i in range(n)], axis=1)
    npd = np.float32 if dtype == "float32" else np.float64
    return net, categorical_update(input_data, em, pm, np.ones(n), npd)
