"""TEST INFRASTRUCTURE ONLY -- CPU restatement of pyhgf's exponential-family state node with a fixed learning rate.

Follows ``add_ef_state`` (pyhgf/model/add_nodes.py:360-470: defaults ``nus = 3``, ``xis = [0, 1]`` for a normal,
``sufficient_statistics_from_parameters(0, I)`` for a multivariate normal), ``set_observation``
(pyhgf/updates/observation.py:10-38) and ``prediction_error_update_exponential_family_fixed``
(pyhgf/updates/prediction_error/exponential.py:13-60), with the sufficient statistics of pyhgf/math.py:90-92 (normal)
and ``MultivariateNormal.sufficient_statistics_from_observations`` (``hstack([x, outer(x, x)[tril_indices(d)]])``).
Plain NumPy in the requested dtype, one rounding per operation like XLA-CPU.

Parity status: PINNED by tests/test_nodes/test_exponential_family.py:34-66 (the bivariate spiral golden,
``xis[-1] = [34.65271, -1.0609777, 1210.3647, -36.398651, 3.3951855]``), asserted in tests/test_oracle_golden.py.
Only tests/ may import this module.
"""
import numpy as np


def initial_xis(dimension: int, distribution: str, dtype=np.float32):
    if distribution == "normal":
        return np.array([0.0, 1.0], dtype=dtype)
    d = dimension
    cov = np.identity(d)
    return np.append(np.zeros(d), cov[np.tril_indices(d)]).astype(dtype)


def sufficient_statistics(x, dimension: int, dtype=np.float32):
    x = np.atleast_1d(np.asarray(x, dtype=dtype))
    if dimension == 1:
        return np.array([x[0], x[0] * x[0]], dtype=dtype)
    return np.hstack([x, np.outer(x, x)[np.tril_indices(dimension)]]).astype(dtype)


def ef_filter(input_data, nus=3.0, dimension: int = 1, distribution: str = "normal", observed=None, xis0=None, dtype=np.float32):
    """-> xis trajectory (T, K) of one node."""
    x = np.asarray(input_data, dtype=dtype).reshape(len(input_data), -1)
    xis = initial_xis(dimension, distribution, dtype) if xis0 is None else np.asarray(xis0, dtype=dtype)
    w = dtype(1.0) / (dtype(1.0) + dtype(nus))
    out = np.empty((x.shape[0], xis.size), dtype=dtype)
    for t in range(x.shape[0]):
        ss = sufficient_statistics(x[t], dimension, dtype)
        new = (xis + w * (ss - xis)).astype(dtype)
        xis = new if (observed is None or observed[t]) else xis
        out[t] = xis
    return out


def spiral_data():
    """The data set of tests/test_nodes/test_exponential_family.py:36-44 (legacy NumPy seeding)."""
    np.random.seed(123)
    n = 1000
    theta = np.sort(np.sqrt(np.random.rand(n)) * 5 * np.pi)
    r_a = -2 * theta - np.pi
    return np.array([np.cos(theta) * r_a, np.sin(theta) * r_a]).T + np.random.randn(n, 2) * 2


# ---------------------------------------------------------------------------------------------------------------
# learning = "hgf-1" | "hgf-2" | "hgf-3": the learning rate is implied by continuous HGFs tracking the statistics
# (add_nodes.py:438-452; prediction_error_update_exponential_family_dynamic, exponential.py:63-120;
# posterior_update_exponential_family_dynamic, pyhgf/updates/posterior/exponential.py).  Parity status of THIS part:
# UNPINNED -- the reference's test of it (tests/test_nodes/test_exponential_family.py:69-95) asserts nothing about the
# hgf-2 network (it re-checks the fixed-rate result), so there is no golden to pin to; the restatement follows the code.
def implied_network(n_statistics: int, learning: str, dtype="float32", **kw):
    """The continuous nodes under a dynamic ef-state node, WITHOUT that node, in the reference's creation order
    (implied index = reference index - 1): per statistic A (fed with t_k(x)), B (value parent of A), then a volatility
    parent of B for hgf-2/3 and one of that for hgf-3.  -> (network, [A...], [B...])."""
    from oracle import network as onet

    net = onet.Network(dtype=dtype, **kw)
    a_nodes, b_nodes = [], []
    for _ in range(n_statistics):
        net.add_nodes()
        a_nodes.append(len(net.edges) - 1)
        net.add_nodes(value_children=len(net.edges) - 1)
        b_nodes.append(len(net.edges) - 1)
        if "-2" in learning or "-3" in learning:
            net.add_nodes(volatility_children=len(net.edges) - 1)
        if "-3" in learning:
            net.add_nodes(volatility_children=len(net.edges) - 1)
    net.input_idxs = tuple(a_nodes)  # explicit: the A nodes keep their own tonic volatility / autoconnection
    return net, a_nodes, b_nodes


def ef_dynamic_filter(input_data, dimension=1, distribution="normal", learning="hgf-2", edit=None, observed=None, dtype=np.float32):
    """-> dict(xis (T, K), nus (T,), network).  ``edit(network)`` may change the implied nodes' attributes first."""
    x = np.asarray(input_data, dtype=dtype).reshape(len(input_data), -1)
    T = x.shape[0]
    ss = np.stack([sufficient_statistics(x[t], dimension, dtype) for t in range(T)])
    K = ss.shape[1]
    net, a_nodes, b_nodes = implied_network(K, learning, dtype="float32" if dtype == np.float32 else "float64")
    if edit is not None:
        edit(net)
    obs = None if observed is None else np.repeat(np.asarray(observed)[:, None], K, axis=1)
    net.input_data(ss.astype(np.float64), observed=obs)
    nt = net.node_trajectories
    xis = initial_xis(dimension, distribution, dtype)
    out = {"xis": np.empty((T, K), dtype), "nus": np.empty(T, dtype), "network": net}
    with np.errstate(all="ignore"):
        for t in range(T):
            pe = np.array([nt[a]["mean"][t] - nt[a]["expected_mean"][t] for a in a_nodes], dtype=dtype)
            ed = np.array([nt[b]["mean"][t] - nt[b]["expected_mean"][t] for b in b_nodes], dtype=dtype)
            nu = dtype((pe / ed).astype(dtype).mean(dtype=dtype))
            new = (xis + (dtype(1) / (dtype(1) + nu)) * (ss[t] - xis)).astype(dtype)
            xis = new if (observed is None or observed[t]) else xis
            out["xis"][t], out["nus"][t] = xis, nu
    return out
