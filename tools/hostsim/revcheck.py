"""The generated reverse sweep (pyhgf_b200/revgen.py) of ANY network the generator covers, compiled for the host in float64 and
compared with the Dual<double> oracle -- no GPU needed.  Development tool and the engine of the CPU tests of the sweep's
update-scheme / topology coverage (tests/test_device_equations_on_host.py).

    python tools/hostsim/revcheck.py          # a few built-in cases
"""
import ctypes as C
import hashlib
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
WORK = os.path.join(ROOT, "build", "hostsim")


def build(network):
    """Generate + compile the sweep of `network` (a pyhgf_b200.network.Network); returns the loaded library."""
    from pyhgf_b200 import revgen

    text = revgen.generate("t", network, network.volatility_updates, network.mean_field_updates, inputs=tuple(network.input_idxs))
    csrc = os.path.join(ROOT, "pyhgf_b200", "csrc")
    h = hashlib.sha1(text.encode())
    for f in ("hgf_math.cuh", "hgf_ops.cuh", "hgf_scan.cuh", "hgf_program.h", "hgf_reverse.cuh"):
        h.update(open(os.path.join(csrc, f), "rb").read())
    src = os.path.join(ROOT, "tools", "hostsim", "revcheck.cpp")
    h.update(open(src, "rb").read())
    work = os.path.join(WORK, "rev_" + h.hexdigest()[:12])
    so = os.path.join(work, "rev.so")
    if not os.path.exists(so):
        os.makedirs(work, exist_ok=True)
        with open(os.path.join(work, "reverse_t.h"), "w") as f:
            f.write(text)
        subprocess.run(["g++", "-O2", "-ffp-contract=off", "-fopenmp", "-shared", "-fPIC", "-Wno-attributes", "-Wno-unused-variable", "-I", csrc,
                        "-I", "/usr/local/cuda/include", "-I", work, "-DREV_HEADER=<reverse_t.h>", "-DREV_TYPE=Rev_t", src, "-o", so], check=True)
    return C.CDLL(so)


def compare(make, keys, values, series, response, resp_inputs=None, time_steps=None, **net_kw):
    """`make(cls, **kw)` builds the network on either Network class.  Returns (logp equal bit for bit, worst gradient error
    relative to each parameter's largest gradient) of the host-compiled sweep against the Dual<double> oracle."""
    from oracle import network as onet
    from pyhgf_b200 import _lib
    from pyhgf_b200.network import Network

    net = make(Network, dtype="float64", **net_kw)
    net.create_belief_propagation_fn()
    lib = build(net)
    fp = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)  # noqa: E731
    n_in = lib.rev_n_inputs()
    plan = net._get_plan(n_in)
    attrs = np.ascontiguousarray(net._flatten(plan), dtype=np.float64)
    assert attrs.size == lib.rev_n_slots()
    slots = np.array([net._resolve_slot(plan, k) for k in keys], dtype=np.int32)
    p = np.ascontiguousarray(np.stack(values), dtype=np.float64)
    n = p.shape[1]
    u = np.ascontiguousarray(np.asarray(series, dtype=np.float64).reshape(len(series), -1))
    y = None if resp_inputs is None else np.ascontiguousarray(resp_inputs, dtype=np.float64)
    ts = None if time_steps is None else np.ascontiguousarray(time_steps, dtype=np.float64)
    logp, grad = np.zeros(n), np.zeros((len(keys), n))
    lib.rev_run.argtypes = [C.c_int, C.c_int] + [C.c_void_p] * 4 + [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_double, C.c_void_p, C.c_void_p]
    lib.rev_run(n, len(u), fp(u), fp(y), fp(ts), fp(attrs), len(keys), fp(slots), fp(p), _lib.RESPONSES[response], float(net.max_posterior_precision),
                fp(logp), fp(grad))
    o = make(onet.Network, dtype="float64", **net_kw).create_belief_propagation_fn()
    oslots = [o.slot(-1, "response_parameter", k[1]) if k[0] == "response_parameter" else o.slot(*k) for k in keys]
    want_s, want_g = o.batch_grad(series, oslots, p, response, resp_inputs=resp_inputs, time_steps=time_steps)
    # instances whose filter blew up (a NaN belief: +inf surprise) have no gradient to compare; the others must all agree
    ok = np.isfinite(want_s)
    assert ok.sum() >= n // 2, "test regime: most instances should be finite"
    worst = max(float(np.abs(grad[i][ok] - want_g[i][ok]).max() / max(np.abs(want_g[i][ok]).max(), 1e-300)) for i in range(len(keys)))
    return bool(np.array_equal(logp, want_s, equal_nan=True)), worst


CASES = {}


def _cases():
    import nets

    rng = np.random.default_rng(11)
    n = 32
    x1 = nets.ar1_series(150, 1, seed=4)[:, 0]
    x2 = nets.ar1_series(150, 2, seed=5)
    usdchf = np.load(os.path.join(ROOT, "pyhgf_b200", "data", "datasets.npz"))["continuous"][:200]
    for scheme in ("eHGF", "standard", "unbounded"):
        CASES[f"continuous3/{scheme}"] = dict(
            make=nets.continuous3, keys=[(1, "tonic_volatility"), (2, "tonic_volatility"), (3, "tonic_volatility"), (2, "volatility_coupling_children", 0), (2, "mean")],
            values=[rng.uniform(-14, -12, n), rng.uniform(-6, -4, n), rng.uniform(-6, -4, n), rng.uniform(0.8, 1.2, n), rng.uniform(0.5, 1.5, n)],
            series=usdchf, response="first_level_gaussian_surprise", volatility_updates=scheme)
        CASES[f"volatile2/{scheme}"] = dict(
            make=nets.volatile2, keys=[(1, "tonic_volatility_vol"), (1, "mean_vol"), (1, "precision_vol"), (1, "autoconnection_strength")],
            values=[rng.uniform(-5, -3, n), rng.uniform(-3, -1, n), np.exp(rng.uniform(-0.3, 0.3, n)), rng.uniform(0.9, 1.0, n)],
            series=x1, response="first_level_gaussian_surprise", volatility_updates=scheme)
        CASES[f"multi_child/{scheme}"] = dict(
            make=nets.multi_child, keys=[(2, "tonic_volatility"), (3, "tonic_volatility"), (4, "tonic_volatility"), (5, "tonic_volatility"),
                                         (2, "value_coupling_children", 1), (4, "volatility_coupling_children", 0)],
            values=[rng.uniform(-5, -4, n), rng.uniform(-6, -5, n), rng.uniform(-5, -4, n), rng.uniform(-6, -5, n), rng.uniform(0.5, 0.9, n), rng.uniform(0.8, 1.2, n)],
            series=x2, response="total_gaussian_surprise", time_steps=rng.uniform(0.8, 1.2, 150), volatility_updates=scheme)
    for name in ("multi_child/unbounded", "multi_child/eHGF", "volatile2/unbounded", "continuous3/standard"):  # the mean-field variants
        CASES[name + "+mean-field"] = dict(CASES[name], mean_field_updates=True)
    # non-linear value couplings (the reference differentiates the user's callable; here g, g', g'' are restated on traced
    # scalars and differentiated by the generator): every enum function, on a continuous parent, a second-level parent
    # (interior Gaussian child) and a volatile-state parent
    def chain(fn_low, fn_high):
        return lambda cls, **kw: (cls(**kw).add_nodes(precision=10.0).add_nodes(value_children=0, coupling_fn=(fn_low,), tonic_volatility=-3.0)
                                  .add_nodes(value_children=1, coupling_fn=(fn_high,), tonic_volatility=-4.0).add_nodes(volatility_children=1, tonic_volatility=-4.0))

    def volatile_parent(fn):
        return lambda cls, **kw: cls(**kw).add_nodes(precision=10.0).add_nodes(kind="volatile-state", value_children=0, coupling_fn=(fn,), autoconnection_strength=1.0)

    xs = 0.5 * x1 + 0.2
    for low, high in (("tanh", "sin"), ("sigmoid", "gelu"), ("relu", "leaky_relu"), ("gelu", "identity"), ("sin", "tanh")):
        CASES[f"coupling/{low}-{high}"] = dict(
            make=chain(low, high), keys=[(1, "tonic_volatility"), (2, "tonic_volatility"), (3, "tonic_volatility"), (1, "value_coupling_children", 0),
                                         (2, "value_coupling_children", 0), (1, "mean"), (2, "mean")],
            values=[rng.uniform(-4, -3, n), rng.uniform(-5, -4, n), rng.uniform(-5, -4, n), rng.uniform(0.8, 1.2, n), rng.uniform(0.8, 1.2, n),
                    rng.uniform(0.1, 0.5, n), rng.uniform(0.1, 0.5, n)],
            series=xs, response="first_level_gaussian_surprise", volatility_updates="eHGF" if low == "sin" else "unbounded")
    for fn in ("tanh", "gelu"):
        CASES[f"coupling/volatile-{fn}"] = dict(
            make=volatile_parent(fn), keys=[(1, "tonic_volatility_vol"), (1, "mean"), (1, "value_coupling_children", 0), (1, "mean_vol")],
            values=[rng.uniform(-5, -3, n), rng.uniform(0.1, 0.5, n), rng.uniform(0.8, 1.2, n), rng.uniform(-3, -1, n)],
            series=xs, response="first_level_gaussian_surprise")
    return CASES


if __name__ == "__main__":
    for name, case in _cases().items():
        equal, worst = compare(**case)
        print(f"{name:28s} logp bit-equal {equal}  worst gradient error {worst:.2e}")
