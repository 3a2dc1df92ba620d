"""What do the shared reciprocals (hgf_math.cuh::Recip) change in the results?

Builds on-demand specialisations of the 3-level binary network from the same sources -- the shipped policy (twice, under
two names: the control), and the same with -DHGF_NO_SHARED_RCP (every quotient of the unbounded update takes its own
MUFU.RCP again) -- runs them on one batch (plain filter, logp + gradient) and compares the results BIT FOR BIT.  Own
process: a registered specialisation serves every later plan of its topology (hgf_select.cu: newest first), so the test
suite runs this as a subprocess.

Measured (profiles/r2_shared_rcp_check.log): the gradient kernels return identical bits.  The plain-filter kernel does
not: the seeds are the same bits (tools/ubench/rcp_scaling.cu), but with the quotients written as seed x numerator nvcc
contracts the OTHER product of (1 - b) / pi1 + b / pi2 into the fma (visible in the two PTX files), one rounding placed
differently per step -- 45 % of the summed logp's move in the last bit (median 1.1e-7 relative), as between any two
compilations of one source.  Exit code 0: control identical, gradient kernels identical, plain filter within 1e-5 on
99 % of the instances (the rest are the ill-conditioned instances tests/parity.py relaxes).

    python tools/shared_rcp_check.py [--build-only]      # prints one JSON line
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

KERNELS = [(0, 0), (2, 1)]
# (the -DHGF_SHARED_RCP_TWIN flags are no-ops that give the objects distinct names: "shared_again" is the A/A control)
BUILDS = (("shared", ["-DHGF_SHARED_RCP_TWIN=0"]), ("shared_again", ["-DHGF_SHARED_RCP_TWIN=1"]), ("separate", ["-DHGF_NO_SHARED_RCP"]))


def main():
    import nets
    from pyhgf_b200 import specialise as sp
    from pyhgf_b200.network import Network

    build_only = "--build-only" in sys.argv
    if build_only:
        for _, flags in BUILDS:
            net = nets.binary3(Network)
            net.create_belief_propagation_fn()
            print(sp.specialise(net, float32=KERNELS, extra_flags=flags, load=False))
        return 0
    import torch

    from pyhgf_b200 import response as R

    rng = np.random.default_rng(11)
    n, t = 8192, 320
    u = (rng.random(t) < 0.5 + 0.3 * np.sin(np.arange(t) / 25.0)).astype(np.float64)
    y = (rng.random(t) < 0.5).astype(np.float64)
    params = {(1, "tonic_volatility"): rng.uniform(-7, -2, n), (2, "tonic_volatility"): rng.uniform(-7, -2, n),
              ("response_parameter", 0): np.exp(rng.uniform(-0.3, 0.8, n))}
    kw = dict(response_function=R.binary_softmax_inverse_temperature, response_function_inputs=y)
    out, names = {}, {}
    what = ["filter logp", "gradient-pass logp"] + [f"d/d{k}" for k in params]

    def run(net, name):
        f = net.batch(u, params, **kw)
        g = net.batch(u, params, gradient="forward", **kw)
        assert f.kernel_name == name and g.kernel_name == name, (f.kernel_name, g.kernel_name, name)
        torch.cuda.synchronize()
        return [f.surprise.cpu().numpy(), g.surprise.cpu().numpy()] + [g.gradient[k].cpu().numpy() for k in params]

    def bits(x, y):
        return {w: int((a.view(np.uint32) != b.view(np.uint32)).sum()) for w, a, b in zip(what, x, y)}

    for label, flags in BUILDS:
        net = nets.binary3(Network)
        names[label] = sp.specialise(net, float32=KERNELS, extra_flags=flags)
        out[label] = run(net, names[label])
        again = run(net, names[label])
        assert not any(bits(out[label], again).values()), ("the same kernel twice", label, bits(out[label], again))
    assert len(set(names.values())) == len(BUILDS)
    control = bits(out["shared"], out["shared_again"])
    diff = bits(out["shared"], out["separate"])
    rel = {w: np.abs(a - b) / np.maximum(np.abs(a), 1e-30) for w, a, b in zip(what, out["shared"], out["separate"]) if diff[w]}
    print(json.dumps({"instances": n, "steps": t, "kernels": names, "control_same_flags_differing": control, "values_differing_in_any_bit": diff,
                      "rel_diff_median_of_differing": {w: float(np.nanmedian(r[r > 0])) for w, r in rel.items()},
                      "rel_diff_p99": {w: float(np.nanquantile(r, 0.99)) for w, r in rel.items()},
                      "rel_diff_max": {w: float(np.nanmax(r)) for w, r in rel.items()}}))
    gradient_kernels_equal = not any(v for w, v in diff.items() if w != "filter logp")
    filter_close = "filter logp" not in rel or float(np.nanquantile(rel["filter logp"], 0.99)) < 1e-5
    return 0 if gradient_kernels_equal and filter_close and not any(control.values()) else 1


if __name__ == "__main__":
    sys.exit(main())
