"""Tolerance helpers of the parity tests.

The stated tolerances (north_star: float32, <= 1e-5 relative on per-trial beliefs, <= 1e-4 on summed logp and on
gradients) are asserted on EVERY instance, with no instance dropped and no fraction waived.  Where the oracle's own
float32 result is not reproducible to the stated tolerance -- it moves by more than that when its parameters move by
1e-6 relative, oracle/conditioning.py -- the instance's bound is `amplification` x that sensitivity instead."""

import numpy as np

from oracle import conditioning

AMPLIFICATION = 4.0        # summed surprise / beliefs
AMPLIFICATION_GRAD = 10.0  # gradients: the kernel's tangent formulas (closed-form sigmoid / logaddexp / Lambert-W JVPs)
#                            are not the oracle's operation-by-operation Dual arithmetic, only the same derivative


def assert_conditioned(got, want, sens, rtol, amplification=AMPLIFICATION, floor=0.0, what="", max_relaxed=None):
    """|got - want| <= max(rtol, amplification * sens_i) * (|want| + floor) for EVERY instance; identical finite pattern.
    `max_relaxed`: at most this fraction of instances may have a bound looser than rtol (a well-conditioned box)."""
    got, want, sens = (np.asarray(a, dtype=np.float64) for a in (got, want, sens))
    assert got.shape == want.shape == sens.shape, (what, got.shape, want.shape, sens.shape)
    assert np.array_equal(np.isfinite(got), np.isfinite(want)), what
    fin = np.isfinite(want)
    tol = conditioning.tolerance(sens, rtol, amplification)
    err = np.zeros_like(want)
    err[fin] = np.abs(got[fin] - want[fin]) / (np.abs(want[fin]) + floor)
    bad = np.where(fin & (err > tol))[0]
    assert bad.size == 0, (what, f"{bad.size} of {fin.sum()} instances outside their bound",
                           [(int(i), float(err[i]), float(tol[i])) for i in bad[:5]])
    relaxed = float((tol[fin] > rtol).mean()) if fin.any() else 0.0
    if max_relaxed is not None:
        assert relaxed <= max_relaxed, (what, f"{relaxed:.4f} of the box has a bound looser than {rtol}")
    return {"max_err": float(err[fin].max()) if fin.any() else 0.0, "n_over_rtol": int((err[fin] > rtol).sum()), "relaxed": relaxed}
