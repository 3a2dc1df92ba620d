"""TEST INFRASTRUCTURE ONLY -- how well-conditioned is an instance, judged by the oracle alone.

The unbounded volatility update and the binary responses contain steps that amplify a rounding-level difference by
orders of magnitude: a saturated belief contradicted by the next observation / decision gives a surprise
``-log(1 - p)`` with ``1 - p ~ 1e-6``, and a belief that jumps across the saturation region within a few trials
carries the jump's own rounding with it.  On such an instance two float32 evaluations of the SAME algorithm that
differ only in legal rounding (libm vs libdevice ``exp``, a contracted multiply-add, an input off by a few ulp)
disagree by 1e-4 .. 1e-2 on the summed surprise, and no float32 implementation can be held to 1e-4 against another.

``sensitivity`` measures exactly that, with the oracle and nothing else: the largest relative change of the oracle's
own float32 result over a small seeded ensemble of relative parameter perturbations of size ``eps`` (default 1e-6,
the size of the kernel's SFU primitive errors).  The parity tests assert EVERY instance against
``max(rtol, amplification * sensitivity_i)``; no instance is dropped and no fraction is waived.
"""

from __future__ import annotations

import numpy as np


def sensitivity(run, params, rows=None, eps: float = 1e-6, n_draws: int = 8, seed: int = 123, floor=0.0):
    """``run(params[P][N]) -> array[..., N]`` (an oracle evaluation); returns, per instance, the largest relative change
    of its result when the rows ``rows`` of ``params`` (default: all) move by ``eps * U(-1, 1)`` relative.  ``floor``
    (scalar or array broadcastable to the result) is added to the denominator for results that cross zero."""
    params = np.asarray(params, dtype=np.float64)
    rows = range(params.shape[0]) if rows is None else rows
    base = np.asarray(run(params), dtype=np.float64)
    rng = np.random.default_rng(seed)
    sens = np.zeros(base.shape)
    for _ in range(n_draws):
        q = params.copy()
        for r in rows:
            q[r] = params[r] * (1.0 + eps * rng.uniform(-1.0, 1.0, params.shape[1]))
        out = np.asarray(run(q), dtype=np.float64)
        with np.errstate(invalid="ignore", divide="ignore"):
            d = np.abs(out - base) / (np.abs(base) + floor)
        d[~np.isfinite(d)] = 0.0
        sens = np.maximum(sens, d)
    return base, sens


def tolerance(sens, rtol: float, amplification: float = 2.0):
    """Per-instance relative tolerance: the stated ``rtol`` wherever the oracle itself is reproducible to it."""
    return np.maximum(rtol, amplification * np.asarray(sens))
