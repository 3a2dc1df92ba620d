"""Generator of the REVERSE SWEEP of the belief-propagation step (north_star: "d logp / d params comes from a hand-written
checkpointed reverse sweep in the same kernel, replacing jax.grad").

The reference differentiates the whole scan in reverse mode (``jax.grad`` over ``lax.scan``, pyhgf/model/network.py:612-624).
Here the adjoint of ONE time step is written out ahead of time, as straight-line CUDA, for a given topology:

  * the update equations of hgf_ops.cuh are restated below on TRACED scalars, one function per opcode, in the same
    operation order (each cites the reference it follows); running the network's schedule on them yields, per scheduled
    operation, an SSA list of primitives (+, -, *, /, exp, log, sigmoid, tanh, sin, cos, erf, logaddexp, lambert_w0, pow,
    select).  Covered: binary, continuous and volatile-state nodes; the unbounded, eHGF and standard volatility updates and
    their mean-field variants; every value-coupling function; every response function;
  * for every operation the generator emits (a) its forward code, which also logs the previous value of every attribute
    slot it overwrites, and (b) its reverse code: restore those slots, recompute the operation's SSA values (so that only
    ONE operation's intermediates are ever live, in registers) and run the adjoint statements in reverse order,
    accumulating into the adjoint attribute vector ``sb``.  The differentiation rules are JAX's: custom JVP/VJP of
    ``lambert_w0`` (pyhgf/math.py:491-498: 1 / (e^w (1 + w))) and ``logaddexp``, and ``where / minimum / maximum /
    clip`` follow the selected branch.

csrc/hgf_reverse.cuh holds the kernel around it: a forward sweep that stores the carried beliefs of every step (the
checkpoints: 2 floats per continuous node), then the backward sweep over time.  Measured, the sweep costs 3.6-3.9 plain
filters whatever the number of parameters; the forward-mode kernels (hgf_math.cuh, Dual<>) cost 1.4-2.5 plain filters per
pass of up to 4 filter parameters + 1 response parameter, so the library takes the sweep as soon as forward mode would
need a second pass (profiles/r2_grad_scaling.log).

    python -m pyhgf_b200.revgen        # (re)writes csrc/generated/reverse_*.h for the BASELINE topologies
"""

from __future__ import annotations

import os
import re

from . import _lib
from .topology import get_update_sequence

NF = _lib.N_FIELDS
F = _lib.FIELD_INDEX
GEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc", "generated")


# ----------------------------------------------------------------------------------------------------------------------
# traced scalars
class V:
    """An SSA value of the traced step.  `diff` False: data or a constant (no adjoint)."""

    __slots__ = ("tr", "name", "diff", "pulls")

    def __init__(self, tr, name, diff):
        self.tr, self.name, self.diff, self.pulls = tr, name, diff, []

    def _bin(self, other, fwd, pull_a, pull_b):
        other = self.tr.lift(other)
        return self.tr.emit(fwd.format(a=self.name, b=other.name), [(self, pull_a), (other, pull_b)])

    def __add__(self, o): return self._bin(o, "{a} + {b}", "{g}", "{g}")
    def __radd__(self, o): return self.tr.lift(o).__add__(self)
    def __sub__(self, o): return self._bin(o, "{a} - {b}", "{g}", "-{g}")
    def __rsub__(self, o): return self.tr.lift(o).__sub__(self)
    def __mul__(self, o): return self._bin(o, "{a} * {b}", "{g} * {b}", "{g} * {a}")
    def __rmul__(self, o): return self.tr.lift(o).__mul__(self)
    def __neg__(self): return self.tr.emit(f"-{self.name}", [(self, "-{g}")])

    def __truediv__(self, o):
        o = self.tr.lift(o)
        # q = a / b:  a_bar += g / b,  b_bar -= g q / b   (one tangent-grade reciprocal, made in the reverse code)
        return self.tr.emit(f"b_div({self.name}, {o.name})", [(self, "{g} * t_rcp({b})"), (o, "-{g} * {r} * t_rcp({b})")], b=o.name)

    def __rtruediv__(self, o): return self.tr.lift(o).__truediv__(self)

    # predicates on the primal (strings of C)
    def __gt__(self, o): return f"({self.name} > {self.tr.lift(o).name})"
    def __ge__(self, o): return f"({self.name} >= {self.tr.lift(o).name})"
    def __lt__(self, o): return f"({self.name} < {self.tr.lift(o).name})"
    def __le__(self, o): return f"({self.name} <= {self.tr.lift(o).name})"


class Tracer:
    def __init__(self):
        self.n = 0
        self.values = []   # (V, forward expression) in creation order, this operation
        self.consts = {}

    def fresh(self, diff):
        self.n += 1
        return V(self, f"v{self.n}", diff)

    def lift(self, x):
        if isinstance(x, V):
            return x
        key = repr(float(x))
        if key not in self.consts:
            v = self.fresh(False)
            self.values.append((v, f"F({key})" if key not in ("inf", "nan", "-inf") else {"inf": "b_inf<F>()", "-inf": "-b_inf<F>()", "nan": "b_nan<F>()"}[key]))
            self.consts[key] = v
        return self.consts[key]

    def emit(self, expr, deps, **names):
        diff = any(d.diff for d, _ in deps)
        v = self.fresh(diff)
        a = deps[0][0].name if deps else ""
        b = names.get("b", deps[1][0].name if len(deps) > 1 else "")
        v.pulls = [(d, p.replace("{a}", a).replace("{b}", b).replace("{r}", v.name)) for d, p in deps if d.diff]
        self.values.append((v, expr))
        return v

    # elementary functions (forward: the primitives of hgf_math.cuh, so that the math policy applies)
    def exp(self, x): return self.emit(f"b_exp({x.name})", [(x, "{g} * {r}")])
    def log(self, x): return self.emit(f"b_log({x.name})", [(x, "{g} * t_rcp({a})")])
    def sigmoid(self, x, exact=False):
        fwd = f"exact_rcp(F(1) + b_exp(-{x.name}))" if exact else f"b_rcp(F(1) + b_exp(-{x.name}))"
        return self.emit(fwd, [(x, "{g} * {r} * (F(1) - {r})")])
    def tanh(self, x): return self.emit(f"b_tanh({x.name})", [(x, "{g} * (F(1) - {r} * {r})")])
    def sin(self, x): return self.emit(f"b_sin({x.name})", [(x, "{g} * b_cos({a})")])
    def cos(self, x): return self.emit(f"b_cos({x.name})", [(x, "-{g} * b_sin({a})")])
    def erf(self, x): return self.emit(f"b_erf({x.name})", [(x, "{g} * F(1.1283791670955126) * b_exp(-{a} * {a})")])
    def lambert_w0(self, z): return self.emit(f"b_lambert_w0<F>({z.name})", [(z, "{g} * t_rcp(b_exp({r}) * (F(1) + {r}))")])
    def logaddexp(self, a, b):
        # jnp.logaddexp VJP: weights exp(a - out), exp(b - out) = the larger argument's 1 / (1 + e), the smaller one's e / (1 + e)
        wa = f"(({a.name} >= {b.name}) ? t_rcp(F(1) + b_exp(-b_abs({a.name} - {b.name}))) : F(1) - t_rcp(F(1) + b_exp(-b_abs({a.name} - {b.name}))))"
        return self.emit(f"b_logaddexp<F>({a.name}, {b.name})", [(a, "{g} * " + wa), (b, "{g} * (F(1) - " + wa + ")")])
    def pow(self, x, y, exact=False):
        fwd = f"exact_pow({x.name}, {y.name})" if exact else f"b_pow({x.name}, {y.name})"
        return self.emit(fwd, [(x, "{g} * {b} * {r} * t_rcp({a})"), (y, "{g} * b_log({a}) * {r}")])
    def select(self, pred, a, b):
        a, b = self.lift(a), self.lift(b)
        return self.emit(f"({pred}) ? {a.name} : {b.name}", [(a, f"(({pred}) ? {{g}} : F(0))"), (b, f"(({pred}) ? F(0) : {{g}})")])
    # jnp.minimum / maximum against a constant bound (hgf_math.cuh: x_min(T, B), x_max(T, B), x_max(B, T))
    def min_c(self, a, c): return self.select(f"{a.name} > {self.lift(c).name}", c, a)
    def max_c(self, a, c): return self.select(f"{a.name} < {self.lift(c).name}", c, a)
    def cmax(self, c, b): return self.select(f"{b.name} <= {self.lift(c).name}", c, b)
    def minimum(self, a, b):  # general x_min: NaN-propagating, the first argument wins ties
        return self.select(f"b_isnan({a.name}) || (!b_isnan({b.name}) && {a.name} <= {b.name})", a, b)
    def maximum(self, a, b):
        return self.select(f"b_isnan({a.name}) || (!b_isnan({b.name}) && {a.name} >= {b.name})", a, b)
    def isfinite(self, x): return f"b_isfinite({x.name})"
    def isnan(self, x): return f"b_isnan({x.name})"


class Op:
    """One scheduled operation being traced: attribute reads become leaves, writes are collected."""

    def __init__(self, layout):
        self.tr, self.layout = Tracer(), layout
        self.reads, self.writes, self.order = {}, {}, []

    def get(self, slot, diff=True):
        if slot in self.writes:
            return self.writes[slot]
        if slot not in self.reads:
            v = self.tr.fresh(diff)
            self.tr.values.append((v, f"s[{slot}]"))
            self.reads[slot] = v
        return self.reads[slot]

    def field(self, node, name, diff=True):
        return self.get(node * NF + F[name], diff)

    def put(self, node, name, v):
        self.writes[node * NF + F[name]] = self.tr.lift(v)

    def data(self, cname):  # a kernel-supplied non-differentiable scalar (dt, log_dt, the response input)
        v = self.tr.fresh(False)
        self.tr.values.append((v, cname))
        return v


class Layout:
    """Slot layout of hgf_plan_create (hgf_program.h) for a network's edges."""

    def __init__(self, edges):
        self.edges, n = edges, len(edges)
        cnt = {k: [len(getattr(e, k) or ()) for e in edges] for k in ("value_parents", "volatility_parents", "value_children", "volatility_children")}
        self.off = {k: [sum(v[:i]) for i in range(n)] for k, v in cnt.items()}
        self.base_cvp = n * NF
        self.base_cvolp = self.base_cvp + sum(cnt["value_parents"])
        self.base_cvc = self.base_cvolp + sum(cnt["volatility_parents"])
        self.base_cvolc = self.base_cvc + sum(cnt["value_children"])
        self.slot_clip = self.base_cvolc + sum(cnt["volatility_children"])
        self.slot_resp = self.slot_clip + 1
        self.n_slots = self.slot_resp + _lib.N_RESPONSE_PARAMETERS

    def cvp(self, node, k): return self.base_cvp + self.off["value_parents"][node] + k
    def cvolp(self, node, k): return self.base_cvolp + self.off["volatility_parents"][node] + k
    def cvc(self, node, k): return self.base_cvc + self.off["value_children"][node] + k
    def cvolc(self, node, k): return self.base_cvolc + self.off["volatility_children"][node] + k


class Unsupported(Exception):
    pass


# ----------------------------------------------------------------------------------------------------------------------
# the update equations on traced scalars -- each mirrors the function of the same name in csrc/hgf_ops.cuh
def _vp(e): return e.value_parents or ()
def _volp(e): return e.volatility_parents or ()
def _vc(e): return e.value_children or ()
def _volc(e): return e.volatility_children or ()


def _coupling_fn(edges, parent, child) -> int:
    """The coupling function (hgf_coupling_fn) of the value edge parent -> child (stored on the parent, add_edges.py)."""
    from .topology import resolve_coupling_fn

    e = edges[parent]
    k = e.value_children.index(child)
    fn = e.coupling_fn[k] if (e.coupling_fn is not None and k < len(e.coupling_fn)) else None
    return resolve_coupling_fn(fn)


CF = _lib.COUPLING_FNS


# g, g', g'' of hgf_ops.cuh (cf_value / cf_prime / cf_second) on traced scalars: the tracer differentiates them like any
# other composition of primitives (the reference applies jax.grad to the user's callable)
def _cf_value(t, fn, x):
    if fn == CF["relu"]:
        return t.max_c(x, 0.0)
    if fn == CF["sigmoid"]:
        return t.sigmoid(x)
    if fn == CF["tanh"]:
        return t.tanh(x)
    if fn == CF["leaky_relu"]:
        return t.select(f"{x.name} >= F(0)", x, x * 0.01)
    if fn == CF["gelu"]:
        return x * 0.5 * (1.0 + t.erf(x / 1.4142135623730951))
    if fn == CF["sin"]:
        return t.sin(x)
    return x


def _cf_prime(t, fn, x):
    if fn == CF["relu"]:
        return t.select(f"{x.name} > F(0)", 1.0, 0.0)
    if fn == CF["sigmoid"]:
        g = t.sigmoid(x)
        return g * (1.0 - g)
    if fn == CF["tanh"]:
        g = t.tanh(x)
        return 1.0 - g * g
    if fn == CF["leaky_relu"]:
        return t.select(f"{x.name} >= F(0)", 1.0, 0.01)
    if fn == CF["gelu"]:
        cdf = 0.5 * (1.0 + t.erf(x / 1.4142135623730951))
        pdf = t.exp(-(x * x) / 2.0) / 2.5066282746310002
        return cdf + x * pdf
    if fn == CF["sin"]:
        return t.cos(x)
    return t.lift(1.0)


def _cf_second(t, fn, x):
    if fn == CF["sigmoid"]:
        g = t.sigmoid(x)
        return g * (1.0 - g) * (1.0 - 2.0 * g)
    if fn == CF["tanh"]:
        g = t.tanh(x)
        return -2.0 * g * (1.0 - g * g)
    if fn == CF["gelu"]:
        pdf = t.exp(-(x * x) / 2.0) / 2.5066282746310002
        return pdf * (2.0 - x * x)
    if fn == CF["sin"]:
        return -t.sin(x)
    return t.lift(0.0)


def predict_binary(op, L, node, dt, log_dt, cap, exact_sigmoid):  # prediction/binary.py:12-93
    t, e = op.tr, L.edges[node]
    mu = t.lift(0.0)
    for p in _vp(e):
        mu = mu + op.field(p, "expected_mean")
    mu = t.sigmoid(mu, exact=exact_sigmoid)
    c = op.get(L.slot_clip)
    mu = t.minimum(t.maximum(mu, c), 1.0 - c)
    op.put(node, "expected_mean", mu)
    op.put(node, "expected_precision", mu * (1.0 - mu))


def _predicted_mean(op, L, node, dt):  # prediction/continuous.py:11-103
    e = L.edges[node]
    lam = op.field(node, "autoconnection_strength")
    if not _vp(e):
        return lam * op.field(node, "mean") + dt * op.field(node, "tonic_drift")
    drift = op.field(node, "tonic_drift")
    for k, p in enumerate(_vp(e)):
        fn = _coupling_fn(L.edges, p, node)
        mu_par = op.field(p, "expected_mean")
        drift = drift + op.get(L.cvp(node, k)) * (mu_par if fn == 0 else _cf_value(op.tr, fn, mu_par))
    return lam * op.field(node, "mean") + dt * drift


def _nan_unless_positive(t, x):
    return t.select(f"{x.name} > F(1e-128)", x, float("nan"))


def predict_continuous(op, L, node, dt, log_dt, cap, exact_sigmoid):  # prediction/continuous.py:106-366
    t, e = op.tr, L.edges[node]
    variance = 1.0 / op.field(node, "precision")
    op.put(node, "temp.current_variance", variance)
    mu_hat = _predicted_mean(op, L, node, dt)
    log_vol = op.field(node, "tonic_volatility")
    for k, p in enumerate(_volp(e)):
        kappa = op.get(L.cvolp(node, k))
        log_vol = log_vol + kappa * op.field(p, "expected_mean")
        if not L.mean_field:  # the MGF term (prediction/continuous.py:207-231; mean-field :369-512 drops it and the Laplace term)
            log_vol = log_vol + (kappa * kappa) / (2.0 * op.field(p, "expected_precision"))
    omega = _nan_unless_positive(t, dt * t.exp(log_vol))
    pi_cond = 1.0 / (variance + omega)
    if not _vp(e) or L.mean_field:
        pi_marg = pi_cond
    else:
        var = t.lift(0.0)
        for k, p in enumerate(_vp(e)):
            fn = _coupling_fn(L.edges, p, node)
            slope = t.lift(1.0) if fn == 0 else _cf_prime(t, fn, op.field(p, "expected_mean"))
            c = dt * op.get(L.cvp(node, k)) * slope
            var = var + (c * c) / op.field(p, "expected_precision")
        pi_marg = 1.0 / (variance + omega + var)
    op.put(node, "temp.effective_precision", omega * pi_marg)
    if not _vc(e) and not _volc(e) and not _volp(e):
        op.put(node, "expected_precision", op.field(node, "precision"))
        op.put(node, "temp.conditional_expected_precision", op.field(node, "precision"))
    else:
        op.put(node, "expected_precision", pi_marg)
        op.put(node, "temp.conditional_expected_precision", pi_cond)
    op.put(node, "expected_mean", mu_hat)


def pe_binary(op, L, node, dt, log_dt, cap, exact_sigmoid):  # prediction_error/binary.py:11-51
    pi_hat = op.field(node, "expected_precision")
    obs = op.field(node, "observed", diff=False)
    op.put(node, "temp.value_prediction_error", ((op.field(node, "mean", diff=False) - op.field(node, "expected_mean")) * obs) / pi_hat)
    op.put(node, "precision", pi_hat)


def pe_continuous(op, L, node, dt, log_dt, cap, exact_sigmoid):  # prediction_error/continuous.py:10-176
    e = L.edges[node]
    is_input = not _vc(e) and not _volc(e)
    delta = op.field(node, "mean", diff=not is_input) - op.field(node, "expected_mean")
    pi_hat = op.field(node, "expected_precision")
    op.put(node, "temp.value_prediction_error", delta)
    vope = pi_hat / op.field(node, "precision") + pi_hat * (delta * delta) - 1.0
    if _volp(e):
        vope = vope / float(len(_volp(e)))
    op.put(node, "temp.volatility_prediction_error", vope)


def _is_gaussian_interior(L, child):
    e = L.edges[child]
    return e.node_type in (2, 6) and bool(_vc(e) or _volc(e))


def _child_precision_weight(op, L, child):  # posterior_update_precision_continuous_node.py:41-80
    pi_hat = op.field(child, "expected_precision")
    if not L.mean_field and _is_gaussian_interior(L, child):
        pi_y = op.field(child, "precision") - pi_hat
        pi_c = op.field(child, "temp.conditional_expected_precision")
        return pi_c * pi_y / (pi_c + pi_y)
    return pi_hat


def _child_gain_precision(op, L, child):  # posterior_update_mean_continuous_node.py:130-144
    pi_hat = op.field(child, "expected_precision")
    if not L.mean_field and _is_gaussian_interior(L, child):
        pi_child = op.field(child, "precision")
        pi_y = pi_child - pi_hat
        pi_c = op.field(child, "temp.conditional_expected_precision")
        return pi_c * pi_child / (pi_c + pi_y)
    return pi_hat


def _posterior_precision(op, L, node, dt, ehgf):  # posterior_update_precision_continuous_node.py:208-365 (eHGF form :627-750)
    t, e = op.tr, L.edges[node]
    obs = [op.field(c, "observed", diff=False) for c in list(_vc(e)) + list(_volc(e))]
    any_observed = " || ".join(f"({o.name} != F(0))" for o in obs)
    # the unobserved branch of the reference's lax.cond, precision_update_missing_values :368-422 (both are evaluated
    # here and one is selected: the generated code is straight-line)
    log_vol = op.field(node, "tonic_volatility")
    for k, p in enumerate(_volp(e)):
        log_vol = log_vol + op.get(L.cvolp(node, k)) * op.field(p, "mean")
    aged = 1.0 / (1.0 / op.field(node, "precision") + dt * t.exp(log_vol))
    from_values = t.lift(0.0)
    for k, c in enumerate(_vc(e)):
        fn = _coupling_fn(L.edges, node, c)
        psi = op.get(L.cvc(node, k))
        if fn == 0:
            curvature = psi * psi
        else:  # posterior_update_precision_continuous_node.py:116-126, at the node's current mean
            g1 = _cf_prime(t, fn, op.field(node, "mean"))
            curvature = (psi * psi) * (g1 * g1) - psi * _cf_second(t, fn, op.field(node, "mean")) * op.field(c, "temp.value_prediction_error")
        from_values = from_values + _child_precision_weight(op, L, c) * curvature * op.field(c, "observed", diff=False)
    from_vol = t.lift(0.0)
    for k, c in enumerate(_volc(e)):
        kappa = op.get(L.cvolc(node, k))
        if ehgf:  # the volatility increment re-predicted from the node's NEW mean, :564-624
            prev_var = op.field(c, "temp.current_variance")
            omega = dt * t.exp(kappa * op.field(node, "mean") + op.field(c, "tonic_volatility"))
            pi_hat = 1.0 / (prev_var + omega)
            gamma = omega * pi_hat
            weight = (omega - prev_var) * pi_hat
            delta = op.field(c, "mean") - op.field(c, "expected_mean")
            vope = (1.0 / op.field(c, "precision") + delta * delta) * pi_hat - 1.0
            inc = t.cmax(0.0, 0.5 * (kappa * kappa) * gamma * (gamma + weight * vope))
            from_vol = from_vol + inc * op.field(c, "observed", diff=False)
        else:
            gamma, vope = op.field(c, "temp.effective_precision"), op.field(c, "temp.volatility_prediction_error")
            kg = kappa * gamma
            from_vol = from_vol + (0.5 * (kg * kg) + (kg * kg) * vope - 0.5 * (kappa * kappa) * gamma * vope) * op.field(c, "observed", diff=False)
    observed = _nan_unless_positive(t, op.field(node, "expected_precision") + (from_values + from_vol))
    return t.select(any_observed, observed, aged)


def _posterior_mean(op, L, node, pi):  # posterior_update_mean_continuous_node.py:10-195
    t, e = op.tr, L.edges[node]
    fv, fvol = t.lift(0.0), t.lift(0.0)
    for k, c in enumerate(_vc(e)):
        fn = _coupling_fn(L.edges, node, c)
        psi = op.get(L.cvc(node, k))
        vape = op.field(c, "temp.value_prediction_error") * op.field(c, "observed", diff=False)
        if fn == 0:
            fv = fv + ((psi * _child_gain_precision(op, L, c)) / pi) * vape
        else:
            fv = fv + ((psi * _child_gain_precision(op, L, c) * _cf_prime(t, fn, op.field(node, "mean"))) / pi) * vape
    for k, c in enumerate(_volc(e)):
        w = op.get(L.cvolc(node, k)) * op.field(c, "temp.effective_precision") * op.field(c, "temp.volatility_prediction_error")
        w = w * (1.0 / (2.0 * pi))
        fvol = fvol + w * op.field(c, "observed", diff=False)
    return op.field(node, "expected_mean") + fv + fvol


def posterior_continuous(op, L, node, dt, log_dt, cap, exact_sigmoid):  # continuous_node_posterior_update.py:20-128: precision, then mean
    pi = op.tr.min_c(_posterior_precision(op, L, node, dt, False), cap)
    op.put(node, "precision", pi)
    op.put(node, "mean", _posterior_mean(op, L, node, pi))


def posterior_continuous_ehgf(op, L, node, dt, log_dt, cap, exact_sigmoid):
    """continuous_node_posterior_update_ehgf.py:20-155: the mean first (with the EXPECTED precision), then the precision
    re-predicted from that mean."""
    op.put(node, "mean", _posterior_mean(op, L, node, op.field(node, "expected_precision")))
    op.put(node, "precision", op.tr.min_c(_posterior_precision(op, L, node, dt, True), cap))


def _unbounded_mixture(t, explicit, prev_variance, c_precision, c_mean, c_expected_mean, c_expected_precision, mu_hat, pi_hat, kappa, omega,
                       log_dt):
    """hgf_ops.cuh::unbounded_mixture<EXPLICIT>: the two-expansion Lambert-W mixture
    (continuous_node_posterior_update_unbounded.py:51-220; volatile_node_posterior_update_unbounded.py:12-178 with kappa = 1 and
    no omega: explicit = False).  Returns (posterior precision, posterior mean)."""
    prev_var = t.max_c(prev_variance, 1e-128)
    delta = c_mean - c_expected_mean
    be = 1.0 / c_precision + delta * delta
    log_prev_var = t.log(prev_var)
    gamma = (log_dt + kappa * mu_hat + omega) if explicit else (log_dt + mu_hat)
    w1 = t.sigmoid(gamma - log_prev_var)
    da1 = c_expected_precision * be - 1.0
    if explicit:
        pi1 = pi_hat + 0.5 * (kappa * kappa) * w1 * (1.0 - w1)
        mu1 = mu_hat + (kappa * w1 / (2.0 * pi1)) * da1
        pi_y = pi_hat / (kappa * kappa)
    else:
        pi1 = pi_hat + 0.5 * w1 * (1.0 - w1)
        mu1 = mu_hat + (w1 / (2.0 * pi1)) * da1
        pi_y = pi_hat
    log_w_arg = t.log(be) - t.log(2.0 * pi_y) + 0.5 / pi_y - gamma
    w_arg = t.exp(t.min_c(log_w_arg, "LOGMAX"))
    y_star = gamma + t.lambert_w0(w_arg) - 0.5 / pi_y
    if explicit:
        x_star = (y_star - log_dt - omega) / kappa
        log_s2 = log_dt + kappa * x_star + omega
    else:
        x_star = y_star - log_dt
        log_s2 = log_dt + x_star
    log_den = t.logaddexp(log_prev_var, log_s2)
    w2 = t.sigmoid(log_s2 - log_prev_var)
    da2 = be * t.exp(-log_den) - 1.0
    if explicit:
        k2 = kappa * kappa
        pi2_full = pi_hat + 0.5 * k2 * w2 * (w2 + (2.0 * w2 - 1.0) * da2)
        pi2_alt = pi_hat + 0.5 * k2 * w2 * (1.0 - w2)
    else:
        pi2_full = pi_hat + 0.5 * w2 * (w2 + (2.0 * w2 - 1.0) * da2)
        pi2_alt = pi_hat + 0.5 * w2 * (1.0 - w2)
    pi2_safe = t.select(f"{pi2_full.name} <= F(0)", pi2_alt, pi2_full)
    if explicit:
        mu2_safe = x_star + (0.5 * kappa * w2 * da2 - pi_hat * (x_star - mu_hat)) / pi2_safe
    else:
        mu2_safe = x_star + (0.5 * w2 * da2 - pi_hat * (x_star - mu_hat)) / pi2_safe
    ok = f"{t.isfinite(pi2_safe)} && {t.isfinite(mu2_safe)}"
    pi2 = t.select(ok, pi2_safe, pi1)
    mu2 = t.select(ok, mu2_safe, mu1)
    arg1 = (log_dt + kappa * mu1 + omega) if explicit else (log_dt + mu1)
    arg2 = (log_dt + kappa * mu2 + omega) if explicit else (log_dt + mu2)
    l1 = t.logaddexp(log_prev_var, arg1)
    e1 = mu1 - mu_hat
    i1 = -0.5 * l1 - 0.5 * be * t.exp(-l1) - 0.5 * pi_hat * (e1 * e1)
    l2 = t.logaddexp(log_prev_var, arg2)
    e2 = mu2 - mu_hat
    i2 = -0.5 * l2 - 0.5 * be * t.exp(-l2) - 0.5 * pi_hat * (e2 * e2)
    b = t.sigmoid(i2 - i1)
    post_mean = (1.0 - b) * mu1 + b * mu2
    dm = mu1 - mu2
    post_precision = 1.0 / ((1.0 - b) / pi1 + b / pi2 + b * (1.0 - b) * (dm * dm))
    return post_precision, post_mean


def posterior_continuous_unbounded(op, L, node, dt, log_dt, cap, exact_sigmoid):
    """continuous_node_posterior_update_unbounded.py:13-48 (reads only the first volatility child, never its `observed` flag)."""
    t, e = op.tr, L.edges[node]
    child = _volc(e)[0]
    pi, mu = _unbounded_mixture(
        t, True, op.field(child, "temp.current_variance"), op.field(child, "precision"), op.field(child, "mean"),
        op.field(child, "expected_mean"), op.field(child, "expected_precision"), op.field(node, "expected_mean"),
        op.field(node, "expected_precision"), op.get(L.cvolc(node, 0)), op.field(child, "tonic_volatility"), log_dt)
    op.put(node, "precision", t.min_c(pi, cap))
    op.put(node, "mean", mu)


# volatile-state nodes (hgf_ops.cuh: predict_volatile, pe_volatile, posterior_volatile)
def predict_volatile(op, L, node, dt, log_dt, cap, exact_sigmoid):  # prediction/volatile.py:230-309
    t, e = op.tr, L.edges[node]
    variance = 1.0 / op.field(node, "precision")
    op.put(node, "temp.current_variance", variance)
    omega_vol = _nan_unless_positive(t, dt * t.exp(op.field(node, "tonic_volatility_vol")))
    pi_hat_vol = 1.0 / (1.0 / op.field(node, "precision_vol") + omega_vol)
    mu_hat_vol = op.field(node, "mean_vol")
    op.put(node, "expected_mean_vol", mu_hat_vol)
    op.put(node, "expected_precision_vol", pi_hat_vol)
    op.put(node, "temp.effective_precision_vol", omega_vol * pi_hat_vol)
    log_vol = mu_hat_vol if L.mean_field else mu_hat_vol + 1.0 / (2.0 * pi_hat_vol)
    omega = _nan_unless_positive(t, dt * t.exp(log_vol))
    pi_cond = 1.0 / (variance + omega)
    if not _vp(e) or L.mean_field:
        pi_marg = pi_cond
    else:
        var = t.lift(0.0)
        for k, p in enumerate(_vp(e)):
            fn = _coupling_fn(L.edges, p, node)
            slope = t.lift(1.0) if fn == 0 else _cf_prime(t, fn, op.field(p, "expected_mean"))
            c = dt * op.get(L.cvp(node, k)) * slope
            var = var + (c * c) / op.field(p, "expected_precision")
        pi_marg = 1.0 / (variance + omega + var)
    lam = op.field(node, "autoconnection_strength")
    drift = t.lift(0.0)  # (a volatile node has no tonic drift, volatile.py:51-108)
    for k, p in enumerate(_vp(e)):
        fn = _coupling_fn(L.edges, p, node)
        mu_par = op.field(p, "expected_mean")
        drift = drift + op.get(L.cvp(node, k)) * (mu_par if fn == 0 else _cf_value(t, fn, mu_par))
    op.put(node, "expected_mean", lam * op.field(node, "mean") + dt * drift)
    if not _vc(e) and not _volc(e) and not _volp(e):
        op.put(node, "expected_precision", op.field(node, "precision"))
        op.put(node, "temp.conditional_expected_precision", op.field(node, "precision"))
        op.put(node, "temp.effective_precision", t.lift(0.0))
    else:
        op.put(node, "expected_precision", pi_marg)
        op.put(node, "temp.conditional_expected_precision", pi_cond)
        op.put(node, "temp.effective_precision", omega * pi_marg)


def pe_volatile(op, L, node, dt, log_dt, cap, exact_sigmoid):  # prediction_error/volatile.py:13-154 + the implicit level's posterior
    t = op.tr
    mean, mu_hat = op.field(node, "mean"), op.field(node, "expected_mean")
    delta = mean - mu_hat
    pi_hat = op.field(node, "expected_precision")
    precision = op.field(node, "precision")
    vope = pi_hat / precision + pi_hat * (delta * delta) - 1.0
    op.put(node, "temp.value_prediction_error", delta)
    op.put(node, "temp.volatility_prediction_error", vope)
    if L.vol_updates == "unbounded":  # volatile_node_posterior_update_unbounded.py:12-178
        pi, mu = _unbounded_mixture(t, False, op.field(node, "temp.current_variance"), precision, mean, mu_hat, pi_hat,
                                    op.field(node, "expected_mean_vol"), op.field(node, "expected_precision_vol"), None, None, log_dt)
        op.put(node, "precision_vol", t.min_c(pi, cap))
        op.put(node, "mean_vol", mu)
    elif L.vol_updates == "eHGF":  # volatile_node_posterior_update_ehgf.py:16-75 + posterior_update_volatility_level.py
        gamma = op.field(node, "temp.effective_precision")
        mean_vol = op.field(node, "expected_mean_vol") + (gamma * vope) / (2.0 * op.field(node, "expected_precision_vol"))
        op.put(node, "mean_vol", mean_vol)
        prev_var = op.field(node, "temp.current_variance")
        omega = dt * t.exp(mean_vol)
        pi_rep = 1.0 / (prev_var + omega)
        gamma_rep = omega * pi_rep
        weight = (omega - prev_var) * pi_rep
        vope_rep = (1.0 / precision + delta * delta) * pi_rep - 1.0
        inc = t.cmax(0.0, 0.5 * gamma_rep * (gamma_rep + weight * vope_rep))
        op.put(node, "precision_vol", t.min_c(op.field(node, "expected_precision_vol") + inc, cap))
    else:  # standard, volatile_node_posterior_update.py:75-107
        gamma = op.field(node, "temp.effective_precision")
        pi_vol = t.min_c(op.field(node, "expected_precision_vol") + (0.5 * (gamma * gamma) + (gamma * gamma) * vope - 0.5 * gamma * vope), cap)
        op.put(node, "precision_vol", pi_vol)
        op.put(node, "mean_vol", op.field(node, "expected_mean_vol") + (gamma * vope) / (2.0 * pi_vol))


def posterior_volatile(op, L, node, dt, log_dt, cap, exact_sigmoid):  # posterior/volatile/posterior_update_value_level.py:11-245
    t, e = op.tr, L.edges[node]
    pi_hat = op.field(node, "expected_precision")
    pi = pi_hat
    for k, c in enumerate(_vc(e)):
        fn = _coupling_fn(L.edges, node, c)
        psi = op.get(L.cvc(node, k))
        w = _child_precision_weight(op, L, c)
        if fn == 0:
            pi = pi + (psi * psi) * w
        else:  # coupling derivatives at the EXPECTED mean (posterior_update_value_level.py:120-125)
            g1 = _cf_prime(t, fn, op.field(node, "expected_mean"))
            g2 = _cf_second(t, fn, op.field(node, "expected_mean"))
            pi = pi + w * ((psi * psi) * (g1 * g1) - psi * g2 * op.field(c, "temp.value_prediction_error"))
    pi = t.min_c(t.maximum(pi, pi_hat), cap)
    op.put(node, "precision", pi)
    gain = t.lift(0.0)
    for k, c in enumerate(_vc(e)):
        fn = _coupling_fn(L.edges, node, c)
        if fn == 0:
            gain = gain + ((op.get(L.cvc(node, k)) * _child_gain_precision(op, L, c)) / pi) * op.field(c, "temp.value_prediction_error")
        else:
            slope = _cf_prime(t, fn, op.field(node, "expected_mean"))
            gain = gain + ((op.get(L.cvc(node, k)) * _child_gain_precision(op, L, c) * slope) / pi) * op.field(c, "temp.value_prediction_error")
    op.put(node, "mean", op.field(node, "expected_mean") + gain)


OPS = {
    _lib.OPS["PRED_BINARY"]: predict_binary, _lib.OPS["PRED_CONT"]: predict_continuous,
    _lib.OPS["PE_BINARY"]: pe_binary, _lib.OPS["PE_CONT"]: pe_continuous,
    _lib.OPS["POST_CONT"]: posterior_continuous, _lib.OPS["POST_CONT_UNBOUNDED"]: posterior_continuous_unbounded,
    _lib.OPS["POST_CONT_EHGF"]: posterior_continuous_ehgf,
    # the mean-field variants are the same functions with Layout.mean_field set (hgf_ops.cuh passes a flag the same way)
    _lib.OPS["PRED_CONT_MF"]: predict_continuous, _lib.OPS["PRED_VOLATILE_MF"]: predict_volatile, _lib.OPS["POST_CONT_MF"]: posterior_continuous,
    _lib.OPS["POST_CONT_EHGF_MF"]: posterior_continuous_ehgf, _lib.OPS["POST_VOLATILE_MF"]: posterior_volatile,
    _lib.OPS["PRED_VOLATILE"]: predict_volatile, _lib.OPS["PE_VOLATILE"]: pe_volatile, _lib.OPS["POST_VOLATILE"]: posterior_volatile,
}


# responses (pyhgf/response.py; hgf_ops.cuh::response_step) -- `y` is the response input, rp the response parameters
def _binary_surprise(t, y_is_one, mu_hat):  # math.py:209-253
    m = t.min_c(t.max_c(mu_hat, 1e-6), 1.0 - 1e-6)
    return t.select(y_is_one, -t.log(m), -t.log(1.0 - m))


def _gaussian_surprise(t, x, mu_hat, pi_hat):  # math.py:256-297
    d = x - mu_hat
    return 0.5 * (t.lift("LOG2PI") - t.log(pi_hat) + pi_hat * (d * d))


def _nan_to_inf(t, x):
    return t.select(t.isnan(x), float("inf"), x)


def response(op, L, fn, y, inputs):
    t = op.tr
    if fn == 1:  # first_level_gaussian_surprise: node 0's observation against node 1's prediction
        return _nan_to_inf(t, _gaussian_surprise(t, op.field(0, "mean", diff=False), op.field(1, "expected_mean"), op.field(1, "expected_precision")))
    if fn == 2:  # total_gaussian_surprise, response.py:51-89: every continuous node (its any-NaN rule is the kernel's)
        if not any(e.node_type == 2 for e in L.edges):
            raise Unsupported("no continuous node")
        total = t.lift(0.0)
        for n, e in enumerate(L.edges):
            if e.node_type == 2:
                total = total + _gaussian_surprise(t, op.field(n, "mean", diff=n not in inputs), op.field(n, "expected_mean"),
                                                   op.field(n, "expected_precision"))
        return total
    if fn == 3:  # first_level_binary_surprise
        total = t.lift(0.0)
        for n in inputs:
            total = total + _binary_surprise(t, f"{op.field(n, 'mean', diff=False).name} != F(0)", op.field(n, "expected_mean"))
        return _nan_to_inf(t, total)
    if fn == 4:  # binary_softmax
        return _nan_to_inf(t, _binary_surprise(t, f"{y.name} != F(0)", op.field(0, "expected_mean")))
    if fn == 5:  # binary_softmax_inverse_temperature (the last-bit arithmetic of hgf_ops.cuh: libm pow, IEEE quotient)
        mu, temp = op.field(0, "expected_mean"), op.get(L.slot_resp)
        p1, p0 = t.pow(mu, temp, exact=True), t.pow(1.0 - mu, temp, exact=True)
        s = p1 + p0
        p = t.emit(f"exact_quot({p1.name}, {s.name})", [(p1, "{g} * t_rcp({b})"), (s, "-{g} * {r} * t_rcp({b})")], b=s.name)
        return _nan_to_inf(t, _binary_surprise(t, f"{y.name} != F(0)", p))
    raise Unsupported(f"response function {fn}")


RESPONSES = (1, 2, 3, 4, 5)


# ----------------------------------------------------------------------------------------------------------------------
# emission
def _emit_forward(op, indent="    "):
    lines = []
    for v, expr in op.tr.values:
        expr = {"LOGMAX": "b_log_max<F>()", "LOG2PI": "b_log(F(2.0) * F(3.141592653589793))"}.get(expr, expr)
        lines.append(f"{indent}const F {v.name} = {expr};")
    return lines


def _emit_reverse(op, out_seeds, indent="    "):
    """Adjoint statements of the operation's SSA list; `out_seeds` {V name: C expression of its incoming adjoint}."""
    lines, has = [], set()
    vals = [v for v, _ in op.tr.values]
    # adjoint variables: every differentiable non-leaf value (leaves accumulate into sb[])
    leaf_of = {v.name: slot for slot, v in op.reads.items()}
    for v in vals:
        if v.diff and v.name not in leaf_of:
            lines.append(f"{indent}F g{v.name[1:]} = {out_seeds.get(v.name, 'F(0)')};")
    for v in reversed(vals):
        if not v.diff or not v.pulls or v.name in leaf_of:
            continue
        g = f"g{v.name[1:]}"
        for d, p in v.pulls:
            term = p.replace("{g}", g)
            if d.name in leaf_of:
                lines.append(f"{indent}sb[{leaf_of[d.name]}] += {term};")
            else:
                lines.append(f"{indent}g{d.name[1:]} += {term};")
    # a leaf that is ALSO written unchanged (s[k] = s[j]) passes its seed straight through
    for name, seed in out_seeds.items():
        if name in leaf_of:
            lines.append(f"{indent}sb[{leaf_of[name]}] += {seed};")
    return lines


def _lift_patch(tr):
    # constants given as names ("LOGMAX", "LOG2PI") become named leaves
    orig = tr.lift

    def lift(x):
        if isinstance(x, str):
            if x not in tr.consts:
                v = tr.fresh(False)
                tr.values.append((v, x))
                tr.consts[x] = v
            return tr.consts[x]
        return orig(x)

    tr.lift = lift


def generate(name, network_like, vol_updates="unbounded", mean_field=False, inputs=None) -> str:
    """The ``Rev_<name>`` struct of a network (anything with ``.edges``); raises Unsupported for node kinds / update
    schemes outside the generated set (then the library keeps to its forward-mode passes)."""
    edges = network_like.edges
    if any(e.node_type not in (1, 2, 6) for e in edges):
        raise Unsupported("the generated reverse sweep covers binary, continuous and volatile state nodes")
    L = Layout(edges)
    L.vol_updates, L.mean_field = vol_updates, bool(mean_field)
    pred, upd = get_update_sequence(network_like, vol_updates, mean_field)
    if inputs is None:
        inputs = tuple(i for i, e in enumerate(edges) if e.value_children is None and e.volatility_children is None)
    steps = [(node, fn.opcode) for node, fn in pred] + [None] + [(node, fn.opcode) for node, fn in upd]
    ops = []
    for st in steps:
        if st is None:
            ops.append(None)  # set_observation happens here (kernel side)
            continue
        node, opcode = st
        if opcode not in OPS:
            raise Unsupported(f"opcode {opcode}")
        op = Op(L)
        _lift_patch(op.tr)
        dt, log_dt, cap = op.data("dt"), op.data("log_dt"), op.data("cap")
        OPS[opcode](op, L, node, dt, log_dt, cap, True)
        ops.append(op)
    # carried slots: read by some operation before any operation of the step (or the observation) wrote them, and
    # written by some operation -- the beliefs that link one step to the next
    written, live_in = set(), set()
    input_slots = {i * NF + F["mean"] for i in inputs} | {i * NF + F["observed"] for i in inputs}
    for op in ops:
        if op is None:
            written |= input_slots
            continue
        for slot in op.reads:
            if slot not in written:
                live_in.add(slot)
        written |= set(op.writes)
    all_written = set().union(*[set(op.writes) for op in ops if op is not None])
    carried = sorted(live_in & all_written)
    out = [f"// GENERATED by pyhgf_b200/revgen.py -- reverse sweep of topology \"{name}\" ({len(edges)} nodes); do not edit.\n",
           "#pragma once\n#include \"hgf_math.cuh\"\n#include \"hgf_program.h\"\n\n#ifndef HGF_LIT\n#define HGF_LIT(c) F(c)\n#endif\n\nnamespace hgf {\n\n",
           f"template <class F> struct Rev_{name} {{\n",
           f"  static constexpr int kSlots = {L.n_slots}, kNodes = {len(edges)}, kCarried = {len(carried)}, kInputs = {len(inputs)};\n",
           f"  static constexpr int kSlotResp = {L.slot_resp};\n"]
    log_sizes = [0 if op is None else len(op.writes) for op in ops]
    out.append(f"  static constexpr int kLog = {sum(log_sizes)};\n")
    out.append("  HGF_D static int carried_slot(int i) { switch (i) { " + " ".join(f"case {i}: return {s};" for i, s in enumerate(carried)) + " default: return 0; } }\n")
    out.append("  HGF_D static int input_node(int k) { switch (k) { " + " ".join(f"case {k}: return {n};" for k, n in enumerate(inputs)) + " default: return 0; } }\n")

    def op_forward(op, log_base, with_log):
        lines = _emit_forward(op)
        for j, (slot, v) in enumerate(sorted(op.writes.items())):
            if with_log:
                lines.append(f"    log[{log_base + j}] = s[{slot}];")
        for slot, v in sorted(op.writes.items()):
            lines.append(f"    s[{slot}] = {v.name};")
        return lines

    # forward halves: predictions, then (after the kernel has set the observations) updates; each in a plain and a
    # logging form
    for with_log in (False, True):
        for part, rng in (("predict", range(0, ops.index(None))), ("update", range(ops.index(None) + 1, len(ops)))):
            sig = f"  HGF_D static void {part}{'_log' if with_log else ''}(F* s, F dt, F log_dt, F cap{', F* log' if with_log else ''}) {{\n"
            out.append(sig)
            for i in rng:
                base = sum(log_sizes[:i])
                out.append("   {\n" + "\n".join(op_forward(ops[i], base, with_log)) + "\n   }\n")
            out.append("  }\n")
    # reverse halves
    for part, rng in (("update", range(len(ops) - 1, ops.index(None), -1)), ("predict", range(ops.index(None) - 1, -1, -1))):
        out.append(f"  HGF_D static void {part}_reverse(F* s, F* sb, const F* log, F dt, F log_dt, F cap) {{\n")
        for i in rng:
            op, base = ops[i], sum(log_sizes[:i])
            lines = []
            seeds = {}
            for j, (slot, v) in enumerate(sorted(op.writes.items())):
                lines.append(f"    s[{slot}] = log[{base + j}];")  # the attribute vector as this operation found it
            for slot, v in sorted(op.writes.items()):
                lines.append(f"    const F o{slot} = sb[{slot}]; sb[{slot}] = F(0);")
                seeds[v.name] = (seeds[v.name] + f" + o{slot}") if v.name in seeds else f"o{slot}"
            lines += _emit_forward(op)
            lines += _emit_reverse(op, seeds)
            out.append("   {\n" + "\n".join(lines) + "\n   }\n")
        out.append("  }\n")
    # responses
    out.append("  HGF_D static F response(int fn, const F* s, F y) {\n    switch (fn) {\n")
    resp_ops = {}
    for fn in RESPONSES:
        op = Op(L)
        _lift_patch(op.tr)
        try:
            r = response(op, L, fn, op.data("y"), inputs)
        except (Unsupported, IndexError):
            continue
        resp_ops[fn] = (op, r)
        out.append(f"    case {fn}: {{\n" + "\n".join(_emit_forward(op, "      ")) + f"\n      return {r.name};\n    }}\n")
    out.append("    default: return F(0);\n    }\n  }\n")
    out.append("  HGF_D static void response_reverse(int fn, const F* s, F y, F seed, F* sb) {\n    switch (fn) {\n")
    for fn, (op, r) in resp_ops.items():
        out.append(f"    case {fn}: {{\n" + "\n".join(_emit_forward(op, "      ")) + "\n" + "\n".join(_emit_reverse(op, {r.name: "seed"}, "      ")) + "\n      break;\n    }\n")
    out.append("    default: break;\n    }\n  }\n};\n\n}  // namespace hgf\n")
    # numeric literals go through HGF_LIT, so that the same straight-line code also compiles with F a SIMD vector type
    # (oracle/hgf_simd.cpp, the instance-vectorised CPU arm of bench.py), where a functional cast from a scalar does not exist
    return re.sub(r"\bF\((-?[0-9][0-9.e+\-]*)\)", r"HGF_LIT(\1)", "".join(out))


def main():
    from . import codegen

    os.makedirs(GEN, exist_ok=True)
    names = []
    for name in ("binary2", "binary3", "continuous2", "continuous3", "ghgf8", "ghgf16"):
        spec = codegen.TOPOLOGIES[name]
        text = generate(name, spec[0](), spec[1], spec[2])
        codegen._write_if_changed(os.path.join(GEN, f"reverse_{name}.h"), text)
        names.append(name)
    return names


if __name__ == "__main__":
    print("reverse sweeps generated for", main())
