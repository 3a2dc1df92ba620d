"""``HGFDistribution`` / ``HGFPointwise`` / ``logp`` / ``hgf_logp`` over the batched CUDA op.

In the reference snapshot these names are tombstones that raise
``NotImplementedError`` (pyhgf/distribution.py:23-66); their surviving call
sites are ``docs/paper.ipynb`` cells 3-4, 19-22, 34-37, 52:

    op = HGFDistribution(n_levels=3, model_type="binary", input_data=u[None, :],
                         response_function=binary_softmax_inverse_temperature,
                         response_function_inputs=y[None, :])
    op(tonic_volatility_2=..., response_function_parameters=...)   # -> sum of log-probabilities

They are re-provided here with that signature: one participant per row of
``input_data``; parameters broadcast over participants (and, with ``n_chains``,
over chains).  The live pattern they stand for is
``jax.vmap(participant_logp)(params).sum()`` and ``jax.grad`` of it
(notebooks 3-Multilevel_HGF cells 16-17, 4-Parameter_recovery cell 11); here that
is ONE kernel launch, the gradient coming from the same pass.

Parity note: the reference holds no implementation or test for these names, so
their results are pinned only through the ``Network`` path they wrap
("parity unpinned" for the wrapper itself).
"""

from __future__ import annotations

from typing import Callable, Optional

import numpy as np
import torch

from .network import Network

_LEVEL_PARAMETERS = ("mean", "precision", "tonic_volatility", "tonic_drift")


def build_hgf(model_type: str = "continuous", n_levels: int = 2, volatility_updates: str = "unbounded",
              continuous_precision: float = 1e4, dtype: str = "float32", device=None, **network_kwargs):
    """The standard 2-/3-level HGF (the removed ``pyhgf.model.HGF``, pyhgf/model/hgf.py:23-30).

    binary: node 0 binary-state (level 1), node 1 its value parent (level 2), node 2 the
    volatility parent (level 3).  continuous: node 0 the input (precision ``continuous_precision``),
    nodes 1..n_levels the levels; level 1 is the input's value parent, each next level the
    volatility parent of the one below.  Returns ``(network, {level: node})``.
    """
    if model_type not in ("binary", "continuous"):
        raise ValueError("Invalid model type. Should be 'binary' or 'continuous'")
    if n_levels not in (2, 3):
        raise ValueError("n_levels should be 2 or 3")
    net = Network(volatility_updates=volatility_updates, dtype=dtype, device=device, **network_kwargs)
    levels = {}
    if model_type == "binary":
        net.add_nodes(kind="binary-state")
        net.add_nodes(kind="continuous-state", value_children=0)
        levels[1], levels[2] = 0, 1
        if n_levels == 3:
            net.add_nodes(kind="continuous-state", volatility_children=1)
            levels[3] = 2
    else:
        net.add_nodes(kind="continuous-state", precision=continuous_precision)
        net.add_nodes(kind="continuous-state", value_children=0)
        net.add_nodes(kind="continuous-state", volatility_children=1)
        levels[1], levels[2] = 1, 2
        if n_levels == 3:
            net.add_nodes(kind="continuous-state", volatility_children=2)
            levels[3] = 3
    return net, levels


class HGFDistribution:
    """Sum of the log-probabilities of N participants' decisions under the HGF (docs/paper.ipynb cells 3, 19, 52).

    Parameters
    ----------
    input_data : ``(N, T)`` observations, one row per participant.
    time_steps : ``(T,)`` or None.
    model_type, n_levels : see :func:`build_hgf`.
    update_type : the network's ``volatility_updates`` ("unbounded" default, "eHGF", "standard").
    response_function : one of ``pyhgf_b200.response``.
    response_function_inputs : ``(N, T)`` decisions, or None.
    n_chains : evaluate ``n_chains`` parameter sets per participant at once; parameters then
        have shape ``(N, n_chains)`` (or broadcast to it) and the op returns ``(n_chains,)`` sums.

    Calling the op with ``mean_k / precision_k / tonic_volatility_k / tonic_drift_k`` (k = level),
    ``volatility_coupling_k`` and ``response_function_parameters`` returns the summed
    log-probability as a device tensor; :meth:`value_and_grad` also returns its gradient.
    """

    pointwise = False

    def __init__(self, input_data=None, time_steps=None, model_type: str = "continuous", update_type: str = "unbounded",
                 n_levels: int = 2, response_function: Optional[Callable] = None, response_function_inputs=None,
                 n_chains: int = 1, dtype: str = "float32", device=None):
        if input_data is None:
            raise ValueError("input_data is required")
        if response_function is None:
            raise ValueError("response_function is required")
        x = np.asarray(input_data)
        if x.ndim == 1:
            x = x[None, :]
        self.n_participants, self.n_steps = x.shape[0], x.shape[1]
        self.n_chains = int(n_chains)
        self.model_type, self.n_levels = model_type, n_levels
        self.response_function = response_function
        self.net, self.levels = build_hgf(model_type, n_levels, update_type, dtype=dtype, device=device)
        self.net.create_belief_propagation_fn()
        dev = self.net.device
        tdt = torch.float32 if self.net.bits == 32 else torch.float64
        # resident on the device across evaluations (a sampler calls the op thousands of times)
        self._input = torch.as_tensor(x).to(device=dev, dtype=tdt).contiguous()
        self._resp = None
        if response_function_inputs is not None and np.size(response_function_inputs) > 0:
            r = np.asarray(response_function_inputs)
            if r.ndim == 1:
                r = r[None, :]
            if r.shape != x.shape:
                raise ValueError("response_function_inputs should have the shape of input_data")
            self._resp = torch.as_tensor(r).to(device=dev, dtype=tdt).contiguous()
        self._time_steps = None if time_steps is None else torch.as_tensor(np.asarray(time_steps)).to(device=dev, dtype=tdt)
        self._group = None
        if self.n_chains > 1:
            self._group = (torch.arange(self.n_participants * self.n_chains, device=dev, dtype=torch.int32) % self.n_chains).contiguous()

    # -- parameter names -> attribute slots ---------------------------------
    def _keys(self, name: str):
        if name == "response_function_parameters":
            return [("response_parameter", 0)]
        stem, _, lvl = name.rpartition("_")
        if not lvl.isdigit():
            raise TypeError(f"unexpected parameter {name!r}")
        k = int(lvl)
        if stem == "volatility_coupling":
            if k not in self.levels or k + 1 not in self.levels or (self.model_type == "binary" and k == 1):
                raise ValueError(f"{name}: the model has no volatility coupling between levels {k} and {k + 1}")
            return [(self.levels[k], "volatility_coupling_parents", 0), (self.levels[k + 1], "volatility_coupling_children", 0)]
        if stem in _LEVEL_PARAMETERS:
            if k not in self.levels:
                raise ValueError(f"{name}: the model has {self.n_levels} levels")
            if self.model_type == "binary" and k == 1:
                raise ValueError(f"{name}: level 1 of a binary HGF is the binary state node, which has no such parameter")
            return [(self.levels[k], stem)]
        raise TypeError(f"unexpected parameter {name!r}")

    def _expand(self, value, dev, tdt):
        n, c = self.n_participants, self.n_chains
        v = torch.as_tensor(np.asarray(value) if not torch.is_tensor(value) else value).to(device=dev, dtype=tdt)
        if v.ndim == 2 and v.shape[-1] == 1 and c == 1:
            v = v[:, 0]
        if c == 1:
            return v.reshape(-1).expand(n) if v.numel() == 1 else v.reshape(n)
        if v.numel() == 1:
            return v.reshape(1, 1).expand(n, c).reshape(-1)
        if v.ndim == 1:
            if v.shape[0] == n:
                return v[:, None].expand(n, c).reshape(-1)
            if v.shape[0] == c:
                return v[None, :].expand(n, c).reshape(-1)
            raise ValueError(f"parameter of length {v.shape[0]} matches neither participants ({n}) nor chains ({c})")
        return v.reshape(n, c).reshape(-1)

    def _run(self, params: dict, gradient: bool, pointwise: bool):
        dev = self._input.device
        tdt = self._input.dtype
        launch_params, names = {}, []
        for name, value in params.items():
            keys = self._keys(name)
            col = self._expand(value, dev, tdt)
            for key in keys:
                launch_params[key] = col
            names.append((name, keys, value))
        n_inst = self.n_participants * self.n_chains
        res = self.net.batch(
            self._input, launch_params, response_function=self.response_function, response_function_inputs=self._resp,
            time_steps=self._time_steps, n_instances=n_inst, input_group_size=self.n_chains,
            response_group_size=self.n_chains, pointwise=pointwise, gradient=gradient and bool(launch_params),
            group=self._group, n_groups=self.n_chains, out_scale=-1.0,
        )
        return res, names

    def _shape_like(self, g, value):
        """Gradient of the summed logp wrt a parameter, in the parameter's own shape."""
        n, c = self.n_participants, self.n_chains
        shape = tuple(np.shape(value)) if not torch.is_tensor(value) else tuple(value.shape)
        g = g.reshape(n, c) if c > 1 else g
        numel = int(np.prod(shape)) if shape else 1
        if numel == 1:
            return (g.sum(0) if c > 1 else g.sum()).reshape(shape if c == 1 else (c,))
        if c > 1 and len(shape) == 1 and shape[0] == n:
            return g  # one value per participant, shared by chains: per-chain gradients (N, C)
        if c > 1 and len(shape) == 1 and shape[0] == c:
            return g.sum(0)
        return g.reshape(shape)

    def __call__(self, **params):
        res, _ = self._run(params, gradient=False, pointwise=self.pointwise)
        if self.pointwise:
            lp = -res.pointwise  # [N*C, T]
            return lp if self.n_chains == 1 else lp.reshape(self.n_participants, self.n_chains, self.n_steps)
        return res.reduced[0] if self.n_chains == 1 else res.reduced

    def value_and_grad(self, **params):
        """``(sum logp, {name: d sum logp / d name})`` from one fused launch (the ``jax.value_and_grad`` of the op)."""
        res, names = self._run(params, gradient=True, pointwise=False)
        grads = {}
        for name, keys, value in names:
            g = sum(res.gradient[k] for k in keys)
            grads[name] = self._shape_like(g, value)  # out_scale = -1: already d logp
        total = res.reduced[0] if self.n_chains == 1 else res.reduced
        return total, grads

    def per_participant(self, **params):
        """Log-probability per participant (and chain): ``(N,)`` or ``(N, n_chains)``."""
        res, _ = self._run(params, gradient=False, pointwise=False)
        lp = res.surprise  # out_scale = -1 -> already logp
        return lp if self.n_chains == 1 else lp.reshape(self.n_participants, self.n_chains)


class HGFPointwise(HGFDistribution):
    """Per-trial log-probabilities ``(N, T)`` (docs/paper.ipynb cells 20, 35), for LOO / WAIC."""

    pointwise = True


def hgf_logp(input_data=None, response_function=None, response_function_inputs=None, model_type: str = "continuous",
             n_levels: int = 2, time_steps=None, update_type: str = "unbounded", dtype: str = "float32", device=None,
             **parameters):
    """``(sum of log-probabilities, per-participant log-probabilities)`` for N participants."""
    op = HGFDistribution(input_data=input_data, time_steps=time_steps, model_type=model_type, update_type=update_type,
                         n_levels=n_levels, response_function=response_function,
                         response_function_inputs=response_function_inputs, dtype=dtype, device=device)
    lp = op.per_participant(**parameters)
    return lp.sum(), lp


def logp(input_data=None, response_function=None, response_function_inputs=None, model_type: str = "continuous",
         n_levels: int = 2, time_steps=None, update_type: str = "unbounded", dtype: str = "float32", device=None,
         **parameters):
    """Log-probability of ONE participant's series under the HGF."""
    total, _ = hgf_logp(input_data=np.asarray(input_data).reshape(1, -1), response_function=response_function,
                        response_function_inputs=None if response_function_inputs is None else np.asarray(
                            response_function_inputs).reshape(1, -1),
                        model_type=model_type, n_levels=n_levels, time_steps=time_steps, update_type=update_type,
                        dtype=dtype, device=device, **parameters)
    return total
