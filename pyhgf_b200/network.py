"""``Network`` -- the model API of the path, backed by the CUDA time loop.

Mirrors the hot-path surface of ``pyhgf.model.Network``
(pyhgf/model/network.py:85-155 constructor, :157-180 ``input_idxs``, :182-234
``create_belief_propagation_fn``, :528-628 ``input_data``, :769-776
``get_network``, :778-945 ``add_nodes``, :986-1022 ``surprise``, :1025-1067
``add_edges``) with the same names, argument meaning and error behaviour.  The
per-observation ``lax.scan`` over ``beliefs_propagation`` is replaced by one
launch of ``libpyhgf_cuda.so`` (``hgf_filter`` / ``hgf_logp_grad``,
include/pyhgf_cuda.h); what the reference obtains with ``jax.vmap`` over
parameters is the native batch dimension of that launch (:meth:`Network.batch`).

There is no CPU path: every method that computes needs the built library and a
CUDA device.
"""

from __future__ import annotations

import os

import ctypes as C
from copy import deepcopy
from typing import Callable, Optional, Sequence, Union

import numpy as np
import torch

from . import _lib, categorical, ef_state
from .topology import AdjacencyLists, Plan, UpdateSequence, get_update_sequence

_COUPLING_KEYS = ("value_coupling_parents", "volatility_coupling_parents", "value_coupling_children",
                  "volatility_coupling_children")
_EDGE_OF_KEY = {"value_coupling_parents": "value_parents", "volatility_coupling_parents": "volatility_parents",
                "value_coupling_children": "value_children", "volatility_coupling_children": "volatility_children"}


# ---------------------------------------------------------------------------
# node defaults (the state layout) -- pyhgf/model/add_nodes.py
# ---------------------------------------------------------------------------
def _continuous_state_defaults() -> dict:  # add_nodes.py:65-85
    return {
        "mean": 0.0, "expected_mean": 0.0, "precision": 1.0, "expected_precision": 1.0,
        "volatility_coupling_children": None, "volatility_coupling_parents": None,
        "value_coupling_children": None, "value_coupling_parents": None,
        "tonic_volatility": -4.0, "tonic_drift": 0.0, "autoconnection_strength": 1.0, "observed": 1,
        "temp": {"effective_precision": 0.0, "conditional_expected_precision": 1.0, "value_prediction_error": 0.0,
                 "volatility_prediction_error": 0.0, "current_variance": 1.0},
    }


def _volatile_state_defaults() -> dict:  # add_nodes.py:159-184
    return {
        "mean": 0.0, "expected_mean": 0.0, "precision": 1.0, "expected_precision": 1.0,
        "autoconnection_strength": 0.0,
        "mean_vol": 0.0, "expected_mean_vol": 0.0, "precision_vol": 1.0, "expected_precision_vol": 1.0,
        "tonic_volatility_vol": -4.0,
        "value_coupling_children": None, "value_coupling_parents": None,
        "temp": {"effective_precision": 0.0, "conditional_expected_precision": 1.0, "value_prediction_error": 0.0,
                 "volatility_prediction_error": 0.0, "effective_precision_vol": 0.0, "current_variance": 1.0},
    }


def _binary_state_defaults() -> dict:  # add_nodes.py:330-340
    return {
        "observed": 1, "mean": 0, "expected_mean": 0.5, "precision": 1.0, "expected_precision": 1.0,
        "value_coupling_parents": None, "temp": {"value_prediction_error": 0.0},
    }


_NODE_KINDS = {"continuous-state": (2, _continuous_state_defaults), "volatile-state": (6, _volatile_state_defaults),
               "binary-state": (1, _binary_state_defaults)}
_OTHER_KINDS = ("dp-state", "constant-state")
_VECTOR_KINDS = ("ef-state", "categorical-state")  # own kernels beside the scan (ef_state.py, categorical.py)


def get_couplings(value_parents, volatility_parents, value_children, volatility_children):
    """Normalise coupling specifications to ``(indexes, strengths)`` tuples (add_nodes.py:603-650)."""
    out = []
    for indexes in (value_parents, volatility_parents, value_children, volatility_children):
        if indexes is None:
            out.append((None, None))
        elif isinstance(indexes, (int, np.integer)):
            out.append(((int(indexes),), (1.0,)))
        elif isinstance(indexes, list):
            out.append((tuple(indexes), tuple([1.0] * len(indexes))))
        elif isinstance(indexes, tuple):
            out.append((tuple(indexes[0]), tuple(indexes[1])))
        else:
            raise TypeError("coupling indexes should be an int, a list of ints or a tuple (indexes, strengths)")
    return out


def _torch_dtype(bits: int):
    return torch.float32 if bits == 32 else torch.float64


def _np_dtype(bits: int):
    return np.float32 if bits == 32 else np.float64


ParamKey = Union[tuple, str]


class BatchResult:
    """Outputs of one batched launch (device tensors).

    ``surprise`` is the per-instance summed surprise ``[N]`` (``logp = -surprise``),
    ``gradient[key]`` its derivative wrt a per-instance parameter ``[N]``
    (``logp_gradient`` flips the sign), ``pointwise`` the per-trial surprise as an
    ``[N, T]`` view, ``trajectories[(node, key)]`` ``[N, T]`` views of the recorded
    leaves and ``last_attributes[(node, key)]`` the final state ``[N]``.
    """

    def __init__(self):
        self.surprise = None
        self.gradient = {}
        self.pointwise = None
        self.trajectories = {}
        self.last_attributes = None
        self.reduced = None
        self.kernel_name = ""

    def relaunch(self, stream=None, consume: bool = True):
        """Enqueue the same evaluation again on the same device buffers (inputs and outputs stay resident
        in HBM; a sampler or a benchmark overwrites the parameter tensors in place between calls).
        Host cost: one C call.  With a fused cross-GPU reduction, ``consume=False`` skips the wait for the other
        ranks' sums (``self.reduced`` is then only valid after ``fused_reduce.result()`` has been asked for it)."""
        s = torch.cuda.current_stream(self._device) if stream is None else stream
        fused = getattr(self, "_fused", None)
        with torch.cuda.device(self._device), torch.cuda.stream(s):  # the zeroing is ordered with the kernel, on ITS stream
            if fused is not None:
                fused.attach(self._call)  # next epoch; the all-reduced sums land in the workspace
            elif self.reduced is not None:
                self.reduced.zero_()
            _lib.check(self._fn(self._plan.handle, C.byref(self._call), C.c_void_p(s.cuda_stream)))
            if fused is not None:
                self.reduced = fused.result(wait=consume, stream=s)
        return self

    @property
    def logp(self):
        return None if self.surprise is None else -self.surprise

    @property
    def logp_gradient(self):
        return {k: -v for k, v in self.gradient.items()}


class ComponentsResult(BatchResult):
    """Outputs of a batched launch that ``Network.batch`` split over the network's connected components
    (one launch per component, on its own stream): the same fields as :class:`BatchResult`, keys in the
    original node numbering."""

    def __init__(self, parts, streams, device):
        super().__init__()
        self._parts, self._streams, self._device = parts, streams, device
        self.kernel_name = f"components({len(parts)}): " + " | ".join(p.kernel_name for p in parts)

    def relaunch(self, stream=None, consume: bool = True):
        cur = torch.cuda.current_stream(self._device) if stream is None else stream
        for part, s in zip(self._parts, self._streams):
            s.wait_stream(cur)
            part.relaunch(stream=s)
        for s in self._streams:
            cur.wait_stream(s)
        return self


class Network:
    """A predictive-coding network whose belief propagation runs on a B200.

    Parameters are those of ``pyhgf.model.Network`` (network.py:85-91);
    ``dtype`` ("float32" default, as JAX; or "float64") and ``device`` are
    additions.
    """

    def __init__(self, volatility_updates: str = "unbounded", max_posterior_precision: float = 1e10,
                 mean_field_updates: bool = False, precision_clipping_value: float = 1e-6, dtype: str = "float32",
                 device: Optional[Union[int, str, torch.device]] = None) -> None:
        if volatility_updates not in ("standard", "eHGF", "unbounded"):
            raise ValueError("Invalid update type. Should be one of the following: 'standard', 'eHGF', 'unbounded'")
        self.edges: tuple = ()
        self.n_nodes: int = 0
        self.node_trajectories: Optional[dict] = None
        self.attributes: dict = {-1: {"time_step": 0.0, "precision_clipping_value": float(precision_clipping_value)}}
        self.last_attributes: Optional[dict] = None
        self.update_sequence = None
        self.scan_fn = None
        self._input_idxs: Optional[tuple] = None
        self.input_dim: list = []
        self.volatility_updates = volatility_updates
        self.max_posterior_precision = float(max_posterior_precision)
        self.mean_field_updates = bool(mean_field_updates)
        name = str(dtype).replace("torch.", "")
        if name not in ("float32", "float64"):
            raise ValueError("dtype should be 'float32' or 'float64'")
        self.bits = 32 if name == "float32" else 64
        self._device = device
        self._plan: Optional[Plan] = None
        self._plan_n_fed = None

    # ------------------------------------------------------------------ graph
    def add_nodes(self, kind: str = "continuous-state", n_nodes: int = 1, node_parameters: dict = {},
                  value_children=None, value_parents=None, volatility_children=None, volatility_parents=None,
                  coupling_fn: tuple = (None,), **additional_parameters):
        """Add state node(s) (network.py:778-945; defaults add_nodes.py:65-85, 159-184, 330-340)."""
        if kind not in _NODE_KINDS and kind not in _OTHER_KINDS and kind not in _VECTOR_KINDS:
            raise ValueError(
                "Invalid node type. Should be one of the following: 'dp-state', 'continuous-state', 'binary-state', "
                "'ef-state', 'categorical-state', 'volatile-state', or 'constant-state'"
            )
        if kind in _OTHER_KINDS:
            raise NotImplementedError(
                f"'{kind}' nodes are outside the belief-propagation path this backend accelerates "
                "(binary-, continuous-, volatile- and categorical-state nodes, ef-state nodes with a fixed learning rate)"
            )
        if kind == "ef-state":
            return self._add_ef_state(n_nodes, node_parameters, additional_parameters, value_children)
        if kind == "categorical-state":
            return self._add_categorical_state(n_nodes, node_parameters, additional_parameters)
        vp, volp, vc, volc = get_couplings(value_parents, volatility_parents, value_children, volatility_children)
        node_type, make_defaults = _NODE_KINDS[kind]
        if node_type == 6:  # the volatility level is implicit (add_nodes.py:106-202)
            volp, volc = (None, None), (None, None)
        params = make_defaults()
        # update_parameters, add_nodes.py:653-697
        unknown = [k for k in additional_parameters if k not in params]
        if unknown:
            raise ValueError(
                f"Some parameter(s) passed as keyword arguments were not found in the default key(s) of a {kind} node "
                f"(i.e. {list(params.keys())}). If you want to create a new key in the node's attributes, "
                "please use the node_parameters argument instead."
            )
        params.update(additional_parameters)
        params.update(node_parameters or {})

        coupling_fn = tuple(coupling_fn)
        if vc[0] is not None and len(vc[0]) != len(coupling_fn):  # insert_nodes, add_nodes.py:746-762
            if len(coupling_fn) == 1:
                coupling_fn = len(vc[0]) * coupling_fn
            else:
                raise ValueError("The number of coupling fn and value children do not match")

        for _ in range(n_nodes):
            idx = self.n_nodes
            self.edges = self.edges + (AdjacencyLists(node_type, None, None, None, None, coupling_fn),)
            self.attributes[idx] = deepcopy(params)
            if vp[0] is not None:
                for j, c in zip(*vp):
                    self.add_edges(parent_idxs=j, children_idxs=idx, kind="value", coupling_strengths=c)
            if vc[0] is not None:
                for j, c in zip(*vc):
                    self.add_edges(parent_idxs=idx, children_idxs=j, kind="value", coupling_strengths=c)
            if volc[0] is not None:
                for j, c in zip(*volc):
                    self.add_edges(parent_idxs=idx, children_idxs=j, kind="volatility", coupling_strengths=c)
            if volp[0] is not None:
                for j, c in zip(*volp):
                    self.add_edges(parent_idxs=j, children_idxs=idx, kind="volatility", coupling_strengths=c)
            self.n_nodes += 1
        self._invalidate()
        return self

    def _add_ef_state(self, n_nodes, node_parameters, additional_parameters, value_children):
        """Exponential-family state node(s), add_nodes.py:360-470 (fixed learning rate only, see ef_state.py)."""
        if value_children is not None:
            raise NotImplementedError("an ef-state node with value children is outside the accelerated path")
        params = ef_state.defaults()
        unknown = [k for k in additional_parameters if k not in params]
        if unknown:
            raise ValueError(
                f"Some parameter(s) passed as keyword arguments were not found in the default key(s) of a ef-state node "
                f"(i.e. {list(params.keys())}). If you want to create a new key in the node's attributes, "
                "please use the node_parameters argument instead."
            )
        params.update(additional_parameters)
        params.update(node_parameters or {})
        learning = str(params["learning"])
        params = ef_state.finish_parameters(params)
        if "hgf" in learning and n_nodes != 1:
            raise NotImplementedError("add ef-state nodes with learning='hgf-*' one at a time")
        for _ in range(n_nodes):
            node = self.n_nodes
            self.edges = self.edges + (AdjacencyLists(ef_state.NODE_TYPE, None, None, None, None, (None,)),)
            self.attributes[node] = deepcopy(params)
            self.n_nodes += 1
            if "hgf" in learning:  # continuous nodes tracking each sufficient statistic (add_nodes.py:438-452)
                for _k in range(ef_state.n_sufficient_statistics(int(params["dimension"]))):
                    self.add_nodes(value_children=node)
                    self.add_nodes(value_children=self.n_nodes - 1)
                    if "-2" in learning or "-3" in learning:
                        self.add_nodes(volatility_children=self.n_nodes - 1)
                    if "-3" in learning:
                        self.add_nodes(volatility_children=self.n_nodes - 1)
        self._invalidate()
        return self

    def _add_categorical_state(self, n_nodes, node_parameters, additional_parameters):
        """Categorical state node and its implied binary HGFs (add_nodes.py:473-545, fill_categorical_state_node.py:13-81)."""
        if n_nodes != 1:
            raise NotImplementedError("add categorical-state nodes one at a time (the reference gives n_nodes > 1 the same binary nodes)")
        node_parameters = dict(node_parameters or {})
        n = int(node_parameters.get("n_categories", additional_parameters.get("n_categories", 4)))
        if not 2 <= n <= _lib.MAX_CATEGORIES:
            raise ValueError(f"n_categories should be between 2 and {_lib.MAX_CATEGORIES}")
        params = categorical.defaults(n, 1 + len(self.edges))
        unknown = [k for k in additional_parameters if k not in params]
        if unknown:
            raise ValueError(
                f"Some parameter(s) passed as keyword arguments were not found in the default key(s) of a categorical-state node "
                f"(i.e. {list(params.keys())}). If you want to create a new key in the node's attributes, "
                "please use the node_parameters argument instead."
            )
        params.update(additional_parameters)
        params.update(node_parameters)
        node = self.n_nodes
        self.edges = self.edges + (AdjacencyLists(categorical.NODE_TYPE, None, None, None, None, (None,)),)
        self.attributes[node] = deepcopy(params)
        self.n_nodes += 1
        binary = list(params["binary_idxs"])
        bp = categorical.default_binary_parameters(n)  # as the reference: the DEFAULTS reach the implied network
        self.add_nodes(kind="binary-state", n_nodes=n, node_parameters={"mean": bp["mean_1"], "precision": bp["precision_1"]})
        edges = list(self.edges)
        edges[node] = AdjacencyLists(categorical.NODE_TYPE, tuple(binary), None, None, None, (None,))
        for b in binary:
            edges[b] = AdjacencyLists(1, None, None, (node,), None, (None,))
        self.edges = tuple(edges)
        for b in binary:
            self.add_nodes(value_children=b, node_parameters={"mean": bp["mean_2"], "precision": bp["precision_2"],
                                                              "tonic_volatility": bp["tonic_volatility_2"]})
        self.add_nodes(volatility_children=[b + n for b in binary],
                       node_parameters={"mean": bp["mean_3"], "precision": bp["precision_3"], "tonic_volatility": bp["tonic_volatility_3"]})
        self._invalidate()
        return self

    def _categorical_nodes(self) -> list:
        return [i for i, e in enumerate(self.edges) if e.node_type == categorical.NODE_TYPE]

    def _vector_nodes(self) -> list:
        """Nodes whose own step runs beside the scan, on the trajectories of the nodes implied under them: categorical
        nodes and ef-state nodes with learning='hgf-*' (those have value parents)."""
        return [i for i, e in enumerate(self.edges)
                if e.node_type == categorical.NODE_TYPE or (e.node_type == ef_state.NODE_TYPE and e.value_parents is not None)]

    def _implied_network(self):
        """(sub-network, {node: sub-network node}): this network without its categorical / dynamic ef-state nodes.  The
        nodes under such an input become the inputs, one column each: the binary nodes get the one-hot observation
        (categorical_state_prediction_error, prediction_error/categorical.py:36-45), the statistic-tracking nodes get t(x)
        (prediction_error_update_exponential_family_dynamic, exponential.py:104-118); attributes are copied at call time."""
        cats = set(self._vector_nodes())
        nodes = [i for i in range(self.n_nodes) if i not in cats]
        local = {g: k for k, g in enumerate(nodes)}
        sub = Network(volatility_updates=self.volatility_updates, max_posterior_precision=self.max_posterior_precision,
                      mean_field_updates=self.mean_field_updates, precision_clipping_value=self.attributes[-1]["precision_clipping_value"],
                      dtype="float32" if self.bits == 32 else "float64", device=self._device)
        remap = lambda t: None if t is None else (tuple(local[j] for j in t if j not in cats) or None)  # noqa: E731
        sub.edges = tuple(AdjacencyLists(self.edges[g].node_type, remap(self.edges[g].value_parents), remap(self.edges[g].volatility_parents),
                                         remap(self.edges[g].value_children), remap(self.edges[g].volatility_children),
                                         self.edges[g].coupling_fn) for g in nodes)
        sub.n_nodes = len(nodes)
        for g in nodes:
            a = deepcopy(self.attributes[g])
            for key, field in _EDGE_OF_KEY.items():  # coupling strengths of the edges that went with the removed nodes
                members = getattr(self.edges[g], field)
                if members is not None and a.get(key) is not None and any(j in cats for j in members):
                    kept = [c for j, c in zip(members, np.atleast_1d(a[key])) if j not in cats]
                    a[key] = np.asarray(kept, dtype=np.float64) if kept else None
            sub.attributes[local[g]] = a
        sub.attributes[-1] = deepcopy(self.attributes[-1])
        inputs = []
        for g in self.input_idxs:
            inputs += [local[b] for b in self.edges[g].value_parents] if g in cats else [local[g]]
        sub._input_idxs = tuple(inputs)
        return sub, local

    def _ef_nodes(self) -> list:
        """The ef-state nodes, after checking that the network is one this backend runs: the hgf_ef_filter path
        serves networks made of ef-state nodes only (they have no parents, so nothing couples them to HGF nodes)."""
        ef = [i for i, e in enumerate(self.edges) if e.node_type == ef_state.NODE_TYPE and e.value_parents is None]
        if ef and len(ef) != len(self.edges):
            raise NotImplementedError("ef-state nodes and HGF nodes in one network: filter them as two networks")
        return ef

    def add_edges(self, parent_idxs, children_idxs, kind: str = "value", coupling_strengths=1.0,
                  coupling_fn: tuple = (None,)):
        """Add value / volatility coupling(s) (network.py:1025-1067, utils/add_edges.py:10-176).

        As in the reference, ``coupling_fn`` is accepted but the function in force
        is the one the parent was created with.
        """
        if kind not in ("value", "volatility"):
            raise ValueError(f"The kind of coupling should be value or volatility, got {kind}")
        children = [children_idxs] if isinstance(children_idxs, (int, np.integer)) else list(children_idxs)
        parents = [parent_idxs] if isinstance(parent_idxs, (int, np.integer)) else list(parent_idxs)
        if isinstance(coupling_strengths, (int, float, np.floating, np.integer)):
            strengths = [float(coupling_strengths)]
        else:
            strengths = [float(c) for c in coupling_strengths]
        edges = list(self.edges)
        child_field, child_key = (("value_children", "value_coupling_children") if kind == "value"
                                  else ("volatility_children", "volatility_coupling_children"))
        parent_field, parent_key = (("value_parents", "value_coupling_parents") if kind == "value"
                                    else ("volatility_parents", "volatility_coupling_parents"))
        for p in parents:
            e = edges[p]._asdict()
            old = self.attributes[p].get(child_key)
            if e[child_field] is None:
                e[child_field] = tuple(children)
                self.attributes[p][child_key] = np.asarray(strengths, dtype=np.float64)
            else:
                e[child_field] = e[child_field] + tuple(children)
                self.attributes[p][child_key] = np.concatenate([np.atleast_1d(old), strengths])
            edges[p] = AdjacencyLists(**e)
        for c in children:
            e = edges[c]._asdict()
            old = self.attributes[c].get(parent_key)
            if e[parent_field] is None:
                e[parent_field] = tuple(parents)
                self.attributes[c][parent_key] = np.asarray(strengths, dtype=np.float64)
            else:
                e[parent_field] = e[parent_field] + tuple(parents)
                self.attributes[c][parent_key] = np.concatenate([np.atleast_1d(old), strengths])
            edges[c] = AdjacencyLists(**e)
        self.edges = tuple(edges)
        self._invalidate()
        return self

    def _invalidate(self):
        self._plan = None
        self.scan_fn = None
        self.update_sequence = None

    def _invalidate_plan(self):
        self._plan = None

    def connected_components(self) -> list:
        """The node sets of the network's connected components (any coupling kind), each sorted, ordered by first node."""
        parent = list(range(self.n_nodes))

        def find(i):
            while parent[i] != i:
                parent[i] = parent[parent[i]]
                i = parent[i]
            return i

        for i, e in enumerate(self.edges):
            for kind in ("value_parents", "volatility_parents", "value_children", "volatility_children"):
                for j in getattr(e, kind) or ():
                    parent[find(i)] = find(j)
        groups: dict = {}
        for i in range(self.n_nodes):
            groups.setdefault(find(i), []).append(i)
        return sorted(groups.values(), key=lambda g: g[0])

    def _component_network(self, nodes):
        """The sub-network of one connected component, nodes renumbered in ascending order; attributes are copied
        from this network at call time (they may have been edited since the last call)."""
        local = {g: k for k, g in enumerate(nodes)}
        sub = Network(volatility_updates=self.volatility_updates, max_posterior_precision=self.max_posterior_precision,
                      mean_field_updates=self.mean_field_updates, precision_clipping_value=self.attributes[-1]["precision_clipping_value"],
                      dtype="float32" if self.bits == 32 else "float64", device=self._device)
        remap = lambda t: None if t is None else tuple(local[j] for j in t)  # noqa: E731
        sub.edges = tuple(AdjacencyLists(self.edges[g].node_type, remap(self.edges[g].value_parents), remap(self.edges[g].volatility_parents),
                                         remap(self.edges[g].value_children), remap(self.edges[g].volatility_children),
                                         self.edges[g].coupling_fn) for g in nodes)
        sub.n_nodes = len(nodes)
        for g in nodes:
            sub.attributes[local[g]] = deepcopy(self.attributes[g])
        sub.attributes[-1] = deepcopy(self.attributes[-1])
        sub._input_idxs = tuple(local[g] for g in self.input_idxs if g in local)
        return sub, local

    def specialise(self, gradient_parameters: int = 0, response_parameter: bool = False, float64: bool = False,
                   verbose: bool = False, reverse: bool = False) -> str:
        """Build (once per topology, cached on disk) and register register-resident kernels for THIS network's
        topology, so that it does not run through the opcode interpreter (``pyhgf_b200/specialise.py``).  The plain
        filter is always instantiated; ``gradient_parameters = k`` adds the logp+grad kernel for k filter parameters
        per pass (<= 4), ``response_parameter`` the variant that also differentiates one response-function parameter.
        ``reverse=True`` also builds the topology's generated reverse sweep (one launch for any number of parameters).
        Returns the kernel name the network's plans report afterwards."""
        from . import specialise as sp

        if self._vector_nodes():  # the scan runs the implied network: that is the topology to specialise
            if self.scan_fn is None:
                self.create_belief_propagation_fn()
            return self._implied_network()[0].specialise(gradient_parameters, response_parameter, float64, verbose, reverse)
        kernels = [(0, 0)]
        if gradient_parameters or response_parameter:
            kernels.append((min(int(gradient_parameters), 4), 1 if response_parameter else 0))
        return sp.specialise(self, float32=() if float64 else kernels, float64=kernels if float64 else (), verbose=verbose, reverse=reverse)

    # ----------------------------------------------------------------- inputs
    @property
    def input_idxs(self) -> tuple:
        """Nodes receiving observations (network.py:157-180).

        Auto-detection (nodes without children) has the reference's side effect:
        continuous input nodes get ``autoconnection_strength = tonic_volatility = 0``.
        """
        if self._input_idxs is not None:
            return self._input_idxs
        auto = tuple(i for i, e in enumerate(self.edges) if e.value_children is None and e.volatility_children is None)
        for idx in auto:
            if self.edges[idx].node_type == 2:
                self.attributes[idx]["autoconnection_strength"] = 0.0
                self.attributes[idx]["tonic_volatility"] = 0.0
            self.attributes[idx].setdefault("observed", 1)
        return auto

    @input_idxs.setter
    def input_idxs(self, value):
        self._input_idxs = tuple(int(v) for v in value)
        self._plan = None

    # --------------------------------------------------------------- lowering
    def create_belief_propagation_fn(self, overwrite: bool = True):
        """Derive the static schedule and lower the network into a CUDA plan (network.py:182-234)."""
        _ = self.input_idxs
        if self._ef_nodes():  # get_input_dimension, network.py:1069-1086; the schedule is one PE step per node
            self.input_dim = [int(self.attributes[i]["dimension"]) for i in self.input_idxs]
            self.update_sequence = ((), tuple((i, "prediction_error_update_exponential_family_fixed") for i in self.input_idxs))
            self.scan_fn = True
            return self
        cats = self._vector_nodes()
        if cats:  # the schedule of the implied nodes, framed by the vector nodes' steps (get_update_sequence.py:282-305, 337-345)
            implied, local = self._implied_network()
            implied.create_belief_propagation_fn()
            back = {k: g for g, k in local.items()}
            seq = implied.update_sequence
            is_cat = lambda c: self.edges[c].node_type == categorical.NODE_TYPE  # noqa: E731
            self.input_dim = [w for _, _, w in self._categorical_columns(None)]
            self.update_sequence = UpdateSequence(
                tuple((back[i], f) for i, f in seq.prediction_steps),
                tuple((c, categorical.categorical_state_prediction_error if is_cat(c)
                       else ef_state.prediction_error_update_exponential_family_dynamic) for c in cats)
                + tuple((back[i], f) for i, f in seq.update_steps)
                + tuple((c, categorical.categorical_state_update if is_cat(c)
                         else ef_state.posterior_update_exponential_family_dynamic) for c in cats))
            self.scan_fn = True
            return self
        self.input_dim = [1 for _ in self.input_idxs]
        if self.update_sequence is None or overwrite:
            self.update_sequence = get_update_sequence(self, self.volatility_updates, self.mean_field_updates)
        self._plan = None
        self._get_plan(len(self.input_idxs))
        self.scan_fn = self._plan  # truthy marker: the "scan function" is the lowered plan
        return self

    def get_network(self):
        """(attributes, edges, update_sequence), network.py:769-776."""
        if self.scan_fn is None:
            self.create_belief_propagation_fn()
        return self.attributes, self.edges, self.update_sequence

    def _get_plan(self, n_fed: int) -> Plan:
        if self.update_sequence is None:
            _ = self.input_idxs
            self.update_sequence = get_update_sequence(self, self.volatility_updates, self.mean_field_updates)
        if self._plan is None or self._plan_n_fed != n_fed:
            self._plan = Plan(self.edges, self.update_sequence, self.input_idxs, n_fed, self.volatility_updates,
                              self.mean_field_updates, self.max_posterior_precision)
            self._plan_n_fed = n_fed
        return self._plan

    @property
    def plan(self) -> Plan:
        if self._plan is None:
            self.create_belief_propagation_fn()
        return self._plan

    @property
    def device(self) -> torch.device:
        if not torch.cuda.is_available():
            raise RuntimeError("pyhgf_b200 needs a CUDA device: the belief-propagation path has no CPU implementation")
        d = self._device
        if d is None:
            return torch.device("cuda", torch.cuda.current_device())
        if isinstance(d, int):
            return torch.device("cuda", d)
        return torch.device(d)

    # ------------------------------------------------------- attribute vector
    def _flatten(self, plan: Plan, attributes: Optional[dict] = None, response_parameters=None) -> np.ndarray:
        """One instance's attribute vector in the library's slot layout (hgf_program.h)."""
        attributes = self.attributes if attributes is None else attributes
        v = np.zeros(plan.n_slots, dtype=np.float64)
        for i in range(self.n_nodes):
            a = attributes[i]
            temp = a.get("temp", {})
            for k, name in enumerate(_lib.FIELDS):
                if name.startswith("temp."):
                    val = temp.get(name[5:], 0.0)
                else:
                    val = a.get(name, 1.0 if name == "observed" else 0.0)
                v[i * _lib.N_FIELDS + k] = float(val)
            for key in _COUPLING_KEYS:
                arr = a.get(key)
                if arr is not None:
                    for pos, c in enumerate(np.atleast_1d(np.asarray(arr, dtype=np.float64))):
                        v[plan.slot(i, key, pos)] = c
        v[plan.slot_clipping] = float(attributes[-1]["precision_clipping_value"])
        if response_parameters is not None:
            rp = np.atleast_1d(np.asarray(response_parameters, dtype=np.float64)).ravel()
            if rp.size > _lib.N_RESPONSE_PARAMETERS:
                raise ValueError(f"at most {_lib.N_RESPONSE_PARAMETERS} response-function parameters are supported")
            v[plan.slot_response: plan.slot_response + rp.size] = rp
        return v.astype(_np_dtype(self.bits))

    def _resolve_slot(self, plan: Plan, key: ParamKey) -> int:
        """``(node, attribute)``, ``(node, coupling_key, position)``, ``"response_parameter"`` or
        ``("response_parameter", j)`` -> slot."""
        if isinstance(key, str):
            key = (key,)
        if key[0] == "response_parameter":
            return plan.slot(-1, "response_parameter", int(key[1]) if len(key) > 1 else 0)
        if key[0] == "precision_clipping_value" or (key[0] == -1 and key[1] == "precision_clipping_value"):
            return plan.slot_clipping
        node, name = int(key[0]), key[1]
        pos = int(key[2]) if len(key) > 2 else 0
        return plan.slot(node, name, pos)

    # ------------------------------------------------------------- the launch
    def _format_series(self, input_data, observed, dev, grouped: bool):
        """-> values [G, T, n_fed] on the device (+ uint8 mask or None)."""
        dt = _torch_dtype(self.bits)
        x = torch.as_tensor(np.asarray(input_data) if not torch.is_tensor(input_data) else input_data)
        x = x.to(device=dev, dtype=dt, non_blocking=True)
        if grouped:
            if x.ndim == 2:
                x = x[:, :, None]
        else:
            if x.ndim == 1:
                x = x[:, None]
            x = x[None]
        if x.ndim != 3:
            raise ValueError("input_data should be (time,), (time, inputs) or, per series, (series, time[, inputs])")
        n_inputs = len(self.input_idxs)
        if x.shape[2] > n_inputs:
            raise ValueError(f"input_data has {x.shape[2]} columns but the network has {n_inputs} input node(s)")
        obs = None
        if observed is not None:
            if isinstance(observed, (tuple, list)) and not grouped:
                observed = np.stack([np.asarray(o) for o in observed], axis=1)
            o = torch.as_tensor(np.asarray(observed) if not torch.is_tensor(observed) else observed).to(dev)
            if grouped:
                if o.ndim == 2:
                    o = o[:, :, None]
            else:
                if o.ndim == 1:
                    o = o[:, None]
                o = o[None]
            n_fed = min(x.shape[2], o.shape[2])
            x = x[:, :, :n_fed]
            obs = (o[:, :, :n_fed] != 0).to(torch.uint8).contiguous()
        return x.contiguous(), obs

    def _launch(self, *, values, stream=None, **kw):
        """Stage, allocate and enqueue one launch.  With ``stream`` given, ALL of it happens in that stream's order
        (it first waits for what the caller's current stream has enqueued, i.e. for the inputs): staging copies and
        output allocations made on another stream would race with the kernel."""
        dev = values.device
        cur = torch.cuda.current_stream(dev)
        if stream is not None and stream != cur:
            stream.wait_stream(cur)
            with torch.cuda.device(dev), torch.cuda.stream(stream):
                return self._launch_impl(values=values, stream=stream, **kw)
        return self._launch_impl(values=values, stream=stream, **kw)

    def _launch_impl(self, *, values, observed, time_steps, input_group_size, n_instances, params: dict,
                     response_fn: int, response_inputs, response_group_size, response_parameters, record_slots,
                     want_last, want_sum, want_pointwise, want_grad, group, n_groups, out_scale=1.0, stream=None,
                     attributes=None, fused_reduce=None, generative_seed=None, plan=None, branch_idx=None):
        L = _lib.lib()
        dev = values.device
        dt = _torch_dtype(self.bits)
        n_fed = values.shape[2]
        plan = self._get_plan(n_fed) if plan is None else plan
        T = int(values.shape[1])
        N = int(n_instances)
        keep = [values]
        call = _lib.Call()
        call.dtype = self.bits
        call.n_instances = N
        call.n_steps = T
        call.input_group_size = int(input_group_size)
        call.input_data = values.data_ptr()
        if observed is not None:
            keep.append(observed)
            call.observed = observed.data_ptr()
        if time_steps is not None:
            ts = torch.as_tensor(np.asarray(time_steps) if not torch.is_tensor(time_steps) else time_steps)
            ts = ts.to(device=dev, dtype=dt).contiguous()
            if ts.numel() != T:
                raise ValueError("time_steps should have one entry per observation")
            keep.append(ts)
            call.time_steps = ts.data_ptr()
        if branch_idx is not None:
            br = torch.as_tensor(np.asarray(branch_idx, dtype=np.int32)).to(dev).contiguous()
            if br.numel() != T:
                raise ValueError("branches_idx should have one entry per observation")
            keep.append(br)
            call.branch_idx = br.data_ptr()
        attr = torch.from_numpy(self._flatten(plan, attributes, response_parameters)).to(dev)
        keep.append(attr)
        call.attributes = attr.data_ptr()
        keys = list(params.keys())
        P = len(keys)
        if P > _lib.MAX_PARAMS:
            raise ValueError(f"at most {_lib.MAX_PARAMS} per-instance parameters per launch")
        slots = (C.c_int32 * max(P, 1))(*[self._resolve_slot(plan, k) for k in keys])
        call.n_params = P
        call.param_slots = C.cast(slots, C.POINTER(C.c_int32))
        if P:
            cols = []
            for k in keys:
                col = params[k]
                col = torch.as_tensor(np.asarray(col) if not torch.is_tensor(col) else col).to(device=dev, dtype=dt).reshape(-1)
                if col.numel() == 1 and N != 1:
                    col = col.expand(N)
                if col.numel() != N:
                    raise ValueError(f"parameter {k!r} has {col.numel()} values for {N} instances")
                cols.append(col)
            pmat = torch.stack(cols, 0).contiguous()
            keep.append(pmat)
            call.params = pmat.data_ptr()
        call.response_fn = int(response_fn)
        call.response_group_size = int(response_group_size)
        if response_inputs is not None:
            ri = torch.as_tensor(np.asarray(response_inputs) if not torch.is_tensor(response_inputs) else response_inputs)
            ri = ri.to(device=dev, dtype=dt).contiguous()
            if ri.shape[-1] != T:
                raise ValueError("response_function_inputs should have one entry per observation")
            keep.append(ri)
            call.response_inputs = ri.data_ptr()
        res = BatchResult()
        res.kernel_name = plan.kernel_name
        R = len(record_slots)
        rec = (C.c_int32 * max(R, 1))(*record_slots)
        traj = None
        if R:
            if R > _lib.MAX_RECORDS:
                raise ValueError(f"at most {_lib.MAX_RECORDS} recorded leaves per launch")
            traj = torch.empty((R, T, N), device=dev, dtype=dt)
            call.n_records = R
            call.record_slots = C.cast(rec, C.POINTER(C.c_int32))
            call.trajectories = traj.data_ptr()
        last = None
        if want_last:
            last = torch.empty((self.n_nodes * _lib.N_FIELDS, N), device=dev, dtype=dt)
            call.last_attributes = last.data_ptr()
        if want_sum:
            res.surprise = torch.empty(N, device=dev, dtype=dt)
            call.surprise_sum = res.surprise.data_ptr()
        if want_pointwise:
            pw = torch.empty((T, N), device=dev, dtype=dt)
            call.surprise_pointwise = pw.data_ptr()
            res.pointwise = pw.t()
        grad = None
        if want_grad:
            grad = torch.empty((P, N), device=dev, dtype=dt)
            call.gradient = grad.data_ptr()
            # "forward" / "reverse" force one of the two gradient kernels (hgf_call_t::gradient_mode); True: the library's choice
            call.gradient_mode = {"forward": 1, "reverse": 2}.get(want_grad, 0) if isinstance(want_grad, str) else 0
        if n_groups:
            if fused_reduce is not None:
                if fused_reduce.n_groups != n_groups or fused_reduce.bits != self.bits:
                    raise ValueError("fused_reduce was built for another n_groups / dtype")
                fused_reduce.attach(call)
            else:
                res.reduced = torch.zeros(n_groups, device=dev, dtype=dt)
                call.reduced = res.reduced.data_ptr()
            if group is not None:
                g = torch.as_tensor(group).to(device=dev, dtype=torch.int32).contiguous()
                keep.append(g)
                call.group = g.data_ptr()
        call.out_scale = float(out_scale)
        if generative_seed is not None:
            call.generative, call.seed = 1, int(generative_seed) & 0xFFFFFFFFFFFFFFFF
        s = torch.cuda.current_stream(dev) if stream is None else stream
        fn = L.hgf_logp_grad if want_grad else L.hgf_filter
        with torch.cuda.device(dev):
            _lib.check(fn(plan.handle, C.byref(call), C.c_void_p(s.cuda_stream)))
        for t in keep:  # inputs staged above must outlive the (asynchronous) kernel
            if t.is_cuda:
                t.record_stream(s)
        # everything a relaunch needs stays referenced by the result (BatchResult.relaunch)
        res._call, res._fn, res._plan, res._keep, res._device = call, fn, plan, (keep, slots, rec), dev
        res._traj = traj
        res._last = last
        res._fused = fused_reduce if n_groups else None
        if res._fused is not None:
            res.reduced = fused_reduce.result(stream=s)
        if grad is not None:
            res.gradient = {k: grad[i] for i, k in enumerate(keys)}
            res._grad = grad
        return res, plan

    # ------------------------------------------------------------ input_data
    def input_data(self, input_data, time_steps=None, observed=None, input_idxs=None, rng_keys=None,
                   record_trajectories: bool = True):
        """Filter one observation series (network.py:528-628).

        ``input_data`` is ``(time,)`` or ``(time, n_inputs)``.  Every call restarts
        from ``self.attributes``.  Populates ``node_trajectories`` (NumPy arrays of
        length ``time`` per node attribute) and ``last_attributes``.
        """
        if rng_keys is not None:
            raise NotImplementedError("rng_keys only serve action functions, which are outside the accelerated path")
        if input_idxs is not None:
            self.input_idxs = input_idxs
        if self.scan_fn is None:
            self.create_belief_propagation_fn()
        if self._ef_nodes():
            return self._ef_input_data(input_data, time_steps, observed, record_trajectories)
        if self._vector_nodes():
            return self._categorical_input_data(input_data, time_steps, observed)
        self._filter_one_series(input_data, time_steps, observed, record_trajectories)
        self._last_inputs = (input_data, time_steps, observed)
        return self

    def input_custom_sequence(self, update_branches, branches_idx, input_data, time_steps=None, observed=None,
                              input_idxs=None, rng_keys=None):
        """Filter one series with an update sequence chosen per observation (network.py:653-767): step ``t`` runs
        ``update_branches[branches_idx[t]]`` (``lax.switch`` in the reference, the interpreter's branch table here).

        ``update_branches`` are ``UpdateSequence``s over this network's nodes -- e.g. ``self.update_sequence`` with some
        steps removed; every step's function must be one of ``pyhgf_b200.topology``'s update functions.
        """
        if rng_keys is not None:
            raise NotImplementedError("rng_keys only serve action functions, which are outside the accelerated path")
        if input_idxs is not None:
            self.input_idxs = input_idxs
        if self.scan_fn is None:
            self.create_belief_propagation_fn()
        if self._ef_nodes() or self._vector_nodes():
            raise NotImplementedError("input_custom_sequence serves binary-, continuous- and volatile-state networks")
        branches = tuple((tuple(b[0]), tuple(b[1])) for b in update_branches)
        if not 1 <= len(branches) <= _lib.MAX_BRANCHES:
            raise ValueError(f"between 1 and {_lib.MAX_BRANCHES} update branches, got {len(branches)}")
        for b in branches:
            for _, fn in b[0] + b[1]:
                if not hasattr(fn, "opcode"):
                    raise NotImplementedError("custom update functions cannot cross the C ABI: build the branches from "
                                              "the steps of Network.update_sequence")
        self._filter_one_series(input_data, time_steps, observed, True, branches=branches, branches_idx=branches_idx)
        return self

    def _filter_one_series(self, input_data, time_steps, observed, record_trajectories, branches=None, branches_idx=None):
        dev = self.device
        values, obs = self._format_series(input_data, observed, dev, grouped=False)
        plan = self._get_plan(values.shape[2])
        if branches is not None and len(branches) == 1:
            plan = Plan(self.edges, branches[0], self.input_idxs, values.shape[2], self.volatility_updates,
                        self.mean_field_updates, self.max_posterior_precision)
            branches_idx = None
        elif branches is not None:
            plan = Plan(self.edges, branches[0], self.input_idxs, values.shape[2], self.volatility_updates,
                        self.mean_field_updates, self.max_posterior_precision, branches=branches)
        T = values.shape[1]
        record = list(range(self.n_nodes * _lib.N_FIELDS)) if record_trajectories else []
        res, plan = self._launch(values=values, observed=obs, time_steps=time_steps, input_group_size=0, n_instances=1,
                                 params={}, response_fn=0, response_inputs=None, response_group_size=0,
                                 response_parameters=None, record_slots=record, want_last=True, want_sum=False,
                                 want_pointwise=False, want_grad=False, group=None, n_groups=0, plan=plan,
                                 branch_idx=branches_idx)
        last = res._last[:, 0].double().cpu().numpy()
        ts = np.ones(T) if time_steps is None else np.asarray(time_steps, dtype=np.float64)
        self.last_attributes = self._unflatten(last, float(ts[-1]) if T else 0.0)
        if record_trajectories:
            traj = res._traj[:, :, 0].cpu().numpy()  # [R, T]
            nt = {}
            for i in range(self.n_nodes):
                a = self.attributes[i]
                d, temp = {}, {}
                for k, name in enumerate(_lib.FIELDS):
                    col = traj[i * _lib.N_FIELDS + k]
                    if name.startswith("temp."):
                        if name[5:] in a.get("temp", {}):
                            temp[name[5:]] = col
                    elif name in a:
                        d[name] = col
                for key in _COUPLING_KEYS:
                    if key in a:
                        d[key] = None if a[key] is None else np.broadcast_to(
                            np.asarray(a[key], dtype=traj.dtype), (T, np.size(a[key])))
                d["temp"] = temp
                nt[i] = d
            nt[-1] = {"time_step": ts.astype(traj.dtype),
                      "precision_clipping_value": np.full(T, self.attributes[-1]["precision_clipping_value"], dtype=traj.dtype)}
            self.node_trajectories = nt
        else:
            self.node_trajectories = None

    # ------------------------------------------------- categorical-state nodes
    def _categorical_columns(self, n_columns: int) -> list:
        """[(input node, first column, columns)] over the input nodes (get_input_dimension, network.py:1069-1086)."""
        out, col = [], 0
        for i in self.input_idxs:
            kind = self.edges[i].node_type
            d = int(self.attributes[i]["n_categories"]) if kind == categorical.NODE_TYPE else (
                int(self.attributes[i]["dimension"]) if kind == ef_state.NODE_TYPE else 1)
            out.append((i, col, d))
            col += d
        if n_columns is not None and n_columns != col:
            raise ValueError(f"input_data has {n_columns} column(s) but the input nodes take {col}")
        return out

    def _categorical_input_data(self, input_data, time_steps, observed):
        dev = self.device
        dt, cast = _torch_dtype(self.bits), _np_dtype(self.bits)
        x = np.asarray(input_data.cpu() if torch.is_tensor(input_data) else input_data, dtype=np.float64)
        if x.ndim == 1:
            x = x[:, None]
        T = x.shape[0]
        cols = self._categorical_columns(x.shape[1])
        masks = None
        if observed is not None:
            if isinstance(observed, (tuple, list)):
                per_node = [np.asarray(o) for o in observed]
            else:
                o = np.asarray(observed)
                per_node = [o] if o.ndim == 1 else [o[:, k] for k in range(o.shape[1])]
            masks = {i: per_node[j] for j, (i, _, _) in enumerate(cols)}
        implied, local = self._implied_network()
        # what the implied inputs are fed with: the raw columns, or t(x) for a dynamic ef-state node
        is_ef = lambda i: self.edges[i].node_type == ef_state.NODE_TYPE  # noqa: E731
        fed = [ef_state.sufficient_statistics(x[:, c0:c0 + d].astype(cast), d).astype(np.float64) if is_ef(i) else x[:, c0:c0 + d]
               for i, c0, d in cols]
        expanded = None if masks is None else np.stack([masks[i] for (i, _, _), f in zip(cols, fed) for _ in range(f.shape[1])], axis=1)
        implied.input_data(np.concatenate(fed, axis=1), time_steps=time_steps, observed=expanded, record_trajectories=True)
        back = {k: g for g, k in local.items()}
        nt = {back[k]: v for k, v in implied.node_trajectories.items() if k != -1}
        nt[-1] = implied.node_trajectories[-1]
        last = deepcopy(self.attributes)
        for k, g in back.items():
            last[g] = implied.last_attributes[k]
        last[-1] = implied.last_attributes[-1]
        for (c, c0, n), ss in zip(cols, fed):
            if is_ef(c):  # posterior_update_exponential_family_dynamic: per-step learning rate from the implied HGFs
                a_nodes = self.edges[c].value_parents
                b_nodes = [self.edges[a].value_parents[0] for a in a_nodes]
                planes = [np.stack([nt[j][key] for j in group]) for group, key in
                          ((a_nodes, "mean"), (a_nodes, "expected_mean"), (b_nodes, "mean"), (b_nodes, "expected_mean"))]
                imp = torch.as_tensor(np.stack(planes)).to(device=dev, dtype=dt)[:, :, :, None].contiguous()  # [4][K][T][1]
                obs = None if masks is None else torch.as_tensor((masks[c] != 0).astype(np.uint8)[None]).to(dev).contiguous()
                series = torch.as_tensor(np.ascontiguousarray(x[None, :, c0:c0 + n])).to(device=dev, dtype=dt)
                traj, fin, nus = ef_state.launch(bits=self.bits, device=dev, series=series, observed=obs, nus=None,
                                                 xis0=self.attributes[c]["xis"], dimension=n, input_group_size=0,
                                                 want_trajectory=True, implied=imp)
                seen = np.ones(T, dtype=np.int64) if masks is None else (masks[c] != 0).astype(np.int64)
                xc = x[:, c0:c0 + n].astype(cast)
                nt[c] = {"dimension": np.full(T, n), "nus": nus[:, 0].cpu().numpy(), "xis": traj[:, :, 0].t().cpu().numpy(),
                         "observed": seen, "mean": xc if n > 1 else xc[:, 0], "observation_ss": ss.astype(cast)}
                if T:
                    for key in ("nus", "xis", "observed", "mean", "observation_ss"):
                        last[c][key] = nt[c][key][-1]
                continue
            if self.edges[c].node_type != categorical.NODE_TYPE:
                continue
            binary = self.edges[c].value_parents
            em = np.stack([nt[b]["expected_mean"] for b in binary])  # [n][T]
            pm = np.stack([nt[self.edges[b].value_parents[0]]["mean"] for b in binary])
            out = categorical.update(bits=self.bits, device=dev,
                                     observations=torch.as_tensor(np.ascontiguousarray(x[None, :, c0:c0 + n])).to(device=dev, dtype=dt),
                                     expected_mean=torch.as_tensor(em).to(device=dev, dtype=dt)[:, :, None].contiguous(),
                                     parent_mean=torch.as_tensor(pm).to(device=dev, dtype=dt)[:, :, None].contiguous(),
                                     alpha0=self.attributes[c]["alpha"], input_group_size=0)
            seen = np.ones(T, dtype=np.int64) if masks is None else (masks[c] != 0).astype(np.int64)
            nt[c] = {"n_categories": np.full(T, n), "binary_idxs": np.broadcast_to(np.asarray(binary), (T, n)),
                     "alpha": out["alpha"][:, :, 0].t().cpu().numpy(), "mean": out["mean"][:, :, 0].t().cpu().numpy(),
                     "kl_divergence": out["kl_divergence"][:, 0].cpu().numpy(), "surprise": out["surprise"][:, 0].cpu().numpy(),
                     "observed": seen}
            if T:
                for key in ("alpha", "mean", "kl_divergence", "surprise", "observed"):
                    last[c][key] = nt[c][key][-1]
        self.node_trajectories = nt
        self.last_attributes = last
        self._last_inputs = (input_data, time_steps, observed)
        return self

    def _categorical_batch(self, input_data, parameters, time_steps, observed, n_instances, input_group_size, record, stream):
        """Runs ``_categorical_batch_impl`` in ``stream``'s order: the implied scan AND the per-node follow-up launches
        (which read the scan's trajectories) are enqueued on the same stream."""
        dev = self.device
        cur = torch.cuda.current_stream(dev)
        if stream is not None and stream != cur:
            stream.wait_stream(cur)
            with torch.cuda.device(dev), torch.cuda.stream(stream):
                return self._categorical_batch_impl(input_data, parameters, time_steps, observed, n_instances, input_group_size, record, stream)
        return self._categorical_batch_impl(input_data, parameters, time_steps, observed, n_instances, input_group_size, record, stream)

    def _categorical_batch_impl(self, input_data, parameters, time_steps, observed, n_instances, input_group_size, record, stream):
        """``batch`` with categorical / dynamic ef-state nodes: the implied network's batch (fed with the one-hot columns
        or with t(x)), then one launch per vector node that is recorded -- ``(node, "kl_divergence" | "surprise" |
        "alpha" | "mean")`` for a categorical node, ``(node, "xis" | "nus")`` for an ef-state node."""
        implied, local = self._implied_network()
        cats = self._vector_nodes()
        is_ef = lambda c: self.edges[c].node_type == ef_state.NODE_TYPE  # noqa: E731
        own = [tuple(k) for k in record if k[0] in cats]
        for k in own:
            allowed = ("xis", "nus") if is_ef(k[0]) else ("kl_divergence", "surprise", "alpha", "mean")
            if k[1] not in allowed:
                raise ValueError(f"node {k[0]} records one of {allowed}")
        params = {}
        for k, v in parameters.items():
            if isinstance(k, (tuple, list)) and isinstance(k[0], (int, np.integer)) and k[0] >= 0:
                if k[0] in cats:
                    raise ValueError("per-instance parameters live on the nodes implied under a categorical / ef-state node")
                params[(local[k[0]],) + tuple(k[1:])] = v
            else:
                params[k] = v
        rec = [(local[k[0]],) + tuple(k[1:]) for k in record if k[0] not in cats]
        needed = {}
        for c in sorted({k[0] for k in own}):
            under = self.edges[c].value_parents
            above = [self.edges[b].value_parents[0] for b in under]
            if is_ef(c):  # hgf_ef_call_t::implied: A.mean, A.expected_mean, B.mean, B.expected_mean
                needed[c] = tuple([(local[j], key) for j in group] for group, key in
                                  ((under, "mean"), (under, "expected_mean"), (above, "mean"), (above, "expected_mean")))
            else:
                needed[c] = ([(local[b], "expected_mean") for b in under], [(local[p], "mean") for p in above])
        extra = [k for groups in needed.values() for group in groups for k in group if k not in rec]
        dev, dt = self.device, _torch_dtype(self.bits)
        grouped = input_group_size > 0
        x = torch.as_tensor(np.asarray(input_data) if not torch.is_tensor(input_data) else input_data).to(device=dev, dtype=dt)
        if not grouped:
            x = (x[:, None] if x.ndim == 1 else x)[None]
        elif x.ndim == 2:
            x = x[:, :, None]
        cols = self._categorical_columns(x.shape[2])
        fed = []
        for i, c0, d in cols:  # what the implied inputs observe: the raw columns, or t(x) for an ef-state node
            part = x[:, :, c0:c0 + d]
            if is_ef(i):
                r_, q_ = np.tril_indices(d)
                part = torch.cat([part, part[:, :, r_.tolist()] * part[:, :, q_.tolist()]], dim=2)
            fed.append(part)
        widths = [f.shape[2] for f in fed]
        fed = torch.cat(fed, dim=2).contiguous()
        masks = None
        if observed is not None:  # one mask column per input node -> one per implied input
            o = torch.as_tensor(np.asarray(observed) if not torch.is_tensor(observed) else observed).to(dev)
            o = o.reshape(x.shape[0], x.shape[1], -1)
            if o.shape[2] != len(cols):
                raise ValueError(f"observed has {o.shape[2]} column(s) for {len(cols)} input node(s)")
            masks = {i: (o[:, :, j] != 0).to(torch.uint8).contiguous() for j, (i, _, _) in enumerate(cols)}
            observed = torch.cat([masks[i][:, :, None].expand(-1, -1, w) for (i, _, _), w in zip(cols, widths)], dim=2)
            observed = observed if grouped else observed[0]
        res = implied.batch(fed if grouped else fed[0], params, time_steps=time_steps, observed=observed, n_instances=n_instances,
                            input_group_size=input_group_size, record=rec + list(dict.fromkeys(extra)), stream=stream)
        traj = res.trajectories
        where = {i: (c0, d) for i, c0, d in cols}
        back = {k: g for g, k in local.items()}
        out = {(back[k[0]],) + tuple(k[1:]): v for k, v in traj.items() if k in rec}
        for c, groups in needed.items():
            c0, n = where[c]
            planes = [torch.stack([traj[k].t() for k in group]) for group in groups]  # each [n or K][T][N]
            if is_ef(c):
                xt, _, nus = ef_state.launch(bits=self.bits, device=dev, series=x[:, :, c0:c0 + n].contiguous(),
                                             observed=None if masks is None else masks[c], nus=None, xis0=self.attributes[c]["xis"],
                                             dimension=n, input_group_size=input_group_size,
                                             want_trajectory=(c, "xis") in own, implied=torch.stack(planes).contiguous())
                if (c, "xis") in own:
                    out[(c, "xis")] = xt.permute(2, 1, 0)  # [N][T][K] view of the kernel's [K][T][N]
                if (c, "nus") in own:
                    out[(c, "nus")] = nus.t()
                continue
            want_vec = any(k[0] == c and k[1] in ("alpha", "mean") for k in own)
            r = categorical.update(bits=self.bits, device=dev, observations=x[:, :, c0:c0 + n].contiguous(),
                                   expected_mean=planes[0].contiguous(), parent_mean=planes[1].contiguous(),
                                   alpha0=self.attributes[c]["alpha"], input_group_size=input_group_size, want_vectors=want_vec)
            for k in own:
                if k[0] == c:
                    out[k] = r[k[1]].t() if k[1] in ("kl_divergence", "surprise") else r[k[1]].permute(2, 1, 0)
        res.trajectories = out
        return res

    # ---------------------------------------------------- ef-state networks
    def _ef_columns(self, n_columns: int) -> list:
        """[(node, first column, dimension)]: the split of network.py:603-604 over the input nodes."""
        out, col = [], 0
        for i in self.input_idxs:
            d = int(self.attributes[i]["dimension"])
            out.append((i, col, d))
            col += d
        if n_columns != col:
            raise ValueError(f"input_data has {n_columns} column(s) but the input nodes take {col}")
        return out

    def _ef_input_data(self, input_data, time_steps, observed, record_trajectories):
        dev = self.device
        dt, cast = _torch_dtype(self.bits), _np_dtype(self.bits)
        x = np.asarray(input_data.cpu() if torch.is_tensor(input_data) else input_data, dtype=cast)
        if x.ndim == 1:
            x = x[:, None]
        T = x.shape[0]
        cols = self._ef_columns(x.shape[1])
        if observed is None:
            masks = [None] * len(cols)
        elif isinstance(observed, (tuple, list)):
            masks = [np.asarray(o) for o in observed]
        else:
            o = np.asarray(observed)
            masks = [o] if o.ndim == 1 else [o[:, k] for k in range(o.shape[1])]
        ts = np.ones(T) if time_steps is None else np.asarray(time_steps, dtype=np.float64)
        last, nt = deepcopy(self.attributes), {}
        for (i, c0, d), m in zip(cols, masks):
            a = self.attributes[i]
            series = torch.as_tensor(np.ascontiguousarray(x[None, :, c0:c0 + d])).to(dev)
            obs = None if m is None else torch.as_tensor((m != 0).astype(np.uint8)[None]).to(dev).contiguous()
            nus = torch.full((1,), float(a["nus"]), dtype=dt, device=dev)
            traj, fin = ef_state.launch(bits=self.bits, device=dev, series=series, observed=obs, nus=nus, xis0=a["xis"],
                                        dimension=d, input_group_size=0, want_trajectory=record_trajectories)
            seen = np.ones(T, dtype=np.int64) if m is None else (m != 0).astype(np.int64)
            ss = ef_state.sufficient_statistics(x[:, c0:c0 + d], d)
            last[i]["xis"] = fin[:, 0].cpu().numpy()
            if T:
                last[i]["mean"] = x[-1, c0:c0 + d] if d > 1 else x[-1, c0]
                last[i]["observation_ss"] = ss[-1]
                last[i]["observed"] = seen[-1]
            if record_trajectories:
                nt[i] = {"dimension": np.full(T, d), "nus": np.full(T, a["nus"], dtype=cast), "xis": traj[:, :, 0].t().cpu().numpy(),
                         "observed": seen, "mean": x[:, c0:c0 + d] if d > 1 else x[:, c0], "observation_ss": ss}
        last[-1]["time_step"] = float(ts[-1]) if T else 0.0
        self.last_attributes = last
        if record_trajectories:
            nt[-1] = {"time_step": ts.astype(cast),
                      "precision_clipping_value": np.full(T, self.attributes[-1]["precision_clipping_value"], dtype=cast)}
            self.node_trajectories = nt
        else:
            self.node_trajectories = None
        self._last_inputs = (input_data, time_steps, observed)
        return self

    def _ef_batch(self, input_data, parameters, observed, n_instances, input_group_size, record, last_attributes):
        """``batch`` for an ef-state network: instances differ in ``(node, "nus")`` and, with groups, in their series."""
        dev = self.device
        dt = _torch_dtype(self.bits)
        grouped = input_group_size > 0
        x = torch.as_tensor(np.asarray(input_data) if not torch.is_tensor(input_data) else input_data).to(device=dev, dtype=dt)
        if not grouped:
            x = x[:, None] if x.ndim == 1 else x
            x = x[None]
        elif x.ndim == 2:
            x = x[:, :, None]
        cols = self._ef_columns(x.shape[2])
        for k in parameters:
            if not (isinstance(k, (tuple, list)) and len(k) == 2 and k[1] == "nus" and self.edges[k[0]].node_type == ef_state.NODE_TYPE):
                raise ValueError(f"unknown parameter {k!r}: an ef-state network takes (node, 'nus')")
        if n_instances is None:
            sizes = {int(v.numel()) if torch.is_tensor(v) else int(np.size(v)) for v in parameters.values()} - {1}
            if len(sizes) > 1:
                raise ValueError(f"per-instance parameters disagree on the batch size: {sorted(sizes)}")
            n_instances = sizes.pop() if sizes else (x.shape[0] * input_group_size if grouped else 1)
        if grouped and x.shape[0] * input_group_size < n_instances:
            raise ValueError("not enough input series for n_instances / input_group_size")
        obs = None
        if observed is not None:
            obs = torch.as_tensor(np.asarray(observed) if not torch.is_tensor(observed) else observed).to(dev)
            obs = obs.reshape(x.shape[0], x.shape[1], -1)
        wanted = {tuple(k) for k in record}
        if any(k[1] != "xis" for k in wanted):
            raise ValueError("an ef-state network records (node, 'xis')")
        res = BatchResult()
        res.n_instances = n_instances
        res.trajectories, lasts = {}, {}
        for j, (i, c0, d) in enumerate(cols):
            v = parameters.get((i, "nus"), self.attributes[i]["nus"])
            nus = torch.as_tensor(v).to(device=dev, dtype=dt).reshape(-1)
            nus = nus.expand(n_instances).contiguous() if nus.numel() == 1 else nus.contiguous()
            m = None if obs is None else (obs[:, :, min(j, obs.shape[2] - 1)] != 0).to(torch.uint8).contiguous()
            traj, fin = ef_state.launch(bits=self.bits, device=dev, series=x[:, :, c0:c0 + d].contiguous(), observed=m, nus=nus,
                                        xis0=self.attributes[i]["xis"], dimension=d, input_group_size=input_group_size,
                                        want_trajectory=(i, "xis") in wanted)
            if traj is not None:
                res.trajectories[(i, "xis")] = traj.permute(2, 1, 0)  # [N][T][K] view of the kernel's [K][T][N]
            lasts[(i, "xis")] = fin.t()
        res.last_attributes = lasts if last_attributes else None
        return res

    def _unflatten(self, vec: np.ndarray, time_step: float) -> dict:
        out = deepcopy(self.attributes)
        cast = _np_dtype(self.bits)
        for i in range(self.n_nodes):
            for k, name in enumerate(_lib.FIELDS):
                val = cast(vec[i * _lib.N_FIELDS + k])
                if name.startswith("temp."):
                    if name[5:] in out[i].get("temp", {}):
                        out[i]["temp"][name[5:]] = val
                elif name in out[i]:
                    out[i][name] = val
        out[-1]["time_step"] = time_step
        return out

    # --------------------------------------------------------------- surprise
    def sample(self, time_steps, n_predictions: int = 1, rng_key=0) -> dict:
        """Sample ``n_predictions`` rollouts forward in generative mode (pyhgf/utils/sample.py:18-116): starting from
        ``last_attributes`` (or the initial attributes), at every step each input node observes a draw from its own
        prediction and the beliefs are updated on it.  One thread per rollout; the random stream is an in-kernel
        Philox4x32-10 keyed by ``(rng_key, rollout, step)`` -- not JAX's threefry, so draws differ from the
        reference's while their distribution is the same.  Returns ``{node: {field: tensor[n_predictions, T]}}``."""
        if self.scan_fn is None:
            self.create_belief_propagation_fn()
        dev = self.device
        ts = np.atleast_1d(np.asarray(time_steps, dtype=np.float64))
        T, n_in = int(ts.size), len(self.input_idxs)
        plan = self._get_plan(n_in)
        start = self.last_attributes if getattr(self, "last_attributes", None) is not None else self.attributes
        record = list(range(self.n_nodes * _lib.N_FIELDS))
        dummy = torch.zeros((1, T, n_in), device=dev, dtype=_torch_dtype(self.bits))  # ignored by the kernel
        res, plan = self._launch(values=dummy, observed=None, time_steps=ts, input_group_size=0, n_instances=int(n_predictions),
                                 params={}, response_fn=0, response_inputs=None, response_group_size=0, response_parameters=None,
                                 record_slots=record, want_last=False, want_sum=False, want_pointwise=False, want_grad=False,
                                 group=None, n_groups=0, attributes=start, generative_seed=int(rng_key))
        out = {}
        for i in range(self.n_nodes):
            a, d, temp = self.attributes[i], {}, {}
            for k, name in enumerate(_lib.FIELDS):
                col = res._traj[i * _lib.N_FIELDS + k].t()
                if name.startswith("temp."):
                    if name[5:] in a.get("temp", {}):
                        temp[name[5:]] = col
                elif name in a:
                    d[name] = col
            d["temp"] = temp
            out[i] = d
        out[-1] = {"time_step": torch.as_tensor(ts, device=dev).expand(int(n_predictions), T)}
        return out

    def to_pandas(self):
        """Export the trajectories and the per-node surprise as a pandas data frame (pyhgf/utils/to_pandas.py:14-136,
        state nodes of type 1, 2 and 6): ``time_steps``, ``time``, every ``x_{i}_{mean|precision|observed...}`` column,
        ``x_{i}_surprise`` (``binary_surprise`` / ``gaussian_surprise`` of node i's ``mean`` under its own prediction,
        pyhgf/math.py:209-297, evaluated in the network's dtype) and ``total_surprise``.  Host-side post-processing
        of ``node_trajectories``, as in the reference."""
        import pandas as pd

        if self.node_trajectories is None:
            raise ValueError("to_pandas needs recorded trajectories: call input_data(..., record_trajectories=True) first")
        nt, ft = self.node_trajectories, _np_dtype(self.bits)
        ts = np.asarray(nt[-1]["time_step"])
        cols = {"time_steps": ts, "time": np.cumsum(ts)}
        for types in ((1, 2), (6,)):
            for i, e in enumerate(self.edges):
                if e.node_type in types:
                    for var, v in nt[i].items():
                        if var != "temp" and v is not None and (("mean" in var) or ("precision" in var) or ("observed" in var)):
                            cols[f"x_{i}_{var}"] = np.asarray(v)
        with np.errstate(divide="ignore", invalid="ignore"):
            for i, e in enumerate(self.edges):
                if e.node_type == 1:  # math.py:270-297
                    x, m = np.asarray(nt[i]["mean"], dtype=ft), np.asarray(nt[i]["expected_mean"], dtype=ft)
                    m = np.clip(m, ft(1e-6), ft(1.0) - ft(1e-6))
                    cols[f"x_{i}_surprise"] = np.where(x != 0, -np.log(m), -np.log(ft(1.0) - m)).astype(ft)
            for i, e in enumerate(self.edges):
                if e.node_type in (2, 6):  # math.py:209-238
                    x, m, p = (np.asarray(nt[i][k], dtype=ft) for k in ("mean", "expected_mean", "expected_precision"))
                    cols[f"x_{i}_surprise"] = (ft(0.5) * (np.log(ft(2.0) * ft(np.pi)) - np.log(p) + p * np.square(x - m))).astype(ft)
        df = pd.DataFrame(cols)
        df["total_surprise"] = df.iloc[:, df.columns.str.contains("_surprise")].sum(axis=1, min_count=1)
        return df

    def surprise(self, response_function: Callable, response_function_inputs=(), response_function_parameters=None):
        """Per-trial surprise under a response function (network.py:986-1022)."""
        return response_function(hgf=self, response_function_inputs=response_function_inputs,
                                 response_function_parameters=response_function_parameters)

    def _fused_response(self, name: str, response_function_inputs=None, response_function_parameters=None) -> np.ndarray:
        """Per-trial surprise of a built-in response function for the series last given to ``input_data``:
        the filter is re-run with the response fused into the time loop (no trajectory is materialised)."""
        if getattr(self, "_last_inputs", None) is None:
            raise RuntimeError("call input_data() before surprise()")
        input_data, time_steps, observed = self._last_inputs
        dev = self.device
        values, obs = self._format_series(input_data, observed, dev, grouped=False)
        ri = None
        if response_function_inputs is not None and np.size(response_function_inputs) > 0:
            ri = np.asarray(response_function_inputs).reshape(1, -1)
        res, _ = self._launch(values=values, observed=obs, time_steps=time_steps, input_group_size=0, n_instances=1,
                              params={}, response_fn=_lib.RESPONSES[name], response_inputs=ri, response_group_size=0,
                              response_parameters=response_function_parameters, record_slots=[], want_last=False,
                              want_sum=False, want_pointwise=True, want_grad=False, group=None, n_groups=0)
        return res.pointwise[0].cpu().numpy()

    # ------------------------------------------------------------------ batch
    def _split_over_components(self, input_data, parameters, response_function, response_function_inputs,
                               response_function_parameters, time_steps, observed, n_instances, input_group_size,
                               response_group_size, record, last_attributes, pointwise, gradient, n_groups, out_scale, stream,
                               fused_reduce):
        """A network made of several disconnected sub-networks IS several independent networks: run one launch per
        connected component, concurrently on separate streams.  Per instance that is more threads with fewer nodes
        each -- a shorter serial chain per time step and half the registers for BASELINE config 5's 16-node network
        (two 8-node components).  Only the plain evaluations are split (no gradient, final attributes or device
        reduction, no response function -- those read across components); everything else, and
        HGF_SPLIT_COMPONENTS=0, takes the single-launch path.  Returns None when the call is not split."""
        if os.environ.get("HGF_SPLIT_COMPONENTS") == "0" or gradient or last_attributes or n_groups or fused_reduce is not None:
            return None
        comps = self.connected_components()
        if len(comps) < 2:
            return None
        if response_function is not None or pointwise or not record:
            return None  # response functions read across components (input nodes, node 1): single launch
        grouped = input_group_size > 0
        x = torch.as_tensor(np.asarray(input_data) if not torch.is_tensor(input_data) else input_data)
        want_dim = 3 if grouped else 2
        if x.ndim == want_dim - 1:
            x = x.unsqueeze(-1)
        inputs = list(self.input_idxs)
        if x.ndim != want_dim or x.shape[-1] != len(inputs):
            return None
        obs = None
        if observed is not None:
            if isinstance(observed, (tuple, list)) and not grouped:
                observed = np.stack([np.asarray(o) for o in observed], axis=1)
            obs = torch.as_tensor(np.asarray(observed) if not torch.is_tensor(observed) else observed)
            if obs.ndim == want_dim - 1:
                obs = obs.unsqueeze(-1)
            if obs.shape != x.shape:
                return None
        if n_instances is None:
            sizes = {int(np.size(v)) if not torch.is_tensor(v) else int(v.numel()) for v in parameters.values()} - {1}
            if len(sizes) > 1:
                raise ValueError(f"per-instance parameters disagree on the batch size: {sorted(sizes)}")
            n_instances = sizes.pop() if sizes else (x.shape[0] * input_group_size if grouped else 1)
        dev = self.device
        fused = self._fused_isomorphic_components(comps, x, obs, parameters, time_steps, n_instances, input_group_size, record, stream)
        if fused is not None:
            return fused
        cur = torch.cuda.current_stream(dev) if stream is None else stream
        parts, streams, merged_traj = [], [], {}
        for nodes in comps:
            sub, local = self._component_network(nodes)
            cols = [c for c, g in enumerate(inputs) if g in local]
            rec_k = [(local[int(n)], f) for n, f in record if int(n) in local]
            if not rec_k:
                continue  # nothing is asked of this component
            p_k = {}
            for key, val in parameters.items():
                if isinstance(key, str) or key[0] == "response_parameter":
                    continue  # no response function on this path
                if key[0] == "precision_clipping_value" or key[0] == -1:
                    p_k[key] = val
                elif int(key[0]) in local:
                    p_k[(local[int(key[0])],) + tuple(key[1:])] = val
            s_k = torch.cuda.Stream(device=dev)
            s_k.wait_stream(cur)
            with torch.cuda.stream(s_k):
                r = sub.batch(x[..., cols], p_k, time_steps=time_steps, observed=None if obs is None else obs[..., cols],
                              n_instances=n_instances, input_group_size=input_group_size, record=rec_k, stream=s_k)
            if getattr(r, "_traj", None) is not None:
                r._traj.record_stream(cur)  # allocated on the side stream, read on the caller's
            back = {v: g for g, v in local.items()}
            for (n_loc, f), t in r.trajectories.items():
                merged_traj[(back[n_loc], f)] = t
            parts.append(r)
            streams.append(s_k)
        if not parts:
            return None
        for s_k in streams:
            cur.wait_stream(s_k)
        out = ComponentsResult(parts, streams, dev)
        out.trajectories = {tuple(key): merged_traj[(int(key[0]), key[1])] for key in record}
        return out

    def _fused_isomorphic_components(self, comps, x, obs, parameters, time_steps, n_instances, input_group_size, record, stream):
        """WIDE networks made of ISOMORPHIC components -- the reference's "wide (x10)" benchmark network, ten 3-level binary
        HGFs side by side (docs/paper.ipynb cell 46, docs/benchmark_results.csv:4), or BASELINE config 5's two 8-node halves
        -- run as ONE launch of the component's kernel over C x N threads: the components of an instance are C lanes'
        worth of work instead of one thread's, thread (c, n) carrying component c of instance n (component-major, so that
        a warp reads ONE component's series: broadcast loads).  Returns None unless every component has the same
        edges, attributes, recorded leaves and per-instance parameter keys (then the per-component launches below run)."""
        if os.environ.get("HGF_FUSE_COMPONENTS") == "0":
            return None
        inputs = list(self.input_idxs)
        subs = [self._component_network(nodes) for nodes in comps]
        sub0, local0 = subs[0]
        sig = lambda net: (net.edges, [sorted((k, repr(v)) for k, v in net.attributes[i].items()) for i in range(net.n_nodes)], net._input_idxs)  # noqa: E731
        if any(sig(sub) != sig(sub0) for sub, _ in subs[1:]):
            return None
        per_comp_rec, per_comp_par = [], []
        for sub, local in subs:
            per_comp_rec.append([(local[int(n)], f) for n, f in record if int(n) in local])
            pk = {}
            for key, val in parameters.items():
                if isinstance(key, str) or key[0] in ("response_parameter", "precision_clipping_value") or key[0] == -1:
                    return None  # shared / response parameters: the per-component path handles them
                if int(key[0]) in local:
                    pk[(local[int(key[0])],) + tuple(key[1:])] = val
            per_comp_par.append(pk)
        if any(r != per_comp_rec[0] for r in per_comp_rec[1:]) or any(set(q) != set(per_comp_par[0]) for q in per_comp_par[1:]):
            return None
        if not per_comp_rec[0]:
            return None
        C, N = len(comps), int(n_instances)
        grouped = input_group_size > 0
        if grouped and x.shape[0] * input_group_size != N:
            return None
        dev, dt = self.device, _torch_dtype(self.bits)
        cols = [[c for c, g in enumerate(inputs) if g in local] for _, local in subs]
        x = x.to(device=dev, dtype=dt)
        series = torch.cat([x[..., c] if grouped else x[..., c][None] for c in cols], dim=0).contiguous()  # [C or C*G, T, n_fed]
        mask = None
        if obs is not None:
            obs = obs.to(dev)
            mask = torch.cat([obs[..., c] if grouped else obs[..., c][None] for c in cols], dim=0).contiguous()
        params = {}
        for key in per_comp_par[0]:
            parts = []
            for pk in per_comp_par:
                v = torch.as_tensor(np.asarray(pk[key]) if not torch.is_tensor(pk[key]) else pk[key]).to(device=dev, dtype=dt).reshape(-1)
                parts.append(v.expand(N) if v.numel() == 1 else v)
            params[key] = torch.cat(parts)
        r = sub0.batch(series, params, time_steps=time_steps, observed=mask, n_instances=C * N, input_group_size=input_group_size if grouped else N,
                       record=per_comp_rec[0], stream=stream)
        traj = {}
        for c, (_, local) in enumerate(subs):
            back = {v: g for g, v in local.items()}
            for (n_loc, f), t in r.trajectories.items():
                traj[(back[n_loc], f)] = t[c * N:(c + 1) * N]
        r.trajectories = {tuple(key): traj[(int(key[0]), key[1])] for key in record}
        r.kernel_name = f"components({C}) in one launch: {r.kernel_name}"
        return r

    def batch(self, input_data, parameters: Optional[dict] = None, *, response_function=None,
              response_function_inputs=None, response_function_parameters=None, time_steps=None, observed=None,
              n_instances: Optional[int] = None, input_group_size: int = 0, response_group_size: int = 0,
              record: Sequence[tuple] = (), last_attributes: bool = False, pointwise: bool = False,
              gradient: bool = False, group=None, n_groups: int = 0, out_scale: float = 1.0, stream=None,
              fused_reduce=None) -> BatchResult:
        """Run N independent instances of this network in one launch.

        What the reference writes as ``jax.vmap(lambda p: Network()...input_data(u).surprise(f).sum())(params)``
        (and ``jax.grad`` of it) -- notebooks 3-Multilevel_HGF / 4-Parameter_recovery.

        parameters : ``{key: array[N]}`` per-instance overrides of the attribute vector; a key is
            ``(node, "tonic_volatility")``, ``(node, "volatility_coupling_children", k)`` or
            ``("response_parameter", j)``.
        input_data : one series ``(T[, n_inputs])`` shared by all instances, or with
            ``input_group_size = g`` a stack ``(G, T[, n_inputs])`` where instance n reads series ``n // g``.
        response_function : one of ``pyhgf_b200.response``'s functions (or its name): fused into the time loop.
        record : ``[(node, attribute), ...]`` leaves to record at every step -> ``result.trajectories``.
        gradient : also return d(surprise)/d(parameters) (``hgf_logp_grad``).  ``True`` lets the library pick between
            forward-mode tangents in the filter's registers and the checkpointed reverse sweep (taken from 5 filter parameters
            on, where the topology has a generated one); ``"forward"`` / ``"reverse"`` force one.
        group / n_groups : sum ``out_scale * surprise`` of the instances of each group (e.g. chain) on the device.
        fused_reduce : a ``pyhgf_b200.distributed.FusedReduce``: the group sums are all-reduced over the ranks
            by the kernel itself (NVLink peer memory); ``result.reduced`` then holds the sums over all ranks.
        """
        parameters = dict(parameters or {})
        if self.scan_fn is None:
            self.create_belief_propagation_fn()
        if self._ef_nodes():
            if response_function is not None or gradient or pointwise or fused_reduce is not None or time_steps is not None:
                raise NotImplementedError("an ef-state network has no response function, gradient or time steps on this path")
            return self._ef_batch(input_data, parameters, observed, n_instances, input_group_size, record, last_attributes)
        if self._vector_nodes():
            if response_function is not None or gradient or pointwise or fused_reduce is not None or last_attributes:
                raise NotImplementedError("a network with categorical nodes batches trajectories (record=[...]) on this path")
            return self._categorical_batch(input_data, parameters, time_steps, observed, n_instances, input_group_size, record, stream)
        dev = self.device
        split = self._split_over_components(input_data, parameters, response_function, response_function_inputs,
                                            response_function_parameters, time_steps, observed, n_instances, input_group_size,
                                            response_group_size, record, last_attributes, pointwise, gradient, n_groups, out_scale,
                                            stream, fused_reduce)
        if split is not None:
            return split
        grouped = input_group_size > 0
        values, obs = self._format_series(input_data, observed, dev, grouped=grouped)
        if n_instances is None:
            sizes = {int(np.size(v)) if not torch.is_tensor(v) else int(v.numel()) for v in parameters.values()} - {1}
            if len(sizes) > 1:
                raise ValueError(f"per-instance parameters disagree on the batch size: {sorted(sizes)}")
            n_instances = sizes.pop() if sizes else (values.shape[0] * input_group_size if grouped else 1)
        if grouped and values.shape[0] * input_group_size < n_instances:
            raise ValueError("not enough input series for n_instances / input_group_size")
        name = response_function if (response_function is None or isinstance(response_function, str)) else getattr(
            response_function, "_hgf_response", None)
        if response_function is not None and name not in _lib.RESPONSES:
            raise NotImplementedError(
                "only the built-in response functions of pyhgf_b200.response are fused into the kernel; for a custom "
                "callable, record the trajectories it needs (record=[...]) and apply it to result.trajectories"
            )
        plan = self._get_plan(values.shape[2])
        if plan.kernel_name == "interpreter" and os.environ.get("HGF_AUTO_SPECIALISE") == "1" and values.shape[2] == len(self.input_idxs):
            # opt-in: build register-resident kernels for this topology on first use (pyhgf_b200/specialise.py)
            n_filter = sum(1 for k in parameters if not (isinstance(k, (tuple, list)) and k[0] == "response_parameter") and k != "response_parameter")
            self.specialise(gradient_parameters=n_filter if gradient else 0, response_parameter=gradient and n_filter < len(parameters),
                            float64=self.bits == 64)
            plan = self._get_plan(values.shape[2])
        rec_slots = [self._resolve_slot(plan, k) for k in record]
        ri = response_function_inputs
        if ri is not None and np.size(ri) == 0:
            ri = None
        if ri is not None:
            ri = torch.as_tensor(np.asarray(ri) if not torch.is_tensor(ri) else ri)
            if ri.ndim == 1:
                ri = ri[None]
        res, plan = self._launch(values=values, observed=obs, time_steps=time_steps, input_group_size=input_group_size,
                                 n_instances=n_instances, params=parameters, response_fn=_lib.RESPONSES[name],
                                 response_inputs=ri, response_group_size=response_group_size,
                                 response_parameters=response_function_parameters, record_slots=rec_slots,
                                 want_last=last_attributes, want_sum=name is not None, want_pointwise=pointwise and name is not None,
                                 want_grad=gradient, group=group, n_groups=n_groups, out_scale=out_scale, stream=stream,
                                 fused_reduce=fused_reduce)
        if rec_slots:
            res.trajectories = {tuple(k): res._traj[i].t() for i, k in enumerate(record)}
        if last_attributes:
            res.last_attributes = _LastAttributes(res._last, plan, self)
        return res


class _LastAttributes:
    """``[(node, key)] -> tensor[N]`` view of the final attribute vectors."""

    def __init__(self, flat, plan, net):
        self.flat, self._plan, self._net = flat, plan, net

    def __getitem__(self, key):
        node, name = key
        f = _lib.FIELD_INDEX[name] if name in _lib.FIELD_INDEX else _lib.FIELD_INDEX["temp." + name]
        return self.flat[node * _lib.N_FIELDS + f]
