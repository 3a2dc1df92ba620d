"""Host-side topology compiler.

Turns a network's edges into (1) the static update schedule the reference
derives in ``pyhgf/utils/get_update_sequence.py:51-351`` and (2) the flat
``hgf_topology_t`` the CUDA library lowers into a kernel program
(``hgf_plan_create``, include/pyhgf_cuda.h).  Runs once per network, on the host.
"""

from __future__ import annotations

import ctypes as C
from typing import NamedTuple, Optional, Sequence

import numpy as np

from . import _lib
from .updates import UPDATE_FUNCTIONS as U

BINARY, CONTINUOUS, VOLATILE = 1, 2, 6
SUPPORTED_NODE_TYPES = (BINARY, CONTINUOUS, VOLATILE)


class AdjacencyLists(NamedTuple):
    """Edges of one node (pyhgf/typing/typing.py:12-52).

    node_type: 1 binary state, 2 continuous state, 6 volatile state (0, 3, 4, 5
    are node families this backend does not run).
    """

    node_type: int
    value_parents: Optional[tuple]
    volatility_parents: Optional[tuple]
    value_children: Optional[tuple]
    volatility_children: Optional[tuple]
    coupling_fn: tuple


class UpdateSequence(NamedTuple):
    """pyhgf/typing/typing.py:65-86 (hot-path members)."""

    prediction_steps: tuple
    update_steps: tuple


def _members(*groups):
    return [j for g in groups if g is not None for j in g]


def get_update_sequence(network, volatility_updates: str, mean_field_updates: bool = False) -> UpdateSequence:
    """Static schedule of one time slice.

    Predictions run parents-before-children; prediction errors and posterior
    updates interleave children-before-parents.  The variant chosen per node
    follows pyhgf/utils/get_update_sequence.py:176-222: a continuous node with
    volatility children takes the ``volatility_updates`` scheme, every other
    continuous node the (relaxed or mean-field) standard update; a volatile
    node's prediction-error step also runs its volatility-level posterior.
    """
    if volatility_updates not in _lib.VOLATILITY_UPDATES:
        raise ValueError("Invalid update type.")
    edges = network.edges
    n = len(edges)
    for i, e in enumerate(edges):
        if e.node_type not in SUPPORTED_NODE_TYPES:
            raise NotImplementedError(
                f"node {i} has node_type {e.node_type}: the CUDA backend runs binary-, continuous- and volatile-state nodes"
            )
    parents = [_members(e.value_parents, e.volatility_parents) for e in edges]
    children = [_members(e.value_children, e.volatility_children) for e in edges]

    # ---- predictions --------------------------------------------------------
    predicted = [False] * n
    prediction_steps = []
    pending = list(range(n))
    while pending:
        progressed = False
        for idx in list(pending):
            if any(not predicted[p] for p in parents[idx]):
                continue
            t = edges[idx].node_type
            if t == BINARY:
                fn = U["binary_state_node_prediction"]
            elif t == CONTINUOUS:
                fn = U["continuous_node_prediction_mean_field" if mean_field_updates else "continuous_node_prediction"]
            else:
                fn = U["volatile_node_prediction_mean_field" if mean_field_updates else "volatile_node_prediction"]
            prediction_steps.append((idx, fn))
            predicted[idx] = True
            pending.remove(idx)
            progressed = True
        if pending and not progressed:
            raise Warning("The structure of the network cannot be updated consistently.")

    # ---- prediction errors / posterior updates --------------------------------
    def posterior_fn(idx):
        if edges[idx].node_type == VOLATILE:
            return U["volatile_node_posterior_update_mean_field" if mean_field_updates else "volatile_node_posterior_update"]
        standard = U["continuous_node_posterior_update_mean_field" if mean_field_updates else "continuous_node_posterior_update"]
        if edges[idx].volatility_children is None or volatility_updates == "standard":
            return standard
        if volatility_updates == "unbounded":
            return U["continuous_node_posterior_update_unbounded"]
        return U["continuous_node_posterior_update_ehgf_mean_field" if mean_field_updates else "continuous_node_posterior_update_ehgf"]

    pe_fn = {BINARY: U["binary_state_node_prediction_error"], CONTINUOUS: U["continuous_node_prediction_error"],
             VOLATILE: U["volatile_node_prediction_error"]}

    needs_posterior = [i for i in range(n) if children[i]]
    needs_pe = list(range(n))
    update_steps = []
    while needs_pe or needs_posterior:
        progressed = False
        for idx in list(needs_posterior):  # a parent updates once all its children have sent their errors
            if all(c not in needs_pe for c in children[idx]):
                update_steps.append((idx, posterior_fn(idx)))
                needs_posterior.remove(idx)
                progressed = True
        for idx in list(needs_pe):
            if not parents[idx] and edges[idx].node_type != VOLATILE:
                needs_pe.remove(idx)  # roots have nobody to send an error to
            elif idx not in needs_posterior:  # leaves, or nodes whose own posterior is done
                update_steps.append((idx, pe_fn[edges[idx].node_type]))
                needs_pe.remove(idx)
                progressed = True
        if (needs_pe or needs_posterior) and not progressed:
            raise Warning("The structure of the network cannot be updated consistently.")
    return UpdateSequence(tuple(prediction_steps), tuple(update_steps))


def get_input_idxs(edges: Sequence[AdjacencyLists]) -> tuple:
    """Nodes without children receive observations (pyhgf/utils/get_input_idxs.py:6-22)."""
    return tuple(i for i, e in enumerate(edges) if e.value_children is None and e.volatility_children is None)


def resolve_coupling_fn(fn) -> int:
    """Map a coupling function (None, name, or a callable known by name) to the library's enum."""
    if fn is None:
        return 0
    name = fn if isinstance(fn, str) else getattr(fn, "__name__", None)
    if name == "logistic":
        name = "sigmoid"
    if name in _lib.COUPLING_FNS:
        return _lib.COUPLING_FNS[name]
    raise NotImplementedError(
        f"coupling function {fn!r} cannot cross the C ABI: use None or one of "
        f"{sorted(k for k in _lib.COUPLING_FNS if k)} (by name, or a callable with that __name__)"
    )


class Plan:
    """A lowered network: owns the library's ``hgf_plan_t`` and knows the attribute-vector layout."""

    def __init__(self, edges, update_sequence: UpdateSequence, input_idxs, n_fed, volatility_updates, mean_field_updates,
                 max_posterior_precision, branches=None):
        L = _lib.lib()
        n = len(edges)
        self.n_nodes = n
        self.edges = edges
        self.input_idxs = tuple(int(i) for i in input_idxs)
        self.n_fed = int(n_fed)

        def csr(name):
            ptr, idx = [0], []
            for e in edges:
                idx.extend(getattr(e, name) or ())
                ptr.append(len(idx))
            return np.asarray(ptr, np.int32), np.asarray(idx or [0], np.int32)

        arrays = {}
        arrays["node_type"] = np.asarray([e.node_type for e in edges], np.int32)
        for name in ("value_parents", "volatility_parents", "value_children", "volatility_children"):
            arrays[name + "_ptr"], arrays[name + "_idx"] = csr(name)
        fns = []
        for e in edges:
            for k in range(len(e.value_children or ())):
                fns.append(resolve_coupling_fn(e.coupling_fn[k] if k < len(e.coupling_fn) else None))
        arrays["value_children_fn"] = np.asarray(fns or [0], np.int32)
        pred, upd = update_sequence
        self.n_branches = 0
        if branches is not None and len(branches) > 1:
            # input_custom_sequence (network.py:653-767): the branches' steps back to back, CSR pointers per branch
            self.n_branches = len(branches)
            pred = tuple(s for b in branches for s in b[0])
            upd = tuple(s for b in branches for s in b[1])
            arrays["branch_prediction_ptr"] = np.cumsum([0] + [len(b[0]) for b in branches]).astype(np.int32)
            arrays["branch_update_ptr"] = np.cumsum([0] + [len(b[1]) for b in branches]).astype(np.int32)
        arrays["prediction_node"] = np.asarray([s[0] for s in pred] or [0], np.int32)
        arrays["prediction_op"] = np.asarray([s[1].opcode for s in pred] or [0], np.int32)
        arrays["update_node"] = np.asarray([s[0] for s in upd] or [0], np.int32)
        arrays["update_op"] = np.asarray([s[1].opcode for s in upd] or [0], np.int32)
        arrays["input_idxs"] = np.asarray(self.input_idxs or (0,), np.int32)

        topo = _lib.Topology()
        topo.n_nodes = n
        for key, arr in arrays.items():
            setattr(topo, key, arr.ctypes.data_as(C.POINTER(C.c_int32)))
        topo.n_branches = self.n_branches
        topo.n_prediction_steps = len(pred)
        topo.n_update_steps = len(upd)
        topo.n_inputs = len(self.input_idxs)
        topo.n_fed = self.n_fed
        topo.volatility_updates = _lib.VOLATILITY_UPDATES[volatility_updates]
        topo.mean_field_updates = int(bool(mean_field_updates))
        topo.max_posterior_precision = float(max_posterior_precision)
        handle = C.c_void_p()
        _lib.check(L.hgf_plan_create(C.byref(topo), C.byref(handle)))
        self._L = L
        self.handle = handle
        self.n_slots = int(L.hgf_plan_n_slots(handle))
        self.slot_clipping = int(L.hgf_plan_slot_clipping(handle))
        self.slot_response = int(L.hgf_plan_slot_response(handle))
        self.kernel_name = L.hgf_plan_kernel_name(handle).decode()

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                self._L.hgf_plan_destroy(self.handle)
                self.handle = None
        except Exception:
            pass

    _COUPLING_KINDS = {"value_coupling_parents": 0, "volatility_coupling_parents": 1, "value_coupling_children": 2,
                       "volatility_coupling_children": 3}

    def slot(self, node: int, key: str, position: int = 0) -> int:
        """Index, in an instance's attribute vector, of ``attributes[node][key]`` (``[position]`` for couplings).

        ``node=-1`` addresses the globals: ``precision_clipping_value`` and ``response_parameter`` (position j).
        """
        if node == -1:
            if key == "precision_clipping_value":
                return self.slot_clipping
            if key == "response_parameter":
                if not 0 <= position < _lib.N_RESPONSE_PARAMETERS:
                    raise KeyError(f"response parameter {position}")
                return self.slot_response + position
            raise KeyError(key)
        if not 0 <= node < self.n_nodes:
            raise KeyError(f"node {node}")
        if key in self._COUPLING_KINDS:
            s = int(self._L.hgf_plan_slot_coupling(self.handle, self._COUPLING_KINDS[key], node, position))
            if s < 0:
                raise KeyError(f"node {node} has no {key}[{position}]")
            return s
        if key.startswith("temp.") or key in _lib.FIELD_INDEX:
            return node * _lib.N_FIELDS + _lib.FIELD_INDEX[key]
        if "temp." + key in _lib.FIELD_INDEX:
            return node * _lib.N_FIELDS + _lib.FIELD_INDEX["temp." + key]
        raise KeyError(key)
