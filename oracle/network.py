"""TEST INFRASTRUCTURE ONLY -- Python front-end of the CPU oracle.

Restates the *host-side* part of the reference path so that golden tests read
like the reference's own tests:

* ``Network`` / ``add_nodes`` / ``input_data`` / ``surprise`` -- the model API of
  ``pyhgf/model/network.py:85-155, 528-628, 778-945, 986-1022``;
* node defaults -- ``pyhgf/model/add_nodes.py:65-85, 159-184, 330-340``;
* edge bookkeeping -- ``pyhgf/utils/add_edges.py:10-176``;
* the static schedule -- ``pyhgf/utils/get_update_sequence.py:51-351``.

The arithmetic itself lives in ``hgf_oracle.cpp`` (``liboracle.so``), driven here
through ctypes.  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
CPU-baseline legs may import this module; the product never does.
"""

from __future__ import annotations

import ctypes
import os
import subprocess
from collections import namedtuple
from copy import deepcopy
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")

AdjacencyLists = namedtuple(
    "AdjacencyLists",
    ["node_type", "value_parents", "volatility_parents", "value_children", "volatility_children", "coupling_fn"],
)

# opcodes / enums: keep in sync with hgf_oracle.cpp
OPS = dict(
    PRED_BINARY=1, PRED_CONT=2, PRED_CONT_MF=3, PRED_VOLATILE=4, PRED_VOLATILE_MF=5,
    PE_BINARY=10, PE_CONT=11, PE_VOLATILE=12,
    POST_CONT=20, POST_CONT_MF=21, POST_CONT_EHGF=22, POST_CONT_EHGF_MF=23,
    POST_CONT_UNBOUNDED=24, POST_VOLATILE=25, POST_VOLATILE_MF=26,
)
VOL_UPDATES = {"unbounded": 0, "eHGF": 1, "standard": 2}
COUPLING_FNS = {None: 0, "linear": 0, "identity": 1, "relu": 2, "sigmoid": 3, "tanh": 4, "leaky_relu": 5, "gelu": 6, "sin": 7}
RESPONSES = {
    "first_level_gaussian_surprise": 1,
    "total_gaussian_surprise": 2,
    "first_level_binary_surprise": 3,
    "binary_softmax": 4,
    "binary_softmax_inverse_temperature": 5,
}
FIELDS = [
    "mean", "expected_mean", "precision", "expected_precision",
    "tonic_volatility", "tonic_drift", "autoconnection_strength", "observed",
    "mean_vol", "expected_mean_vol", "precision_vol", "expected_precision_vol", "tonic_volatility_vol",
    "temp.effective_precision", "temp.conditional_expected_precision", "temp.value_prediction_error",
    "temp.volatility_prediction_error", "temp.current_variance", "temp.effective_precision_vol",
]
NF = len(FIELDS)
FIELD_INDEX = {name: i for i, name in enumerate(FIELDS)}


def build_library(force: bool = False) -> str:
    """Compile liboracle.so (g++, no FMA contraction so float32 rounds once per op)."""
    src = os.path.join(_HERE, "hgf_oracle.cpp")
    hdr = os.path.join(_HERE, "hgf_scalar.hpp")
    if (
        force
        or not os.path.exists(_LIB_PATH)
        or os.path.getmtime(_LIB_PATH) < max(os.path.getmtime(src), os.path.getmtime(hdr))
    ):
        cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-fopenmp", "-ffp-contract=off", "-o", _LIB_PATH, src]
        subprocess.run(cmd, check=True, cwd=_HERE)
    return _LIB_PATH


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build_library()
        L = ctypes.CDLL(_LIB_PATH)
        ip = ctypes.POINTER(ctypes.c_int)
        dp = ctypes.POINTER(ctypes.c_double)
        L.oracle_net_create.restype = ctypes.c_void_p
        L.oracle_net_create.argtypes = [ctypes.c_int] + [ip] * 10 + [ctypes.c_int, ctypes.c_int, ctypes.c_double]
        L.oracle_net_destroy.argtypes = [ctypes.c_void_p]
        L.oracle_net_set_schedule.argtypes = [ctypes.c_void_p, ctypes.c_int, ip, ip, ctypes.c_int, ip, ip]
        L.oracle_net_set_inputs.argtypes = [ctypes.c_void_p, ctypes.c_int, ip, ctypes.c_int]
        L.oracle_n_slots.argtypes = [ctypes.c_void_p]
        L.oracle_slot_coupling.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int]
        L.oracle_slot_clip.argtypes = [ctypes.c_void_p]
        L.oracle_slot_resp.argtypes = [ctypes.c_void_p]
        L.oracle_run.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, dp, dp, dp, dp, dp]
        L.oracle_apply_op.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_double, dp]
        L.oracle_response.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, dp, dp, dp, dp]
        L.oracle_batch_surprise.argtypes = [
            ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_int, ctypes.c_longlong, dp, dp, dp, dp,
            ctypes.c_int, ip, dp, ctypes.c_int, dp, dp, dp, ctypes.c_int,
        ]
        L.oracle_batch_grad.argtypes = [
            ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_int, ctypes.c_longlong, dp, dp, dp, dp,
            ctypes.c_int, ip, dp, ctypes.c_int, dp, dp, dp, ctypes.c_int,
        ]
        L.oracle_binary_surprise.restype = ctypes.c_double
        L.oracle_binary_surprise.argtypes = [ctypes.c_int, ctypes.c_double, ctypes.c_double]
        L.oracle_gaussian_surprise.restype = ctypes.c_double
        L.oracle_gaussian_surprise.argtypes = [ctypes.c_int, ctypes.c_double, ctypes.c_double, ctypes.c_double]
        L.oracle_lambert_w0.restype = ctypes.c_double
        L.oracle_lambert_w0.argtypes = [ctypes.c_int, ctypes.c_double]
        _lib = L
    return _lib


def _ip(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_int))


def _dp(a):
    return None if a is None else a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


# ---------------------------------------------------------------------------
# node defaults -- pyhgf/model/add_nodes.py
# ---------------------------------------------------------------------------
def _continuous_defaults():  # add_nodes.py:65-85
    return {
        "mean": 0.0, "expected_mean": 0.0, "precision": 1.0, "expected_precision": 1.0,
        "volatility_coupling_children": None, "volatility_coupling_parents": None,
        "value_coupling_children": None, "value_coupling_parents": None,
        "tonic_volatility": -4.0, "tonic_drift": 0.0, "autoconnection_strength": 1.0, "observed": 1,
        "temp": {
            "effective_precision": 0.0, "conditional_expected_precision": 1.0, "value_prediction_error": 0.0,
            "volatility_prediction_error": 0.0, "current_variance": 1.0,
        },
    }


def _volatile_defaults():  # add_nodes.py:159-184
    return {
        "mean": 0.0, "expected_mean": 0.0, "precision": 1.0, "expected_precision": 1.0,
        "autoconnection_strength": 0.0,
        "mean_vol": 0.0, "expected_mean_vol": 0.0, "precision_vol": 1.0, "expected_precision_vol": 1.0,
        "tonic_volatility_vol": -4.0,
        "value_coupling_children": None, "value_coupling_parents": None,
        "temp": {
            "effective_precision": 0.0, "conditional_expected_precision": 1.0, "value_prediction_error": 0.0,
            "volatility_prediction_error": 0.0, "effective_precision_vol": 0.0, "current_variance": 1.0,
        },
    }


def _binary_defaults():  # add_nodes.py:330-340
    return {
        "observed": 1, "mean": 0, "expected_mean": 0.5, "precision": 1.0, "expected_precision": 1.0,
        "value_coupling_parents": None, "temp": {"value_prediction_error": 0.0},
    }


def _get_couplings(*specs):  # add_nodes.py:603-650
    out = []
    for indexes in specs:
        if indexes is None:
            out.append((None, None))
        elif isinstance(indexes, int):
            out.append(((indexes,), (1.0,)))
        elif isinstance(indexes, list):
            out.append((tuple(indexes), tuple([1.0] * len(indexes))))
        elif isinstance(indexes, tuple):
            out.append((tuple(indexes[0]), tuple(indexes[1])))
        else:
            raise TypeError("invalid coupling specification")
    return out


class Network:
    """CPU-oracle twin of ``pyhgf.model.Network`` (hot-path subset)."""

    def __init__(self, volatility_updates="unbounded", max_posterior_precision=1e10, mean_field_updates=False,
                 precision_clipping_value=1e-6, dtype="float32"):
        if volatility_updates not in VOL_UPDATES:
            raise ValueError("Invalid update type.")
        self.edges: tuple = ()
        self.n_nodes = 0
        self.attributes: dict = {-1: {"time_step": 0.0, "precision_clipping_value": float(precision_clipping_value)}}
        self.volatility_updates = volatility_updates
        self.mean_field_updates = bool(mean_field_updates)
        self.max_posterior_precision = float(max_posterior_precision)
        self.dtype = 32 if str(dtype) in ("float32", "32", "f32") else 64
        self.update_sequence = None
        self._input_idxs = None
        self._handle = None
        self.node_trajectories = None
        self.last_attributes = None
        self._traj_flat = None

    # -- graph building ------------------------------------------------------
    def add_nodes(self, kind="continuous-state", n_nodes=1, node_parameters=None, value_children=None,
                  value_parents=None, volatility_children=None, volatility_parents=None, coupling_fn=(None,),
                  **additional_parameters):
        node_parameters = dict(node_parameters or {})
        if kind not in ("continuous-state", "binary-state", "volatile-state"):
            raise ValueError("Invalid node type (oracle supports continuous-, binary- and volatile-state).")
        vp, volp, vc, volc = _get_couplings(value_parents, volatility_parents, value_children, volatility_children)
        if kind == "continuous-state":
            node_type, defaults = 2, _continuous_defaults()
        elif kind == "volatile-state":
            node_type, defaults = 6, _volatile_defaults()
            volp, volc = (None, None), (None, None)  # add_nodes.py:154-157
        else:
            node_type, defaults = 1, _binary_defaults()
        invalid = [k for k in additional_parameters if k not in defaults]  # add_nodes.py:673-689
        if invalid:
            raise ValueError(f"Some parameter(s) passed as keyword arguments were not found ({invalid}).")
        defaults.update(additional_parameters)
        defaults.update(node_parameters)
        coupling_fn = tuple(coupling_fn)
        if vc[0] is not None and len(vc[0]) != len(coupling_fn):  # add_nodes.py:746-762
            if len(coupling_fn) == 1:
                coupling_fn = len(vc[0]) * coupling_fn
            else:
                raise ValueError("The number of coupling fn and value children do not match")
        for _ in range(n_nodes):
            node_idx = self.n_nodes
            self.edges = self.edges + (AdjacencyLists(node_type, None, None, None, None, coupling_fn),)
            self.attributes[node_idx] = deepcopy(defaults)
            if vp[0] is not None:
                for idx, cpl in zip(*vp):
                    self.add_edges("value", idx, node_idx, cpl)
            if vc[0] is not None:
                for idx, cpl in zip(*vc):
                    self.add_edges("value", node_idx, idx, cpl)
            if volc[0] is not None:
                for idx, cpl in zip(*volc):
                    self.add_edges("volatility", node_idx, idx, cpl)
            if volp[0] is not None:
                for idx, cpl in zip(*volp):
                    self.add_edges("volatility", idx, node_idx, cpl)
            self.n_nodes += 1
        self._handle = None
        self.update_sequence = None
        return self

    def add_edges(self, kind, parent_idx, child_idx, coupling_strength=1.0):  # utils/add_edges.py:10-176
        edges = list(self.edges)
        key_c = "value_coupling_children" if kind == "value" else "volatility_coupling_children"
        key_p = "value_coupling_parents" if kind == "value" else "volatility_coupling_parents"
        fld_c = "value_children" if kind == "value" else "volatility_children"
        fld_p = "value_parents" if kind == "value" else "volatility_parents"
        e = edges[parent_idx]._asdict()
        pa = self.attributes[parent_idx]
        if e[fld_c] is None:
            e[fld_c] = (child_idx,)
            pa[key_c] = np.array([coupling_strength], dtype=np.float64)
        else:
            e[fld_c] = e[fld_c] + (child_idx,)
            pa[key_c] = np.concatenate([pa[key_c], [coupling_strength]])
        edges[parent_idx] = AdjacencyLists(**e)
        e = edges[child_idx]._asdict()
        ca = self.attributes[child_idx]
        if e[fld_p] is None:
            e[fld_p] = (parent_idx,)
            ca[key_p] = np.array([coupling_strength], dtype=np.float64)
        else:
            e[fld_p] = e[fld_p] + (parent_idx,)
            ca[key_p] = np.concatenate([ca[key_p], [coupling_strength]])
        edges[child_idx] = AdjacencyLists(**e)
        self.edges = tuple(edges)
        return self

    # -- inputs ----------------------------------------------------------------
    @property
    def input_idxs(self):  # network.py:157-176
        auto = tuple(
            i for i in range(len(self.edges))
            if self.edges[i].value_children is None and self.edges[i].volatility_children is None
        )
        if self._input_idxs is not None:
            return self._input_idxs
        for idx in auto:
            if self.edges[idx].node_type == 2:
                self.attributes[idx]["autoconnection_strength"] = 0.0
                self.attributes[idx]["tonic_volatility"] = 0.0
            self.attributes[idx].setdefault("observed", 1)
        return auto

    @input_idxs.setter
    def input_idxs(self, value):
        self._input_idxs = tuple(value)

    # -- schedule -- utils/get_update_sequence.py:51-351 -----------------------
    def get_update_sequence(self):
        edges, n = self.edges, len(self.edges)
        vu, mf = self.volatility_updates, self.mean_field_updates
        pred, upd = [], []
        no_pe = list(range(n))
        no_pred = list(range(n))
        no_post = [
            i for i in range(n)
            if not (edges[i].value_children is None and edges[i].volatility_children is None) and edges[i].node_type != 0
        ]

        def parents(i):
            return [j for grp in (edges[i].value_parents, edges[i].volatility_parents) if grp is not None for j in grp]

        def children(i):
            return [j for grp in (edges[i].value_children, edges[i].volatility_children) if grp is not None for j in grp]

        while True:  # :106-151
            no_update = True
            for idx in list(no_pred):
                if not any(p in no_pred for p in parents(idx)):
                    no_update = False
                    no_pred.remove(idx)
                    t = edges[idx].node_type
                    if t == 1:
                        pred.append((idx, OPS["PRED_BINARY"]))
                    elif t == 2:
                        pred.append((idx, OPS["PRED_CONT_MF"] if mf else OPS["PRED_CONT"]))
                    elif t == 6:
                        pred.append((idx, OPS["PRED_VOLATILE_MF"] if mf else OPS["PRED_VOLATILE"]))
            if not no_pred:
                break
            if no_update:
                raise Warning("The structure of the network cannot be updated consistently.")

        while True:  # :156-329
            no_update = True
            for idx in list(no_post):
                if all(c not in no_pe for c in children(idx)):
                    no_update = False
                    t = edges[idx].node_type
                    op = None
                    if t == 2:
                        has_vol_children = edges[idx].volatility_children is not None
                        if vu == "unbounded":
                            op = "POST_CONT_UNBOUNDED" if has_vol_children else ("POST_CONT_MF" if mf else "POST_CONT")
                        elif vu == "eHGF":
                            if has_vol_children:
                                op = "POST_CONT_EHGF_MF" if mf else "POST_CONT_EHGF"
                            else:
                                op = "POST_CONT_MF" if mf else "POST_CONT"
                        else:
                            op = "POST_CONT_MF" if mf else "POST_CONT"
                    elif t == 6:
                        op = "POST_VOLATILE_MF" if mf else "POST_VOLATILE"
                    if op is not None:
                        upd.append((idx, OPS[op]))
                    no_post.remove(idx)
            for idx in list(no_pe):
                if len(parents(idx)) == 0 and edges[idx].node_type != 6:
                    no_pe.remove(idx)
                elif idx not in no_post:
                    t = edges[idx].node_type
                    if t == 1:
                        upd.append((idx, OPS["PE_BINARY"]))
                    elif t == 2:
                        upd.append((idx, OPS["PE_CONT"]))
                    elif t == 6:
                        upd.append((idx, OPS["PE_VOLATILE"]))
                    no_update = False
                    no_pe.remove(idx)
            if not no_pe and not no_post:
                break
            if no_update:
                raise Warning("The structure of the network cannot be updated consistently.")
        return tuple(pred), tuple(upd)

    def create_belief_propagation_fn(self, overwrite: bool = True):
        _ = self.input_idxs  # triggers the input-node attribute override
        if self.update_sequence is None:
            self.update_sequence = self.get_update_sequence()
        self._build_handle()
        return self

    def get_network(self):
        if self._handle is None:
            self.create_belief_propagation_fn()
        return self.attributes, self.edges, self.update_sequence

    # -- flattening to the C oracle ---------------------------------------------
    def _build_handle(self):
        L = lib()
        n = len(self.edges)

        def csr(name):
            ptr, idx = [0], []
            for e in self.edges:
                lst = getattr(e, name) or ()
                idx.extend(lst)
                ptr.append(len(idx))
            return np.asarray(ptr, dtype=np.int32), np.asarray(idx if idx else [0], dtype=np.int32)

        vp_ptr, vp_idx = csr("value_parents")
        volp_ptr, volp_idx = csr("volatility_parents")
        vc_ptr, vc_idx = csr("value_children")
        volc_ptr, volc_idx = csr("volatility_children")
        vc_fn = []
        for e in self.edges:
            if e.value_children is not None:
                fns = e.coupling_fn
                for k in range(len(e.value_children)):
                    vc_fn.append(COUPLING_FNS[fns[k]])
        vc_fn = np.asarray(vc_fn if vc_fn else [0], dtype=np.int32)
        node_type = np.asarray([e.node_type for e in self.edges], dtype=np.int32)
        h = L.oracle_net_create(
            n, _ip(node_type), _ip(vp_ptr), _ip(vp_idx), _ip(volp_ptr), _ip(volp_idx), _ip(vc_ptr), _ip(vc_idx),
            _ip(vc_fn), _ip(volc_ptr), _ip(volc_idx), VOL_UPDATES[self.volatility_updates],
            int(self.mean_field_updates), self.max_posterior_precision,
        )
        pred, upd = self.update_sequence
        pn = np.asarray([s[0] for s in pred] or [0], dtype=np.int32)
        po = np.asarray([s[1] for s in pred] or [0], dtype=np.int32)
        un = np.asarray([s[0] for s in upd] or [0], dtype=np.int32)
        uo = np.asarray([s[1] for s in upd] or [0], dtype=np.int32)
        L.oracle_net_set_schedule(h, len(pred), _ip(pn), _ip(po), len(upd), _ip(un), _ip(uo))
        ii = np.asarray(self.input_idxs, dtype=np.int32)
        self._n_fed = len(ii)
        L.oracle_net_set_inputs(h, len(ii), _ip(ii), self._n_fed)
        self._handle = h
        self.n_slots = L.oracle_n_slots(h)
        return h

    def slot(self, node: int, key: str, pos: int = 0) -> int:
        """Flat slot index of an attribute (node=-1 for globals / response parameters)."""
        L = lib()
        if node == -1:
            if key == "precision_clipping_value":
                return L.oracle_slot_clip(self._handle)
            if key == "response_parameter":
                return L.oracle_slot_resp(self._handle) + pos
            raise KeyError(key)
        kinds = {"value_coupling_parents": 0, "volatility_coupling_parents": 1, "value_coupling_children": 2,
                 "volatility_coupling_children": 3}
        if key in kinds:
            return L.oracle_slot_coupling(self._handle, kinds[key], node, pos)
        return node * NF + FIELD_INDEX[key]

    def flatten(self, attributes=None, response_parameters=None) -> np.ndarray:
        attributes = self.attributes if attributes is None else attributes
        slots = np.zeros(self.n_slots, dtype=np.float64)
        for i in range(len(self.edges)):
            a = attributes[i]
            for k, name in enumerate(FIELDS):
                if name.startswith("temp."):
                    v = a.get("temp", {}).get(name[5:], 0.0)
                else:
                    v = a.get(name, 1.0 if name == "observed" else 0.0)
                slots[i * NF + k] = float(v)
            for key in ("value_coupling_parents", "volatility_coupling_parents", "value_coupling_children",
                        "volatility_coupling_children"):
                arr = a.get(key)
                if arr is not None:
                    for pos, c in enumerate(np.atleast_1d(arr)):
                        slots[self.slot(i, key, pos)] = float(c)
        slots[self.slot(-1, "precision_clipping_value")] = attributes[-1]["precision_clipping_value"]
        if response_parameters is not None:
            for pos, c in enumerate(np.atleast_1d(np.asarray(response_parameters, dtype=np.float64))):
                slots[self.slot(-1, "response_parameter", pos)] = float(c)
        if self.dtype == 32:
            slots = slots.astype(np.float32).astype(np.float64)
        return slots

    def unflatten(self, slots: np.ndarray, time_step=None) -> dict:
        out = deepcopy(self.attributes)
        for i in range(len(self.edges)):
            for k, name in enumerate(FIELDS):
                v = slots[i * NF + k]
                if name.startswith("temp."):
                    if name[5:] in out[i].get("temp", {}):
                        out[i]["temp"][name[5:]] = v
                elif name in out[i]:
                    out[i][name] = v
        if time_step is not None:
            out[-1]["time_step"] = time_step
        return out

    def _cast(self, x):
        x = np.asarray(x, dtype=np.float64)
        return x.astype(np.float32).astype(np.float64) if self.dtype == 32 else x

    def _format_inputs(self, input_data, time_steps, observed):
        x = self._cast(input_data)
        if x.ndim == 1:
            x = x[:, None]
        n_steps = x.shape[0]
        ts = None if time_steps is None else np.ascontiguousarray(self._cast(time_steps))
        if observed is None:
            obs = None
        else:
            if isinstance(observed, tuple):
                observed = np.stack([np.asarray(o) for o in observed], axis=1)
            obs = np.asarray(observed, dtype=np.float64)
            if obs.ndim == 1:
                obs = obs[:, None]
            obs = np.ascontiguousarray(obs)
        return np.ascontiguousarray(x), ts, obs, n_steps

    # -- Network.input_data -- network.py:528-628 ----------------------------------
    def input_data(self, input_data, time_steps=None, observed=None, input_idxs=None, record_trajectories=True):
        if input_idxs is not None:
            self.input_idxs = input_idxs
        if self._handle is None:
            self.create_belief_propagation_fn()
        x, ts, obs, n_steps = self._format_inputs(input_data, time_steps, observed)
        # the reference zips (values, observed, input_idxs): extra auto-detected input
        # nodes beyond the supplied columns are simply not fed (beliefs_propagation.py:128)
        n_fed = min(x.shape[1], len(self.input_idxs)) if obs is None else min(x.shape[1], obs.shape[1], len(self.input_idxs))
        if x.shape[1] > len(self.input_idxs):
            raise ValueError("input_data has more columns than input nodes")
        if n_fed != self._n_fed:
            ii = np.asarray(self.input_idxs, dtype=np.int32)
            lib().oracle_net_set_inputs(self._handle, len(ii), _ip(ii), n_fed)
            self._n_fed = n_fed
        x = np.ascontiguousarray(x[:, :n_fed])
        obs = None if obs is None else np.ascontiguousarray(obs[:, :n_fed])
        slots = self.flatten()
        traj = np.zeros((n_steps, len(self.edges) * NF), dtype=np.float64) if record_trajectories else None
        lib().oracle_run(self._handle, self.dtype, n_steps, _dp(x), _dp(obs), _dp(ts), _dp(slots), _dp(traj))
        last_ts = 1.0 if ts is None else ts[-1]
        self.last_attributes = self.unflatten(slots, last_ts)
        if record_trajectories:
            self._traj_flat = traj
            self._n_steps = n_steps
            nt = {}
            for i in range(len(self.edges)):
                d, temp = {}, {}
                for k, name in enumerate(FIELDS):
                    col = traj[:, i * NF + k]
                    if name.startswith("temp."):
                        if name[5:] in self.attributes[i].get("temp", {}):
                            temp[name[5:]] = col
                    elif name in self.attributes[i]:
                        d[name] = col
                d["temp"] = temp
                nt[i] = d
            nt[-1] = {"time_step": np.ones(n_steps) if ts is None else ts}
            self.node_trajectories = nt
        else:
            self.node_trajectories = None
            self._traj_flat = None
        return self

    # -- Network.surprise -- network.py:986-1022 ----------------------------------
    def to_pandas(self):
        """pyhgf/utils/to_pandas.py:14-136 (state nodes of type 1, 2, 6): sufficient statistics, per-node surprise
        (binary_surprise / gaussian_surprise of each node's `mean` under its own prediction) and their row sum."""
        import pandas as pd

        nt, bits = self.node_trajectories, 32 if self.dtype == np.float32 else 64
        ts = np.asarray(nt[-1]["time_step"])
        cols = {"time_steps": ts, "time": np.cumsum(ts)}
        for types in ((1, 2), (6,)):
            for i, e in enumerate(self.edges):
                if e.node_type in types:
                    for var, v in nt[i].items():
                        if ("mean" in var) or ("precision" in var) or ("observed" in var):
                            cols[f"x_{i}_{var}"] = np.asarray(v)
        for i, e in enumerate(self.edges):
            if e.node_type == 1:
                cols[f"x_{i}_surprise"] = np.array([binary_surprise(x, m, bits) for x, m in zip(nt[i]["mean"], nt[i]["expected_mean"])])
        for i, e in enumerate(self.edges):
            if e.node_type in (2, 6):
                cols[f"x_{i}_surprise"] = np.array([gaussian_surprise(x, m, p, bits) for x, m, p in
                                                    zip(nt[i]["mean"], nt[i]["expected_mean"], nt[i]["expected_precision"])])
        df = pd.DataFrame(cols)
        df["total_surprise"] = df.iloc[:, df.columns.str.contains("_surprise")].sum(axis=1, min_count=1)
        return df

    def surprise(self, response_function, response_function_inputs=(), response_function_parameters=None):
        return response_function(
            hgf=self, response_function_inputs=response_function_inputs,
            response_function_parameters=response_function_parameters,
        )

    def _response(self, name, inputs=None, parameters=None):
        slots0 = self.flatten(response_parameters=parameters)
        out = np.zeros(self._n_steps, dtype=np.float64)
        ri = None
        if inputs is not None and np.size(inputs) > 0:
            ri = np.ascontiguousarray(np.asarray(inputs, dtype=np.float64))
        lib().oracle_response(self._handle, self.dtype, RESPONSES[name], self._n_steps, _dp(self._traj_flat),
                              _dp(slots0), _dp(ri), _dp(out))
        return out

    # -- Network.input_custom_sequence -- network.py:653-767 ---------------------
    def _set_schedule(self, seq):
        pred, upd = seq
        pn = np.asarray([s[0] for s in pred] or [0], dtype=np.int32)
        po = np.asarray([s[1] for s in pred] or [0], dtype=np.int32)
        un = np.asarray([s[0] for s in upd] or [0], dtype=np.int32)
        uo = np.asarray([s[1] for s in upd] or [0], dtype=np.int32)
        lib().oracle_net_set_schedule(self._handle, len(pred), _ip(pn), _ip(po), len(upd), _ip(un), _ip(uo))

    def input_custom_sequence(self, update_branches, branches_idx, input_data, time_steps=None, observed=None):
        """``lax.switch(idx, branches_fn, attributes, data)`` under ``scan`` (network.py:746-758): every step runs the
        belief propagation of the branch its index names (clamped, as lax.switch does).  Runs of equal index are
        filtered with that branch's schedule, the attribute vector carried from one run to the next."""
        if self._handle is None:
            self.create_belief_propagation_fn()
        x, ts, obs, n_steps = self._format_inputs(input_data, time_steps, observed)
        n_fed = x.shape[1]
        if n_fed != self._n_fed:
            ii = np.asarray(self.input_idxs, dtype=np.int32)
            lib().oracle_net_set_inputs(self._handle, len(ii), _ip(ii), n_fed)
            self._n_fed = n_fed
        idx = np.clip(np.asarray(branches_idx, dtype=np.int64), 0, len(update_branches) - 1)
        assert idx.shape == (n_steps,)
        slots = self.flatten()
        traj = np.zeros((n_steps, len(self.edges) * NF), dtype=np.float64)
        t0 = 0
        while t0 < n_steps:
            t1 = t0 + 1
            while t1 < n_steps and idx[t1] == idx[t0]:
                t1 += 1
            self._set_schedule(update_branches[idx[t0]])
            seg = np.zeros((t1 - t0, traj.shape[1]), dtype=np.float64)
            lib().oracle_run(self._handle, self.dtype, t1 - t0, _dp(np.ascontiguousarray(x[t0:t1])),
                             _dp(None if obs is None else np.ascontiguousarray(obs[t0:t1])),
                             _dp(None if ts is None else np.ascontiguousarray(ts[t0:t1])), _dp(slots), _dp(seg))
            traj[t0:t1] = seg
            t0 = t1
        self._set_schedule(self.update_sequence)
        self.last_attributes = self.unflatten(slots, 1.0 if ts is None else ts[-1])
        nt = {}
        for i in range(len(self.edges)):
            d, temp = {}, {}
            for k, name in enumerate(FIELDS):
                col = traj[:, i * NF + k]
                if name.startswith("temp."):
                    if name[5:] in self.attributes[i].get("temp", {}):
                        temp[name[5:]] = col
                elif name in self.attributes[i]:
                    d[name] = col
            d["temp"] = temp
            nt[i] = d
        nt[-1] = {"time_step": np.ones(n_steps) if ts is None else ts}
        self.node_trajectories = nt
        return self

    def apply_op(self, op: str, node: int, attributes: dict, time_step: Optional[float] = None) -> dict:
        """Run one update step on an attributes dict (isolated-kernel goldens)."""
        if self._handle is None:
            self.create_belief_propagation_fn()
        slots = self.flatten(attributes)
        ts = attributes[-1]["time_step"] if time_step is None else time_step
        lib().oracle_apply_op(self._handle, self.dtype, OPS[op], node, float(ts), _dp(slots))
        out = deepcopy(attributes)
        for i in range(len(self.edges)):
            for k, name in enumerate(FIELDS):
                v = slots[i * NF + k]
                if name.startswith("temp."):
                    if name[5:] in out[i].get("temp", {}):
                        out[i]["temp"][name[5:]] = v
                elif name in out[i]:
                    out[i][name] = v
        return out

    # -- batched drivers (CPU baseline / bulk parity / gradient oracle) -----------
    def batch_surprise(self, values, param_slots, params, response, resp_inputs=None, response_parameters=None,
                       group_size=0, time_steps=None, observed=None, n_threads=0, return_last=False):
        """values [G][T][n_in] (or [T] / [T][n_in] shared); params [P][N] -> surprise_sum [N]."""
        if self._handle is None:
            self.create_belief_propagation_fn()
        v = self._cast(values)
        n_in = len(self.input_idxs)
        if v.ndim == 1:
            v = v[None, :, None]
        elif v.ndim == 2:
            v = v[None, :, :] if v.shape[1] == n_in and group_size == 0 else v[:, :, None]
        v = np.ascontiguousarray(v)
        n_steps = v.shape[1]
        params = np.ascontiguousarray(self._cast(np.atleast_2d(params)))
        n_params, n_inst = params.shape
        ps = np.asarray(list(param_slots) or [0], dtype=np.int32)
        slots0 = self.flatten(response_parameters=response_parameters)
        ri = None if resp_inputs is None else np.ascontiguousarray(np.atleast_2d(np.asarray(resp_inputs, dtype=np.float64)))
        ts = None if time_steps is None else np.ascontiguousarray(self._cast(time_steps))
        obs = None if observed is None else np.ascontiguousarray(np.asarray(observed, dtype=np.float64))
        out = np.zeros(n_inst, dtype=np.float64)
        last = np.zeros((n_inst, len(self.edges) * NF), dtype=np.float64) if return_last else None
        rc = lib().oracle_batch_surprise(
            self._handle, self.dtype, n_inst, n_steps, group_size, _dp(v), _dp(obs), _dp(ts), _dp(slots0),
            len(param_slots), _ip(ps), _dp(params), RESPONSES[response], _dp(ri), _dp(out), _dp(last), n_threads,
        )
        assert rc == 0
        return (out, last) if return_last else out

    def batch_grad(self, values, param_slots, params, response, resp_inputs=None, response_parameters=None,
                   group_size=0, time_steps=None, observed=None, n_threads=0, dtype=64):
        """Forward-mode gradient oracle (Dual<double>; dtype=32: the Dual<float> twin): returns
        (surprise_sum [N], grad [P][N])."""
        if self._handle is None:
            self.create_belief_propagation_fn()
        v = np.asarray(values, dtype=np.float64)
        n_in = len(self.input_idxs)
        if v.ndim == 1:
            v = v[None, :, None]
        elif v.ndim == 2:
            v = v[None, :, :] if v.shape[1] == n_in and group_size == 0 else v[:, :, None]
        v = np.ascontiguousarray(v)
        n_steps = v.shape[1]
        params = np.ascontiguousarray(np.atleast_2d(np.asarray(params, dtype=np.float64)))
        n_params, n_inst = params.shape
        ps = np.asarray(list(param_slots), dtype=np.int32)
        save = self.dtype
        self.dtype = 64
        slots0 = self.flatten(response_parameters=response_parameters)
        self.dtype = save
        ri = None if resp_inputs is None else np.ascontiguousarray(np.atleast_2d(np.asarray(resp_inputs, dtype=np.float64)))
        ts = None if time_steps is None else np.ascontiguousarray(np.asarray(time_steps, dtype=np.float64))
        obs = None if observed is None else np.ascontiguousarray(np.asarray(observed, dtype=np.float64))
        out = np.zeros(n_inst, dtype=np.float64)
        grad = np.zeros((n_params, n_inst), dtype=np.float64)
        if dtype == 32:
            v = v.astype(np.float32).astype(np.float64)
            params = params.astype(np.float32).astype(np.float64)
            slots0 = slots0.astype(np.float32).astype(np.float64)
        rc = lib().oracle_batch_grad(
            self._handle, int(dtype), n_inst, n_steps, group_size, _dp(v), _dp(obs), _dp(ts), _dp(slots0), n_params, _ip(ps),
            _dp(params), RESPONSES[response], _dp(ri), _dp(out), _dp(grad), n_threads,
        )
        assert rc == 0
        return out, grad


# ---------------------------------------------------------------------------
# response functions -- pyhgf/response.py:15-215 (same call signature)
# ---------------------------------------------------------------------------
def first_level_gaussian_surprise(hgf, response_function_inputs=None, response_function_parameters=None):
    return hgf._response("first_level_gaussian_surprise")


def total_gaussian_surprise(hgf, response_function_inputs=None, response_function_parameters=None):
    return hgf._response("total_gaussian_surprise")


def first_level_binary_surprise(hgf, response_function_inputs=None, response_function_parameters=None):
    return hgf._response("first_level_binary_surprise")


def binary_softmax(hgf, response_function_inputs=None, response_function_parameters=None):
    return hgf._response("binary_softmax", response_function_inputs)


def binary_softmax_inverse_temperature(hgf, response_function_inputs=None, response_function_parameters=None):
    return hgf._response("binary_softmax_inverse_temperature", response_function_inputs, response_function_parameters)


# scalar maths -- pyhgf/math.py:209-297, 450-488
def binary_surprise(x, expected_mean, dtype=32):
    return lib().oracle_binary_surprise(dtype, float(x), float(expected_mean))


def gaussian_surprise(x, expected_mean, expected_precision, dtype=32):
    return lib().oracle_gaussian_surprise(dtype, float(x), float(expected_mean), float(expected_precision))


def lambert_w0(z, dtype=64):
    return lib().oracle_lambert_w0(dtype, float(z))
