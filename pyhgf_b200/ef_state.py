"""Exponential-family state nodes with a fixed learning rate, on ``hgf_ef_filter``.

Mirrors ``add_ef_state`` (pyhgf/model/add_nodes.py:360-470) and the update the schedule gives such a node
(``prediction_error_update_exponential_family_fixed``, pyhgf/updates/prediction_error/exponential.py:13-60;
``get_update_sequence`` schedules nothing else for it: no prediction, no posterior step).  The node has no
parents, so every node is its own time loop over its ``dimension`` input columns; one launch filters a batch
of instances that differ in ``nus`` (and optionally in their series).

``learning="hgf-1|2|3"`` (add_nodes.py:438-452) puts continuous HGFs under the node, one per sufficient statistic; they
run on the ordinary scan as its inputs, fed with t(x) (``Network._implied_network``), and ``hgf_ef_filter`` takes their
trajectories to form the per-step learning rate (``implied``).
"""

from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib

NODE_TYPE = 3
MAX_DIMENSION = 8  # template instances of ef_fixed_kernel (csrc/hgf_ef.cu)


def prediction_error_update_exponential_family_dynamic(*args, **kwargs):
    """Schedule marker (pyhgf/updates/prediction_error/exponential.py:63-120): t(x) goes to the continuous nodes that
    track the statistics -- here they are the scan's inputs, fed with t(x)."""
    raise RuntimeError("schedule marker: the step runs inside the CUDA time loop")


def posterior_update_exponential_family_dynamic(*args, **kwargs):
    """Schedule marker (pyhgf/updates/posterior/exponential.py): runs inside hgf_ef_filter (per-step learning rate)."""
    raise RuntimeError("schedule marker: the step runs as hgf_ef_filter")


def n_sufficient_statistics(d: int) -> int:
    return d + d * (d + 1) // 2


def defaults() -> dict:  # add_nodes.py:389-396
    return {"dimension": 1, "distribution": "normal", "learning": "generalised-filtering", "nus": 3.0,
            "xis": np.array([0.0, 1.0]), "observed": 1}


def finish_parameters(params: dict) -> dict:
    """The derived entries of add_nodes.py:402-415 and the checks of :430-437."""
    d = int(params["dimension"])
    dist = params["distribution"]
    if dist not in ("normal", "multivariate-normal"):
        raise ValueError("The distribution should be either 'normal' or 'multivariate-normal'.")
    if str(params["learning"]) not in ("generalised-filtering", "hgf-1", "hgf-2", "hgf-3"):
        raise ValueError("learning should be 'generalised-filtering', 'hgf-1', 'hgf-2' or 'hgf-3'")
    if dist == "normal" and d != 1:
        raise ValueError("a 'normal' ef-state node has dimension 1; use 'multivariate-normal'")
    if not 1 <= d <= MAX_DIMENSION:
        raise NotImplementedError(f"ef-state dimension {d} is outside the built kernels (1..{MAX_DIMENSION})")
    k = n_sufficient_statistics(d)
    params["mean"] = np.zeros(d) if d > 1 else 0.0
    params["observation_ss"] = np.zeros(k)
    if dist == "normal":
        params["xis"] = np.array([0.0, 1.0])
    else:  # MultivariateNormal.sufficient_statistics_from_parameters(0, I): [mean, tril(cov + mean mean^T)]
        params["xis"] = np.append(np.zeros(d), np.identity(d)[np.tril_indices(d)])
    params.pop("distribution")
    params.pop("learning")
    return params


def sufficient_statistics(x: np.ndarray, d: int) -> np.ndarray:
    """t(x) for every row of ``x`` (T, d): math.py:90-92 (normal), MultivariateNormal (tril of the outer product)."""
    r, q = np.tril_indices(d)
    return np.concatenate([x, x[:, r] * x[:, q]], axis=1)


def launch(*, bits: int, device, series: torch.Tensor, observed, nus, xis0: np.ndarray, dimension: int,
           input_group_size: int, want_trajectory: bool, implied=None):
    """One ``hgf_ef_filter`` launch. series [G][T][d], nus [N] (device tensors) -> (trajectory [K][T][N] | None, last [K][N]).
    With ``implied`` [4][K][T][N] (learning="hgf-*") the learning rate is per step and a third result, nus [T][N], is returned."""
    L = _lib.lib()
    tdt = torch.float32 if bits == 32 else torch.float64
    G, T, d = series.shape
    N = int(nus.shape[0]) if implied is None else int(implied.shape[3])
    K = n_sufficient_statistics(dimension)
    if d != dimension or len(xis0) != K:
        raise ValueError("series / xis do not match the node's dimension")
    x0 = torch.as_tensor(np.asarray(xis0, dtype=np.float64), dtype=tdt).to(device)
    traj = torch.empty((K, T, N), dtype=tdt, device=device) if want_trajectory else None
    last = torch.empty((K, N), dtype=tdt, device=device)
    c = _lib.EfCall()
    c.dtype = bits
    c.n_instances = N
    c.n_steps = T
    c.dimension = dimension
    c.input_group_size = int(input_group_size)
    c.input_data = series.data_ptr()
    c.observed = observed.data_ptr() if observed is not None else None
    c.nus = nus.data_ptr() if implied is None else None
    nus_t = None
    if implied is not None:
        if tuple(implied.shape) != (4, K, T, N) or not implied.is_contiguous():
            raise ValueError("implied should be a contiguous [4][K][T][N] tensor")
        nus_t = torch.empty((T, N), dtype=tdt, device=device)
        c.implied, c.nus_trajectory = implied.data_ptr(), nus_t.data_ptr()
    c.xis0 = x0.data_ptr()
    c.xis_trajectory = traj.data_ptr() if traj is not None else None
    c.xis_last = last.data_ptr()
    with torch.cuda.device(device):  # the library launches on the CURRENT device: make it the tensors' one
        s = torch.cuda.current_stream(device)  # the temporaries above are freed in this stream's order
        _lib.check(L.hgf_ef_filter(C.byref(c), C.c_void_p(s.cuda_stream)))
    return (traj, last) if implied is None else (traj, last, nus_t)
