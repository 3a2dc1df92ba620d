"""Categorical state nodes: the implied binary HGFs on the scan, the node's own step on ``hgf_categorical_update``.

Mirrors ``add_categorical_state`` (pyhgf/model/add_nodes.py:473-545) and ``fill_categorical_state_node``
(pyhgf/utils/fill_categorical_state_node.py:13-81): under a categorical node sit one binary node per category, one
continuous value parent per binary node and ONE volatility parent shared by the continuous nodes.  The schedule
(get_update_sequence.py:301-305, 337-345) hands the observation vector to the binary nodes first
(``categorical_state_prediction_error``) and runs ``categorical_state_update`` last, so the time loop is the ordinary
one over the implied nodes -- with the binary nodes as its inputs -- and the categorical node is a per-step function of
their trajectories (include/pyhgf_cuda.h: ``hgf_categorical_update``).

As in the reference, ``node_parameters["binary_parameters"]`` is stored on the node but the implied network is built
from the DEFAULT binary parameters (add_nodes.py:532-537 passes the local defaults to ``fill_categorical_state_node``);
the reference's golden (tests/test_categorical.py:81-88) depends on it.
"""

from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib

NODE_TYPE = 5


def categorical_state_prediction_error(*args, **kwargs):
    """Schedule marker (pyhgf/updates/prediction_error/categorical.py:12-47): the observation vector goes to the binary
    nodes -- here they ARE the scan's inputs, so the step is the kernel's set_observation."""
    raise RuntimeError("schedule marker: the step runs inside the CUDA time loop")


def categorical_state_update(*args, **kwargs):
    """Schedule marker (pyhgf/updates/posterior/categorical.py:12-91): runs as hgf_categorical_update."""
    raise RuntimeError("schedule marker: the step runs as hgf_categorical_update")


def default_binary_parameters(n_categories: int) -> dict:  # add_nodes.py:496-506
    return {"n_categories": n_categories, "precision_1": 1.0, "precision_2": 1.0, "precision_3": 1.0,
            "mean_1": 1.0 / n_categories, "mean_2": -float(np.log(n_categories - 1)), "mean_3": 0.0,
            "tonic_volatility_2": -4.0, "tonic_volatility_3": -4.0}


def defaults(n_categories: int, first_binary_idx: int) -> dict:  # add_nodes.py:507-517
    return {"binary_idxs": [first_binary_idx + i for i in range(n_categories)], "n_categories": n_categories,
            "surprise": 0.0, "kl_divergence": 0.0, "alpha": np.ones(n_categories), "observed": 1,
            "mean": np.array([1.0 / n_categories] * n_categories), "binary_parameters": default_binary_parameters(n_categories)}


def update(*, bits: int, device, observations: torch.Tensor, expected_mean: torch.Tensor, parent_mean: torch.Tensor,
           alpha0, input_group_size: int, want_vectors: bool = True) -> dict:
    """One ``hgf_categorical_update`` launch.  observations [G][T][n], expected_mean / parent_mean [n][T][N] (device)
    -> {"kl_divergence": [T][N], "surprise": [T][N], "alpha": [n][T][N], "mean": [n][T][N]}."""
    L = _lib.lib()
    tdt = torch.float32 if bits == 32 else torch.float64
    n, T, N = expected_mean.shape
    if observations.shape[1:] != (T, n) or parent_mean.shape != expected_mean.shape:
        raise ValueError("observations / trajectories do not match the node's categories")
    a0 = torch.as_tensor(np.asarray(alpha0, dtype=np.float64), dtype=tdt).to(device)
    out = {"kl_divergence": torch.empty((T, N), dtype=tdt, device=device), "surprise": torch.empty((T, N), dtype=tdt, device=device)}
    if want_vectors:
        out["alpha"] = torch.empty((n, T, N), dtype=tdt, device=device)
        out["mean"] = torch.empty((n, T, N), dtype=tdt, device=device)
    c = _lib.CategoricalCall()
    c.dtype, c.n_instances, c.n_steps, c.n_categories = bits, N, T, n
    c.input_group_size = int(input_group_size)
    c.observations = observations.data_ptr()
    c.expected_mean = expected_mean.data_ptr()
    c.parent_mean = parent_mean.data_ptr()
    c.alpha0 = a0.data_ptr()
    c.alpha = out["alpha"].data_ptr() if want_vectors else None
    c.mean = out["mean"].data_ptr() if want_vectors else None
    c.kl_divergence = out["kl_divergence"].data_ptr()
    c.surprise = out["surprise"].data_ptr()
    with torch.cuda.device(device):  # the library launches on the CURRENT device: make it the tensors' one
        s = torch.cuda.current_stream(device)  # the temporaries above are freed in this stream's order
        _lib.check(L.hgf_categorical_update(C.byref(c), C.c_void_p(s.cuda_stream)))
    return out
