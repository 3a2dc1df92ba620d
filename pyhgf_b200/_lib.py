"""ctypes binding of ``libpyhgf_cuda.so`` (declared in ``include/pyhgf_cuda.h``).

There is no fallback: if the library has not been built, importing anything
that computes raises with the build command.
"""

from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# HGF_MATH=accurate loads the validation build (IEEE division, libdevice transcendentals, no FMA contraction:
# `HGF_MATH=accurate python -m pyhgf_b200.build`); the shipped library is the default.
LIB_PATH = os.path.join(_HERE, "lib", "libpyhgf_cuda_accurate.so" if os.environ.get("HGF_MATH") == "accurate" else "libpyhgf_cuda.so")

N_FIELDS = 19
N_RESPONSE_PARAMETERS = 4
MAX_PARAMS = 16
MAX_BRANCHES = 8  # HGF_MAX_BRANCHES
MAX_CATEGORIES = 16  # HGF_MAX_CATEGORIES
MAX_RECORDS = 640

#: order of the per-node attribute fields inside an instance's attribute vector
FIELDS = (
    "mean", "expected_mean", "precision", "expected_precision",
    "tonic_volatility", "tonic_drift", "autoconnection_strength", "observed",
    "mean_vol", "expected_mean_vol", "precision_vol", "expected_precision_vol", "tonic_volatility_vol",
    "temp.effective_precision", "temp.conditional_expected_precision", "temp.value_prediction_error",
    "temp.volatility_prediction_error", "temp.current_variance", "temp.effective_precision_vol",
)
FIELD_INDEX = {name: i for i, name in enumerate(FIELDS)}

OPS = dict(
    PRED_BINARY=1, PRED_CONT=2, PRED_CONT_MF=3, PRED_VOLATILE=4, PRED_VOLATILE_MF=5,
    PE_BINARY=10, PE_CONT=11, PE_VOLATILE=12,
    POST_CONT=20, POST_CONT_MF=21, POST_CONT_EHGF=22, POST_CONT_EHGF_MF=23,
    POST_CONT_UNBOUNDED=24, POST_VOLATILE=25, POST_VOLATILE_MF=26,
)
VOLATILITY_UPDATES = {"unbounded": 0, "eHGF": 1, "standard": 2}
COUPLING_FNS = {None: 0, "linear": 0, "identity": 1, "relu": 2, "sigmoid": 3, "tanh": 4, "leaky_relu": 5, "gelu": 6, "sin": 7}
RESPONSES = {
    None: 0,
    "first_level_gaussian_surprise": 1,
    "total_gaussian_surprise": 2,
    "first_level_binary_surprise": 3,
    "binary_softmax": 4,
    "binary_softmax_inverse_temperature": 5,
}

_i32p = C.POINTER(C.c_int32)


class Topology(C.Structure):
    """``hgf_topology_t``"""

    _fields_ = [
        ("n_nodes", C.c_int32),
        ("node_type", _i32p),
        ("value_parents_ptr", _i32p), ("value_parents_idx", _i32p),
        ("volatility_parents_ptr", _i32p), ("volatility_parents_idx", _i32p),
        ("value_children_ptr", _i32p), ("value_children_idx", _i32p), ("value_children_fn", _i32p),
        ("volatility_children_ptr", _i32p), ("volatility_children_idx", _i32p),
        ("n_prediction_steps", C.c_int32), ("prediction_node", _i32p), ("prediction_op", _i32p),
        ("n_update_steps", C.c_int32), ("update_node", _i32p), ("update_op", _i32p),
        ("n_inputs", C.c_int32), ("input_idxs", _i32p), ("n_fed", C.c_int32),
        ("volatility_updates", C.c_int32), ("mean_field_updates", C.c_int32),
        ("max_posterior_precision", C.c_double),
        ("n_branches", C.c_int32), ("branch_prediction_ptr", _i32p), ("branch_update_ptr", _i32p),
    ]


class Call(C.Structure):
    """``hgf_call_t``"""

    _fields_ = [
        ("dtype", C.c_int32),
        ("n_instances", C.c_int64),
        ("n_steps", C.c_int32),
        ("input_group_size", C.c_int64),
        ("input_data", C.c_void_p),
        ("observed", C.c_void_p),
        ("time_steps", C.c_void_p),
        ("attributes", C.c_void_p),
        ("n_params", C.c_int32),
        ("param_slots", _i32p),
        ("params", C.c_void_p),
        ("response_fn", C.c_int32),
        ("response_group_size", C.c_int64),
        ("response_inputs", C.c_void_p),
        ("n_records", C.c_int32),
        ("record_slots", _i32p),
        ("trajectories", C.c_void_p),
        ("last_attributes", C.c_void_p),
        ("surprise_sum", C.c_void_p),
        ("surprise_pointwise", C.c_void_p),
        ("gradient", C.c_void_p),
        ("reduced", C.c_void_p),
        ("group", C.c_void_p),
        ("out_scale", C.c_double),
        ("n_groups", C.c_int32),
        ("world_size", C.c_int32),
        ("rank", C.c_int32),
        ("epoch", C.c_uint32),
        ("peer_workspace", C.POINTER(C.c_void_p)),
        ("multicast_workspace", C.c_void_p),
        ("generative", C.c_int32),
        ("seed", C.c_uint64),
        ("branch_idx", C.c_void_p),
        ("gradient_mode", C.c_int32),
    ]


#: every symbol include/pyhgf_cuda.h declares
EXPORTS = (
    "hgf_abi_version", "hgf_last_error", "hgf_plan_create", "hgf_plan_destroy", "hgf_plan_n_slots",
    "hgf_plan_slot_coupling", "hgf_plan_slot_clipping", "hgf_plan_slot_response", "hgf_plan_kernel_name",
    "hgf_filter", "hgf_logp_grad", "hgf_filter_host", "hgf_logp_grad_host", "hgf_launch_count",
    "hgf_fused_reduce_bytes", "hgf_fused_reduce_result_offset", "hgf_fused_reduce_wait", "hgf_register_specialisation", "hgf_ef_filter",
    "hgf_categorical_update", "hgf_logp_grad_reverse_noalloc",
)

_lib = None


class EfCall(C.Structure):
    """``hgf_ef_call_t``"""

    _fields_ = [
        ("dtype", C.c_int32),
        ("n_instances", C.c_int64),
        ("n_steps", C.c_int32),
        ("dimension", C.c_int32),
        ("input_group_size", C.c_int64),
        ("input_data", C.c_void_p),
        ("observed", C.c_void_p),
        ("nus", C.c_void_p),
        ("xis0", C.c_void_p),
        ("xis_trajectory", C.c_void_p),
        ("xis_last", C.c_void_p),
        ("implied", C.c_void_p),
        ("nus_trajectory", C.c_void_p),
    ]


class CategoricalCall(C.Structure):
    """``hgf_categorical_call_t``"""

    _fields_ = [
        ("dtype", C.c_int32),
        ("n_instances", C.c_int64),
        ("n_steps", C.c_int32),
        ("n_categories", C.c_int32),
        ("input_group_size", C.c_int64),
        ("observations", C.c_void_p),
        ("expected_mean", C.c_void_p),
        ("parent_mean", C.c_void_p),
        ("alpha0", C.c_void_p),
        ("alpha", C.c_void_p),
        ("mean", C.c_void_p),
        ("kl_divergence", C.c_void_p),
        ("surprise", C.c_void_p),
    ]


class BackendError(RuntimeError):
    """The CUDA library reported an error (invalid arguments or a CUDA failure)."""


def lib() -> C.CDLL:
    """Load the library (once).  Raises if it has not been built: there is no CPU path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m pyhgf_b200.build` "
            "(nvcc, sm_100a). pyhgf_b200 has no CPU fallback."
        )
    L = C.CDLL(LIB_PATH)
    L.hgf_abi_version.restype = C.c_int
    L.hgf_last_error.restype = C.c_char_p
    L.hgf_plan_create.argtypes = [C.POINTER(Topology), C.POINTER(C.c_void_p)]
    L.hgf_plan_destroy.argtypes = [C.c_void_p]
    L.hgf_plan_destroy.restype = None
    L.hgf_plan_n_slots.argtypes = [C.c_void_p]
    L.hgf_plan_n_slots.restype = C.c_int32
    L.hgf_plan_slot_coupling.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32]
    L.hgf_plan_slot_coupling.restype = C.c_int32
    L.hgf_plan_slot_clipping.argtypes = [C.c_void_p]
    L.hgf_plan_slot_clipping.restype = C.c_int32
    L.hgf_plan_slot_response.argtypes = [C.c_void_p]
    L.hgf_plan_slot_response.restype = C.c_int32
    L.hgf_plan_kernel_name.argtypes = [C.c_void_p]
    L.hgf_plan_kernel_name.restype = C.c_char_p
    for name in ("hgf_filter", "hgf_logp_grad", "hgf_logp_grad_reverse_noalloc"):
        f = getattr(L, name)
        f.argtypes = [C.c_void_p, C.POINTER(Call), C.c_void_p]
        f.restype = C.c_int
    for name in ("hgf_filter_host", "hgf_logp_grad_host"):
        f = getattr(L, name)
        f.argtypes = [C.c_void_p, C.POINTER(Call), C.c_int32]
        f.restype = C.c_int
    L.hgf_launch_count.restype = C.c_int64
    L.hgf_ef_filter.argtypes = [C.POINTER(EfCall), C.c_void_p]
    L.hgf_ef_filter.restype = C.c_int
    L.hgf_categorical_update.argtypes = [C.POINTER(CategoricalCall), C.c_void_p]
    L.hgf_categorical_update.restype = C.c_int
    L.hgf_register_specialisation.argtypes = [C.c_char_p]
    L.hgf_register_specialisation.restype = C.c_int
    L.hgf_fused_reduce_bytes.argtypes = [C.c_int32, C.c_int32]
    L.hgf_fused_reduce_bytes.restype = C.c_int64
    L.hgf_fused_reduce_result_offset.argtypes = [C.c_int32, C.c_int32, C.c_uint32]
    L.hgf_fused_reduce_result_offset.restype = C.c_int64
    L.hgf_fused_reduce_wait.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_uint32, C.c_void_p]
    L.hgf_fused_reduce_wait.restype = C.c_int
    if L.hgf_abi_version() != 5:
        raise ImportError("libpyhgf_cuda.so ABI version mismatch; rebuild with `python -m pyhgf_b200.build --force`")
    _lib = L
    return L


def check(rc: int) -> None:
    if rc != 0:
        raise BackendError(lib().hgf_last_error().decode())


def launch_count() -> int:
    return int(lib().hgf_launch_count())
