"""Names of the update steps.

In the reference each schedule entry is ``(node_idx, python_function)``
(pyhgf/updates/**); here the arithmetic lives in CUDA
(pyhgf_b200/csrc/hgf_ops.cuh), so a step is a named marker carrying the opcode
the kernels dispatch on.  ``__name__`` is the reference function's name, so
code that inspects ``update_sequence`` keeps working.
"""

from __future__ import annotations

from ._lib import OPS


class UpdateStep:
    __slots__ = ("__name__", "opcode", "source")

    def __init__(self, name: str, opcode: int, source: str):
        self.__name__ = name
        self.opcode = opcode
        self.source = source

    def __repr__(self):
        return f"<cuda update step {self.__name__} (op {self.opcode})>"

    def __call__(self, *args, **kwargs):
        raise TypeError(
            f"{self.__name__} is executed inside the CUDA time loop (hgf_ops.cuh); it is not callable from Python"
        )


_TABLE = [
    ("binary_state_node_prediction", "PRED_BINARY", "pyhgf/updates/prediction/binary.py:12-93"),
    ("continuous_node_prediction", "PRED_CONT", "pyhgf/updates/prediction/continuous.py:282-366"),
    ("continuous_node_prediction_mean_field", "PRED_CONT_MF", "pyhgf/updates/prediction/continuous.py:455-512"),
    ("volatile_node_prediction", "PRED_VOLATILE", "pyhgf/updates/prediction/volatile.py:230-309"),
    ("volatile_node_prediction_mean_field", "PRED_VOLATILE_MF", "pyhgf/updates/prediction/volatile.py:312-429"),
    ("binary_state_node_prediction_error", "PE_BINARY", "pyhgf/updates/prediction_error/binary.py:11-51"),
    ("continuous_node_prediction_error", "PE_CONT", "pyhgf/updates/prediction_error/continuous.py:123-176"),
    ("volatile_node_prediction_error", "PE_VOLATILE", "pyhgf/updates/prediction_error/volatile.py:85-154"),
    ("continuous_node_posterior_update", "POST_CONT",
     "pyhgf/updates/posterior/continuous/continuous_node_posterior_update.py:20-70"),
    ("continuous_node_posterior_update_mean_field", "POST_CONT_MF",
     "pyhgf/updates/posterior/continuous/continuous_node_posterior_update.py:73-128"),
    ("continuous_node_posterior_update_ehgf", "POST_CONT_EHGF",
     "pyhgf/updates/posterior/continuous/continuous_node_posterior_update_ehgf.py:20-83"),
    ("continuous_node_posterior_update_ehgf_mean_field", "POST_CONT_EHGF_MF",
     "pyhgf/updates/posterior/continuous/continuous_node_posterior_update_ehgf.py:86-155"),
    ("continuous_node_posterior_update_unbounded", "POST_CONT_UNBOUNDED",
     "pyhgf/updates/posterior/continuous/continuous_node_posterior_update_unbounded.py:13-48"),
    ("volatile_node_posterior_update", "POST_VOLATILE",
     "pyhgf/updates/posterior/volatile/volatile_node_posterior_update.py:23-72"),
    ("volatile_node_posterior_update_mean_field", "POST_VOLATILE_MF",
     "pyhgf/updates/posterior/volatile/volatile_node_posterior_update.py:110-160"),
]

UPDATE_FUNCTIONS = {name: UpdateStep(name, OPS[op], src) for name, op, src in _TABLE}
globals().update(UPDATE_FUNCTIONS)
__all__ = list(UPDATE_FUNCTIONS) + ["UPDATE_FUNCTIONS", "UpdateStep"]
