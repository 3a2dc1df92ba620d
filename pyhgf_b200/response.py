"""Response (surprise) functions -- same names and call signature as ``pyhgf/response.py:15-215``.

``f(hgf, response_function_inputs=None, response_function_parameters=None) -> array(T,)``.
Each is a marker for the reduction fused into the CUDA time loop
(``response_step`` in csrc/hgf_ops.cuh): called on a filtered ``Network`` it
returns the per-trial surprise of that network's series; passed to
``Network.batch`` / ``HGFDistribution`` it selects the in-kernel reduction.
"""

from __future__ import annotations


def _fused(name: str, doc: str):
    def fn(hgf, response_function_inputs=None, response_function_parameters=None):
        return hgf._fused_response(name, response_function_inputs, response_function_parameters)

    fn.__name__ = fn.__qualname__ = name
    fn.__doc__ = doc
    fn._hgf_response = name
    return fn


first_level_gaussian_surprise = _fused(
    "first_level_gaussian_surprise",
    "Gaussian surprise of node 0's observations under node 1's prediction; NaN -> inf (response.py:15-48).",
)
total_gaussian_surprise = _fused(
    "total_gaussian_surprise",
    "Sum of the Gaussian surprises of every continuous-state node; inf everywhere if node 1's mean "
    "was ever NaN (response.py:51-89).",
)
first_level_binary_surprise = _fused(
    "first_level_binary_surprise",
    "Binary surprise of the observations under the input nodes' predictions; NaN -> inf (response.py:92-130).",
)
binary_softmax = _fused(
    "binary_softmax",
    "Surprise of the decisions `response_function_inputs` under node 0's expected mean (response.py:133-164).",
)
binary_softmax_inverse_temperature = _fused(
    "binary_softmax_inverse_temperature",
    "As binary_softmax with the belief sharpened by the inverse temperature "
    "`response_function_parameters`: mu^t / (mu^t + (1-mu)^t) (response.py:167-215).",
)

__all__ = [
    "first_level_gaussian_surprise", "total_gaussian_surprise", "first_level_binary_surprise", "binary_softmax",
    "binary_softmax_inverse_temperature",
]
