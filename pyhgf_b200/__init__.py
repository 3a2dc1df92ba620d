"""pyhgf_b200 -- the batched belief-propagation path of pyhgf on B200 (sm_100a).

Drop-in surface for that path: ``Network`` (``add_nodes``, ``input_data``,
``surprise``), the response functions, ``HGFDistribution`` / ``HGFPointwise`` /
``logp``; all arithmetic runs in ``libpyhgf_cuda.so`` (include/pyhgf_cuda.h).
"""

import os as _os

import numpy as _np

__version__ = "0.1.0"


def load_data(dataset: str):
    """The reference's example series (pyhgf/__init__.py:14-66): ``"continuous"`` -> USD-CHF (614,),
    ``"binary"`` -> ``(u, y)`` of Iglesias et al. (320,) each."""
    d = _np.load(_os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "data", "datasets.npz"))
    if dataset == "continuous":
        return d["continuous"]
    if dataset == "binary":
        return d["binary_u"], d["binary_y"]
    raise ValueError("Invalid dataset argument. Should be 'continous' or 'binary'.")


def __getattr__(name):  # lazy: importing the package must not need torch / the built library
    if name in ("Network", "BatchResult"):
        from . import network

        return getattr(network, name)
    if name in ("HGFDistribution", "HGFPointwise", "logp", "hgf_logp"):
        from . import distribution

        return getattr(distribution, name)
    raise AttributeError(name)
