"""Regenerate pyhgf_b200/data/datasets.npz from the reference's data files.

The three series are the reference's own fixtures (pyhgf/data/binary_input.txt,
binary_response.txt -- Iglesias et al. 2013, 320 trials; usdchf.txt -- 614
USD-CHF rates), read exactly as `pyhgf.load_data` does (pyhgf/__init__.py:14-66).
Run in the authoring container only (the GPU box has no /root/reference).
"""
import os

import numpy as np

REF = "/root/reference/pyhgf/data"
OUT = os.path.join(os.path.dirname(__file__), "..", "..", "pyhgf_b200", "data", "datasets.npz")

if __name__ == "__main__":
    u = np.loadtxt(os.path.join(REF, "binary_input.txt"))
    y = np.loadtxt(os.path.join(REF, "binary_response.txt"))
    x = np.loadtxt(os.path.join(REF, "usdchf.txt"))
    assert u.shape == (320,) and y.shape == (320,) and x.shape == (614,)
    np.savez_compressed(OUT, binary_u=u, binary_y=y, continuous=x)
    print("wrote", os.path.abspath(OUT), u.sum(), y.sum(), x[:3])
