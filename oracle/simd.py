"""TEST / BASELINE INFRASTRUCTURE ONLY -- Python side of oracle/hgf_simd.cpp, the instance-vectorised float32 CPU arm.

The reference's CPU path is ``jit(vmap(...))`` on XLA-CPU: float32, vectorised across the batch, reverse-mode gradients.
The scalar oracle (hgf_oracle.cpp) is the parity checker; THIS is the CPU implementation bench.py times beside the GPU
(`cpu_baseline`, `--impl reference`): 16 (AVX-512) or 8 (AVX2) instances per SIMD vector, libmvec, OpenMP, and the generated
reverse sweep for logp + gradient.  It covers the networks of the BASELINE configurations (the 3-level binary, the 4-node
continuous and the 16-node generalised HGF; default "unbounded" scheme, unit or given time steps, every observation present).

Only tests/ and bench.py's CPU legs may import it.
"""

from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

from . import network as onet

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)


class _Call(ctypes.Structure):
    _fields_ = [
        ("n_instances", ctypes.c_longlong), ("n_steps", ctypes.c_int),
        ("input_group_size", ctypes.c_longlong), ("response_group_size", ctypes.c_longlong),
        ("input_data", ctypes.c_void_p), ("time_steps", ctypes.c_void_p), ("response_inputs", ctypes.c_void_p),
        ("attributes", ctypes.c_void_p), ("n_params", ctypes.c_int), ("param_slots", ctypes.c_void_p), ("params", ctypes.c_void_p),
        ("response_fn", ctypes.c_int), ("max_posterior_precision", ctypes.c_float),
        ("surprise", ctypes.c_void_p), ("gradient", ctypes.c_void_p),
    ]


def _isa() -> tuple:
    """(tag, compiler flags) of the widest vector ISA this machine's CPU has (the library is built where it runs)."""
    flags = set()
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("flags"):
                flags = set(line.split(":", 1)[1].split())
                break
    except OSError:
        pass
    if {"avx512f", "avx512vl", "avx512dq", "avx512bw"} <= flags:
        return "avx512", ["-mavx512f", "-mavx512vl", "-mavx512dq", "-mavx512bw", "-mavx2", "-mfma"]
    if {"avx2", "fma"} <= flags:
        return "avx2", ["-mavx2", "-mfma"]
    raise RuntimeError("the instance-vectorised CPU arm needs AVX2 or AVX-512")


def library_path() -> str:
    return os.path.join(_HERE, f"libhgf_simd_{_isa()[0]}.so")


def build_library(force: bool = False) -> str:
    path = library_path()
    src = os.path.join(_HERE, "hgf_simd.cpp")
    gen = os.path.join(_ROOT, "pyhgf_b200", "csrc", "generated")
    deps = [src] + [os.path.join(gen, f) for f in ("reverse_binary3.h", "reverse_continuous3.h", "reverse_ghgf16.h")]
    if force or not os.path.exists(path) or os.path.getmtime(path) < max(os.path.getmtime(d) for d in deps):
        cmd = ["g++", "-O3", "-std=c++17", "-fopenmp", "-shared", "-fPIC", "-Wno-attributes", "-Wno-unused-variable", "-Wno-ignored-attributes",
               *_isa()[1], "-I", os.path.join(_ROOT, "pyhgf_b200", "csrc"), "-I", "/usr/local/cuda/include", src, "-o", path, "-lmvec", "-lm"]
        subprocess.run(cmd, check=True, cwd=_HERE)
    return path


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        L = ctypes.CDLL(build_library())
        L.hgf_simd_batch.argtypes = [ctypes.c_int, ctypes.POINTER(_Call), ctypes.c_int]
        L.hgf_simd_n_slots.argtypes = [ctypes.c_int]
        _lib = L
    return _lib


def width() -> int:
    return int(lib().hgf_simd_width())


def _topology(net: "onet.Network") -> int:
    """0: the 3-level binary HGF, 1: the 4-node continuous HGF, 2: the 16-node generalised HGF -- recognised by their edges."""
    kinds = [e.node_type for e in net.edges]
    if net.volatility_updates != "unbounded" or net.mean_field_updates:
        raise NotImplementedError("the SIMD CPU arm covers the default scheme")
    e = net.edges
    if kinds == [1, 2, 2] and e[1].value_children == (0,) and e[2].volatility_children == (1,):
        return 0
    if kinds == [2, 2, 2, 2] and e[1].value_children == (0,) and e[2].volatility_children == (1,) and e[3].volatility_children == (2,):
        return 1
    if kinds == [2] * 12 + [6] * 4 and all(e[4 + i].value_children == (i,) and e[8 + i].volatility_children == (4 + i,) for i in range(4)) \
            and [e[i].value_children for i in range(12, 16)] == [(4, 5), (6, 7), (12,), (13,)]:
        return 2
    raise NotImplementedError("the SIMD CPU arm covers the 3-level binary, the 4-node continuous and the 16-node generalised HGF")


def batch(net: "onet.Network", values, param_slots, params, response: str, resp_inputs=None, response_parameters=None,
          group_size: int = 0, response_group_size: int | None = None, time_steps=None, gradient: bool = False, n_threads: int = 0):
    """``params [P][N] -> surprise_sum [N]`` (and ``grad [P][N]`` with ``gradient=True``), float32; arguments as
    ``oracle.network.Network.batch_surprise / batch_grad``."""
    if net._handle is None:
        net.create_belief_propagation_fn()
    topo = _topology(net)
    n_in = len(net.input_idxs)
    v = np.asarray(values, dtype=np.float32)
    if v.ndim == 1:
        v = v[None, :, None]
    elif v.ndim == 2:
        v = v[None, :, :] if v.shape[1] == n_in and group_size == 0 else v[:, :, None]
    v = np.ascontiguousarray(v)
    params = np.ascontiguousarray(np.atleast_2d(np.asarray(params, dtype=np.float32)))
    n_params, n_inst = params.shape
    save = net.dtype
    net.dtype = 64
    slots0 = np.ascontiguousarray(net.flatten(response_parameters=response_parameters).astype(np.float32))
    net.dtype = save
    assert slots0.size == lib().hgf_simd_n_slots(topo), "slot layout out of sync with the generated step"
    ps = np.ascontiguousarray(np.asarray(list(param_slots) or [0], dtype=np.int32))
    ri = None if resp_inputs is None else np.ascontiguousarray(np.atleast_2d(np.asarray(resp_inputs, dtype=np.float32)))
    ts = None if time_steps is None else np.ascontiguousarray(np.asarray(time_steps, dtype=np.float32))
    out = np.zeros(n_inst, dtype=np.float32)
    grad = np.zeros((n_params, n_inst), dtype=np.float32) if gradient else None
    c = _Call()
    c.n_instances, c.n_steps = n_inst, v.shape[1]
    c.input_group_size = int(group_size)
    c.response_group_size = int(group_size if response_group_size is None else response_group_size)
    c.input_data = v.ctypes.data
    c.time_steps = None if ts is None else ts.ctypes.data
    c.response_inputs = None if ri is None else ri.ctypes.data
    c.attributes = slots0.ctypes.data
    c.n_params, c.param_slots, c.params = len(list(param_slots)), ps.ctypes.data, params.ctypes.data
    c.response_fn = 0 if response is None else onet.RESPONSES[response]  # None: the filter alone
    c.max_posterior_precision = float(net.max_posterior_precision)
    c.surprise = out.ctypes.data
    c.gradient = None if grad is None else grad.ctypes.data
    rc = lib().hgf_simd_batch(topo, ctypes.byref(c), int(n_threads))
    if rc != 0:
        raise RuntimeError(f"hgf_simd_batch failed ({rc})")
    return (out, grad) if gradient else out
