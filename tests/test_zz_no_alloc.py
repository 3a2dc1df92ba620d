"""hgf_logp_grad_reverse_noalloc (include/pyhgf_cuda.h): the reverse sweep of the 3-level binary HGF with its checkpoint
array in the library's static device scratch instead of cudaMallocAsync memory.  Called through the C ABI on the SAME
call structure as the existing entry point (hgf_logp_grad, gradient_mode = reverse) and compared with its outputs, twice
in a row (the second call finds the first call's checkpoints in the shared scratch).

Collected last (test_zz_*): a failure here must not hide the parity tests.
"""

import ctypes as C

import numpy as np
import pytest

import nets

pytestmark = pytest.mark.gpu

# float32: the suite's bound on summed logp and on gradients (tests/test_gpu_parity.py: RTOL_SUM); the two entry points run
# the same per-instance code, so this is generous
RTOL = 1e-4

KEYS = [(1, "tonic_volatility"), (2, "tonic_volatility"), ("response_parameter", 0), (1, "mean"), (2, "precision")]


def _params(n, seed):
    rng = np.random.default_rng(seed)
    return [rng.uniform(-6, -3.5, n), rng.uniform(-6, -3.5, n), np.exp(rng.uniform(-0.3, 0.8, n)), rng.uniform(-0.5, 0.5, n),
            np.exp(rng.uniform(-0.5, 0.5, n))]


@pytest.fixture(scope="module")
def G():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from pyhgf_b200 import _lib

    _lib.lib()  # no fallback: the CUDA library must be there
    import pyhgf_b200.network as g

    return g


def _existing(G, datasets, n, seed, n_chains=8):
    """One evaluation through the existing entry point; returns the result (which keeps the call structure and the device
    buffers, BatchResult.relaunch) and copies of its outputs."""
    from pyhgf_b200 import response

    u, y = datasets["u"], datasets["y"]
    net = nets.binary3(G.Network, dtype="float32")
    res = net.batch(u, dict(zip(KEYS, _params(n, seed))), response_function=response.binary_softmax_inverse_temperature,
                    response_function_inputs=y, gradient="reverse", group=np.arange(n) % n_chains, n_groups=n_chains)
    assert res.kernel_name == "static:binary3"
    assert res._call.gradient_mode == 2
    return res, res.surprise.clone(), res._grad.clone(), res.reduced.clone()


def _close(got, want, what):
    got, want = got.double().cpu().numpy(), want.double().cpu().numpy()
    assert np.isfinite(want).all(), what
    scale = np.abs(want).max(axis=-1, keepdims=True)
    assert (np.abs(got - want) <= RTOL * np.maximum(np.abs(want), 1e-3 * scale)).all(), (what, float(np.abs(got - want).max()))


def test_noalloc_reverse_sweep_matches_existing_entry_point_twice(G, datasets):
    import torch

    from pyhgf_b200 import _lib

    L = _lib.lib()
    dev = torch.device("cuda", 0)
    stream = torch.cuda.current_stream(dev)
    free_after_first = None
    for call_no, (n, seed) in enumerate([(4096, 11), (1000, 12)]):  # a full batch, then a ragged one over the used scratch
        res, want_s, want_g, want_r = _existing(G, datasets, n, seed)
        # the outputs are the call's own buffers: poison them so that only the new entry point can have filled them
        res.surprise.fill_(float("nan"))
        res._grad.fill_(float("nan"))
        res.reduced.zero_()  # (accumulated into: the caller zeroes it, include/pyhgf_cuda.h)
        torch.cuda.synchronize(dev)
        before = _lib.launch_count()
        with torch.cuda.device(dev):
            _lib.check(L.hgf_logp_grad_reverse_noalloc(res._plan.handle, C.byref(res._call), C.c_void_p(stream.cuda_stream)))
        torch.cuda.synchronize(dev)
        assert _lib.launch_count() - before == 1
        _close(res.surprise, want_s, f"surprise_sum, call {call_no}")
        _close(res._grad, want_g, f"gradient, call {call_no}")
        _close(res.reduced, want_r, f"reduced, call {call_no}")
        # nothing was allocated on the device by the second call (the first one may load the module that holds the scratch)
        free = torch.cuda.mem_get_info(dev)[0]
        if call_no == 0:
            free_after_first = free
            res0 = res
    # the first call once more, on the scratch the ragged call left behind
    _, want_s, want_g, want_r = _existing(G, datasets, 4096, 11)
    res0.surprise.fill_(float("nan"))
    res0._grad.fill_(float("nan"))
    res0.reduced.zero_()
    torch.cuda.synchronize(dev)
    free_before = torch.cuda.mem_get_info(dev)[0]
    with torch.cuda.device(dev):
        _lib.check(L.hgf_logp_grad_reverse_noalloc(res0._plan.handle, C.byref(res0._call), C.c_void_p(stream.cuda_stream)))
    torch.cuda.synchronize(dev)
    assert torch.cuda.mem_get_info(dev)[0] == free_before, "the call changed the device's free memory"
    assert free_after_first is not None
    _close(res0.surprise, want_s, "surprise_sum, repeated call")
    _close(res0._grad, want_g, "gradient, repeated call")
    _close(res0.reduced, want_r, "reduced, repeated call")


def test_noalloc_rejects_what_it_cannot_run(G, datasets):
    import torch

    from pyhgf_b200 import _lib, response

    L = _lib.lib()
    dev = torch.device("cuda", 0)
    s = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    max_cells = 2097152  # HGF_NOALLOC_MAX_STEP_INSTANCES
    T = len(datasets["u"])
    res, *_ = _existing(G, datasets, 64, 3)
    call = res._call
    # beyond the static scratch: refused before anything is launched (the pointers are never read)
    call.n_instances = max_cells // T + 1
    before = _lib.launch_count()
    with pytest.raises(_lib.BackendError, match="HGF_NOALLOC_MAX_STEP_INSTANCES"):
        _lib.check(L.hgf_logp_grad_reverse_noalloc(res._plan.handle, C.byref(call), s))
    assert _lib.launch_count() == before
    call.n_instances = 64
    # float64 and other topologies have no static-scratch variant
    r64 = nets.binary3(G.Network, dtype="float64").batch(
        datasets["u"], dict(zip(KEYS, _params(16, 4))), response_function=response.binary_softmax_inverse_temperature,
        response_function_inputs=datasets["y"], gradient="reverse")
    with pytest.raises(_lib.BackendError, match="float32 only"):
        _lib.check(L.hgf_logp_grad_reverse_noalloc(r64._plan.handle, C.byref(r64._call), s))
    rc = nets.continuous3(G.Network, dtype="float32").batch(
        datasets["x"][:100], {(1, "tonic_volatility"): np.full(16, -10.0)}, response_function=response.first_level_gaussian_surprise,
        gradient="reverse")
    with pytest.raises(_lib.BackendError, match="static:binary3"):
        _lib.check(L.hgf_logp_grad_reverse_noalloc(rc._plan.handle, C.byref(rc._call), s))
    torch.cuda.synchronize(dev)
