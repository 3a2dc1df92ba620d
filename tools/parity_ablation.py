"""Per-primitive ablation of the float32 math policy (csrc/hgf_math.cuh): what each knob costs and buys.

For every policy variant the 3-level binary HGF kernels (plain filter and logp+grad of omega_2, omega_3, inverse
temperature) and the 3-level continuous filter are built as on-demand specialisations with the variant's flags
(pyhgf_b200/specialise.py), run through the C ABI, and compared with the float32 CPU oracle on EVERY instance:

  * "smoke box":   4 096 instances, omega_2, omega_3 ~ U[-6, -3], the reference's binary series (what smoke() runs);
  * "config4 box": BASELINE config 4's synthetic batch (1 024 participants x 64 chains x 320 trials, bench.py);
  * "config2 box": 4 096 samples of BASELINE config 2's prior on the USD-CHF series.

    python tools/parity_ablation.py --build          # CPU box: compile every variant's object (travels with gpurun)
    python tools/parity_ablation.py [names...]       # GPU box: run, one JSON object per variant on stdout
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

#: name -> extra nvcc flags (knobs of csrc/hgf_math.cuh; -fmad is nvcc's contraction switch)
VARIANTS = {
    "sfu": [],                                                   # what round 1 shipped
    "accurate": ["-DHGF_MATH_ACCURATE", "-fmad=false"],          # the validation build
    "div1": ["-DHGF_DIV_MODE=1"],
    "div2": ["-DHGF_DIV_MODE=2"],
    "exp2": ["-DHGF_EXP_MODE=2"],
    "log1": ["-DHGF_LOG_MODE=1"],
    "log2": ["-DHGF_LOG_MODE=2"],
    "pow2": ["-DHGF_POW_MODE=2"],
    "noftz": ["-DHGF_FTZ=0"],
    "nofmad": ["-fmad=false"],
    "exp2_log1": ["-DHGF_EXP_MODE=2", "-DHGF_LOG_MODE=1"],
    "div1_exp2": ["-DHGF_DIV_MODE=1", "-DHGF_EXP_MODE=2"],
    "div1_exp2_log1": ["-DHGF_DIV_MODE=1", "-DHGF_EXP_MODE=2", "-DHGF_LOG_MODE=1"],
    "div1_exp2_log1_nofmad": ["-DHGF_DIV_MODE=1", "-DHGF_EXP_MODE=2", "-DHGF_LOG_MODE=1", "-fmad=false"],
    "div2_exp2_log2": ["-DHGF_DIV_MODE=2", "-DHGF_EXP_MODE=2", "-DHGF_LOG_MODE=2"],
    "div2_exp2_log2_pow2_nofmad": ["-DHGF_DIV_MODE=2", "-DHGF_EXP_MODE=2", "-DHGF_LOG_MODE=2", "-DHGF_POW_MODE=2", "-fmad=false"],
    "pow2_div1": ["-DHGF_POW_MODE=2", "-DHGF_DIV_MODE=1"],
    "pow2_div2": ["-DHGF_POW_MODE=2", "-DHGF_DIV_MODE=2"],
    "pow2_div2_nofmad": ["-DHGF_POW_MODE=2", "-DHGF_DIV_MODE=2", "-fmad=false"],
    "pow2_div2_log2": ["-DHGF_POW_MODE=2", "-DHGF_DIV_MODE=2", "-DHGF_LOG_MODE=2"],
    "pow2_div2_exp2": ["-DHGF_POW_MODE=2", "-DHGF_DIV_MODE=2", "-DHGF_EXP_MODE=2"],
    # last-bit arithmetic only where a probability's complement is taken (response pow + quotient, binary sigmoid)
    "resp1": ["-DHGF_RESP_MODE=1"],
    "resp2": ["-DHGF_RESP_MODE=2"],
    "resp1_div1": ["-DHGF_RESP_MODE=1", "-DHGF_DIV_MODE=1"],
    "resp2_div1": ["-DHGF_RESP_MODE=2", "-DHGF_DIV_MODE=1"],
    "resp2_exp2": ["-DHGF_RESP_MODE=2", "-DHGF_EXP_MODE=2"],
    "resp2_log1": ["-DHGF_RESP_MODE=2", "-DHGF_LOG_MODE=1"],
    "resp2_nofmad": ["-DHGF_RESP_MODE=2", "-fmad=false"],
    "resp2_div1_exp2": ["-DHGF_RESP_MODE=2", "-DHGF_DIV_MODE=1", "-DHGF_EXP_MODE=2"],
    "resp2_div1_log1": ["-DHGF_RESP_MODE=2", "-DHGF_DIV_MODE=1", "-DHGF_LOG_MODE=1"],
    "resp2_exp2_log1": ["-DHGF_RESP_MODE=2", "-DHGF_EXP_MODE=2", "-DHGF_LOG_MODE=1"],
    "resp2_div1_exp2_log1": ["-DHGF_RESP_MODE=2", "-DHGF_DIV_MODE=1", "-DHGF_EXP_MODE=2", "-DHGF_LOG_MODE=1"],
    "resp2_div1_exp2_log1_nofmad": ["-DHGF_RESP_MODE=2", "-DHGF_DIV_MODE=1", "-DHGF_EXP_MODE=2", "-DHGF_LOG_MODE=1", "-fmad=false"],
}
for _k, _v in json.loads(os.environ.get("HGF_ABLATION_EXTRA", "{}")).items():
    VARIANTS[_k] = list(_v)

KEYS = [(1, "tonic_volatility"), (2, "tonic_volatility"), ("response_parameter", 0)]


def networks():
    import nets
    from pyhgf_b200.network import Network

    return nets.binary3(Network), nets.continuous3(Network)


def build_all(names):
    from pyhgf_b200 import specialise as sp

    b3, c3 = networks()
    for name in names:
        sp.specialise(b3, float32=[(0, 0), (2, 1)], extra_flags=VARIANTS[name], load=False, verbose=False)
        sp.specialise(c3, float32=[(0, 0)], extra_flags=VARIANTS[name], load=False, verbose=False)
        print("built", name, flush=True)


def stats(err, ok):
    import numpy as np

    e = err[ok]
    return {"max": float(e.max()), "p999": float(np.quantile(e, 0.999)), "p99": float(np.quantile(e, 0.99)), "p50": float(np.median(e)),
            "n_gt_1e-4": int((e > 1e-4).sum()), "n_gt_1e-5": int((e > 1e-5).sum()), "n": int(ok.sum())}


def main(names):
    import numpy as np
    import torch

    import bench
    import nets
    from oracle import network as onet
    from pyhgf_b200 import load_data, response
    from pyhgf_b200 import specialise as sp
    from pyhgf_b200.network import Network

    u_bin, y_bin = load_data("binary")
    x_cont = load_data("continuous")
    rng = np.random.default_rng(0)
    n = int(os.environ.get("HGF_ABLATION_BOX", "4096"))  # 4 096 = smoke()'s box; larger: the same box, more samples
    p_smoke = np.stack([rng.uniform(-6, -3, n), rng.uniform(-6, -3, n), np.exp(rng.uniform(-0.3, 0.8, n))]).astype(np.float32)
    u4, y4, p4 = bench.synth_config4(0)
    rng2 = np.random.default_rng(2)
    n = 4096
    p2 = np.stack([np.clip(rng2.normal(m, 1.0, n), m - 3, m + 3) for m in (-13.0, -2.0, -2.0)]).astype(np.float32)

    # the float32 oracle, once
    o32 = nets.binary3(onet.Network).create_belief_propagation_fn()
    slots = [o32.slot(1, "tonic_volatility"), o32.slot(2, "tonic_volatility"), o32.slot(-1, "response_parameter", 0)]
    rf = "binary_softmax_inverse_temperature"
    want = {}
    want["smoke"] = (o32.batch_surprise(u_bin, slots, p_smoke.astype(np.float64), rf, resp_inputs=y_bin),
                     o32.batch_grad(u_bin, slots, p_smoke.astype(np.float64), rf, resp_inputs=y_bin, dtype=32)[1])
    s4, g4 = o32.batch_grad(u4.astype(np.float64), slots, p4.astype(np.float64), rf, resp_inputs=y4.astype(np.float64),
                            group_size=bench.N_CHAINS, dtype=32)
    want["config4"] = (s4, g4)
    oc = nets.continuous3(onet.Network).create_belief_propagation_fn()
    cslots = [oc.slot(1, "tonic_volatility"), oc.slot(2, "tonic_volatility"), oc.slot(3, "tonic_volatility")]
    want["config2"] = oc.batch_surprise(x_cont, cslots, p2.astype(np.float64), "first_level_gaussian_surprise")
    ckeys = [(1, "tonic_volatility"), (2, "tonic_volatility"), (3, "tonic_volatility")]

    dev = torch.device("cuda", 0)
    d_u4, d_y4, d_p4 = (torch.from_numpy(a).to(dev) for a in (u4, y4, p4))

    def grad_err(got, ref):
        # relative to max(|g|, 1e-3 max|g|): a gradient that crosses zero has no relative error of its own
        floor = 1e-3 * np.abs(ref[np.isfinite(ref)]).max()
        return np.abs(got - ref) / np.maximum(np.abs(ref), floor)

    dump = {"p_smoke": p_smoke, "p4": p4, "smoke/oracle_logp": want["smoke"][0], "smoke/oracle_grad": want["smoke"][1],
            "config4/oracle_logp": s4, "config4/oracle_grad": g4}
    for name in names:
        flags = VARIANTS[name]
        b3, c3 = networks()
        k3 = sp.specialise(b3, float32=[(0, 0), (2, 1)], extra_flags=flags)
        kc = sp.specialise(c3, float32=[(0, 0)], extra_flags=flags)
        out = {"variant": name, "flags": flags}
        # smoke box
        r = b3.batch(u_bin, dict(zip(KEYS, p_smoke)), response_function=response.binary_softmax_inverse_temperature,
                     response_function_inputs=y_bin, gradient=True)
        assert r.kernel_name == k3, (r.kernel_name, k3)
        for box, res, (ws, wg) in (("smoke", r, want["smoke"]),):
            got = res.surprise.double().cpu().numpy()
            dump[f"{box}/{name}/logp"] = got.astype(np.float32)
            dump[f"{box}/{name}/grad"] = np.stack([res.gradient[k].cpu().numpy() for k in KEYS])
            fin = np.isfinite(ws)
            out[box] = {"finite_pattern_equal": bool(np.array_equal(np.isfinite(got), fin)), "logp": stats(np.abs(got / ws - 1), fin & np.isfinite(got))}
            for i, k in enumerate(KEYS):
                gg = res.gradient[k].double().cpu().numpy()
                ok = fin & np.isfinite(wg[i]) & np.isfinite(gg)
                out[box][f"grad{i}"] = stats(grad_err(gg, wg[i]), ok)
        rfw = b3.batch(u_bin, dict(zip(KEYS, p_smoke)), response_function=response.binary_softmax_inverse_temperature,
                       response_function_inputs=y_bin)
        gotf = rfw.surprise.double().cpu().numpy()
        dump[f"smoke/{name}/logp_fwd"] = gotf.astype(np.float32)
        fin = np.isfinite(want["smoke"][0])
        out["smoke"]["logp_fwd_kernel"] = stats(np.abs(gotf / want["smoke"][0] - 1), fin & np.isfinite(gotf))
        # config 4 box (+ timing of the resident launch)
        r4 = b3.batch(d_u4, {k: d_p4[i] for i, k in enumerate(KEYS)}, response_function=response.binary_softmax_inverse_temperature,
                      response_function_inputs=d_y4, input_group_size=bench.N_CHAINS, response_group_size=bench.N_CHAINS, gradient=True)
        got = r4.surprise.double().cpu().numpy()
        dump[f"config4/{name}/logp"] = got.astype(np.float32)
        dump[f"config4/{name}/grad"] = np.stack([r4.gradient[k].cpu().numpy() for k in KEYS])
        ws, wg = want["config4"]
        fin = np.isfinite(ws)
        out["config4"] = {"finite_pattern_equal": bool(np.array_equal(np.isfinite(got), fin)), "logp": stats(np.abs(got / ws - 1), fin & np.isfinite(got))}
        for i, k in enumerate(KEYS):
            gg = r4.gradient[k].double().cpu().numpy()
            ok = fin & np.isfinite(wg[i]) & np.isfinite(gg)
            out["config4"][f"grad{i}"] = stats(grad_err(gg, wg[i]), ok)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        for _ in range(5):
            r4.relaunch()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(20)]
        for a, b in ev:
            flush.zero_()
            a.record(); r4.relaunch(); b.record()
        torch.cuda.synchronize()
        ms = [a.elapsed_time(b) for a, b in ev]
        out["config4"]["logp_grad_ms"] = {"mean": float(np.mean(ms)), "min": float(np.min(ms))}
        # forward-only timing on the same batch (plain filter kernel)
        rf4 = b3.batch(d_u4, {k: d_p4[i] for i, k in enumerate(KEYS)}, response_function=response.binary_softmax_inverse_temperature,
                       response_function_inputs=d_y4, input_group_size=bench.N_CHAINS, response_group_size=bench.N_CHAINS)
        gotf = rf4.surprise.double().cpu().numpy()
        dump[f"config4/{name}/logp_fwd"] = gotf.astype(np.float32)
        out["config4"]["logp_fwd_kernel"] = stats(np.abs(gotf / want["config4"][0] - 1), np.isfinite(want["config4"][0]) & np.isfinite(gotf))
        for _ in range(5):
            rf4.relaunch()
        for a, b in ev:
            flush.zero_()
            a.record(); rf4.relaunch(); b.record()
        torch.cuda.synchronize()
        ms = [a.elapsed_time(b) for a, b in ev]
        out["config4"]["forward_ms"] = {"mean": float(np.mean(ms)), "min": float(np.min(ms))}
        # config 2 box
        rc = c3.batch(x_cont, dict(zip(ckeys, p2)), response_function=response.first_level_gaussian_surprise)
        assert rc.kernel_name == kc, (rc.kernel_name, kc)
        got = rc.surprise.double().cpu().numpy()
        ws = want["config2"]
        fin = np.isfinite(ws)
        out["config2"] = {"finite_pattern_equal": bool(np.array_equal(np.isfinite(got), fin)),
                          "surprise": stats(np.abs(got - ws) / (np.abs(ws) + 1.0), fin & np.isfinite(got))}
        print(json.dumps(out), flush=True)
    if os.environ.get("HGF_ABLATION_DUMP"):
        np.savez_compressed(os.environ["HGF_ABLATION_DUMP"], **dump)


if __name__ == "__main__":
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    names = args or list(VARIANTS)
    if "--build" in sys.argv:
        build_all(names)
    else:
        main(names)
