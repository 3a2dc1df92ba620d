"""Driver of the host numerics test-bed (tools/hostsim/hostsim.cpp): builds the device equations for the host with a
math policy's flags and compares the result with the float32 oracle on the smoke box / config-4 box, like
tools/parity_ablation.py does on the GPU -- minutes of CPU instead of a GPU call.  Development tool.

    python tools/hostsim/run.py NAME=-DFLAG,-DFLAG ...     e.g.  sfu=  resp2=-DHGF_RESP_MODE=2  ieee=@
(`@` alone builds without the SFU emulation: IEEE / libm primitives.)
"""
import ctypes as C
import hashlib
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
WORK = os.path.join(ROOT, "build", "hostsim")


def build(flags):
    os.makedirs(WORK, exist_ok=True)
    emulate = "@" not in flags
    flags = [f for f in flags if f and f != "@"]
    src = os.path.join(ROOT, "tools", "hostsim", "hostsim.cpp")
    h = hashlib.sha1(repr((flags, emulate)).encode())
    csrc = os.path.join(ROOT, "pyhgf_b200", "csrc")
    deps = [os.path.join(csrc, x) for x in ("hgf_math.cuh", "hgf_ops.cuh", "hgf_scan.cuh", "hgf_program.h", "hgf_reverse.cuh")]
    deps += sorted(os.path.join(csrc, "generated", x) for x in os.listdir(os.path.join(csrc, "generated")) if x.endswith(".h"))
    for f in [src] + deps:
        h.update(open(f, "rb").read())
    so = os.path.join(WORK, f"sim_{h.hexdigest()[:12]}.so")
    if not os.path.exists(so):
        cmd = ["g++", "-O2", "-ffp-contract=off", "-fopenmp", "-shared", "-fPIC", "-Wno-attributes", "-Wno-unused-variable", "-I", os.path.join(ROOT, "pyhgf_b200", "csrc"),
               "-I", "/usr/local/cuda/include"] + (["-DHGF_HOST_EMULATE_SFU"] if emulate else []) + flags + [src, "-o", so]
        subprocess.run(cmd, check=True)
    return C.CDLL(so)


def fp(a):
    return a.ctypes.data_as(C.c_void_p)


class Box:
    def __init__(self, name, u, y, params, group):
        import nets
        from oracle import network as onet
        from pyhgf_b200.network import Network

        self.name, self.group = name, group
        self.u = np.ascontiguousarray(u, dtype=np.float32)
        self.y = np.ascontiguousarray(y, dtype=np.float32)
        self.p = np.ascontiguousarray(params, dtype=np.float32)
        net = nets.binary3(Network)
        net.create_belief_propagation_fn()
        plan = net._get_plan(1)
        self.attrs = np.ascontiguousarray(net._flatten(plan), dtype=np.float32)
        self.slots = np.array([plan.slot(1, "tonic_volatility"), plan.slot(2, "tonic_volatility"), plan.slot(-1, "response_parameter", 0)], dtype=np.int32)
        o32 = nets.binary3(onet.Network).create_belief_propagation_fn()
        os_ = [o32.slot(1, "tonic_volatility"), o32.slot(2, "tonic_volatility"), o32.slot(-1, "response_parameter", 0)]
        self.want, self.want_g = o32.batch_grad(self.u.astype(np.float64), os_, self.p.astype(np.float64), "binary_softmax_inverse_temperature",
                                                resp_inputs=self.y.astype(np.float64), group_size=group, dtype=32)

    def run(self, lib, entry="sim_binary3_grad"):
        n = self.p.shape[1]
        T = self.u.shape[-1]
        logp = np.zeros(n, dtype=np.float32)
        grad = np.zeros((3, n), dtype=np.float32)
        getattr(lib, entry)(n, T, self.group, fp(self.u), fp(self.y), fp(self.attrs), fp(self.slots), fp(self.p), fp(logp), fp(grad))
        return logp.astype(np.float64), grad.astype(np.float64)


def boxes(which=("smoke", "config4")):
    import bench
    from pyhgf_b200 import load_data

    out = []
    if "smoke" in which:
        u, y = load_data("binary")
        rng = np.random.default_rng(0)
        n = 4096
        p = np.stack([rng.uniform(-6, -3, n), rng.uniform(-6, -3, n), np.exp(rng.uniform(-0.3, 0.8, n))])
        out.append(Box("smoke", u, y, p, 0))
    if "config4" in which:
        u4, y4, p4 = bench.synth_config4(0, n_participants=int(os.environ.get("HOSTSIM_PARTICIPANTS", "256")))
        out.append(Box("config4", u4, y4, p4, bench.N_CHAINS))
    return out


def report(name, box, logp, grad):
    fin = np.isfinite(box.want) & np.isfinite(logp)
    e = np.abs(logp / box.want - 1)[fin]
    line = f"{name:28s} {box.name:8s} logp max {e.max():.2e} n>1e-4 {int((e > 1e-4).sum()):5d} n>1e-5 {int((e > 1e-5).sum()):5d}  finite== {np.array_equal(np.isfinite(logp), np.isfinite(box.want))}"
    for i in range(3):
        ref = box.want_g[i]
        ok = fin & np.isfinite(ref) & np.isfinite(grad[i])
        floor = 1e-3 * np.abs(ref[ok]).max()
        ge = (np.abs(grad[i] - ref) / np.maximum(np.abs(ref), floor))[ok]
        line += f" | g{i} {ge.max():.1e} {int((ge > 1e-4).sum()):4d}"
    print(line, flush=True)


if __name__ == "__main__":
    bx = boxes()
    for arg in sys.argv[1:]:
        name, _, fl = arg.partition("=")
        lib = build(fl.split(","))
        for b in bx:
            report(name, b, *b.run(lib))
