"""Build libpyhgf_cuda.so (sm_100a) in-tree with nvcc.

One kernel per translation unit, compiled in parallel; objects go to
``build/obj`` (ignored), the library to ``pyhgf_b200/lib/libpyhgf_cuda.so`` (git-
ignored, but it travels to the GPU box with the snapshot).  No JIT at import.

    python -m pyhgf_b200.build [--force] [-j N]
"""

from __future__ import annotations

import argparse
import concurrent.futures as cf
import hashlib
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "pyhgf_b200", "csrc")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
#: "sfu" (default): MUFU-based float32 division/exp/log, FMA contraction; "accurate": IEEE division, libdevice
#: transcendentals, one rounding per primitive (the validation build) -- see csrc/hgf_math.cuh
MATH = os.environ.get("HGF_MATH", "sfu")
if MATH not in ("sfu", "accurate"):
    raise SystemExit("HGF_MATH must be 'sfu' or 'accurate'")
OBJ = os.path.join(ROOT, "build", "obj" if MATH == "sfu" else "obj_accurate")
LIB = os.path.join(ROOT, "pyhgf_b200", "lib", "libpyhgf_cuda.so" if MATH == "sfu" else "libpyhgf_cuda_accurate.so")
COMMON = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O3", "-lineinfo",
    "-Xcompiler", "-fPIC", "-diag-suppress", "221", "-I", CSRC, "-I", os.path.join(ROOT, "include"),
] + (["-DHGF_MATH_ACCURATE", "-fmad=false"] if MATH == "accurate" else [])

# interpreter kernels: (tag, scalar type)
INTERP_SCALARS = [
    ("f32", "float", 0), ("f32d1", "float", 1), ("f32d2", "float", 2), ("f32d3", "float", 3),
    ("f32d4", "float", 4), ("f64", "double", 0), ("f64d4", "double", 4),
]


def _headers():
    return [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".h", ".cuh", ".inc"))] + [
        os.path.join(ROOT, "include", "pyhgf_cuda.h")
    ]


def jobs():
    """(source, object, extra flags) of every translation unit."""
    out = []
    for name in ("hgf_capi", "hgf_select", "hgf_interp", "hgf_static", "hgf_ef", "hgf_categorical", "hgf_fused", "hgf_noalloc"):
        out.append((os.path.join(CSRC, name + ".cu"), os.path.join(OBJ, name + ".o"), []))
    for tag, base, ntan in INTERP_SCALARS:
        out.append((os.path.join(CSRC, "hgf_interp_inst.cu"), os.path.join(OBJ, f"hgf_interp_{tag}.o"),
                    [f"-DHGF_BASE={base}", f"-DHGF_NTAN={ntan}"]))
    gen = os.path.join(CSRC, "generated")
    if os.path.isdir(gen):
        manifest = os.path.join(gen, "instances.txt")
        if os.path.exists(manifest):
            for line in open(manifest):
                line = line.strip()
                if not line or line.startswith("#"):
                    continue
                topo, base, ntan, nresp = line.split()
                out.append((os.path.join(CSRC, "hgf_static_inst.cu"), os.path.join(OBJ, f"hgf_static_{topo}_{base}_d{ntan}r{nresp}.o"),
                            [f"-DHGF_TOPO_HEADER=<generated/topo_{topo}.h>", f"-DHGF_TOPO_TYPE=Topo_{topo}",
                             f"-DHGF_BASE={base}", f"-DHGF_NTAN={ntan}", f"-DHGF_NRESP={nresp}"] + os.environ.get("HGF_EXTRA_STATIC_FLAGS", "").split()))
        rev = os.path.join(gen, "reverse.txt")
        if os.path.exists(rev):
            for name in open(rev).read().split():
                out.append((os.path.join(CSRC, "hgf_reverse_inst.cu"), os.path.join(OBJ, f"hgf_reverse_{name}.o"),
                            [f"-DHGF_REV_HEADER=<generated/reverse_{name}.h>", f"-DHGF_REV_TYPE=Rev_{name}", f"-DHGF_REV_FN=reverse_{name}"]))
    return out


def _stamp(src, flags, dep_hash):
    """Content stamp of one translation unit: flags + source + every header (mtimes do not survive the copy
    to the GPU box, so staleness is decided on content)."""
    h = hashlib.sha1(" ".join(COMMON + flags).encode())
    h.update(open(src, "rb").read())
    h.update(dep_hash.encode())
    return h.hexdigest()


def _stale(src, obj, flags, dep_hash):
    if not os.path.exists(obj):
        return True
    stamp = obj + ".flags"
    return not os.path.exists(stamp) or open(stamp).read() != _stamp(src, flags, dep_hash)


def _compile(job, verbose, dep_hash):
    src, obj, flags = job
    t0 = time.time()
    cmd = [NVCC] + COMMON + flags + ["-c", src, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {os.path.basename(obj)}:\n{r.stdout}\n{r.stderr}")
    with open(obj + ".flags", "w") as f:
        f.write(_stamp(src, flags, dep_hash))
    if verbose:
        print(f"  [{time.time() - t0:6.1f}s] {os.path.basename(obj)}", flush=True)
    return obj


def build(force: bool = False, n_jobs: int | None = None, verbose: bool = True) -> str:
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    gen_script = os.path.join(ROOT, "pyhgf_b200", "codegen.py")
    if os.path.exists(gen_script):
        subprocess.run([sys.executable, "-m", "pyhgf_b200.codegen"], check=True, cwd=ROOT)
    deps = _headers()
    gen = os.path.join(CSRC, "generated")
    if os.path.isdir(gen):
        deps += [os.path.join(gen, f) for f in sorted(os.listdir(gen))]
    dh = hashlib.sha1()
    for d in deps:
        dh.update(open(d, "rb").read())
    dep_hash = dh.hexdigest()
    all_jobs = jobs()
    todo = [j for j in all_jobs if force or _stale(j[0], j[1], j[2], dep_hash)]
    if verbose:
        print(f"pyhgf_b200.build: {len(todo)} of {len(all_jobs)} translation units to compile", flush=True)
    n_jobs = n_jobs or os.cpu_count() or 4
    if todo:
        with cf.ThreadPoolExecutor(max_workers=n_jobs) as ex:
            list(ex.map(lambda j: _compile(j, verbose, dep_hash), todo))
    objs = [j[1] for j in all_jobs]
    link_stamp = LIB + ".stamp"
    want_link = hashlib.sha1("".join(open(o + ".flags").read() for o in objs).encode()).hexdigest()
    if todo or not os.path.exists(LIB) or not os.path.exists(link_stamp) or open(link_stamp).read() != want_link:
        cmd = [NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + objs
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
        with open(link_stamp, "w") as f:
            f.write(want_link)
        if verbose:
            print(f"pyhgf_b200.build: linked {LIB}", flush=True)
    return LIB


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("-j", type=int, default=None)
    a = ap.parse_args()
    build(force=a.force, n_jobs=a.j)
