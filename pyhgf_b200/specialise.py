"""Register-resident kernels for ANY topology, built on demand.

The BASELINE topologies get their generated specialisations at library build time
(``codegen.py`` -> ``csrc/generated``); every other network runs through the opcode
interpreter, whose per-thread attribute vector lives in local memory (2-6x slower,
profiles/r1_configs_interpreter.log).  ``specialise(network)`` closes that gap for a
user network: it generates the same ``Topo_<hash>`` accessor for the network's own edges
and schedule, compiles the kernel template for it with nvcc (sm_100a, the library's own
headers and flags, one translation unit per scalar type, in parallel), links a small shared
object next to the library and registers it (``hgf_register_specialisation``).  Plans
created afterwards dispatch to it; the object is cached on disk by content hash, so the
cost (tens of seconds for a 16-node network) is paid once per topology and build.

This is the reference's "trace once, run many" (``jax.jit`` of ``lax.scan``,
pyhgf/model/network.py:182-233, 612-624) done ahead of time with hand-written CUDA: there
is no tracing and no code generation beyond the table of compile-time topology answers.
"""

from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import subprocess
from typing import Iterable, Sequence

from . import _lib, build, codegen

ROOT = build.ROOT
OUT_DIR = os.path.join(ROOT, "pyhgf_b200", "lib", "specialised")
WORK_DIR = os.path.join(ROOT, "build", "specialised")
_registered: dict = {}


def _variant_bit(pf: int, pr: int) -> int:
    return 1 << (pf + 5 * pr)


def header_for(network, salt: str = "") -> tuple:
    """(name, header text) of the network's topology accessor; the name is a content hash (salted with any extra
    compiler flags, so that differently-built kernels of one topology never share a symbol name)."""
    inputs = network.input_idxs
    body = codegen.generate_header("@", network, network.volatility_updates, network.mean_field_updates, inputs=inputs)
    name = "h" + hashlib.sha1((body + salt).encode()).hexdigest()[:14]
    return name, body.replace("Topo_@", "Topo_" + name).replace('"@"', f'"{name}"')


def specialise(network, float32: Iterable[Sequence[int]] = ((0, 0),), float64: Iterable[Sequence[int]] = (),
               verbose: bool = False, extra_flags: Sequence[str] = (), load: bool = True, reverse: bool = False) -> str:
    """Build (or load from the cache) and register the specialisation of ``network``'s topology.

    float32 / float64 : the kernels to instantiate, as ``(PF, PR)`` pairs -- PF filter-parameter tangent
        directions (0 = the plain filter, up to 4 per pass), PR = 1 adds a response-parameter tangent.
        A launch that needs a kernel that was not instantiated runs through the interpreter.
    extra_flags : further nvcc flags for the kernel translation units (e.g. a math policy, ``-DHGF_EXP_MODE=2``:
        tools/parity_ablation.py builds one object per policy this way).
    load : False only builds the object (no GPU or loaded library needed: the CPU box pre-builds what the GPU box runs).
    reverse : also generate and compile the topology's checkpointed reverse sweep (``revgen.py``; float32 and float64), so
        that ``gradient=True`` with many parameters / ``gradient="reverse"`` costs one launch whatever their number.  Raises
        ``revgen.Unsupported`` for node kinds / update schemes the generator does not restate.
    Returns the kernel name the network's plans will report (``"specialised:<hash>"``).
    """
    extra_flags = [str(f) for f in extra_flags]
    name, header = header_for(network, salt=" ".join(extra_flags))
    f32 = sorted({(int(a), int(b)) for a, b in float32})
    f64 = sorted({(int(a), int(b)) for a, b in float64})
    for pf, pr in f32 + f64:
        if not (0 <= pf <= 4 and pr in (0, 1)):
            raise ValueError("kernels are (PF, PR) with 0 <= PF <= 4 and PR in (0, 1)")
    # content tag: kernels, math policy, every source the objects are compiled from (no paths: the cache travels)
    rev_header = None
    if reverse:
        from . import revgen

        rev_header = revgen.generate(name, network, network.volatility_updates, network.mean_field_updates, inputs=tuple(network.input_idxs))
    tag = hashlib.sha1(repr((f32, f64, build.MATH, _source_hash(), header, extra_flags, rev_header)).encode()).hexdigest()[:12]
    so = os.path.join(OUT_DIR, f"libhgf_{name}_{tag}.so")
    label = f"specialised:{name}"
    if not os.path.exists(so):
        _compile(name, header, f32, f64, so, label, verbose, extra_flags, rev_header)
    if not load:
        return label
    L = _lib.lib()
    if _registered.get(name) != so:
        _lib.check(L.hgf_register_specialisation(so.encode()))
        _registered[name] = so
    if hasattr(network, "_invalidate_plan"):
        network._invalidate_plan()
    return label


def _source_hash() -> str:
    h = hashlib.sha1()
    for path in build._headers() + [os.path.join(build.CSRC, f) for f in ("hgf_static_inst.cu", "hgf_specialisation.cu", "hgf_reverse_inst.cu")]:
        h.update(open(path, "rb").read())
    return h.hexdigest()


def _compile(name, header, f32, f64, so, label, verbose, extra_flags=(), rev_header=None):
    work = os.path.join(WORK_DIR, name)
    os.makedirs(work, exist_ok=True)
    os.makedirs(OUT_DIR, exist_ok=True)
    with open(os.path.join(work, f"topo_{name}.h"), "w") as f:
        f.write(header)
    topo_flags = [f"-DHGF_TOPO_HEADER=<topo_{name}.h>", f"-DHGF_TOPO_TYPE=Topo_{name}", "-I", work]
    jobs = []
    for base, kernels in (("float", f32), ("double", f64)):
        for pf, pr in kernels:
            jobs.append((os.path.join(build.CSRC, "hgf_static_inst.cu"), os.path.join(work, f"inst_{base}_d{pf}r{pr}.o"),
                         topo_flags + [f"-DHGF_BASE={base}", f"-DHGF_NTAN={pf}", f"-DHGF_NRESP={pr}"] + list(extra_flags)))
    m32 = sum(_variant_bit(pf, pr) for pf, pr in f32)
    m64 = sum(_variant_bit(pf, pr) for pf, pr in f64)
    rev_flags = []
    if rev_header is not None:
        with open(os.path.join(work, f"reverse_{name}.h"), "w") as f:
            f.write(rev_header)
        rev_flags = [f"-DHGF_REV_FN=reverse_{name}"]
        jobs.append((os.path.join(build.CSRC, "hgf_reverse_inst.cu"), os.path.join(work, "reverse.o"),
                     ["-I", work, f"-DHGF_REV_HEADER=<reverse_{name}.h>", f"-DHGF_REV_TYPE=Rev_{name}"] + rev_flags + list(extra_flags)))
    jobs.append((os.path.join(build.CSRC, "hgf_specialisation.cu"), os.path.join(work, "table.o"),
                 topo_flags + [f'-DHGF_TOPO_LABEL="{label}"', f"-DHGF_MASK32={m32}u", f"-DHGF_MASK64={m64}u"] + rev_flags))

    def run(job):
        src, obj, flags = job
        r = subprocess.run([build.NVCC] + build.COMMON + flags + ["-c", src, "-o", obj], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {os.path.basename(obj)}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            print(f"  specialise: {os.path.basename(obj)}", flush=True)
        return obj

    with cf.ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
        objs = list(ex.map(run, jobs))
    lib_dir = os.path.dirname(_lib.LIB_PATH)
    lib_name = os.path.basename(_lib.LIB_PATH)[3:-3]
    tmp = so + f".tmp{os.getpid()}"
    r = subprocess.run([build.NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", tmp] + objs +
                       ["-L", lib_dir, f"-l{lib_name}", "-Xlinker", "-rpath=$ORIGIN/.."], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    os.replace(tmp, so)
