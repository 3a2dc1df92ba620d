"""Generate the register-resident kernel specialisations.

For each topology listed in ``TOPOLOGIES`` (the BASELINE configurations) this
writes ``csrc/generated/topo_<name>.h``: a ``Topo_<name>`` accessor struct that
answers every topology query of ``hgf_ops.cuh`` with a compile-time constant.
With it the schedule and edge loops of the time step unroll completely, every
attribute index is static, and the compiler keeps the whole instance in
registers (the opcode interpreter keeps it in local memory).  A specialisation
is used when ``hgf_plan_create`` finds that a lowered network answers every
query identically (``topo_matches`` in hgf_static_run.h) -- the generator never
has to reproduce the library's byte layout.

    python -m pyhgf_b200.codegen        # (re)writes csrc/generated/*
"""

from __future__ import annotations

import os

from . import _lib
from .topology import AdjacencyLists, get_update_sequence

HERE = os.path.dirname(os.path.abspath(__file__))
GEN = os.path.join(HERE, "csrc", "generated")
NF = _lib.N_FIELDS


class _Graph:
    """Just enough of ``Network`` to describe edges (no torch, no library)."""

    def __init__(self):
        self.edges = ()

    def add(self, node_type, value_children=(), volatility_children=(), coupling_fn=None):
        idx = len(self.edges)
        fns = tuple(coupling_fn or [None] * len(value_children))
        self.edges += (AdjacencyLists(node_type, None, None, None, None, fns),)
        for c in value_children:
            self._edge("value", idx, c)
        for c in volatility_children:
            self._edge("volatility", idx, c)
        return idx

    def _edge(self, kind, parent, child):
        e = list(self.edges)
        cf, pf = ("value_children", "value_parents") if kind == "value" else ("volatility_children", "volatility_parents")
        d = e[parent]._asdict()
        d[cf] = (d[cf] or ()) + (child,)
        e[parent] = AdjacencyLists(**d)
        d = e[child]._asdict()
        d[pf] = (d[pf] or ()) + (parent,)
        e[child] = AdjacencyLists(**d)
        self.edges = tuple(e)


def _binary(levels):
    g = _Graph()
    g.add(1)
    g.add(2, value_children=[0])
    if levels == 3:
        g.add(2, volatility_children=[1])
    return g


def _continuous(levels):
    g = _Graph()
    g.add(2)
    g.add(2, value_children=[0])
    for k in range(1, levels):
        g.add(2, volatility_children=[k])
    return g


def _ghgf16():
    g = _Graph()
    for _ in range(4):
        g.add(2)
    for i in range(4):
        g.add(2, value_children=[i])
    for i in range(4):
        g.add(2, volatility_children=[4 + i])
    g.add(6, value_children=[4, 5])
    g.add(6, value_children=[6, 7])
    g.add(6, value_children=[12])
    g.add(6, value_children=[13])
    return g


def _ghgf8():
    """One connected component of the 16-node gHGF above (its two components are isomorphic): what
    ``Network.batch`` launches per component when it splits a disconnected network (network.py)."""
    g = _Graph()
    g.add(2)
    g.add(2)
    g.add(2, value_children=[0])
    g.add(2, value_children=[1])
    g.add(2, volatility_children=[2])
    g.add(2, volatility_children=[3])
    g.add(6, value_children=[2, 3])
    g.add(6, value_children=[6])
    return g


#: name -> (graph factory, volatility_updates, mean_field, float32 kernels, float64 kernels); a kernel is
#: (PF, PR): PF filter tangent directions, PR response-parameter-only tangent directions (0 or 1)
_ALL = tuple((pf, pr) for pr in (0, 1) for pf in range(5))
TOPOLOGIES = {
    "binary2": (lambda: _binary(2), "unbounded", False, _ALL, ((0, 0),)),
    "binary3": (lambda: _binary(3), "unbounded", False, _ALL, ((0, 0), (4, 0))),
    "continuous2": (lambda: _continuous(2), "unbounded", False, ((0, 0), (3, 0)), ((0, 0),)),
    "continuous3": (lambda: _continuous(3), "unbounded", False, ((0, 0), (3, 0), (4, 0)), ((0, 0),)),
    "ghgf16": (_ghgf16, "unbounded", False, ((0, 0),), ()),
    "ghgf8": (_ghgf8, "unbounded", False, ((0, 0),), ()),
}


def _switch(name, args, table, default=0):
    """``HGF_D int name(args) const { switch ... }`` over a dict {key tuple: value}."""
    if not table:
        return f"  HGF_D int {name}({', '.join('int ' + a for a in args)}) const {{ return {default}; }}\n"
    if len(args) == 1:
        body = " ".join(f"case {k[0]}: return {v};" for k, v in sorted(table.items()))
        return f"  HGF_D int {name}(int {args[0]}) const {{ switch ({args[0]}) {{ {body} default: return {default}; }} }}\n"
    key = f"{args[0]} * 256 + {args[1]}"
    body = " ".join(f"case {k[0] * 256 + k[1]}: return {v};" for k, v in sorted(table.items()))
    return (f"  HGF_D int {name}(int {args[0]}, int {args[1]}) const {{ switch ({key}) {{ {body} default: return {default}; }} }}\n")


def generate_topology(name, spec) -> str:
    factory, vol_updates, mean_field, _, _ = spec
    return generate_header(name, factory(), vol_updates, mean_field)


def generate_header(name, g, vol_updates, mean_field, inputs=None) -> str:
    """The ``Topo_<name>`` accessor of a network description ``g`` (anything with ``.edges``): the BASELINE
    topologies above at build time, any user network on demand (pyhgf_b200/specialise.py)."""
    edges = g.edges
    n = len(edges)
    pred, upd = get_update_sequence(g, vol_updates, mean_field)
    if inputs is None:
        inputs = tuple(i for i, e in enumerate(edges) if e.value_children is None and e.volatility_children is None)
    inputs = tuple(int(i) for i in inputs)
    # slot layout of hgf_plan_create (hgf_program.h); checked at run time by topo_matches
    counts = {k: [len(getattr(e, k) or ()) for e in edges] for k in ("value_parents", "volatility_parents", "value_children", "volatility_children")}
    off = {k: [sum(v[:i]) for i in range(n)] for k, v in counts.items()}
    base_cvp = n * NF
    base_cvolp = base_cvp + sum(counts["value_parents"])
    base_cvc = base_cvolp + sum(counts["volatility_parents"])
    base_cvolc = base_cvc + sum(counts["value_children"])
    slot_clip = base_cvolc + sum(counts["volatility_children"])
    slot_resp = slot_clip + 1
    n_slots = slot_resp + _lib.N_RESPONSE_PARAMETERS

    def fn_of(parent, child):
        e = edges[parent]
        k = e.value_children.index(child)
        f = e.coupling_fn[k] if (e.coupling_fn is not None and k < len(e.coupling_fn)) else None
        from .topology import resolve_coupling_fn

        return resolve_coupling_fn(f)

    t = {k: {} for k in ("vp", "vp_fn", "volp", "vc", "vc_fn", "volc", "slot_cvp", "slot_cvolp", "slot_cvc", "slot_cvolc")}
    for i, e in enumerate(edges):
        for k, p in enumerate(e.value_parents or ()):
            t["vp"][(i, k)] = p
            t["vp_fn"][(i, k)] = fn_of(p, i)
            t["slot_cvp"][(i, k)] = base_cvp + off["value_parents"][i] + k
        for k, p in enumerate(e.volatility_parents or ()):
            t["volp"][(i, k)] = p
            t["slot_cvolp"][(i, k)] = base_cvolp + off["volatility_parents"][i] + k
        for k, c in enumerate(e.value_children or ()):
            t["vc"][(i, k)] = c
            t["vc_fn"][(i, k)] = fn_of(i, c)
            t["slot_cvc"][(i, k)] = base_cvc + off["value_children"][i] + k
        for k, c in enumerate(e.volatility_children or ()):
            t["volc"][(i, k)] = c
            t["slot_cvolc"][(i, k)] = base_cvolc + off["volatility_children"][i] + k

    out = [
        f"// GENERATED by pyhgf_b200/codegen.py -- topology \"{name}\" ({n} nodes, volatility_updates={vol_updates}).\n",
        "// Compile-time answers to the topology queries of hgf_ops.cuh; do not edit.\n",
        "#pragma once\n#include \"hgf_ops.cuh\"\n\nnamespace hgf {\n\n",
        f"struct Topo_{name} {{\n",
        "  static constexpr bool is_static = true;\n  static constexpr bool kConstAttributes = false;\n  static constexpr int kUnroll = 256;\n",
        f"  static constexpr int kSlots = {n_slots};\n  static constexpr int kNodes = {n};\n",
        "  const Program& p;\n",
        f"  HGF_D explicit Topo_{name}(const Program& prog) : p(prog) {{}}\n",
        "  HGF_D double max_posterior_precision() const { return p.max_posterior_precision; }\n",
    ]

    def const(fn, v):
        out.append(f"  HGF_D constexpr int {fn}() const {{ return {v}; }}\n")

    const("n_nodes", n)
    const("n_slots", n_slots)
    const("n_pred", len(pred))
    const("n_upd", len(upd))
    const("n_inputs", len(inputs))
    const("n_fed", len(inputs))
    const("vol_updates", _lib.VOLATILITY_UPDATES[vol_updates])
    const("slot_clip", slot_clip)
    const("slot_resp", slot_resp)
    out.append(_switch("pred_node", ["i"], {(i,): s[0] for i, s in enumerate(pred)}))
    out.append(_switch("pred_op", ["i"], {(i,): s[1].opcode for i, s in enumerate(pred)}))
    out.append(_switch("upd_node", ["i"], {(i,): s[0] for i, s in enumerate(upd)}))
    out.append(_switch("upd_op", ["i"], {(i,): s[1].opcode for i, s in enumerate(upd)}))
    out.append(_switch("input_idx", ["k"], {(k,): v for k, v in enumerate(inputs)}))
    out.append(_switch("node_type", ["n"], {(i,): e.node_type for i, e in enumerate(edges)}))
    for fn, key in (("n_vp", "value_parents"), ("n_volp", "volatility_parents"), ("n_vc", "value_children"), ("n_volc", "volatility_children")):
        out.append(_switch(fn, ["n"], {(i,): c for i, c in enumerate(counts[key]) if c}))
    for fn in ("vp", "vp_fn", "volp", "vc", "vc_fn", "volc", "slot_cvp", "slot_cvolp", "slot_cvc", "slot_cvolc"):
        out.append(_switch(fn, ["n", "k"], t[fn]))
    # the schedule as straight-line code: every op and node index is a literal, so the switch in apply_op and
    # the edge loops inside the update functions resolve at compile time without relying on loop unrolling
    # (`self` is the accessor the kernel was instantiated with: this struct or a wrapper such as ConstAttributes<>)
    out.append("  template <class T, class Self> HGF_D void run_predictions(const Self& self, T* s, const Step<T>& st) const {\n")
    for node, fn in pred:
        out.append(f"    apply_op(self, s, st, {fn.opcode}, {node});  // {fn.__name__}\n")
    out.append("  }\n  template <class T, class Self> HGF_D void run_updates(const Self& self, T* s, const Step<T>& st) const {\n")
    for node, fn in upd:
        out.append(f"    apply_op(self, s, st, {fn.opcode}, {node});  // {fn.__name__}\n")
    out.append("  }\n};\n\n}  // namespace hgf\n")
    return "".join(out)


def _write_if_changed(path, text):
    if os.path.exists(path) and open(path).read() == text:
        return False
    with open(path, "w") as f:
        f.write(text)
    return True


def main():
    os.makedirs(GEN, exist_ok=True)
    from . import revgen

    includes, table, manifest = [], [], ["# topology base PF PR  (one translation unit each; written by codegen.py)\n"]
    reverse = []
    for name, spec in TOPOLOGIES.items():
        _write_if_changed(os.path.join(GEN, f"topo_{name}.h"), generate_topology(name, spec))
        includes.append(f"#include <generated/topo_{name}.h>\n")
        m32 = sum(1 << (pf + 5 * pr) for pf, pr in spec[3])
        m64 = sum(1 << (pf + 5 * pr) for pf, pr in spec[4])
        try:  # the reverse sweep of the topology, where the generator covers its node kinds (revgen.py)
            _write_if_changed(os.path.join(GEN, f"reverse_{name}.h"), revgen.generate(name, spec[0](), spec[1], spec[2]))
            reverse.append(name)
            rev = f"kReverse_{name}"
        except revgen.Unsupported:
            rev = "nullptr"
        table.append(f"HGF_STATIC_TOPO({name}, {m32}u, {m64}u, {rev})\n")
        for pf, pr in spec[3]:
            manifest.append(f"{name} float {pf} {pr}\n")
        for pf, pr in spec[4]:
            manifest.append(f"{name} double {pf} {pr}\n")
    _write_if_changed(os.path.join(GEN, "static_includes.inc"), "".join(includes))
    _write_if_changed(os.path.join(GEN, "static_table.inc"), "".join(table))
    _write_if_changed(os.path.join(GEN, "instances.txt"), "".join(manifest))
    _write_if_changed(os.path.join(GEN, "reverse_table.inc"), "".join(f"HGF_STATIC_REVERSE({n})\n" for n in reverse))
    _write_if_changed(os.path.join(GEN, "reverse.txt"), "".join(f"{n}\n" for n in reverse))
    keep = ({f"topo_{n}.h" for n in TOPOLOGIES} | {f"reverse_{n}.h" for n in reverse} |
            {"static_includes.inc", "static_table.inc", "instances.txt", "reverse_table.inc", "reverse.txt"})
    for f in os.listdir(GEN):
        if f not in keep:
            os.remove(os.path.join(GEN, f))


if __name__ == "__main__":
    main()
