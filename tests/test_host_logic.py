"""CPU-side checks: the C-ABI library loads and exports what include/pyhgf_cuda.h declares, the host
topology compiler derives the reference's schedule, and plan creation validates its input.  No kernel
is launched here (no GPU in the authoring container)."""

import ctypes as C
import os
import re

import numpy as np
import pytest

import nets
from oracle import network as onet

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from pyhgf_b200 import _lib, build

    build.build(verbose=False)
    return _lib


def test_library_exports_every_declared_symbol(lib):
    header = open(os.path.join(ROOT, "include", "pyhgf_cuda.h")).read()
    declared = set(re.findall(r"\b(hgf_[a-z_0-9]+)\s*\(", header))
    assert declared == set(lib.EXPORTS), declared ^ set(lib.EXPORTS)
    L = lib.lib()
    for name in declared:
        assert hasattr(L, name), name
    assert L.hgf_abi_version() == 5
    assert lib.launch_count() == 0


def test_struct_layouts_match_the_header(lib):
    """ctypes mirrors of hgf_topology_t / hgf_call_t: same field order as the header."""
    header = open(os.path.join(ROOT, "include", "pyhgf_cuda.h")).read()

    def fields(struct_name):
        end = header.index("} " + struct_name + ";")
        body = header[header.rindex("typedef struct {", 0, end) + len("typedef struct {"):end]
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        return [re.search(r"(\w+)\s*$", d.strip()).group(1) for d in body.split(";") if d.strip()]

    assert fields("hgf_topology_t") == [f[0] for f in lib.Topology._fields_]
    assert fields("hgf_call_t") == [f[0] for f in lib.Call._fields_]
    assert fields("hgf_ef_call_t") == [f[0] for f in lib.EfCall._fields_]
    assert fields("hgf_categorical_call_t") == [f[0] for f in lib.CategoricalCall._fields_]


def test_ef_state_node_defaults(lib):
    """add_ef_state, pyhgf/model/add_nodes.py:360-470: defaults, derived entries, the keys it pops, its errors."""
    from pyhgf_b200.network import Network

    n = Network().add_nodes(kind="ef-state")
    a = n.attributes[0]
    assert n.edges[0].node_type == 3 and a["nus"] == 3.0 and a["dimension"] == 1 and a["observed"] == 1
    assert np.array_equal(a["xis"], [0.0, 1.0]) and a["mean"] == 0.0 and a["observation_ss"].shape == (2,)
    assert "distribution" not in a and "learning" not in a
    m = Network().add_nodes(kind="ef-state", nus=5.0, dimension=2, distribution="multivariate-normal")
    assert np.array_equal(m.attributes[0]["xis"], [0, 0, 1, 0, 1]) and m.attributes[0]["observation_ss"].shape == (5,)
    m.create_belief_propagation_fn()
    assert m.input_idxs == (0,) and m.input_dim == [2]
    with pytest.raises(ValueError, match="either 'normal' or 'multivariate-normal'"):
        Network().add_nodes(kind="ef-state", distribution="gamma")
    with pytest.raises(ValueError, match="generalised-filtering"):
        Network().add_nodes(kind="ef-state", learning="adam")
    # learning="hgf-2" (add_nodes.py:438-452): per statistic a fed node, its value parent and a volatility parent of that;
    # the node indexes the reference's test edits (tests/test_nodes/test_exponential_family.py:78-83) mean the same nodes
    h = Network().add_nodes(kind="ef-state", learning="hgf-2", distribution="multivariate-normal", dimension=2)
    assert len(h.edges) == 16 and h.edges[0].value_parents == (1, 4, 7, 10, 13)
    assert h.edges[1].value_children == (0,) and h.edges[1].value_parents == (2,) and h.edges[2].volatility_parents == (3,)
    assert len(Network().add_nodes(kind="ef-state", learning="hgf-1").edges) == 5
    assert len(Network().add_nodes(kind="ef-state", learning="hgf-3").edges) == 9
    h.create_belief_propagation_fn()
    assert h.input_idxs == (0,) and h.input_dim == [2]
    steps = [(i, f.__name__) for i, f in h.update_sequence.update_steps]
    assert steps[0] == (0, "prediction_error_update_exponential_family_dynamic")
    assert steps[-1] == (0, "posterior_update_exponential_family_dynamic")
    implied, _ = h._implied_network()
    assert implied.input_idxs == (0, 3, 6, 9, 12) and implied.attributes[0]["tonic_volatility"] == -4.0
    with pytest.raises(ValueError, match="not found in the default"):
        Network().add_nodes(kind="ef-state", tonic_volatility=-2.0)
    mixed = Network().add_nodes(kind="ef-state").add_nodes()
    with pytest.raises(NotImplementedError, match="two networks"):
        mixed.create_belief_propagation_fn()


BUILDERS = [nets.binary3, nets.binary2, nets.continuous2, nets.continuous3, nets.nonlinear, nets.tanh_coupling,
            nets.volatile2, nets.multi_child, nets.ghgf16]


@pytest.mark.parametrize("vu", ["unbounded", "eHGF", "standard"])
@pytest.mark.parametrize("mf", [False, True])
@pytest.mark.parametrize("build", BUILDERS, ids=[b.__name__ for b in BUILDERS])
def test_schedule_and_layout_match_the_oracle(lib, build, vu, mf):
    from pyhgf_b200.network import Network

    g = build(Network, volatility_updates=vu, mean_field_updates=mf).create_belief_propagation_fn()
    o = build(onet.Network, volatility_updates=vu, mean_field_updates=mf).create_belief_propagation_fn()
    assert tuple((i, f.opcode) for i, f in g.update_sequence.prediction_steps) == o.update_sequence[0]
    assert tuple((i, f.opcode) for i, f in g.update_sequence.update_steps) == o.update_sequence[1]
    assert g.input_idxs == o.input_idxs
    assert g.plan.n_slots == o.n_slots
    for i, e in enumerate(g.edges):
        assert tuple(e)[:5] == tuple(o.edges[i])[:5]
        for key, field in [("value_coupling_parents", e.value_parents), ("volatility_coupling_parents", e.volatility_parents),
                           ("value_coupling_children", e.value_children), ("volatility_coupling_children", e.volatility_children)]:
            for pos in range(len(field or ())):
                assert g.plan.slot(i, key, pos) == o.slot(i, key, pos)
    assert np.array_equal(g._flatten(g.plan).astype(np.float64), o.flatten())
    assert g.plan.kernel_name  # a variant was selected


def test_schedule_counts_of_the_reference(lib):  # tests/test_utils.py:70-92, tests/test_model.py:26-41
    from pyhgf_b200.network import Network

    n1 = Network().add_nodes(kind="binary-state").add_nodes(value_children=0).create_belief_propagation_fn()
    assert len(n1.update_sequence.prediction_steps) == 2 and len(n1.update_sequence.update_steps) == 2
    n2 = (Network(volatility_updates="standard").add_nodes().add_nodes(value_children=0).add_nodes(volatility_children=1)
          .create_belief_propagation_fn())
    assert len(n2.update_sequence.prediction_steps) == 3 and len(n2.update_sequence.update_steps) == 4
    names = [f.__name__ for _, f in nets.binary3(Network).create_belief_propagation_fn().update_sequence.update_steps]
    assert names == ["binary_state_node_prediction_error", "continuous_node_posterior_update",
                     "continuous_node_prediction_error", "continuous_node_posterior_update_unbounded"]


def test_api_errors_match_the_reference(lib):
    from pyhgf_b200.network import Network

    with pytest.raises(ValueError, match="Invalid update type"):
        Network(volatility_updates="fancy")
    with pytest.raises(ValueError, match="Invalid node type"):
        Network().add_nodes(kind="quantum-state")
    with pytest.raises(ValueError, match="not found in the default key"):
        Network().add_nodes(volatility=3.0)
    with pytest.raises(ValueError, match="coupling fn and value children do not match"):
        Network().add_nodes(n_nodes=3).add_nodes(value_children=[0, 1, 2], coupling_fn=(None, None))
    with pytest.raises(NotImplementedError):
        Network().add_nodes(kind="dp-state")
    net = Network().add_nodes().add_nodes(value_children=0)
    assert net.input_idxs == (0,)
    assert net.attributes[0]["autoconnection_strength"] == 0.0 and net.attributes[0]["tonic_volatility"] == 0.0
    net2 = Network().add_nodes().add_nodes(value_children=0)
    net2.input_idxs = (0,)  # a manual assignment returns before the mutation (network.py:160-176)
    assert net2.input_idxs == (0,) and net2.attributes[0]["tonic_volatility"] == -4.0


def test_plan_create_rejects_bad_topologies(lib):
    L = lib.lib()
    topo = lib.Topology()
    handle = C.c_void_p()
    assert L.hgf_plan_create(None, C.byref(handle)) != 0
    topo.n_nodes = 0
    assert L.hgf_plan_create(C.byref(topo), C.byref(handle)) != 0
    assert b"n_nodes" in L.hgf_last_error()
    from pyhgf_b200.network import Network
    from pyhgf_b200.topology import AdjacencyLists, get_update_sequence

    class Fake:
        edges = (AdjacencyLists(3, None, None, None, None, (None,)),)

    with pytest.raises(NotImplementedError):
        get_update_sequence(Fake(), "unbounded")
    with pytest.raises(NotImplementedError, match="cannot cross the C ABI"):
        Network().add_nodes().add_nodes(value_children=0, coupling_fn=(lambda x: x**2,)).create_belief_propagation_fn()


def test_no_cpu_fallback(lib):
    """The product fails loudly without a device instead of computing on the host."""
    import torch

    from pyhgf_b200.network import Network

    if torch.cuda.is_available():
        pytest.skip("a device is present")
    with pytest.raises(RuntimeError, match="needs a CUDA device"):
        nets.binary3(Network).input_data(np.ones(4))


def test_product_never_imports_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "pyhgf_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text and "liboracle" not in text, f


def test_on_demand_specialisation_builds_and_is_selected(lib):
    """pyhgf_b200/specialise.py: a topology without a built-in specialisation gets register-resident kernels
    compiled on demand (nvcc cross-compiles here), registered through hgf_register_specialisation and selected by
    hgf_plan_create for exactly that topology -- other networks keep their own kernels."""
    from pyhgf_b200.network import Network

    net = nets.tanh_coupling(Network)
    net.create_belief_propagation_fn()
    assert net.plan.kernel_name == "interpreter"
    name = net.specialise()
    assert name.startswith("specialised:")
    assert net.plan.kernel_name == name
    assert nets.tanh_coupling(Network).plan.kernel_name == name          # same topology, another instance
    assert nets.volatile2(Network).plan.kernel_name == "interpreter"     # a different one is untouched
    assert nets.binary3(Network).plan.kernel_name == "static:binary3"
    assert net.specialise() == name                                      # cached: no rebuild
    L = lib.lib()
    assert L.hgf_register_specialisation(b"/nonexistent.so") != 0 and b"nonexistent" in L.hgf_last_error()


def test_switching_plan_lowering(lib):
    """input_custom_sequence (network.py:653-767): several update sequences in one plan, run by the interpreter."""
    from pyhgf_b200.network import Network
    from pyhgf_b200.topology import Plan

    net = nets.binary3(Network).create_belief_propagation_fn()
    seq = net.update_sequence
    short = (seq.prediction_steps, tuple(s for s in seq.update_steps if s[0] != 2))
    args = (net.edges, seq, net.input_idxs, 1, "unbounded", False, 1e10)
    plan = Plan(*args, branches=(tuple(seq), short))
    assert plan.n_branches == 2 and plan.kernel_name == "interpreter" and plan.n_slots == net.plan.n_slots
    assert Plan(*args, branches=(tuple(seq),)).n_branches == 0  # one branch is an ordinary plan
    with pytest.raises(lib.BackendError, match="branches"):
        Plan(*args, branches=(tuple(seq),) * 9)
    with pytest.raises(ValueError, match="update branches"):
        net.input_custom_sequence((), [0], [1.0])


def test_categorical_state_structure(lib):
    """tests/test_categorical.py:19-72: the implied network of a categorical node and its coupling bookkeeping."""
    from pyhgf_b200.network import Network

    net = Network().add_nodes(kind="categorical-state", node_parameters={"n_categories": 3, "binary_parameters": {"tonic_volatility_2": -2.0}})
    assert [e.node_type for e in net.edges] == [5, 1, 1, 1, 2, 2, 2, 2]
    assert net.edges[0].value_parents == (1, 2, 3) and net.edges[2].value_children == (0,) and net.edges[2].value_parents == (5,)
    assert net.edges[7].volatility_children == (4, 5, 6) and net.edges[5].volatility_parents == (7,)
    for i, e in enumerate(net.edges):
        if e.node_type == 2:
            for field, key in (("value_parents", "value_coupling_parents"), ("value_children", "value_coupling_children"),
                               ("volatility_parents", "volatility_coupling_parents"), ("volatility_children", "volatility_coupling_children")):
                if getattr(e, field):
                    assert len(getattr(e, field)) == len(net.attributes[i][key])
                else:
                    assert getattr(e, field) is None and net.attributes[i][key] is None
    a = net.attributes[0]
    assert a["binary_idxs"] == [1, 2, 3] and np.array_equal(a["alpha"], np.ones(3)) and np.allclose(a["mean"], 1 / 3)
    # the reference hands the DEFAULT binary parameters to the implied network (add_nodes.py:532-537)
    assert net.attributes[4]["tonic_volatility"] == -4.0 and a["binary_parameters"] == {"tonic_volatility_2": -2.0}
    assert np.isclose(net.attributes[4]["mean"], -np.log(2.0)) and net.attributes[1]["mean"] == 1 / 3
    net.create_belief_propagation_fn()
    assert net.input_idxs == (0,) and net.input_dim == [3]
    steps = [(i, f.__name__) for i, f in net.update_sequence.update_steps]
    assert steps[0] == (0, "categorical_state_prediction_error") and steps[-1] == (0, "categorical_state_update")
    assert (7, "continuous_node_posterior_update_unbounded") in steps
    implied, local = net._implied_network()
    assert implied.input_idxs == (0, 1, 2) and [e.node_type for e in implied.edges] == [1, 1, 1, 2, 2, 2, 2]
    assert implied.edges[0].value_children is None
    with pytest.raises(ValueError, match="n_categories"):
        Network().add_nodes(kind="categorical-state", node_parameters={"n_categories": 40})
