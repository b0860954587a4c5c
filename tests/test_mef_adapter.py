"""The reference-side binding, compiled and run: canopy::lower_model (include/canopy-cpp.h, -DCANOPY_WITH_MEF_OPENPSA) on
stand-ins that have the reference's signatures (tests/cpp/mef_shim/; the reference headers need Boost, which this image lacks).

The model of tests/cpp/mef_adapter.cpp holds what the reference's Initializer leaves behind for the adapter to meet:
Model::basic_events() incl. an unused row, house events (one set, one unset), nested gates, complemented arguments, an
atleast gate, two fault trees, and a CCF group applied with CcfGroup::ApplyModel (reference ccf_group.h:215-260) whose
CcfEvents are owned by the group and are NOT rows of Model::basic_events().

CPU: the lowered descriptor against an independent Python construction of the same model (exact enumeration, both CCF
settings).  GPU: canopy::quantify through the C++ layer against the scalar oracle, bit for bit."""
import itertools
import json
import os
import subprocess

import numpy as np
import pytest

from canopy_b200 import graph as G
from helpers import ROOT
from oracle import oracle as O


@pytest.fixture(scope="module")
def exe(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("adapter") / "mef_adapter")
    lib_dir = os.path.join(ROOT, "canopy_b200")
    subprocess.check_call(["g++", "-std=c++17", "-Wall", "-Wextra", "-Werror", "-DCANOPY_WITH_MEF_OPENPSA",
                           "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "tests", "cpp", "mef_shim"),
                           os.path.join(ROOT, "tests", "cpp", "mef_adapter.cpp"), "-o", out,
                           os.path.join(lib_dir, "libcanopy.so"), f"-Wl,-rpath,{lib_dir}"])
    return out


def _graph(d):
    return G.Graph(p=d["p"], op=d["op"], kmin=d["kmin"], kmax=d["kmax"], child_off=d["child_off"], child=d["child"], roots=d["roots"])


def _expected(ccf: bool):
    """The same model, written down independently with the Python builder; CCF expansion restated from ccf_group.h:215-260:
    each member becomes an OR over the CCF events of every combination (of each level) it belongs to."""
    b = G.GraphBuilder()
    pa, pb, pc = (b.basic_event(0.05) for _ in range(3))
    valve, power = b.basic_event(0.02), b.basic_event(0.003)
    b.basic_event(0.5)   # unused
    maintenance, bypass = b.constant(True), b.constant(False)
    members = [pa, pb, pc]
    if ccf:
        uses = [[] for _ in members]
        for level, prob in ((1, 0.04), (2, 0.006), (3, 0.004)):
            for comb in itertools.combinations(range(3), level):
                e = b.basic_event(prob)
                for i in comb:
                    uses[i].append(e)
        members = [b.gate(G.OR, u) for u in uses]
    two = b.gate(G.ATLEAST, members, 2)
    iso = b.gate(G.OR, [valve, bypass])
    cooling = b.gate(G.OR, [two, iso])
    supply = b.gate(G.AND, [maintenance, power])
    b.root(b.gate(G.OR, [cooling, supply]))
    b.root(b.gate(G.AND, [members[0], (members[1], True), (two, True)]))
    return b.build()


@pytest.mark.parametrize("ccf", [0, 1])
def test_adapter_lowers_the_model(exe, ccf):
    d = json.loads(subprocess.check_output([exe, str(ccf)], text=True))
    g = _graph(d)
    assert d["n_basic"] == (13 if ccf else 6) and d["n_gates"] == (9 if ccf else 6)
    # the model's own rows come first and in table order; CCF events follow in walk order
    assert d["p"][:6] == [0.05, 0.05, 0.05, 0.02, 0.003, 0.5]
    if ccf:
        assert sorted(d["p"][6:]) == sorted([0.04] * 3 + [0.006] * 3 + [0.004])
    want = O.exact(_expected(bool(ccf)), False, 64)
    got = O.exact(g, False, 64)
    assert np.allclose(got, want, rtol=1e-13, atol=0)
    if ccf:   # the members themselves are no longer sampled: replaced by their proxy gates
        from canopy_b200.engine import Plan
        with Plan(g) as plan:
            assert not set(int(e) for e in plan.view()["event_orig"]) & {0, 1, 2, 5}


def test_adapter_without_a_gpu_fails_loudly(exe):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    d = json.loads(subprocess.check_output([exe, "1", "run", "3", "1000"], text=True))
    assert d["run_status"] == 5 and "tallies" not in d      # CANOPY_E_CUDA: no CPU fallback


@pytest.mark.gpu
@pytest.mark.parametrize("ccf", [0, 1])
def test_gpu_adapter_quantify_matches_oracle(exe, ccf):
    n = 200_000
    d = json.loads(subprocess.check_output([exe, str(ccf), "run", "11", str(n)], text=True))
    g = _graph(d)
    assert d["seed"] == 11 and d["n"] == n       # Settings::seed() / num_trials() -> McSettings (settings.h:118,178)
    want = O.mc_scalar(g, 11, 0, n, True)
    assert np.array_equal(np.array(d["tallies"], dtype=np.uint64), want)
    # and the law: top events within 4.5 sigma of exact enumeration with the quantised probabilities
    ex = O.exact(g, True, 32)
    assert np.all(np.abs(want - n * ex) <= 4.5 * np.sqrt(n * ex * (1 - ex)) + 1)
