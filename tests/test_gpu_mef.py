"""GPU parity for the MEF front rows of SURVEY.md 8 (a6 CCF groups, a8 / f2 event trees): models go through the library's
MEF reader (canopy_model_load_mef_string_ex), are lowered and run on the device, EVERY sequence and EVERY gate tallied, and
compared bit for bit with the scalar oracle on the loader's graph.

Reference anchors: CcfGroup::ApplyModel ccf_group.h:215-260 and the four model formulas :330-455; EventTree / Fork / Path /
Sequence event_tree.h:65-197; CollectFormula / SetHouseEvent / Link instruction.h:38-83,157-167."""
import os

import numpy as np
import pytest

from helpers import O
from canopy_b200 import synth
from canopy_b200.engine import Plan, load_mef
from test_mef_model import ET_XML, ccf_xml

pytestmark = pytest.mark.gpu
CORES = len(os.sched_getaffinity(0))

CCF_CASES = [
    ("beta-factor", 3, 0.1, [(3, 0.2)]),
    ("MGL", 4, 0.05, [(2, 0.3), (3, 0.4), (4, 0.5)]),
    ("alpha-factor", 3, 0.08, [(1, 0.7), (2, 0.2), (3, 0.1)]),
    ("phi-factor", 3, 0.06, [(1, 0.8), (2, 0.15), (3, 0.05)]),
]


@pytest.mark.parametrize("ccf_analysis", [False, True], ids=["ccf_off", "ccf_on"])
@pytest.mark.parametrize("model_name,n,q,factors", CCF_CASES, ids=[c[0] for c in CCF_CASES])
@pytest.mark.parametrize("spec", [2, 4])
def test_ccf_models_on_gpu(model_name, n, q, factors, ccf_analysis, spec):
    g = load_mef(text=ccf_xml(model_name, n, q, factors), ccf_analysis=ccf_analysis)
    assert g.n_basic == (n if not ccf_analysis else g.n_basic) and (ccf_analysis or g.n_gates == 1)
    ns, begin = 3 * 8192 + 1234, 4096 * 7 + 9
    with Plan(g, tally_all_gates=True, sampler=spec) as plan:
        got, rep = plan.run(ns, seed=0xCCF, sample_begin=begin)
        assert rep.gate_launches >= 1
    want = O.mc_scalar(g, 0xCCF, begin, ns, True, spec=spec)
    assert np.array_equal(got, want)
    # the model is not degenerate: "at least two members fail" happens, and not always
    assert 0 < int(got[0]) < ns
    # law check against exact enumeration of the loader's graph (quantised probabilities), 4 sigma
    exact = O.exact(g, True)
    with Plan(g, tally_all_gates=True, sampler=spec) as plan:
        big, _ = plan.run(1 << 22, seed=11)
    sigma = np.sqrt(np.maximum(exact * (1 - exact), 1e-12) / (1 << 22))
    assert np.all(np.abs(big / float(1 << 22) - exact) <= 4 * sigma + 1e-9)


def test_ccf_members_replaced_by_proxy_gates_on_gpu():
    """With ccf_analysis the member events are no longer sampled: their rows are the proxy OR gates over the CCF events
    (ccf_group.h:215-260).  The per-gate tallies of the proxies must show the members' total probability q."""
    n, q = 3, 0.1
    g = load_mef(text=ccf_xml("beta-factor", n, q, [(3, 0.2)]), ccf_analysis=True)
    S = 1 << 22
    with Plan(g, tally_all_gates=True) as plan:
        got, _ = plan.run(S, seed=5)
    proxies = [i for i, nm in enumerate(g.gate_names) if nm.startswith("m")]
    assert len(proxies) == n
    for gi in proxies:
        ph = int(got[g.n_roots + gi]) / S
        # P(member fails) = 1 - (1 - 0.8 q)(1 - 0.2 q)
        want = 1 - (1 - O.quantised_p(0.8 * q)) * (1 - O.quantised_p(0.2 * q))
        assert abs(ph - want) < 5 * (want / S) ** 0.5


@pytest.mark.parametrize("spec", [2, 4])
def test_event_tree_model_on_gpu(spec):
    """The hand-written event tree of tests/test_mef_model.py (forks, a named branch, a rule with
    test-functional-event, set-house-event clones, a link into a second tree): every sequence and every gate."""
    g = load_mef(text=ET_XML)
    names = g.meta["root_names"]
    assert names == ["G1", "G2", "IE.OK", "IE.CD", "IE.PART", "IE.S3"]
    ns, begin = 5 * 8192 + 77, (1 << 33) + 4000
    with Plan(g, tally_all_gates=True, sampler=spec) as plan:
        got, _ = plan.run(ns, seed=31, sample_begin=begin)
        big, _ = plan.run(1 << 22, seed=32)
    assert np.array_equal(got, O.mc_scalar(g, 31, begin, ns, True, spec=spec))
    exact = O.exact(g, True)
    sigma = np.sqrt(np.maximum(exact * (1 - exact), 1e-12) / (1 << 22))
    assert np.all(np.abs(big / float(1 << 22) - exact) <= 4 * sigma + 1e-9)
    t = dict(zip(names, (int(x) for x in big[: len(names)])))
    assert t["IE.OK"] + t["G1"] == 1 << 22          # OK collects not G1 and nothing else
    assert 0 < t["IE.CD"] < t["G1"] and 0 < t["IE.S3"]


def test_c3_event_tree_200_sequences_over_one_20k_gate_dag():
    """BASELINE.json config 3 as its text states it: 200 sequences sharing one 20k-gate DAG, per-sequence and per-gate
    tallies -- written out as Open-PSA MEF (synth.mef_xml), read back through the library's reader and event-tree
    walker, lowered, run on the GPU; compared bit for bit with the scalar oracle (short range) and with the packed port
    (longer ragged range) on the loader's graph."""
    base = synth.config_graph("c3")
    g = load_mef(text=synth.mef_xml(base, n_seq=200, seed=3))
    names = g.meta["root_names"]
    seqs = [n for n in names if n.startswith("IE.")]
    assert len(seqs) == 201 and seqs[0] == "IE.S0" and seqs[-1] == "IE.REST"
    assert g.n_basic == base.n_basic and g.n_gates >= base.n_gates + 201
    with Plan(g, tally_all_gates=True) as plan:
        st = plan.stats
        assert st["n_gates_live"] > 19000 and st["n_tallies"] == g.n_roots + g.n_gates
        n1, b1 = 600, 8192 * 3 - 300                     # straddles a slab boundary
        got1, _ = plan.run(n1, seed=0xE7, sample_begin=b1)
        n2, b2 = 2 * 8192 + 4096 + 77, (1 << 40) + 5
        got2, _ = plan.run(n2, seed=0xE7, sample_begin=b2)
        S = 1 << 22
        big, _ = plan.run(S, seed=0xE8)
    assert np.array_equal(got1, O.mc_scalar(g, 0xE7, b1, n1, True, spec=plan.spec))
    assert np.array_equal(got2, O.mc_packed(g, 0xE7, b2, n2, True, nthreads=CORES, spec=plan.spec))
    # the leaves of the fork tree partition the sample space: the sequence tallies sum to the number of samples
    first = names.index("IE.S0")
    assert int(big[first: first + 201].sum()) == S
    assert int((big[first: first + 201] > 0).sum()) >= 20   # and a fair number of the sequences do occur
