"""MEF XML loader (C++ in libcanopy) and the world_size-2 sharding logic over gloo.  CPU only."""
import ctypes as C
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from helpers import G, O, ROOT, eval_lowered, random_graph
from canopy_b200 import _lib
from canopy_b200.engine import Plan, load_mef


def test_c1_example_loads_and_matches_structure():
    g = load_mef(os.path.join(ROOT, "examples", "c1_small_ft.xml"))
    assert (g.n_basic, g.n_gates, g.n_edges) == (50, 16, 65)
    assert [g.gate_names[r - g.n_basic] for r in g.roots] == ["TOP"]      # the only gate nobody references
    assert set(int(o) for o in g.op) == {G.AND, G.OR}
    exact_top = 1.0
    # TOP = OR_s AND_t OR_e: independent events, so the exact value has a closed form
    p = {n: q for n, q in zip(g.basic_names, g.p)}
    k = 0
    for s in range(5):
        sub = 1.0
        for t in range(2):
            tr = 1.0
            for e in range(5):
                k += 1
                tr *= 1 - O.quantised_p(p[f"E{k}"])
            sub *= 1 - tr
        exact_top *= 1 - sub
    exact_top = 1 - exact_top
    n = 1 << 18
    t = O.mc_packed(g, 1, 0, n, nthreads=4)[0]
    sigma = (exact_top * (1 - exact_top) / n) ** 0.5
    assert abs(t / n - exact_top) < 4 * sigma


XML = """<?xml version="1.0"?>
<!-- every connective spelling of kConnectiveToString (reference event.h:164-166) -->
<opsa-mef>
  <define-fault-tree name="FT">
    <define-gate name="top"><or><gate name="g_and"/><event name="g_vote"/><not><gate name="g_card"/></not></or></define-gate>
    <define-gate name="g_and"><and><basic-event name="a"/><event name="b" type="basic-event"/><house-event name="h_on"/></and></define-gate>
    <define-gate name="g_vote"><atleast min="2"><event name="a"/><event name="c"/><event name="d"/></atleast></define-gate>
    <define-gate name="g_card"><cardinality min="1" max="2"><event name="a"/><event name="b"/><event name="c"/></cardinality></define-gate>
    <define-gate name="g_x"><xor><event name="a"/><event name="b"/></xor></define-gate>
    <define-gate name="g_n"><not><event name="c"/></not></define-gate>
    <define-gate name="g_nand"><nand><event name="a"/><event name="b"/></nand></define-gate>
    <define-gate name="g_nor"><nor><event name="a"/><constant value="false"/></nor></define-gate>
    <define-gate name="g_null"><event name="d"/></define-gate>
    <define-gate name="g_iff"><iff><event name="a"/><event name="h_off"/></iff></define-gate>
    <define-gate name="g_imp"><imply><event name="a"/><event name="b"/></imply></define-gate>
    <define-component name="sub">
      <define-basic-event name="d"><float value="0.4"/></define-basic-event>
    </define-component>
  </define-fault-tree>
  <model-data>
    <define-basic-event name="a"><float value="0.1"/></define-basic-event>
    <define-basic-event name="b"><float value="2.0e-1"/></define-basic-event>
    <define-basic-event name="c"><float value="0.3"/></define-basic-event>
    <define-house-event name="h_on"><constant value="true"/></define-house-event>
    <define-house-event name="h_off"/>
  </model-data>
</opsa-mef>
"""


def test_all_connectives_from_xml():
    g = load_mef(text=XML)
    names = g.gate_names
    assert g.basic_names == ["d", "a", "b", "c"]
    op = {n: int(o) for n, o in zip(names, g.op)}
    assert op == dict(top=G.OR, g_and=G.AND, g_vote=G.ATLEAST, g_card=G.CARDINALITY, g_x=G.XOR, g_n=G.NOT, g_nand=G.NAND,
                      g_nor=G.NOR, g_null=G.NULL, g_iff=G.IFF, g_imp=G.IMPLY)
    top_children = g.child[int(g.child_off[0]):int(g.child_off[1])]
    assert int(top_children[2]) >> 31 == 1                                 # <not> sets the complement flag
    and_children = [int(c) for c in g.child[int(g.child_off[1]):int(g.child_off[2])]]
    assert and_children[2] == G.CONST_TRUE                                 # house event state
    roots = sorted(names[r - g.n_basic] for r in g.roots)
    assert roots == sorted(["top", "g_x", "g_n", "g_nand", "g_nor", "g_null", "g_iff", "g_imp"])
    exact = O.exact(g, tally_all=True)
    pa, pb, pc, pd = (O.quantised_p(x) for x in (0.1, 0.2, 0.3, 0.4))
    by = {n: exact[g.n_roots + i] for i, n in enumerate(names)}
    assert abs(by["g_and"] - pa * pb) < 1e-12
    assert abs(by["g_iff"] - (1 - pa)) < 1e-12                             # a <-> false
    assert abs(by["g_imp"] - (1 - pa * (1 - pb))) < 1e-12
    assert abs(by["g_null"] - pd) < 1e-12
    with Plan(g, tally_all_gates=True) as plan:
        assert np.array_equal(eval_lowered(plan.view(), g, 9, 0, 3000), O.mc_scalar(g, 9, 0, 3000, True))


BAD_XML = [
    ("<opsa-mef><define-fault-tree name='t'><define-gate name='g'><and><event name='x'/><event name='y'/></and></define-gate></define-fault-tree></opsa-mef>",
     _lib.E_VALIDITY, "Undefined Element"),
    ("<opsa-mef><model-data><define-basic-event name='a'><float value='0.1'/></define-basic-event>"
     "<define-basic-event name='a'><float value='0.2'/></define-basic-event></model-data></opsa-mef>", _lib.E_VALIDITY, "Duplicate"),
    ("<opsa-mef><model-data><define-basic-event name='a'><float value='1.5'/></define-basic-event></model-data></opsa-mef>",
     _lib.E_DOMAIN, "probability"),
    ("<opsa-mef><model-data><define-basic-event name='a'><extern-function name='f'/></define-basic-event></model-data></opsa-mef>",
     _lib.E_UNSUPPORTED, "not supported"),
    ("<opsa-mef><define-fault-tree name='t'></opsa-mef>", _lib.E_IO, "mismatched"),
    ("<other/>", _lib.E_IO, "opsa-mef"),
]


@pytest.mark.parametrize("xml,status,msg", BAD_XML, ids=[str(i) for i in range(len(BAD_XML))])
def test_loader_errors(xml, status, msg):
    with pytest.raises(_lib.CanopyError) as ei:
        load_mef(text=xml)
    assert ei.value.status == status and msg in ei.value.message


def test_missing_file():
    with pytest.raises(_lib.CanopyError) as ei:
        load_mef("/nonexistent/model.xml")
    assert ei.value.status == _lib.E_IO


# ---- world_size 2 over gloo: shard by counter range + one all-reduce ---------------------------------------------
def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from canopy_b200.dist import run_sharded
    g = random_graph(3, n_basic=12, n_gates=30, n_roots=2)
    seen = []

    class OracleShard:
        """Stands in for a Plan on this CPU-only box (the CUDA engine refuses to run here): same run_device signature, the
        oracle computes ONE rank's shard into the tensor run_sharded hands to the all-reduce."""
        n_tallies = g.n_tallies(True)

        def run_device(self, tallies_ptr, n, seed=0, sample_begin=0, device=0, stream=None):
            seen.append((sample_begin, n))
            out = np.ctypeslib.as_array(C.cast(tallies_ptr, C.POINTER(C.c_int64)), shape=(self.n_tallies,))
            out[:] = O.mc_packed(g, seed, sample_begin, n, True).astype(np.int64)
    total = run_sharded(OracleShard(), 3 * 4096 + 100, seed=11, sample_begin=50, device="cpu")
    q.put((rank, seen[0], total.tolist()))
    dist.destroy_process_group()


def test_two_rank_sharding_matches_single_range():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    g = random_graph(3, n_basic=12, n_gates=30, n_roots=2)
    want = O.mc_scalar(g, 11, 50, 3 * 4096 + 100, True).astype(np.int64).tolist()
    (r0, s0, t0), (r1, s1, t1) = res
    assert t0 == want and t1 == want
    assert s0[0] == 50 and s0[0] + s0[1] == s1[0] and s1[0] + s1[1] == 50 + 3 * 4096 + 100 and s1[0] % 4096 == 0
