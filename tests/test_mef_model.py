"""MEF loader rows of SURVEY.md 8f: expressions (mean value), CCF groups, event trees.  CPU only.

The expected numbers are computed here from the formulas the reference states (ccf_group.h:330-455,
exponential.h:15-103, random_deviate.h:44-264) and from hand-derived sequence probabilities; the graph the
loader emits is evaluated by the oracle's exact 2^B enumeration (oracle/, quantised probabilities)."""
import math
from itertools import combinations

import numpy as np
import pytest

from canopy_b200 import _lib
from canopy_b200.engine import Plan, load_mef
from oracle import oracle as O
from helpers import eval_lowered

Q = O.quantised_p


def model(body: str) -> str:
    return f"<?xml version='1.0'?><opsa-mef>{body}</opsa-mef>"


def be(name, expr):
    return f"<define-basic-event name='{name}'>{expr}</define-basic-event>"


def fl(v):
    return f"<float value='{v!r}'/>"


# ---- expressions -----------------------------------------------------------------------------------------------------
def test_expression_means_and_parameters():
    lam, mu, gam, t = 1e-4, 2e-3, 0.05, 8760.0
    xml = model(
        "<define-fault-tree name='ft'><define-gate name='top'><or>"
        + "".join(f"<basic-event name='e{i}'/>" for i in range(13)) + "</or></define-gate></define-fault-tree><model-data>"
        + "<define-parameter name='lambda'><float value='1e-4'/></define-parameter>"
        + "<define-parameter name='twice'><mul><parameter name='lambda'/><int value='2'/></mul></define-parameter>"
        + be("e0", "<exponential><parameter name='lambda'/><system-mission-time/></exponential>")
        + be("e1", "<GLM><float value='0.05'/><parameter name='lambda'/><float value='2e-3'/><system-mission-time/></GLM>")
        + be("e2", "<Weibull><float value='5000'/><float value='1.5'/><float value='100'/><system-mission-time/></Weibull>")
        + be("e3", "<sub><float value='1'/><div><float value='1'/><add><int value='1'/><int value='3'/></add></div></sub>")
        + be("e4", "<uniform-deviate><float value='0.1'/><float value='0.3'/></uniform-deviate>")
        + be("e5", "<lognormal-deviate><float value='0.02'/><float value='3'/><float value='0.95'/></lognormal-deviate>")
        + be("e6", "<beta-deviate><float value='2'/><float value='6'/></beta-deviate>")
        + be("e7", "<ite><lt><parameter name='twice'/><float value='1'/></lt><float value='0.125'/><float value='0.9'/></ite>")
        + be("e8", "<min><float value='0.4'/><pow><float value='0.5'/><int value='3'/></pow><abs><neg><float value='0.3'/></neg></abs></min>")
        + be("e9", "<switch><case><bool value='false'/><float value='0.7'/></case><case><geq><pi/><int value='3'/></geq><float value='0.0625'/></case><float value='0.9'/></switch>")
        + be("e10", "<gamma-deviate><float value='2'/><float value='0.01'/></gamma-deviate>")
        + be("e11", "<mul><normal-deviate><float value='0.5'/><float value='0.1'/></normal-deviate><exp><neg><parameter name='twice'/></neg></exp></mul>")
        + be("e12", "<histogram><float value='0'/><bin><float value='0.1'/><float value='3'/></bin>"
                    "<bin><float value='0.5'/><float value='1'/></bin></histogram>")
        + "</model-data>")
    g = load_mef(text=xml)
    r = lam + mu
    want = [1 - math.exp(-lam * t), (lam - (lam - gam * r) * math.exp(-r * t)) / r, 1 - math.exp(-(((t - 100) / 5000) ** 1.5)),
            0.75, 0.2, 0.02, 0.25, 0.125, 0.125, 0.0625, 0.02, 0.5 * math.exp(-2e-4),
            (3 * 0.05 + 1 * 0.3) / 4]  # histogram: bin masses 3/4 and 1/4 at midpoints 0.05 and 0.3
    assert g.p == pytest.approx(want, rel=1e-14, abs=0)
    # mission time is a loader option (Settings::mission_time)
    g2 = load_mef(text=xml, mission_time=100.0)
    assert g2.p[0] == pytest.approx(1 - math.exp(-lam * 100.0), rel=1e-14)
    assert g2.p[2] == 0.0  # Weibull before its time shift


@pytest.mark.parametrize("expr,status,msg", [
    ("<parameter name='nope'/>", _lib.E_VALIDITY, "Undefined Element"),
    ("<parameter name='loop'/>", _lib.E_CYCLE, "Cycle"),
    ("<div><float value='1'/><float value='0'/></div>", _lib.E_DOMAIN, "Division by 0"),
    ("<exponential><float value='-1'/><float value='1'/></exponential>", _lib.E_DOMAIN, "rate of failure"),
    ("<exponential/>", _lib.E_VALIDITY, "takes 2"),
    ("<periodic-test/>", _lib.E_VALIDITY, "Invalid number of arguments for Periodic Test"),
    ("<periodic-test>" + "<float value=\"1\"/>" * 11 + "</periodic-test>", _lib.E_UNSUPPORTED, "not supported"),
    ("<histogram><float value='0.2'/><bin><float value='0.1'/><float value='1'/></bin></histogram>", _lib.E_VALIDITY, "strictly increasing"),
    ("<histogram><float value='0'/><bin><float value='0.1'/><float value='-1'/></bin></histogram>", _lib.E_VALIDITY, "negative"),
    ("<add><float value='0.8'/><float value='0.8'/></add>", _lib.E_DOMAIN, "probability"),
])
def test_expression_errors(expr, status, msg):
    xml = model("<model-data><define-parameter name='loop'><add><parameter name='loop'/><int value='1'/></add></define-parameter>"
                + be("a", expr) + "</model-data>")
    with pytest.raises(_lib.CanopyError) as ei:
        load_mef(text=xml)
    assert ei.value.status == status and msg in ei.value.message


# ---- CCF groups ------------------------------------------------------------------------------------------------------
def ccf_xml(model_name, n, q, factors, top="<atleast min='2'>{args}</atleast>"):
    members = "".join(f"<basic-event name='m{i}'/>" for i in range(n))
    fac = "".join(f"<factor level='{lv}'>{fl(v)}</factor>" for lv, v in factors)
    args = "".join(f"<basic-event name='m{i}'/>" for i in range(n))
    return model(f"<define-fault-tree name='ft'><define-gate name='top'>{top.format(args=args)}</define-gate>"
                 f"<define-CCF-group name='grp' model='{model_name}'><members>{members}</members>"
                 f"<distribution>{fl(q)}</distribution><factors>{fac}</factors></define-CCF-group></define-fault-tree>")


def expected_ccf_probs(model_name, n, q, factors):
    """level -> probability of one CCF event of that level (ccf_group.h:330-455)."""
    f = dict(factors)
    if model_name == "beta-factor":
        return {1: (1 - f[n]) * q, n: f[n] * q}
    if model_name == "MGL":
        mx = max(f)
        out = {}
        for i in range(mx):
            v = 1.0 / math.comb(n - 1, i)
            for j in range(i):
                v *= f[j + 2]
            if i < mx - 1:
                v *= 1 - f[i + 2]
            out[i + 1] = v * q
        return out
    if model_name == "alpha-factor":
        s = sum(k * v for k, v in f.items())
        return {k: k / math.comb(n - 1, k - 1) * (f[k] / s) * q for k in f}
    return {k: v * q for k, v in f.items()}


@pytest.mark.parametrize("model_name,n,q,factors", [
    ("beta-factor", 3, 0.1, [(3, 0.2)]),
    ("MGL", 4, 0.05, [(2, 0.3), (3, 0.4), (4, 0.5)]),
    ("alpha-factor", 3, 0.08, [(1, 0.7), (2, 0.2), (3, 0.1)]),
    ("phi-factor", 3, 0.06, [(1, 0.8), (2, 0.15), (3, 0.05)]),
])
def test_ccf_models(model_name, n, q, factors):
    xml = ccf_xml(model_name, n, q, factors)
    # without CCF analysis the members are plain basic events with the group distribution (ccf_group.h:127-130)
    g0 = load_mef(text=xml)
    assert g0.n_basic == n and np.all(g0.p == q)
    # with it: one new basic event per combination and level, members replaced by proxy OR gates (ccf_group.h:215-260)
    g = load_mef(text=xml, ccf_analysis=True)
    probs = expected_ccf_probs(model_name, n, q, factors)
    want_names, want_p = [], []
    for level in sorted(probs):
        for comb in combinations(range(n), level):
            want_names.append("[" + " ".join(f"m{i}" for i in comb) + "]")
            want_p.append(probs[level])
    assert g.basic_names[n:] == want_names
    assert g.p[n:] == pytest.approx(want_p, rel=1e-13)
    proxies = {g.gate_names[i]: i for i in range(g.n_gates) if g.gate_names[i].startswith("m")}
    assert sorted(proxies) == [f"m{i}" for i in range(n)]
    for i in range(n):
        gi = proxies[f"m{i}"]
        kids = [int(c) for c in g.child[int(g.child_off[gi]):int(g.child_off[gi + 1])]]
        assert g.op[gi] == 1 and [g.basic_names[c] for c in kids] == [nm for nm in want_names if f"m{i}" in nm[1:-1].split()]
    # exact probability of "at least 2 members fail": enumerate the CCF events by hand
    ev = [(set(comb), Q(probs[level])) for level in sorted(probs) for comb in combinations(range(n), level)]
    exact = 0.0
    for mask in range(1 << len(ev)):
        pr, failed = 1.0, set()
        for k, (members, pk) in enumerate(ev):
            if mask >> k & 1:
                pr *= pk; failed |= members
            else:
                pr *= 1 - pk
        if len(failed) >= 2:
            exact += pr
    assert O.exact(g)[0] == pytest.approx(exact, rel=1e-12)
    with Plan(g) as plan:  # and the lowered form evaluates like the oracle on the loader's graph
        assert np.array_equal(eval_lowered(plan.view(), g, 3, 0, 2048), O.mc_scalar(g, 3, 0, 2048))


@pytest.mark.parametrize("model_name,n,factors,msg", [
    ("beta-factor", 3, [(2, 0.2)], "less than the minimum level"),
    ("MGL", 3, [(2, 0.2), (4, 0.1)], "more than the number of members"),
    ("MGL", 4, [(2, 0.2), (4, 0.1)], "Missing some CCF factors"),
    ("alpha-factor", 3, [(1, 0.2), (1, 0.1)], "Redefinition of CCF factor"),
    ("phi-factor", 3, [(1, 0.5), (2, 0.2)], "must sum to 1"),
    ("beta-factor", 1, [(1, 0.2)], "at least 2 members"),
])
def test_ccf_validation(model_name, n, factors, msg):
    with pytest.raises(_lib.CanopyError) as ei:
        load_mef(text=ccf_xml(model_name, n, 0.1, factors, top="<null>{args}</null>" if n == 1 else "<or>{args}</or>"), ccf_analysis=True)
    assert ei.value.status == _lib.E_VALIDITY and msg in ei.value.message


# ---- event trees -----------------------------------------------------------------------------------------------------
ET_XML = model("""
 <define-initiating-event name="IE" event-tree="ET"/>
 <define-initiating-event name="unused"/>
 <define-rule name="cd-factor">
   <if><test-functional-event name="F1" state="fails"/><collect-expression><float value="0.25"/></collect-expression>
       <collect-expression><float value="0.0"/></collect-expression></if>
   <if><test-initiating-event name="IE"/><block/></if>
 </define-rule>
 <define-event-tree name="ET">
  <define-functional-event name="F1"/><define-functional-event name="F2"/>
  <define-sequence name="OK"/>
  <define-sequence name="CD"><rule name="cd-factor"/></define-sequence>
  <define-sequence name="PART"><collect-expression><float value="0.5"/></collect-expression></define-sequence>
  <define-sequence name="TO2"><event-tree name="ET2"/></define-sequence>
  <define-branch name="B2">
    <fork functional-event="F2">
      <path state="fails"><collect-formula><gate name="G2"/></collect-formula><sequence name="CD"/></path>
      <path state="works"><collect-formula><not><gate name="G2"/></not></collect-formula><sequence name="PART"/></path>
      <path state="bypass"><collect-formula><basic-event name="d"/></collect-formula><sequence name="TO2"/></path>
    </fork>
  </define-branch>
  <initial-state>
   <fork functional-event="F1">
    <path state="works"><collect-formula><not><gate name="G1"/></not></collect-formula><sequence name="OK"/></path>
    <path state="fails"><collect-formula><gate name="G1"/></collect-formula>
        <set-house-event name="H"><constant value="true"/></set-house-event><branch name="B2"/></path>
   </fork>
  </initial-state>
 </define-event-tree>
 <define-event-tree name="ET2">
  <define-functional-event name="F3"/>
  <define-sequence name="S3"/>
  <initial-state>
   <fork functional-event="F3">
    <path state="a"><collect-formula><and><basic-event name="a"/><basic-event name="c"/></and></collect-formula><sequence name="S3"/></path>
    <path state="b"><collect-formula><gate name="G2"/></collect-formula><sequence name="CD"/></path>
   </fork>
  </initial-state>
 </define-event-tree>
 <define-fault-tree name="FT">
  <define-gate name="G1"><or><basic-event name="a"/><basic-event name="b"/></or></define-gate>
  <define-gate name="G2"><and><basic-event name="c"/><gate name="G3"/></and></define-gate>
  <define-gate name="G3"><or><house-event name="H"/><basic-event name="e"/></or></define-gate>
 </define-fault-tree>
 <model-data>
  <define-basic-event name="a"><float value="0.1"/></define-basic-event>
  <define-basic-event name="b"><float value="0.2"/></define-basic-event>
  <define-basic-event name="c"><float value="0.3"/></define-basic-event>
  <define-basic-event name="d"><float value="0.4"/></define-basic-event>
  <define-basic-event name="e"><float value="0.05"/></define-basic-event>
  <define-house-event name="H"><constant value="false"/></define-house-event>
 </model-data>
""")


def test_event_tree_sequences():
    g = load_mef(text=ET_XML)
    names = g.meta["root_names"]
    # fault-tree top events first (G1 and G2 are not used inside FT; G3 is), then the sequences in the order reached
    assert names == ["G1", "G2", "IE.OK", "IE.CD", "IE.PART", "IE.S3"]
    pa, pb, pc, pd, pe = (Q(x) for x in (0.1, 0.2, 0.3, 0.4, 0.05))
    p_g1 = 1 - (1 - pa) * (1 - pb)
    got = dict(zip(names, O.exact(g)))
    assert got["G1"] == pytest.approx(p_g1, rel=1e-12)
    assert got["G2"] == pytest.approx(pc * pe, rel=1e-12)                  # declared state of H is false
    assert got["IE.OK"] == pytest.approx(1 - p_g1, rel=1e-12)
    # after <set-house-event H true> G3 is certain, so G2 reads as c on the paths below the first fork
    # CD is reached twice: F2 fails (x 0.25 from the rule: F1 is "fails" there), and through the link
    # TO2 -> ET2 path b, where the functional-event context of ET does not carry over (rule takes its else: x 0)
    assert got["IE.CD"] == pytest.approx(p_g1 * pc * Q(0.25), rel=1e-12)
    assert got["IE.PART"] == pytest.approx(p_g1 * (1 - pc) * Q(0.5), rel=1e-12)
    assert got["IE.S3"] == pytest.approx(pa * pc * pd, rel=1e-12)          # G1 & d & (a & c) = a & c & d
    # the house-event override produced clones of G2/G3, the originals are untouched
    assert any(n.startswith("G2@") for n in g.gate_names) and any(n.startswith("G3@") for n in g.gate_names)
    with Plan(g, tally_all_gates=True) as plan:
        assert np.array_equal(eval_lowered(plan.view(), g, 5, 64, 4096), O.mc_scalar(g, 5, 64, 4096, True))


@pytest.mark.parametrize("patch,status,msg", [
    (("<sequence name=\"OK\"/>", "<sequence name=\"NOPE\"/>"), _lib.E_VALIDITY, "Undefined Element Error: sequence NOPE"),
    (("<fork functional-event=\"F3\">", "<fork functional-event=\"F9\">"), _lib.E_VALIDITY, "functional event F9"),
    (("<path state=\"b\">", "<path state=\"a\">"), _lib.E_VALIDITY, "Duplicate state"),
    (("<branch name=\"B2\"/></path>", "<branch name=\"B9\"/></path>"), _lib.E_VALIDITY, "branch B9"),
    (("<sequence name=\"TO2\"/>", "<branch name=\"B2\"/>"), _lib.E_CYCLE, "branch B2"),
    (("<define-sequence name=\"S3\"/>", "<define-sequence name=\"S3\"><event-tree name=\"ET\"/></define-sequence>"), _lib.E_CYCLE, "links to itself"),
    (("<set-house-event name=\"H\">", "<set-house-event name=\"a\">"), _lib.E_VALIDITY, "house event a"),
    (("event-tree=\"ET\"", "event-tree=\"ETX\""), _lib.E_VALIDITY, "event tree ETX"),
    (("<collect-expression><float value=\"0.5\"/>", "<collect-expression><float value=\"1.5\"/>"), _lib.E_DOMAIN, "collected on a path"),
])
def test_event_tree_errors(patch, status, msg):
    assert patch[0] in ET_XML
    with pytest.raises(_lib.CanopyError) as ei:
        load_mef(text=ET_XML.replace(patch[0], patch[1], 1))
    assert ei.value.status == status and msg in ei.value.message


# ---- reader robustness ------------------------------------------------------------------------------------
def test_xml_nesting_depth_is_capped_and_reported():
    deep = "<a>" * 600 + "</a>" * 600
    with pytest.raises(_lib.CanopyError) as ei:
        load_mef(text=f"<?xml version='1.0'?><opsa-mef>{deep}</opsa-mef>")
    assert ei.value.status == _lib.E_IO and "nested deeper" in ei.value.message
    # a very deep input must come back as a status too, not as a stack overflow of the process
    with pytest.raises(_lib.CanopyError):
        load_mef(text="<?xml version='1.0'?><opsa-mef>" + "<x>" * 200000)


def test_numeric_character_references_are_utf8():
    xml = model("<define-fault-tree name='ft'><define-gate name='top'><or><basic-event name='pump&#xE9;'/>"
                "<basic-event name='&#26085;&#x672C;'/></or></define-gate></define-fault-tree><model-data>"
                + be("pump&#233;", fl(0.1)) + be("日本", fl(0.2)) + "</model-data>")
    g = load_mef(text=xml)
    assert g.basic_names == ["pumpé", "日本"] and g.n_gates == 1
    with pytest.raises(_lib.CanopyError) as ei:
        load_mef(text=model("<define-fault-tree name='f&#x110000;'/>"))
    assert ei.value.status == _lib.E_IO


# ---- roles and scopes (element.h:300-379; initializer.cpp:42-57, 712-758, 1378-1437) -----------------------------------
def test_private_roles_and_scoped_lookup():
    """Two private events of the same name in different components are two nodes; a reference resolves in its local scope
    first, then as a public id, then as a full path; a private element is not visible by its bare name elsewhere."""
    xml = model(
        "<define-fault-tree name='FT'>"
        "<define-gate name='top'><and><gate name='A.g'/><gate name='B.g'/><basic-event name='shared'/></and></define-gate>"
        "<define-component name='A' role='private'>"
        "  <define-gate name='g'><or><basic-event name='pump'/><basic-event name='shared'/></or></define-gate>"
        "  <define-basic-event name='pump'><parameter name='rate'/></define-basic-event>"
        "  <define-parameter name='rate'><float value='0.125'/></define-parameter>"
        "</define-component>"
        "<define-component name='B' role='private'>"
        "  <define-gate name='g'><or><basic-event name='pump'/><basic-event name='FT.A.pump'/></or></define-gate>"
        "  <define-basic-event name='pump'><parameter name='rate'/></define-basic-event>"
        "  <define-parameter name='rate'><float value='0.25'/></define-parameter>"
        "</define-component>"
        "<define-basic-event name='shared'><float value='0.5'/></define-basic-event>"
        "</define-fault-tree>")
    g = load_mef(text=xml)
    assert sorted(g.basic_names) == ["FT.A.pump", "FT.B.pump", "shared"]          # ids: private = full path
    assert sorted(g.gate_names) == ["FT.A.g", "FT.B.g", "top"]
    p = dict(zip(g.basic_names, g.p))
    assert p == {"FT.A.pump": 0.125, "FT.B.pump": 0.25, "shared": 0.5}           # each pump reads the rate of ITS scope
    assert g.meta["root_names"] == ["top"]
    # top = (pa | s) & (pb | pa) & s = s & (pa | pb)
    ex = O.exact(g, False, 32)[0]
    assert ex == pytest.approx(Q(0.5) * (1 - (1 - Q(0.125)) * (1 - Q(0.25))), rel=1e-12)
    # a private element is not reachable by its bare name from another scope
    bad = xml.replace("<basic-event name='FT.A.pump'/>", "<basic-event name='nosuch'/>").replace(
        "<gate name='A.g'/>", "<gate name='g'/>")
    with pytest.raises(_lib.CanopyError) as ei:
        load_mef(text=bad)
    assert ei.value.status == _lib.E_VALIDITY and "Undefined Element" in ei.value.message
    # role attribute on a single element overrides the container's; public elements collide across components
    dup = model("<define-fault-tree name='FT'><define-gate name='top'><or><basic-event name='e'/><basic-event name='C.e'/></or></define-gate>"
                "<define-basic-event name='e'><float value='0.1'/></define-basic-event>"
                "<define-component name='C'><define-basic-event name='e'><float value='0.2'/></define-basic-event></define-component>"
                "</define-fault-tree>")
    with pytest.raises(_lib.CanopyError) as ei:
        load_mef(text=dup)
    assert "Duplicate Element" in ei.value.message
    ok = dup.replace("<define-basic-event name='e'><float value='0.2'/>", "<define-basic-event name='e' role='private'><float value='0.2'/>")
    g2 = load_mef(text=ok)
    assert sorted(g2.basic_names) == ["FT.C.e", "e"]


def test_scoped_parameters_and_ccf_group_in_a_component():
    """Parameters resolve like events (local scope, public id, full path); a CCF group declared in a private component
    resolves its distribution / factors in the component's scope.  Its members are added to the model only, not to the
    full-path table (reference initializer.cpp:1324-1339), so members of a PRIVATE group cannot be referred to at all --
    kept as the reference has it."""
    def xml(group_role):
        return model(
            "<define-fault-tree name='FT'>"
            "<define-gate name='top'><or><basic-event name='x'/><basic-event name='y'/><gate name='C.two'/></or></define-gate>"
            "<define-basic-event name='x'><parameter name='FT.C.q'/></define-basic-event>"          # full path into a private scope
            "<define-basic-event name='y'><parameter name='q'/></define-basic-event>"               # the public q of model-data
            "<define-component name='C' role='private'>"
            "  <define-parameter name='q'><float value='0.125'/></define-parameter>"
            "  <define-parameter name='beta'><float value='0.25'/></define-parameter>"
            "  <define-gate name='two'><and><basic-event name='m1'/><basic-event name='m2'/></and></define-gate>"
            f"  <define-CCF-group name='grp' model='beta-factor'{group_role}><members><basic-event name='m1'/><basic-event name='m2'/></members>"
            "    <distribution><parameter name='q'/></distribution><factor level='2'><parameter name='beta'/></factor></define-CCF-group>"
            "</define-component></define-fault-tree>"
            "<model-data><define-parameter name='q'><float value='0.5'/></define-parameter></model-data>")
    g = load_mef(text=xml(" role='public'"), ccf_analysis=True)
    p = dict(zip(g.basic_names, g.p))
    assert p["x"] == 0.125 and p["y"] == 0.5
    # the group's q and beta are the component's (0.125, 0.25), not the public q
    assert p["[m1]"] == pytest.approx(0.75 * 0.125) and p["[m2]"] == pytest.approx(0.75 * 0.125)
    assert p["[m1 m2]"] == pytest.approx(0.25 * 0.125)
    assert "FT.C.two" in g.gate_names and "m1" in g.gate_names            # the members' proxy OR gates
    assert np.array_equal(g.meta["psource"].mean(), g.p)                  # the sampled source sees the same scoping
    with pytest.raises(_lib.CanopyError) as ei:                           # private group: its members have no reachable name
        load_mef(text=xml(""), ccf_analysis=True)
    assert ei.value.status == _lib.E_VALIDITY and "Undefined Element" in ei.value.message


_UNUSED_PARAM_MODEL = """<opsa-mef>
<define-fault-tree name="FT"><define-gate name="top"><or><basic-event name="a"/><basic-event name="b"/></or></define-gate></define-fault-tree>
<model-data>
  <define-basic-event name="a"><float value="0.1"/></define-basic-event>
  <define-basic-event name="b"><parameter name="ok"/></define-basic-event>
  <define-parameter name="ok"><float value="0.2"/></define-parameter>
  %s
</model-data></opsa-mef>"""


@pytest.mark.parametrize("extra,status,msg", [
    ('<define-parameter name="u"><parameter name="nope"/></define-parameter>', _lib.E_VALIDITY, "Undefined Element Error"),
    ('<define-parameter name="u"><parameter name="v"/></define-parameter>'
     '<define-parameter name="v"><parameter name="u"/></define-parameter>', _lib.E_CYCLE, "Cycle Error"),
    ('<define-parameter name="u"><div><float value="1"/><float value="0"/></div></define-parameter>', _lib.E_DOMAIN, "Division by 0"),
])
def test_parameters_nothing_refers_to_are_still_defined_and_validated(extra, status, msg):
    """The reference defines every registered parameter, cycle-checks the whole table and validates every expression
    (initializer.cpp:668-681, 1860-1874): an invalid parameter is an error even when no event uses it."""
    with pytest.raises(_lib.CanopyError) as e:
        load_mef(text=_UNUSED_PARAM_MODEL % extra)
    assert e.value.status == status and msg in e.value.message


def test_a_valid_unused_parameter_changes_nothing():
    g0 = load_mef(text=_UNUSED_PARAM_MODEL % "")
    g1 = load_mef(text=_UNUSED_PARAM_MODEL % '<define-parameter name="u"><add><parameter name="ok"/><float value="1"/></add></define-parameter>')
    assert np.array_equal(g0.p, g1.p) and np.array_equal(g0.child, g1.child)
    assert np.array_equal(g0.meta["psource"].sample(3, 5), g1.meta["psource"].sample(3, 5))


def test_typed_reference_is_not_shadowed_by_a_local_element_of_another_kind():
    """<gate name='x'/> is looked up in the gate tables only (GetGate -> GetEntity, initializer.cpp:1372-1403): a private
    basic event FT.x in the referring scope does not hide the public gate x of another tree.  The untyped <event name='x'/>
    takes the local element (GetEvent, :1416-1437)."""
    xml = model(
        "<define-fault-tree name='Other'><define-gate name='x'><and><basic-event name='a'/><basic-event name='b'/></and></define-gate>"
        "<define-basic-event name='a'><float value='0.5'/></define-basic-event>"
        "<define-basic-event name='b'><float value='0.25'/></define-basic-event></define-fault-tree>"
        "<define-fault-tree name='FT'>"
        "<define-gate name='top'><or><gate name='x'/><basic-event name='c'/></or></define-gate>"
        "<define-basic-event name='x' role='private'><float value='0.0625'/></define-basic-event>"
        "<define-basic-event name='c'><float value='0.125'/></define-basic-event>"
        "</define-fault-tree>")
    g = load_mef(text=xml)
    top = g.meta["root_names"].index("top")
    assert O.exact(g, False, 32)[top] == pytest.approx(1 - (1 - Q(0.5) * Q(0.25)) * (1 - Q(0.125)), rel=1e-12)
    g = load_mef(text=xml.replace("<gate name='x'/>", "<event name='x'/>"))       # untyped: the local basic event FT.x
    top = g.meta["root_names"].index("top")
    assert O.exact(g, False, 32)[top] == pytest.approx(1 - (1 - Q(0.0625)) * (1 - Q(0.125)), rel=1e-12)
    with pytest.raises(_lib.CanopyError) as ei:                                   # typed and of the wrong kind everywhere
        load_mef(text=xml.replace("<gate name='x'/>", "<house-event name='x'/>"))
    assert ei.value.status == _lib.E_VALIDITY and "is not a house-event" in ei.value.message
