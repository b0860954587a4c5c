"""Deterministic synthetic fault-tree / event-tree DAG generator for BASELINE.json configs c2..c5.

The reference ships no models, examples or fixtures (tests/, examples/, benchmarks/ hold one empty
CMakeLists.txt each), so the benchmark inputs are authored here, to the law written in
SURVEY.md 8d:
  seed 0xC0FFEE + config id (Python random.Random); gates created layer by layer; fan-in from a truncated
  discrete Pareto(alpha 1.5, min 2, max F); children without replacement, 70 % from the layer directly
  below and 30 % uniform over all lower nodes; connective mix 45 % or / 33 % and / 10 % atleast (k uniform
  in [2, n-1], n >= 3) / 4 % not / 2 % xor / 3 % nand / 3 % nor; 2 % complemented edges; p log-uniform in
  [1e-4, 1e-1].
Two additions keep the result a useful benchmark instead of dead code or a constant:
  (1) nodes left without a parent are attached to a random higher gate with spare fan-in, so (almost)
      nothing is removed by dead-code elimination;
  (2) and<->or is flipped when an independence estimate of the gate probability would leave
      [1e-5, 0.6], so tallies stay non-degenerate through 30+ layers.
"""
from __future__ import annotations

import math
import random

import numpy as np

from .graph import (AND, ATLEAST, COMPLEMENT, NAND, NOR, NOT, OR, XOR, Graph)

CONFIGS = {
    # name: (config id, B, G, layers, max fan-in, samples, roots)
    "c2": dict(cid=2, n_basic=10_000, n_gates=5_000, n_layers=12, max_fanin=8, n_samples=1 << 26, n_seq=0),
    "c3": dict(cid=3, n_basic=30_000, n_gates=20_000, n_layers=16, max_fanin=8, n_samples=1 << 28, n_seq=200),
    "c4": dict(cid=4, n_basic=100_000, n_gates=60_000, n_layers=36, max_fanin=256, n_samples=1 << 30, n_seq=0),
    "c5": dict(cid=4, n_basic=100_000, n_gates=60_000, n_layers=36, max_fanin=256, n_samples=1 << 36, n_seq=0),
}


def _layer_sizes(n_gates: int, n_layers: int) -> list[int]:
    """Geometric shrink from a wide bottom layer to a single top gate, summing to n_gates."""
    if n_layers == 1:
        return [n_gates]
    lo, hi = 1.0, float(n_gates)
    for _ in range(80):
        n1 = 0.5 * (lo + hi)
        r = n1 ** (-1.0 / (n_layers - 1))
        tot = sum(n1 * r ** i for i in range(n_layers))
        if tot < n_gates:
            lo = n1
        else:
            hi = n1
    r = n1 ** (-1.0 / (n_layers - 1))
    sizes = [max(1, int(round(n1 * r ** i))) for i in range(n_layers)]
    sizes[-1] = 1
    sizes[0] += n_gates - sum(sizes)
    assert sizes[0] >= 1 and sum(sizes) == n_gates
    return sizes


def _pareto_fanin(rng: random.Random, fmax: int, alpha: float = 1.5) -> int:
    # inverse CDF of a Pareto with x_min = 2, truncated at fmax, floored to an integer
    u = rng.random()
    tail = (2.0 / (fmax + 1.0)) ** alpha
    x = 2.0 / ((1.0 - u * (1.0 - tail)) ** (1.0 / alpha))
    return max(2, min(fmax, int(x)))


def _atleast_prob(qs: list[float], k: int) -> float:
    # Poisson-binomial tail P(count >= k) under independence
    dist = [1.0] + [0.0] * k  # dist[j] = P(count == j), j == k absorbs ">= k"
    for q in qs:
        for j in range(k, 0, -1):
            if j == k:
                dist[j] = dist[j] + dist[j - 1] * q
            else:
                dist[j] = dist[j] * (1 - q) + dist[j - 1] * q
        dist[0] *= (1 - q)
    return dist[k]


def generate(n_basic: int, n_gates: int, n_layers: int, max_fanin: int, seed: int, n_seq: int = 0) -> Graph:
    rng = random.Random(seed)
    p = [10.0 ** (-4.0 + 3.0 * rng.random()) for _ in range(n_basic)]
    sizes = _layer_sizes(n_gates, n_layers)
    # node ids: events [0,B), gates B + index; gate index increases with layer
    layer_nodes: list[list[int]] = [list(range(n_basic))]
    q = list(p)                       # independence estimate of P(node)
    ops: list[int] = []
    kmins: list[int] = []
    children: list[list[int]] = []    # encoded child codes per gate
    has_parent = bytearray(n_basic + n_gates)
    gate_layer: list[int] = []
    n_lower = n_basic
    for li, size in enumerate(sizes, start=1):
        below = layer_nodes[li - 1]
        this_layer = []
        for _ in range(size):
            g = len(ops)
            u = rng.random()
            if u < 0.45: op = OR
            elif u < 0.78: op = AND
            elif u < 0.88: op = ATLEAST
            elif u < 0.92: op = NOT
            elif u < 0.94: op = XOR
            elif u < 0.97: op = NAND
            else: op = NOR
            n = 1 if op == NOT else 2 if op == XOR else _pareto_fanin(rng, max_fanin)
            n = min(n, n_lower)
            if op == ATLEAST and n < 3:
                op = OR
            if op in (OR, AND, NAND, NOR) and n < 2:
                op, n = NOT, 1
            chosen: list[int] = []
            seen = set()
            guard = 0
            while len(chosen) < n and guard < 64 * n:
                guard += 1
                if rng.random() < 0.7:
                    c = below[int(rng.random() * len(below))]
                else:
                    c = int(rng.random() * n_lower)  # events and all earlier gates share one contiguous id space
                if c in seen:
                    continue
                seen.add(c)
                chosen.append(c)
            n = len(chosen)
            if op == XOR and n < 2:
                op = NOT
            if op == ATLEAST and n < 3:
                op = OR if n >= 2 else NOT
            if op in (OR, AND, NAND, NOR) and n < 2:
                op = NOT
            comp = [False if op == NOT else (rng.random() < 0.02) for _ in chosen]
            qs = [(1 - q[c]) if cm else q[c] for c, cm in zip(chosen, comp)]
            k = 0
            if op == ATLEAST:
                k = 2 + int(rng.random() * (n - 2))  # uniform in [2, n-1]
            # independence estimate, and the and<->or flip that keeps probabilities in band
            def est(o):
                if o == AND: return math.prod(qs)
                if o == OR: return 1 - math.prod(1 - x for x in qs)
                if o == NAND: return 1 - math.prod(qs)
                if o == NOR: return math.prod(1 - x for x in qs)
                if o == XOR: return qs[0] * (1 - qs[1]) + qs[1] * (1 - qs[0])
                if o == NOT: return 1 - qs[0]
                return _atleast_prob(qs, k)
            pr = est(op)
            if op == AND and pr < 1e-5: op = OR
            elif op == OR and pr > 0.6: op = AND
            elif op == NAND and pr > 0.6 and est(NOR) < pr: op = NOR
            pr = est(op)
            ops.append(op); kmins.append(k)
            children.append([c | (COMPLEMENT if cm else 0) for c, cm in zip(chosen, comp)])
            for c in chosen:
                has_parent[c] = 1
            q.append(min(max(pr, 0.0), 1.0))
            gate_layer.append(li)
            this_layer.append(n_basic + g)
        layer_nodes.append(this_layer)
        n_lower += size
    # (1) attach orphans (all nodes except the top gate) to a random higher gate with spare fan-in
    first_gate_of_layer = [0]
    for s in sizes:
        first_gate_of_layer.append(first_gate_of_layer[-1] + s)
    top = n_gates - 1
    for node in range(n_basic + n_gates - 1):
        if has_parent[node]:
            continue
        lyr = 0 if node < n_basic else gate_layer[node - n_basic]
        lo_g = first_gate_of_layer[lyr]  # first gate index strictly above `lyr`
        for _ in range(200):
            g = lo_g + int(rng.random() * (n_gates - lo_g))
            if ops[g] in (NOT, XOR) or len(children[g]) >= max_fanin:
                continue
            if any((c & 0x7FFFFFFF) == node for c in children[g]):
                continue
            children[g].append(node)
            has_parent[node] = 1
            break
    roots = [n_basic + top]
    # c3: event-tree sequences = AND over 3..6 (possibly complemented) mid-level gates of the shared DAG
    if n_seq:
        mid_lo = first_gate_of_layer[len(sizes) // 3]
        mid_hi = first_gate_of_layer[2 * len(sizes) // 3 + 1]
        for _ in range(n_seq):
            n = 3 + int(rng.random() * 4)
            seen = set()
            while len(seen) < n:
                seen.add(mid_lo + int(rng.random() * (mid_hi - mid_lo)))
            ops.append(AND); kmins.append(0)
            children.append([(n_basic + gi) | (COMPLEMENT if rng.random() < 0.3 else 0) for gi in sorted(seen)])
            roots.append(n_basic + len(ops) - 1)
    off = [0]
    flat: list[int] = []
    for ch in children:
        flat.extend(ch)
        off.append(len(flat))
    g = Graph(p=np.array(p), op=np.array(ops, dtype=np.uint8), kmin=np.array(kmins, dtype=np.uint16),
              kmax=np.zeros(len(ops), dtype=np.uint16), child_off=np.array(off, dtype=np.uint64),
              child=np.array(flat, dtype=np.uint32), roots=np.array(roots, dtype=np.uint32))
    g.meta = dict(seed=seed, n_layers_requested=n_layers, max_fanin=max_fanin, est_p_top=q[n_basic + top],
                  est_q=np.array(q[n_basic:n_basic + n_gates]))
    return g


def config_graph(name: str, scale: float = 1.0) -> Graph:
    """Graph of a BASELINE.json config; scale < 1 shrinks B and G proportionally (parity-test sizes)."""
    c = CONFIGS[name]
    nb = max(8, int(c["n_basic"] * scale))
    ng = max(4, int(c["n_gates"] * scale))
    nl = c["n_layers"] if scale == 1.0 else max(3, min(c["n_layers"], int(math.log2(ng)) + 2))
    g = generate(nb, ng, nl, c["max_fanin"], 0xC0FFEE + c["cid"], int(c["n_seq"] * (scale if scale < 1 else 1)) if c["n_seq"] else 0)
    g.meta.update(config=name, scale=scale, n_samples=c["n_samples"])
    return g


# ---- config 3 as BASELINE.json words it: an EVENT TREE whose sequences share one gate DAG, as Open-PSA MEF text ----------
_MEF_CONNECTIVE = {AND: "and", OR: "or", ATLEAST: "atleast", XOR: "xor", NOT: "not", NAND: "nand", NOR: "nor"}


def mef_xml(g: Graph, n_seq: int = 0, seed: int = 1, ft_name: str = "FT") -> str:
    """Open-PSA MEF text of a generated DAG: one fault tree (gate i = "g<i>", basic event j = "e<j>", complemented
    edges as <not>), and -- with n_seq > 0 -- an event tree "ET" with initiating event "IE": a binary tree of forks over
    functional events F0, F1, ...; the path taken at depth d collects (state "fails") a mid-level gate of the DAG or
    (state "works") its negation; each of the first n_seq leaves ends in its own sequence S<k>, the others in "REST".
    A sequence formula is therefore the conjunction of 8 or so (possibly negated) gates of the one shared DAG, which the
    library's event-tree walker (csrc/mef_model.cpp; reference data model event_tree.h:65-197, instruction.h:38-83)
    has to derive.  Only the graph's own roots argument is ignored: the loader finds the top events itself."""
    nb = g.n_basic
    out = ["<?xml version='1.0'?>\n<opsa-mef>\n"]

    def arg(code: int) -> str:
        node = code & 0x7FFFFFFF
        ref = f"<basic-event name='e{node}'/>" if node < nb else f"<gate name='g{node - nb}'/>"
        return f"<not>{ref}</not>" if code & COMPLEMENT else ref

    if n_seq:
        rng = random.Random(seed)
        depth = max(1, math.ceil(math.log2(n_seq)))
        lo, hi = g.n_gates // 3, 2 * g.n_gates // 3
        # one collected gate per functional event; gates whose independence estimate is moderate, so that both
        # branches of every fork are taken and many sequences do occur
        est = g.meta.get("est_q")
        pool = [i for i in range(lo, hi) if est is None or 0.1 <= est[i] <= 0.9] or list(range(lo, hi))
        fork_gate = rng.sample(pool, depth) if len(pool) >= depth else [pool[int(rng.random() * len(pool))] for _ in range(depth)]
        out.append("<define-initiating-event name='IE' event-tree='ET'/>\n<define-event-tree name='ET'>\n")
        out += [f"<define-functional-event name='F{d}'/>\n" for d in range(depth)]
        out += [f"<define-sequence name='S{k}'/>\n" for k in range(n_seq)]
        out.append("<define-sequence name='REST'/>\n<initial-state>\n")
        leaf = [0]

        def fork(d: int) -> str:
            if d == depth:
                k = leaf[0]; leaf[0] += 1
                return f"<sequence name='{'S%d' % k if k < n_seq else 'REST'}'/>"
            gate = f"<gate name='g{fork_gate[d]}'/>"
            return (f"<fork functional-event='F{d}'>"
                    f"<path state='fails'><collect-formula>{gate}</collect-formula>{fork(d + 1)}</path>"
                    f"<path state='works'><collect-formula><not>{gate}</not></collect-formula>{fork(d + 1)}</path></fork>")
        out.append(fork(0))
        out.append("\n</initial-state>\n</define-event-tree>\n")
    out.append(f"<define-fault-tree name='{ft_name}'>\n")
    off = g.child_off
    for i in range(g.n_gates):
        op = int(g.op[i])
        kids = "".join(arg(int(c)) for c in g.child[int(off[i]):int(off[i + 1])])
        tag = _MEF_CONNECTIVE[op]
        attr = f" min='{int(g.kmin[i])}'" if op == ATLEAST else ""
        out.append(f"<define-gate name='g{i}'><{tag}{attr}>{kids}</{tag}></define-gate>\n")
    out.append("</define-fault-tree>\n<model-data>\n")
    out += [f"<define-basic-event name='e{j}'><float value='{float(g.p[j])!r}'/></define-basic-event>\n" for j in range(nb)]
    out.append("</model-data>\n</opsa-mef>\n")
    return "".join(out)
