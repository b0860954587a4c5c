"""Test-only helpers: random small graphs, and a numpy evaluator of the LOWERED form (so the host lowering
pass can be checked against the oracle on a CPU box, without ever running the product on a CPU)."""
from __future__ import annotations

import os
import random
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from canopy_b200 import graph as G  # noqa: E402
from oracle import oracle as O  # noqa: E402


def random_graph(seed: int, n_basic: int = 12, n_gates: int = 20, max_fanin: int = 6, consts: bool = True,
                 p_lo: float = 0.05, p_hi: float = 0.6, n_roots: int = 2, extreme_p: bool = True) -> G.Graph:
    """Random valid DAG exercising all 11 connectives, complements, constants and p in {0, 1}."""
    rng = random.Random(seed)
    p = [p_lo + (p_hi - p_lo) * rng.random() for _ in range(n_basic)]
    if extreme_p and n_basic >= 4:
        p[rng.randrange(n_basic)] = 0.0
        p[rng.randrange(n_basic)] = 1.0
    gates = []
    for g in range(n_gates):
        pool = list(range(n_basic + g))
        op = rng.choice([G.AND, G.OR, G.ATLEAST, G.XOR, G.NOT, G.NAND, G.NOR, G.NULL, G.IFF, G.IMPLY, G.CARDINALITY,
                         G.AND, G.OR, G.ATLEAST])
        if op in (G.NOT, G.NULL):
            n = 1
        elif op in (G.XOR, G.IFF, G.IMPLY):
            n = 2
        elif op == G.ATLEAST:
            n = rng.randint(3, max(3, max_fanin))
        elif op == G.CARDINALITY:
            n = rng.randint(1, max_fanin)
        else:
            n = rng.randint(2, max_fanin)
        n = min(n, len(pool))
        if op == G.ATLEAST and n < 3:
            op = G.OR
        if op in (G.AND, G.OR, G.NAND, G.NOR) and n < 2:
            op, n = G.NULL, 1
        ids = rng.sample(pool, n)
        ch = []
        for i, c in enumerate(ids):
            code = c
            if consts and op != G.NOT and rng.random() < 0.06:
                other = G.CONST_TRUE if rng.random() < 0.5 else G.CONST_FALSE
                if other not in [x & G.ID_MASK for x in ch]:
                    code = other
            if op not in (G.NOT, G.NULL) and rng.random() < 0.2:
                code |= G.COMPLEMENT
            ch.append(code)
        lo = hi = 0
        if op == G.ATLEAST:
            lo = rng.randint(2, n - 1)
        if op == G.CARDINALITY:
            hi = rng.randint(0, n)
            lo = rng.randint(0, hi)
        gates.append((op, ch, lo, hi))
    roots = [n_basic + n_gates - 1] + [rng.randrange(n_basic + n_gates) for _ in range(n_roots - 1)]
    return G.from_lists(p, gates, roots)


def eval_lowered(view: dict, graph: G.Graph, seed: int, begin: int, n: int, bits: int = 32, spec: int = O.SPEC) -> np.ndarray:
    """Evaluate the lowered CSR arrays with numpy on 64-bit words; event words come from the oracle sampler."""
    first_word, last_word = begin // 64, (begin + n - 1) // 64
    nw = last_word - first_word + 1
    n_ev, n_g = view["n_events"], view["n_gates"]
    vals = np.zeros((n_ev + n_g, nw), dtype=np.uint64)
    if n_ev:
        vals[:n_ev] = O.sample_words(graph, seed, view["event_orig"], first_word, nw, bits, spec, fast=True)
    mask = np.full(nw, ~np.uint64(0), dtype=np.uint64)
    s0 = first_word * 64
    if begin > s0:
        mask[0] &= ~np.uint64(0) << np.uint64(begin - s0)
    tail = (begin + n) - last_word * 64
    if tail < 64:
        mask[-1] &= (np.uint64(1) << np.uint64(tail)) - np.uint64(1)
    ones = ~np.uint64(0)
    gate_t = np.zeros(n_g, dtype=np.uint64)
    off, ch = view["gate_child_off"], view["gate_child"]
    lay = view["layer_off"]
    assert lay[0] == 0 and (n_g == 0 or lay[-1] == n_g)
    for l in range(view["n_layers"]):
        for g in range(lay[l], lay[l + 1]):
            xs = []
            for code in ch[off[g]:off[g + 1]]:
                node = int(code) & 0x7FFFFFFF
                assert node < n_ev or node - n_ev < lay[l], "child must come from an earlier layer"
                v = vals[node]
                xs.append(~v if int(code) >> 31 else v)
            op = view["gate_op"][g]
            if op == 0:
                r = np.full(nw, ones, dtype=np.uint64)
                for x in xs:
                    r &= x
            elif op == 1:
                r = np.zeros(nw, dtype=np.uint64)
                for x in xs:
                    r |= x
            elif op == 2:
                r = np.zeros(nw, dtype=np.uint64)
                for x in xs:
                    r ^= x
            else:
                lo, hi = int(view["gate_kmin"][g]), int(view["gate_kmax"][g])
                cnt = np.zeros((nw, 64), dtype=np.int32)
                for x in xs:
                    cnt += np.unpackbits(x.view(np.uint8).reshape(nw, 8), axis=1, bitorder="little").reshape(nw, 64)
                ok = (cnt >= lo) & (cnt <= hi)
                r = np.packbits(ok.astype(np.uint8), axis=1, bitorder="little").view(np.uint64).reshape(nw)
            vals[n_ev + g] = r
            gate_t[g] = sum(bin(int(w)).count("1") for w in (r & mask))
    out = np.zeros(view["n_tallies"], dtype=np.uint64)
    for i in range(view["n_tallies"]):
        k, gi = view["tally_kind"][i], view["tally_gate"][i]
        out[i] = 0 if k == 0 else n if k == 1 else gate_t[gi] if k == 2 else n - gate_t[gi]
    return out
