"""A SECOND, independent restatement of the sampling specs and of the connective semantics, in plain Python integers.

Written from the text of DESIGN.md section 1 only (per-sample DEFINITIONS, no lazy or streaming forms) and sharing no code
with oracle/canopy_oracle.c: the C oracle, this file and the CUDA kernels are three implementations of one written spec.
tests/test_restatement.py checks the oracle's words and tallies against this file; it is slow on purpose (every sample of a
group is rebuilt from Philox calls) and only ever runs on small cases.  TEST INFRASTRUCTURE ONLY.
"""
from __future__ import annotations

from functools import lru_cache

M0, M1, W0, W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85
MASK32 = 0xFFFFFFFF

H = 7          # CANOPY_SLICED_HEAD_BITS
Z = 5          # CANOPY_SPARSE_MIN_ZEROS
KMAX = 13      # CANOPY_SPARSE_MAX_LOG2
CDF_LEN = 32   # CANOPY_SPARSE_CDF_LEN


@lru_cache(maxsize=1 << 16)
def philox(c0: int, c1: int, c2: int, c3: int, k0: int, k1: int):
    """Philox4x32-10: (c0,c1,c2,c3) <- (hi(M1 c2)^c1^k0, lo(M1 c2), hi(M0 c0)^c3^k1, lo(M0 c0)); keys += (W0, W1); 10 rounds."""
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2
        c0, c1, c2, c3 = (p1 >> 32) ^ c1 ^ k0, p1 & MASK32, (p0 >> 32) ^ c3 ^ k1, p0 & MASK32
        k0, k1 = (k0 + W0) & MASK32, (k1 + W1) & MASK32
    return c0, c1, c2, c3


def call(seed: int, e: int, g: int, c3: int):
    """One Philox call of the specs: key = (seed lo, seed hi), counter = (g lo, g hi, event, c3)."""
    return philox(g & MASK32, (g >> 32) & MASK32, e, c3, seed & MASK32, (seed >> 32) & MASK32)


def threshold(p: float, bits: int):
    """('false' | 'true' | 'sampled', thr): thr = floor(p 2^bits); p <= 0 false, p >= 1 true, thr == 0 can never fire."""
    if p <= 0.0:
        return "false", 0
    if p >= 1.0:
        return "true", 0
    num, den = p.as_integer_ratio()          # exact
    return "sampled", (num << bits) // den


def bit_spec1(seed, e, s, thr, bits):
    if bits == 32:
        return int(call(seed, e, s >> 2, 0)[s & 3] < thr)
    o = call(seed, e, s >> 1, 1)
    j = 2 * (s & 1)
    return int(((o[j + 1] << 32) | o[j]) < thr)


def bit_spec2(seed, e, s, thr, bits):
    """u(e, s) assembled MSB first: bit (bits-1-k) of u is bit t of word j of call 0x100+k, for k < H; the samples of the
    group whose top H bits equal the threshold's take their remaining bits, in ascending sample order, from the tail calls."""
    g, i = s >> 7, s & 127
    j, t = i >> 5, i & 31

    def head(ii):
        jj, tt = ii >> 5, ii & 31
        v = 0
        for k in range(H):
            v = (v << 1) | ((call(seed, e, g, 0x100 + k)[jj] >> tt) & 1)
        return v
    thr_head = thr >> (bits - H)
    mine = head(i)
    if mine != thr_head:
        return int(mine < thr_head)
    rank = sum(1 for ii in range(i) if head(ii) == thr_head)
    low_mask = (1 << (bits - H)) - 1
    if bits == 32:
        v = call(seed, e, g, 0x200 + (rank >> 2))[rank & 3]
    else:
        o = call(seed, e, g, 0x300 + (rank >> 1))
        v = (o[2 * (rank & 1) + 1] << 32) | o[2 * (rank & 1)]
    return int((v >> H) < (thr & low_mask))


def class_k(thr64: int) -> int:
    """0 for a dense event; else the group exponent k = min(KMAX, z + 2), z = leading zero bits of the 64-bit threshold."""
    if thr64 == 0:
        return 0
    z = 64 - thr64.bit_length()
    return 0 if z < Z else min(KMAX, z + 2)


@lru_cache(maxsize=4096)
def count_table(thr64: int, k: int):
    q = (1 << 64) - thr64
    r = (thr64 << 64) // q
    a = q
    for _ in range(k):
        a = (a * a) >> 64
    acc, out = 0, []
    for j in range(CDF_LEN):
        acc += a
        out.append(min(acc, (1 << 64) - 1))
        a = ((((a * r) >> 51) * ((1 << k) - j)) // (j + 1)) >> 13
    return tuple(out)


def field(seed, e, g, k, i):
    fpw = 32 // k
    if i < 2 * fpw:
        word = call(seed, e, g, 0x400)[2 + i // fpw]
    else:
        i2 = i - 2 * fpw
        word = call(seed, e, g, 0x401 + i2 // (4 * fpw))[(i2 // fpw) % 4]
    return (word >> (k * (i % fpw))) & ((1 << k) - 1)


@lru_cache(maxsize=1 << 14)
def group_ones(seed, e, g, k, table):
    """The set of positions of the ones of group g: the first N distinct fields, N = #{j : U >= cdf_j}."""
    a = call(seed, e, g, 0x400)
    u = (a[1] << 32) | a[0]
    n = sum(1 for c in table if u >= c)
    ones, i = [], 0
    while len(ones) < n:
        pos = field(seed, e, g, k, i)
        i += 1
        if pos not in ones:
            ones.append(pos)
    return frozenset(ones)


def bit_spec4(seed, e, s, thr, bits, table=None):
    thr64 = thr << 32 if bits == 32 else thr
    k = class_k(thr64)
    if k == 0:
        return bit_spec2(seed, e, s, thr, bits)
    return int((s & ((1 << k) - 1)) in group_ones(seed, e, s >> k, k, table or count_table(thr64, k)))


def sample_bit(seed, e, s, p, bits=32, spec=4):
    kind, thr = threshold(p, bits)
    if kind != "sampled":
        return int(kind == "true")
    if thr == 0:
        return 0
    return {1: bit_spec1, 2: bit_spec2, 4: bit_spec4}[spec](seed, e, s, thr, bits)


def sample_word(seed, e, w, p, bits=32, spec=4):
    return sum(sample_bit(seed, e, 64 * w + b, p, bits, spec) << b for b in range(64))


# ---- connectives (reference event/event.h:143-157) and a scalar Monte Carlo on the descriptor -----------
def connective(op, x, kmin, kmax):
    c, n = sum(x), len(x)
    return [c == n, c > 0, c >= kmin, x[0] != x[1] if n > 1 else False, not x[0], c != n, c == 0, bool(x[0]),
            x[0] == x[1] if n > 1 else False, (not x[0]) or x[1] if n > 1 else False, kmin <= c <= kmax][op]


def mc(graph, seed, begin, n, tally_all=False, bits=32, spec=4):
    """tallies[i] = #{s in [begin, begin+n) : output_i(s)}; outputs = roots, then (tally_all) gates, descriptor order."""
    B, G = graph.n_basic, graph.n_gates
    off, ch = [int(x) for x in graph.child_off], [int(x) for x in graph.child]
    p = [float(x) for x in graph.p]
    roots = [int(r) for r in graph.roots]
    out = [0] * (len(roots) + (G if tally_all else 0))
    for s in range(begin, begin + n):
        val = {}

        def node(i):
            if i == 0x7FFFFFFF:
                return True
            if i == 0x7FFFFFFE:
                return False
            if i not in val:
                if i < B:
                    val[i] = bool(sample_bit(seed, i, s, p[i], bits, spec))
                else:
                    g = i - B
                    xs = [node(c & 0x7FFFFFFF) != bool(c >> 31) for c in ch[off[g]:off[g + 1]]]
                    val[i] = bool(connective(int(graph.op[g]), xs, int(graph.kmin[g]), int(graph.kmax[g])))
            return val[i]
        for r, rid in enumerate(roots):
            out[r] += node(rid)
        if tally_all:
            for g in range(G):
                out[len(roots) + g] += node(B + g)
    return out
