"""The C oracle against the independent pure-Python restatement of the specs (tests/spec_restatement.py), and the Philox
rounds against cuRAND's own `curand_Philox4x32_10` compiled from the toolkit header (SURVEY.md 8c ladder item 1)."""
import os
import random
import subprocess

import numpy as np
import pytest

import spec_restatement as R
from helpers import G, O, ROOT, random_graph


@pytest.mark.parametrize("ctr,key", [((0, 0, 0, 0), (0, 0)), ((0xFFFFFFFF,) * 4, (0xFFFFFFFF,) * 2),
                                     ((0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344), (0xA4093822, 0x299F31D0))])
def test_philox_restatement_kat(ctr, key):
    assert list(R.philox(*ctr, *key)) == O.philox4x32_10(ctr, key)


def test_philox_random_counters_three_ways():
    rng = random.Random(5)
    for _ in range(200):
        ctr = [rng.getrandbits(32) for _ in range(4)]
        key = [rng.getrandbits(32) for _ in range(2)]
        assert list(R.philox(*ctr, *key)) == O.philox4x32_10(ctr, key)


CURAND_SRC = r"""
#include <cstdio>
#include <cstdlib>
#define QUALIFIERS static inline __host__ __device__   /* the header defaults to __device__ only */
#include <curand_philox4x32_x.h>
int main(int argc, char** argv) {
    // stdin-free: counters and keys from argv, one call per 6 numbers; prints the 4 output words
    for (int i = 1; i + 5 < argc; i += 6) {
        uint4 c = make_uint4(strtoul(argv[i], 0, 0), strtoul(argv[i + 1], 0, 0), strtoul(argv[i + 2], 0, 0), strtoul(argv[i + 3], 0, 0));
        uint2 k = make_uint2(strtoul(argv[i + 4], 0, 0), strtoul(argv[i + 5], 0, 0));
        uint4 r = curand_Philox4x32_10(c, k);
        printf("%u %u %u %u\n", r.x, r.y, r.z, r.w);
    }
    return 0;
}
"""


def test_philox_against_curand_header(tmp_path):
    """curand_Philox4x32_10(uint4, uint2) is __host__ __device__ in the toolkit's curand_philox4x32_x.h: compiled with
    nvcc and run on the host, it is an implementation of the rounds that nobody in this repository wrote."""
    src = tmp_path / "kat.cu"
    src.write_text(CURAND_SRC)
    exe = tmp_path / "kat"
    try:
        subprocess.check_call(["nvcc", "-Wno-deprecated-gpu-targets", "-O1", "-o", str(exe), str(src)], timeout=300)
    except (FileNotFoundError, subprocess.CalledProcessError):
        pytest.skip("nvcc / curand header not usable here")
    rng = random.Random(11)
    cases = [([0, 0, 0, 0], [0, 0]), ([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2)] + \
            [([rng.getrandbits(32) for _ in range(4)], [rng.getrandbits(32) for _ in range(2)]) for _ in range(64)]
    args = [str(x) for c, k in cases for x in c + k]
    out = subprocess.check_output([str(exe)] + args, text=True, timeout=60).split("\n")
    for (c, k), line in zip(cases, out):
        assert [int(x) for x in line.split()] == O.philox4x32_10(c, k) == list(R.philox(*c, *k))


PS = [1e-9, 3e-7, 1.3e-4, 2.0 ** -12, 6e-4, 0.0021, 2.0 ** -8, 0.0078, 0.02, 2.0 ** -5, 0.04, 0.3, 0.5, 0.97]


@pytest.mark.parametrize("bits", [32, 64])
@pytest.mark.parametrize("spec", [1, 2, 4])
def test_oracle_words_equal_restatement(spec, bits):
    """Packed words of events across every class (dense, k = 7 .. 13) from the oracle's definition form, its lazy form,
    and the Python restatement."""
    b = G.GraphBuilder()
    b.root(b.gate(G.OR, [b.basic_event(p) for p in PS]))
    g = b.build()
    seed, first, nw = 0xABCDEF0123456789, 127 * 3, 3     # words 381..383: straddles a 8192-sample group boundary (word 384)
    ev = np.arange(len(PS), dtype=np.uint32)
    slow = O.sample_words(g, seed, ev, first, nw, bits, spec, fast=False)
    fast = O.sample_words(g, seed, ev, first, nw, bits, spec, fast=True)
    assert np.array_equal(slow, fast)
    n_check = len(PS) if spec != 2 else 6    # spec 2's definition form is slow in Python (rank of the undecided samples)
    for e in range(n_check):
        for w in range(nw if spec != 2 else 1):
            assert int(slow[e, w]) == R.sample_word(seed, e, first + w, PS[e], bits, spec), (e, w)


@pytest.mark.parametrize("thr64,k", [(1, 13), (12345, 13), (1 << 40, 13), ((1 << 51) - 1, 13), (1 << 51, 13), (3 << 51, 13), (1 << 53, 12), (1 << 54, 11), (1 << 55, 10),
                                     ((1 << 57) - 12345, 9), (1 << 57, 8), ((1 << 59) - 1, 7), (429496 << 32, 13)])
def test_class_and_table_oracle_vs_restatement(thr64, k):
    assert O.sparse_class_k(thr64) == R.class_k(thr64) == k
    assert O.sparse_cdf(thr64, k) == list(R.count_table(thr64, k))


@pytest.mark.parametrize("seed", range(4))
def test_oracle_tallies_equal_restatement_on_random_graphs(seed):
    g = random_graph(500 + seed, n_basic=8, n_gates=14, max_fanin=5, p_lo=0.0005, p_hi=0.4, n_roots=3)
    spec = (4, 2, 4, 1)[seed]
    n, begin = 96, 8192 * seed + 8150          # crosses a slab boundary
    want = R.mc(g, seed, begin, n, True, 32, spec)
    assert [int(x) for x in O.mc_scalar(g, seed, begin, n, True, spec=spec)] == want
    assert [int(x) for x in O.mc_packed(g, seed, begin, n, True, nthreads=3, spec=spec)] == want
