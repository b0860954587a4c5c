"""ctypes wrapper around oracle/liboracle.so.  TEST INFRASTRUCTURE ONLY (see canopy_oracle.c header).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
PARITY UNPINNED: the reference ships no implementation, tests or golden vectors for this path.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "canopy_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        L = _LIB
        u32p, u64p, dblp = C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.POINTER(C.c_double)
        L.oracle_philox4x32_10.argtypes = [u32p, u32p, u32p]
        L.oracle_threshold.argtypes = [C.c_double, C.c_int, u64p]
        L.oracle_quantised_p.argtypes = [C.c_double, C.c_int]
        L.oracle_quantised_p.restype = C.c_double
        L.oracle_sample_bit.argtypes = [C.c_uint64, C.c_uint32, C.c_uint64, C.c_int, C.c_uint64, C.c_int, C.c_int]
        L.oracle_mc_scalar.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_int, u64p]
        L.oracle_mc_packed.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int, u64p]
        L.oracle_exact.argtypes = [C.c_void_p, C.c_int, C.c_int, dblp]
        for fn in (L.oracle_sample_words, L.oracle_sample_words_fast):
            fn.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_int, u32p, C.c_uint32, C.c_uint64, C.c_uint64, u64p]
    return _LIB


_RC = {-1: "bad node id", -3: "cycle", -4: "too many basic events for enumeration", -8: "probability out of domain",
       -9: "unknown sampling spec"}

SPEC = 4  # default sampling spec of the build (include/canopy.h CANOPY_SPEC_VERSION); 1 = direct, 2 = sliced, 4 = classed sparse


def _check(rc):
    if rc != 0:
        raise ValueError(f"oracle: {_RC.get(rc, rc)}")


def philox4x32_10(ctr, key):
    c = (C.c_uint32 * 4)(*[int(x) & 0xFFFFFFFF for x in ctr])
    k = (C.c_uint32 * 2)(*[int(x) & 0xFFFFFFFF for x in key])
    o = (C.c_uint32 * 4)()
    lib().oracle_philox4x32_10(c, k, o)
    return [int(x) for x in o]


def philox_batch8(a0: int, a1: int, a2: int, c3, key):
    """The eight-lane Philox of the timed CPU port: calls (a0, a1, a2, c3[i]); returns eight 4-word outputs."""
    c = (C.c_uint32 * 8)(*[int(x) & 0xFFFFFFFF for x in c3])
    k = (C.c_uint32 * 2)(*[int(x) & 0xFFFFFFFF for x in key])
    o = (C.c_uint32 * 32)()
    lib().oracle_philox_batch8(C.c_uint32(a0), C.c_uint32(a1), C.c_uint32(a2), c, k, o)
    return [[int(o[w * 8 + i]) for w in range(4)] for i in range(8)]


def threshold(p: float, bits: int = 32):
    t = C.c_uint64()
    kind = lib().oracle_threshold(p, bits, C.byref(t))
    return kind, int(t.value)


def quantised_p(p: float, bits: int = 32) -> float:
    return float(lib().oracle_quantised_p(p, bits))


def sparse_class_k(thr64: int) -> int:
    """Spec 4: group exponent k of an event with 64-bit threshold thr64 (0 = dense event, spec 2 planes)."""
    lib().oracle_sparse_class_k.argtypes = [C.c_uint64]
    lib().oracle_sparse_class_k.restype = C.c_uint
    return int(lib().oracle_sparse_class_k(thr64))


def sparse_cdf(thr64: int, k: int) -> list:
    """Spec 4 count table of one sparse event (CANOPY_SPARSE_CDF_LEN words of 0.64 fixed point)."""
    out = (C.c_uint64 * int(lib().oracle_sparse_cdf_len()))()
    lib().oracle_sparse_cdf.argtypes = [C.c_uint64, C.c_uint, C.POINTER(C.c_uint64)]
    lib().oracle_sparse_cdf.restype = None
    lib().oracle_sparse_cdf(thr64, k, out)
    return [int(x) for x in out]


def sparse_words_with_table(seed: int, e: int, k: int, table, first_word: int, n_words: int, stream: bool = True) -> np.ndarray:
    """Test hook (spec 4): packed words of sparse event `e` with group exponent k under an explicit count table."""
    tab = np.ascontiguousarray(table, dtype=np.uint64)
    out = np.zeros(n_words, dtype=np.uint64)
    fn = lib().oracle_sparse_words_with_table
    fn.argtypes = [C.c_uint64, C.c_uint32, C.c_uint, C.POINTER(C.c_uint64), C.c_int, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64)]
    _check(fn(seed, e, k, tab.ctypes.data_as(C.POINTER(C.c_uint64)), int(stream), first_word, n_words,
              out.ctypes.data_as(C.POINTER(C.c_uint64))))
    return out


def sample_bit(seed: int, e: int, s: int, p: float, bits: int = 32, spec: int = SPEC) -> int:
    """One basic-event bit straight from the written definition of the variate (slow; the pin for everything else)."""
    kind, thr = threshold(p, bits)
    return int(lib().oracle_sample_bit(seed, e, s, kind, thr, bits, spec))


def mc_scalar(graph, seed: int, begin: int, n: int, tally_all: bool = False, bits: int = 32, spec: int = SPEC) -> np.ndarray:
    d = graph.c_desc()
    out = np.zeros(graph.n_tallies(tally_all), dtype=np.uint64)
    _check(lib().oracle_mc_scalar(C.addressof(d), seed, begin, n, int(tally_all), bits, spec,
                                  out.ctypes.data_as(C.POINTER(C.c_uint64))))
    return out


def mc_packed(graph, seed: int, begin: int, n: int, tally_all: bool = False, bits: int = 32, nthreads: int = 1,
              spec: int = SPEC) -> np.ndarray:
    d = graph.c_desc()
    out = np.zeros(graph.n_tallies(tally_all), dtype=np.uint64)
    _check(lib().oracle_mc_packed(C.addressof(d), seed, begin, n, int(tally_all), bits, spec, nthreads,
                                  out.ctypes.data_as(C.POINTER(C.c_uint64))))
    return out


def exact(graph, tally_all: bool = False, bits: int = 32) -> np.ndarray:
    d = graph.c_desc()
    out = np.zeros(graph.n_tallies(tally_all), dtype=np.float64)
    _check(lib().oracle_exact(C.addressof(d), int(tally_all), bits, out.ctypes.data_as(C.POINTER(C.c_double))))
    return out


def sample_words(graph, seed: int, event_ids, first_word: int, n_words: int, bits: int = 32, spec: int = SPEC,
                 fast: bool = False) -> np.ndarray:
    """Packed basic-event words. fast=False builds every bit from the definition (oracle_sample_bit);
    fast=True uses the lazily evaluated group routine that mc_scalar / mc_packed use."""
    d = graph.c_desc()
    ev = np.ascontiguousarray(event_ids, dtype=np.uint32)
    out = np.zeros((len(ev), n_words), dtype=np.uint64)
    fn = lib().oracle_sample_words_fast if fast else lib().oracle_sample_words
    _check(fn(C.addressof(d), seed, bits, spec, ev.ctypes.data_as(C.POINTER(C.c_uint32)), len(ev),
                                     first_word, n_words, out.ctypes.data_as(C.POINTER(C.c_uint64))))
    return out
