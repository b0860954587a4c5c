"""canopy_mc_run_static: the entry point for hosts that cannot let this library allocate device memory.

Some processes that load libcanopy do not own the GPU's memory: another allocator took it all at start-up, or a quota is
already committed, or the steady-state loop bans cudaMalloc / cudaFree because they stall the whole device.  For them the
library offers one call that runs the fused engine over a fixed scratch array compiled into the library (include/canopy.h
explains the sharing rules).  It must give exactly the tallies of the ordinary entry point, canopy_mc_run -- tallies are
integers, so "the tolerance the other tests use" is bit equality.

What is checked here, through the C ABI only:
  - two calls in a row on DIFFERENT plans (the second smaller than the first, another spec, threshold width, seed and
    range), each equal to canopy_mc_run on the same plan.  The array is shared and keeps the first call's bytes, so a
    counter that is not cleared or a table that is not rebuilt shows up in the second result;
  - the same plan twice (leftover counters would double the tallies);
  - output into device memory the caller owns;
  - refusal of a model that is too large, with a message that names the size and the limit.

This file is named test_zz_* so that pytest collects it last: `pytest -x` stops at the first failure and a failure here must
not hide the rest of the suite.  It passes both alone in a fresh process (where its first call is also the first CUDA call
of the process) and after all the other tests.
"""
import ctypes as C

import numpy as np
import pytest

from helpers import random_graph
from canopy_b200 import _lib, synth
from canopy_b200.engine import Plan

pytestmark = pytest.mark.gpu


def test_static_scratch_run_equals_ordinary_run_twice_in_a_row():
    big = synth.generate(120, 70, 6, 8, seed=5)       # 190 nodes before folding: close to the fused engine's limit
    small = random_graph(3, n_basic=14, n_gates=30, n_roots=3)
    with Plan(big, tally_all_gates=True, threshold_bits=64, sampler=4) as pa, \
            Plan(small, tally_all_gates=True, threshold_bits=32, sampler=2) as pb:
        # the new entry point first: in a fresh process nothing else has touched CUDA yet
        got_a = pa.run_static(50_000, seed=0xABCDEF, sample_begin=777)
        got_b = pb.run_static(8_192 * 3 + 11, seed=42, sample_begin=8_191)
        got_a2 = pa.run_static(50_000, seed=0xABCDEF, sample_begin=777)
        want_a, _ = pa.run(50_000, seed=0xABCDEF, sample_begin=777)
        want_b, _ = pb.run(8_192 * 3 + 11, seed=42, sample_begin=8_191)
    assert want_a.sum() > 0 and want_b.sum() > 0
    assert np.array_equal(got_a, want_a)
    assert np.array_equal(got_b, want_b)
    assert np.array_equal(got_a2, want_a)


def test_static_scratch_run_into_caller_owned_device_memory():
    torch = pytest.importorskip("torch")
    g = random_graph(9, n_basic=12, n_gates=24, n_roots=2)
    with Plan(g, tally_all_gates=True) as plan:
        want, _ = plan.run(20_000, seed=7)
        dev = torch.full((plan.n_tallies,), -1, dtype=torch.int64, device="cuda:0")   # the caller's memory, not the library's
        st = torch.cuda.Stream()
        torch.cuda.synchronize()                        # `dev` was filled on torch's default stream
        for _ in range(2):
            plan.run_static(20_000, seed=7, stream=st.cuda_stream, tallies_ptr=dev.data_ptr(), host=False)
            st.synchronize()
            assert np.array_equal(dev.cpu().numpy().astype(np.uint64), want)
            dev.fill_(-1)
            torch.cuda.synchronize()
        # an empty range: the counters of the previous call must not leak into it
        assert np.array_equal(plan.run_static(0, seed=7), plan.run(0, seed=7)[0])


def test_static_scratch_run_refuses_what_does_not_fit():
    g = synth.generate(400, 300, 6, 8, seed=2)          # far more rows than one slab holds in shared memory
    with Plan(g) as plan:
        with pytest.raises(_lib.CanopyError) as e:
            plan.run_static(1000)
    assert e.value.status == _lib.E_UNSUPPORTED
    assert "live nodes" in e.value.message and str(plan.stats["n_basic_live"] + plan.stats["n_gates_live"]) in e.value.message
    with Plan(random_graph(1)) as plan:                  # and both outputs missing is an argument error, not a crash
        o = plan._opts(0, 0, 10, 0, 0, None, False, False)
        assert _lib.lib().canopy_mc_run_static(plan._h, C.byref(o), None, None) == _lib.E_INVALID
