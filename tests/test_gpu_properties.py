"""Size-independent properties at BASELINE sizes, and the edge cases of the boundary.  GPU only."""
import os

import numpy as np
import pytest
import torch

from helpers import G, O, random_graph
from canopy_b200 import _lib, synth
from canopy_b200.engine import Plan, quantify, Settings

pytestmark = pytest.mark.gpu
CORES = len(os.sched_getaffinity(0))


@pytest.fixture(scope="module")
def c2_plan():
    g = synth.config_graph("c2")
    with Plan(g, tally_all_gates=True) as plan:
        yield plan


def test_c2_full_size_decomposition_and_batching_invariance(c2_plan):
    """2^26 samples of config 2: tallies of [a,c) == [a,b) + [b,c); independent of batch size and of overlap."""
    S = 1 << 26
    whole, rep = c2_plan.run(S, seed=9)
    cut = 4096 * 5000 + 1234
    a, _ = c2_plan.run(cut, seed=9)
    b, _ = c2_plan.run(S - cut, seed=9, sample_begin=cut)
    assert np.array_equal(whole, a + b)
    small_batches, rep2 = c2_plan.run(S, seed=9, batch_slabs=148, overlap=False)
    assert np.array_equal(whole, small_batches) and rep2.n_batches > rep.n_batches
    assert rep.n_batches * rep.batch_slabs * rep.slab_samples >= S
    # statistical sanity of the top event against its own binomial error (no oracle at this size)
    top = int(whole[0]); p = top / S
    assert 0 < top < S and (p * (1 - p) / S) ** 0.5 < 1e-3


def test_c2_prefix_matches_oracle_inside_a_long_run(c2_plan):
    """The first 2^14 samples of a long run are the same samples as a short run (counter-based RNG)."""
    n = 1 << 14
    got, _ = c2_plan.run(n, seed=9)
    assert np.array_equal(got, O.mc_packed(c2_plan.graph, 9, 0, n, True, nthreads=CORES))


def test_complement_and_alias_identities():
    b = G.GraphBuilder()
    e = [b.basic_event(0.3), b.basic_event(0.6), b.basic_event(0.1)]
    x = b.gate(G.AND, [e[0], e[1]])
    for h in (x, b.gate(G.NOT, [x]), b.gate(G.NAND, [e[0], e[1]]), b.gate(G.NULL, [x]),
              b.gate(G.OR, [e[2], e[0]]), b.gate(G.NOR, [e[2], e[0]])):
        b.root(h)
    n = 100_000
    with Plan(b.build()) as plan:
        t, _ = plan.run(n, seed=4, sample_begin=17)
    t = [int(v) for v in t]
    assert t[0] + t[1] == n and t[1] == t[2] and t[3] == t[0] and t[4] + t[5] == n


@pytest.mark.parametrize("begin,n", [(0, 0), (0, 1), (63, 2), (4095, 2), (5, 4091), (4096 * 3, 4096), ((1 << 44) + 4097, 9000),
                                     ((1 << 63) - 5000, 4999)])
def test_ragged_ranges(begin, n):
    g = random_graph(21, n_basic=10, n_gates=18)
    with Plan(g, tally_all_gates=True) as plan:
        got, _ = plan.run(n, seed=8, sample_begin=begin)
    assert np.array_equal(got, O.mc_scalar(g, 8, begin, n, True))


def test_constant_and_empty_outputs():
    b = G.GraphBuilder()
    e0, e1 = b.basic_event(0.0), b.basic_event(1.0)
    b.root(b.gate(G.OR, [e0, G.GraphBuilder.constant(False)]))
    b.root(b.gate(G.AND, [e1, G.GraphBuilder.constant(True)]))
    b.root(e1)
    with Plan(b.build()) as plan:
        t, rep = plan.run(12345, seed=1)
        assert [int(v) for v in t] == [0, 12345, 12345] and rep.gate_launches == 0
    with Plan(G.from_lists([], [], [])) as plan:
        t, _ = plan.run(1000)
        assert len(t) == 0


def test_device_output_on_torch_stream_matches_host_output():
    g = synth.config_graph("c2", 0.1)
    with Plan(g, tally_all_gates=True) as plan:
        host, _ = plan.run(1 << 18, seed=3)
        side = torch.cuda.Stream()
        out = torch.zeros(plan.n_tallies, dtype=torch.int64, device="cuda")
        with torch.cuda.stream(side):
            plan.run_device(out.data_ptr(), 1 << 18, seed=3, stream=side.cuda_stream)   # no host sync inside
            doubled = out * 2                                                         # ordered after the run on `side`
        side.synchronize()
    assert np.array_equal(out.cpu().numpy().astype(np.uint64), host)
    assert np.array_equal(doubled.cpu().numpy().astype(np.uint64), host * 2)


def test_run_multi_equals_single_device():
    """canopy_mc_run_multi (ncclCommInitAll + ncclAllReduce over every visible device) == one device.  Needs two
    devices: on a 1-GPU lease this test is SKIPPED, not passed -- mc_run_multi's n_devices == 1 branch never touches NCCL."""
    if torch.cuda.device_count() < 2:
        pytest.skip("NEEDS >= 2 GPUS: canopy_mc_run_multi's NCCL all-reduce is not exercised on a single-device lease "
                    "(bench.py's tally_check covers the N-way == 1-way sum on the driver's multi-GPU runs)")
    g = synth.config_graph("c2", 0.1)
    n = 4096 * 37 + 5
    devs = list(range(torch.cuda.device_count()))
    with Plan(g, tally_all_gates=True) as plan:
        one, _ = plan.run(n, seed=6, sample_begin=11)
        multi, rep = plan.run_multi(devs, n, seed=6, sample_begin=11)      # NCCL all-reduce when len(devs) > 1
        assert np.array_equal(one, multi)
        two_way = plan.run(n // 2, seed=6, sample_begin=11)[0] + plan.run(n - n // 2, seed=6, sample_begin=11 + n // 2)[0]
        assert np.array_equal(one, two_way)


def test_quantify_front_end_and_three_sigma_vs_exact():
    g = random_graph(60, n_basic=10, n_gates=16, n_roots=2, extreme_p=False)
    res = quantify(g, Settings(num_trials=1 << 20, seed=12, tally_all_gates=True))
    exact = O.exact(g, True)
    sigma = np.sqrt(np.maximum(exact * (1 - exact), 1e-12) / (1 << 20))
    assert np.all(np.abs(res["p"] - exact) <= 4 * sigma + 1e-9)
    assert np.all(res["ci_lo"] <= res["p"]) and np.all(res["p"] <= res["ci_hi"])


def test_bad_run_arguments():
    with Plan(random_graph(2)) as plan:
        with pytest.raises(_lib.CanopyError) as ei:
            plan.run(100, device=99)
        assert ei.value.status == _lib.E_INVALID
        with pytest.raises(_lib.CanopyError):
            plan.run(1 << 20, sample_begin=(1 << 64) - 10)
        with pytest.raises(_lib.CanopyError):
            plan.sample_words(1, 3, 64)


@pytest.mark.gpu
def test_session_chunks_equal_one_run_and_resume_from_disk(tmp_path):
    from canopy_b200.engine import Session
    g = synth.generate(n_basic=400, n_gates=250, n_layers=7, max_fanin=9, seed=21)
    with Plan(g, tally_all_gates=True) as plan:
        whole, _ = plan.run(3 * 4096 + 1000, seed=5, sample_begin=100)
        s = Session(plan, seed=5, sample_begin=100)
        s.advance(777)
        s.advance(4096)
        f = str(tmp_path / "ck.npz")
        s.save(f)
        r = Session.load(f, plan)                      # "restart": a fresh record from disk
        r.advance(3 * 4096 + 1000 - 777 - 4096)
        assert r.n_done == 3 * 4096 + 1000 and np.array_equal(r.tallies[: plan.n_tallies], whole)
        e = r.estimates()[0]
        assert e[0] == int(whole[0]) / r.n_done and e[2] <= e[0] <= e[3]


@pytest.mark.gpu
def test_update_p_equals_a_fresh_plan_and_uncertainty_loop():
    from canopy_b200.engine import Settings, uncertainty
    g = synth.generate(n_basic=300, n_gates=200, n_layers=6, max_fanin=8, seed=22)
    rng = np.random.default_rng(0)
    with Plan(g, tally_all_gates=True) as plan:
        plan.run(4096, seed=1)                          # make the plan resident before the update
        for k in range(3):
            p = np.clip(g.p * rng.lognormal(0.0, 0.5, g.n_basic), 1e-6, 0.9)
            plan.update_p(p)
            got, _ = plan.run(2 * 4096 + 99, seed=9, sample_begin=k * 10000)
            g2 = G.Graph(p=p, op=g.op, kmin=g.kmin, kmax=g.kmax, child_off=g.child_off, child=g.child, roots=g.roots)
            assert np.array_equal(got, O.mc_scalar(g2, 9, k * 10000, 2 * 4096 + 99, True, spec=plan.spec))
    u = uncertainty(g, Settings(num_trials=8192, seed=3), lambda k: np.clip(g.p * (1 + 0.1 * k), 0, 0.9), n_outer=4)
    assert u["p"].shape == (4, len(g.roots)) and np.all(np.diff(u["p"][:, 0]) >= -0.05) and u["q05"][0] <= u["q95"][0]


@pytest.mark.gpu
def test_parked_value_buffers_are_reused_correctly():
    """Buffers of destroyed plans are parked and handed to the next plan (dirty, other graph, other size)."""
    from canopy_b200.engine import trim_device_cache
    for seed, nb in ((31, 500), (32, 200), (33, 800), (34, 500)):
        g = synth.generate(n_basic=nb, n_gates=nb // 2, n_layers=6, max_fanin=7, seed=seed)
        with Plan(g, tally_all_gates=True) as plan:
            got, _ = plan.run(5 * 4096 + 17, seed=seed, sample_begin=4096 * seed)
        assert np.array_equal(got, O.mc_packed(g, seed, 4096 * seed, 5 * 4096 + 17, True, nthreads=CORES, spec=plan.spec))
        if seed == 32:
            trim_device_cache()


def test_pipelined_run_is_the_same_on_every_kind_of_stream_and_pipeline():
    """The pipelined run (sampler of batch b+1 launched as a programmatic dependent of gate walk b) on the library's
    stream, on the legacy default stream (handle 1), on the per-thread default stream (handle 2), serialised, and the
    oracle: identical tallies.  Many small batches so that every buffer is rewritten several times."""
    g = random_graph(17, n_basic=300, n_gates=160)
    n = 20 * 3 * 8192 + 123
    want = O.mc_packed(g, 3, 77, n, True, nthreads=CORES)
    with Plan(g, tally_all_gates=True) as plan:
        for kw in (dict(), dict(stream=1), dict(stream=2), dict(overlap=False)):
            got, rep = plan.run(n, seed=3, sample_begin=77, batch_slabs=3, **kw)
            assert rep.n_batches >= 20 and np.array_equal(got, want), kw


def test_run_sharded_and_run_device_on_torchs_default_stream():
    """canopy_b200.dist.run_sharded passes torch's CURRENT stream; on the default stream that handle is 0, which the C ABI
    would read as "use the library's own stream" and leave the run unordered with the torch ops around it.  The Python
    layer maps 0 to cudaStreamLegacy: fills before and reads after the run, on the default stream, must see the tallies."""
    from canopy_b200.dist import run_sharded
    g = synth.config_graph("c2", 0.1)
    n = 1 << 20
    with Plan(g, tally_all_gates=True) as plan:
        want, _ = plan.run(n, seed=3, sample_begin=5)
        assert torch.cuda.current_stream().cuda_stream == 0
        for _ in range(5):
            got = run_sharded(plan, n, seed=3, sample_begin=5)
            doubled = got * 2                                    # default stream, right behind the run
            assert np.array_equal(got.cpu().numpy().astype(np.uint64), want)
            assert np.array_equal(doubled.cpu().numpy().astype(np.uint64), want * 2)
        out = torch.full((plan.n_tallies,), -1, dtype=torch.int64, device="cuda")   # fill on the default stream ...
        plan.run_device(out.data_ptr(), n, seed=3, sample_begin=5, stream=torch.cuda.current_stream().cuda_stream)
        assert np.array_equal((out + 0).cpu().numpy().astype(np.uint64), want)     # ... and read right behind it


def test_two_runs_of_one_plan_on_different_streams_do_not_overlap():
    """Per-plan scratch (value buffers, gate tallies) is reused by every run: a second run on ANOTHER stream must wait for
    the first (the library records a completion event per device state and makes the next run's stream wait on it)."""
    g = synth.config_graph("c2", 0.2)
    n = 1 << 22
    with Plan(g, tally_all_gates=True) as plan:
        want_a, _ = plan.run(n, seed=1)
        want_b, _ = plan.run(n, seed=2)
        sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
        a = torch.zeros(plan.n_tallies, dtype=torch.int64, device="cuda")
        b = torch.zeros(plan.n_tallies, dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        for _ in range(3):
            plan.run_device(a.data_ptr(), n, seed=1, stream=sa.cuda_stream)    # asynchronous: returns with kernels in flight
            plan.run_device(b.data_ptr(), n, seed=2, stream=sb.cuda_stream)
            torch.cuda.synchronize()
            assert np.array_equal(a.cpu().numpy().astype(np.uint64), want_a)
            assert np.array_equal(b.cpu().numpy().astype(np.uint64), want_b)
        # canopy_plan_update_p while a run is in flight on a user stream must not change that run's thresholds
        plan.run_device(a.data_ptr(), n, seed=1, stream=sa.cuda_stream)
        plan.update_p(np.clip(g.p * 1.5, 0, 0.9))
        torch.cuda.synchronize()
        assert np.array_equal(a.cpu().numpy().astype(np.uint64), want_a)


def test_pipeline_hazard_on_a_gate_bound_graph():
    """on a graph whose gate walk outlasts the sampler (c4-like: deep, high fan-in) the pipelined run with many
    small batches must equal the serialised run, repeatedly (S(b+2) must not overwrite the buffer G(b) still reads)."""
    g = synth.config_graph("c4", 0.05)
    n = 24 * 4 * 8192 + 99
    with Plan(g, tally_all_gates=True) as plan:
        want, _ = plan.run(n, seed=7, batch_slabs=4, overlap=False)
        assert np.array_equal(want, O.mc_packed(g, 7, 0, n, True, nthreads=CORES, spec=plan.spec))
        for rep in range(12):
            got, r = plan.run(n, seed=7, batch_slabs=4, overlap=True)
            assert r.n_batches >= 24 and np.array_equal(got, want), rep


@pytest.mark.gpu
@pytest.mark.parametrize("spec,bits", [(4, 32), (4, 64), (2, 32), (2, 64)])
def test_fused_engine_equals_streamed_and_oracle(spec, bits):
    """The fused small-model kernel (csrc/small.cuh) against the streamed pair and the scalar oracle: all connectives,
    constants, complements, ragged sample ranges, every class of spec 4."""
    from canopy_b200.engine import ENGINE_FUSED, ENGINE_STREAMED
    from helpers import random_graph
    rng = np.random.default_rng(5)
    for seed in range(6):
        g = random_graph(100 + seed, n_basic=40, n_gates=60, max_fanin=9)
        g.p[:] = np.where(g.p > 0, 10.0 ** rng.uniform(-4.5, -0.3, g.n_basic), 0.0) * (g.p < 1) + (g.p >= 1)   # spread over the classes, keep p = 0 / 1
        n, begin = 2 * 8192 + 777 + seed, 4096 * seed + 13
        with Plan(g, tally_all_gates=True, threshold_bits=bits, sampler=spec, engine=ENGINE_FUSED) as pf, \
             Plan(g, tally_all_gates=True, threshold_bits=bits, sampler=spec, engine=ENGINE_STREAMED) as ps:
            tf, rf = pf.run(n, seed=21, sample_begin=begin)
            ts, rs = ps.run(n, seed=21, sample_begin=begin)
            assert rf.sampler_launches == 0 and rf.gate_launches == 1 and rs.sampler_launches >= 1
        assert np.array_equal(tf, ts)
        assert np.array_equal(tf, O.mc_scalar(g, 21, begin, n, True, bits=bits, spec=spec))


@pytest.mark.gpu
def test_fused_engine_limits_and_update_p():
    from canopy_b200 import _lib
    from canopy_b200.engine import ENGINE_FUSED
    big = synth.generate(n_basic=300, n_gates=160, n_layers=6, max_fanin=12, seed=3)
    with pytest.raises(_lib.CanopyError) as ei:
        Plan(big, engine=ENGINE_FUSED)
    assert ei.value.status == _lib.E_UNSUPPORTED
    g = synth.generate(n_basic=120, n_gates=70, n_layers=5, max_fanin=8, seed=4)
    with pytest.raises(_lib.CanopyError):
        Plan(g, sampler=1, engine=ENGINE_FUSED)                     # the direct sampler has no fused form
    rng = np.random.default_rng(1)
    with Plan(g, tally_all_gates=True, engine=ENGINE_FUSED) as plan:
        plan.run(4096, seed=1)
        for k in range(3):                                          # classes (hence the row order) change with p
            p = np.clip(g.p * rng.lognormal(0.0, 1.5, g.n_basic), 1e-7, 0.9)
            plan.update_p(p)
            got, rep = plan.run(8192 + 5, seed=3, sample_begin=k * 50000)
            assert rep.sampler_launches == 0
            g2 = G.Graph(p=p, op=g.op, kmin=g.kmin, kmax=g.kmax, child_off=g.child_off, child=g.child, roots=g.roots)
            assert np.array_equal(got, O.mc_scalar(g2, 3, k * 50000, 8192 + 5, True))


def test_update_p_from_another_thread_while_the_plan_runs():
    """canopy_plan_update_p and runs of the same plan from different host threads (include/canopy.h): the update waits for
    the host side of a run, a run waits for the update, so every run's tallies are those of one of the two probability
    sets, whole -- never thresholds of one set with the sampler order or count tables of the other."""
    import threading
    g = synth.config_graph("c2", 0.05)
    p0 = np.array(g.p, dtype=np.float64)
    p1 = np.where((p0 > 0) & (p0 < 1), np.clip(p0 * 3.0, 0.0, 0.9), p0)      # moves many events across class boundaries (dense <-> sparse, k <-> k - 1)
    n, seed = 40_000, 11
    with Plan(g, tally_all_gates=True, sampler=4) as plan:
        want = []
        for p in (p0, p1):
            plan.update_p(p)
            want.append(plan.run(n, seed=seed)[0].copy())
        assert not np.array_equal(want[0], want[1])
        stop = threading.Event()
        errors = []

        def updater():
            i = 0
            try:
                while not stop.is_set():
                    plan.update_p(p0 if i & 1 else p1)
                    i += 1
            except Exception as e:   # noqa: BLE001
                errors.append(e)

        t = threading.Thread(target=updater)
        t.start()
        try:
            seen = [0, 0]
            for _ in range(60):
                got = plan.run(n, seed=seed)[0]
                which = [np.array_equal(got, w) for w in want]
                assert any(which), "a run mixed two probability sets"
                seen[which.index(True)] += 1
        finally:
            stop.set()
            t.join()
        assert not errors, errors
