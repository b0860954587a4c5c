"""Host-side mirror of the quantification interface, over the C ABI (include/canopy.h).

Vocabulary follows the reference: `Settings` carries the Monte Carlo knobs of
reference src/mef/openpsa/settings.h:118-187,324-327 (num_trials, seed, batch_size), widened to 64 bit;
`Plan` is the lowered fault-tree DAG; `run()` returns tallies per top event / per gate.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _lib
from .graph import CGraphDesc, Graph


class SettingsError(ValueError):
    """reference src/mef/openpsa/error.h:66 (SettingsError)."""


@dataclass
class Settings:
    """Monte Carlo subset of the reference Settings builder (settings.h:118-187)."""
    num_trials: int = 1000      # settings.h:325 default 1e3
    seed: int = 0               # settings.h:324
    batch_size: int = 0         # slabs per batch here (settings.h:326 is 1 trial); 0 = engine default
    sample_begin: int = 0       # first global sample index (restart / shard by counter range)
    tally_all_gates: bool = False
    threshold_bits: int = 32
    sampler: int = 0            # sampling spec: 0 = library default (4), 1 = "direct", 2 = "sliced", 4 = "classed sparse"
    overlap: bool = True
    profile: bool = False

    def validate(self):
        if self.num_trials < 1:
            raise SettingsError("The number of trials cannot be less than 1.")      # settings.h:129
        if self.seed < 0:
            raise SettingsError("The seed for PRNG cannot be negative.")            # settings.h:186
        if self.batch_size < 0:
            raise SettingsError("The batch size cannot be negative.")
        if self.threshold_bits not in (32, 64):
            raise SettingsError("threshold_bits must be 32 or 64")
        if self.sampler not in (0, 1, 2, 4):
            raise SettingsError("sampler must be 0 (default), 1 (direct), 2 (sliced) or 4 (classed sparse)")
        return self


@dataclass
class Report:
    total_ms: float
    sampler_ms: float
    gate_ms: float
    sampler_launches: int
    gate_launches: int
    other_launches: int
    n_batches: int
    slab_samples: int
    batch_slabs: int
    h2d_bytes: int
    d2h_bytes: int

    @property
    def launches(self) -> int:
        return self.sampler_launches + self.gate_launches + self.other_launches


def _report(r: _lib.CMcReport) -> Report:
    return Report(r.total_ms, r.sampler_ms, r.gate_ms, r.sampler_launches, r.gate_launches, r.other_launches,
                  r.n_batches, r.slab_samples, r.batch_slabs, r.h2d_bytes, r.d2h_bytes)


def _arr(ptr, n, dtype):
    if n == 0:
        return np.zeros(0, dtype=dtype)
    return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dtype, copy=True)


ENGINE_AUTO, ENGINE_STREAMED, ENGINE_FUSED = 0, 1, 2   # CANOPY_ENGINE_* (include/canopy.h)
DEFAULT_ENGINE = ENGINE_AUTO   # what Plan(engine=None) asks for; the GPU tests run under both AUTO and STREAMED (tests/conftest.py)


class Plan:
    """Lowered DAG (host) + lazily created per-device state (owned by libcanopy)."""

    def __init__(self, graph: Graph, tally_all_gates: bool = False, threshold_bits: int = 32, sampler: int = 0,
                 plan_cache: bool = True, engine: int | None = None):
        """engine: ENGINE_AUTO (fused kernel for models whose rows fit in shared memory), ENGINE_STREAMED, ENGINE_FUSED."""
        engine = DEFAULT_ENGINE if engine is None else engine
        self._h = C.c_void_p()
        self.graph = graph
        self.tally_all_gates = bool(tally_all_gates)
        self.threshold_bits = threshold_bits
        desc = graph.c_desc()
        opts = _lib.CPlanOpts(tally_all_gates=int(tally_all_gates), threshold_bits=threshold_bits, sampler=int(sampler),
                              no_plan_cache=int(not plan_cache), engine=int(engine))
        _lib.check(_lib.lib().canopy_plan_create(C.byref(desc), C.byref(opts), C.byref(self._h)))
        st = _lib.CPlanStats()
        _lib.check(_lib.lib().canopy_plan_info(self._h, C.byref(st)))
        self.stats = {f: getattr(st, f) for f, _ in _lib.CPlanStats._fields_}
        self.spec = int(st.spec_version)  # the sampling spec the oracle must be asked for

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            _lib.lib().canopy_plan_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def n_tallies(self) -> int:
        return int(self.stats["n_tallies"])

    def view(self) -> dict:
        """Copies of the lowered arrays (for the host-logic tests)."""
        v = _lib.CPlanView()
        _lib.check(_lib.lib().canopy_plan_view_get(self._h, C.byref(v)))
        return dict(
            n_events=v.n_events, n_gates=v.n_gates, n_layers=v.n_layers, n_tallies=v.n_tallies, n_edges=v.n_edges,
            event_orig=_arr(v.event_orig, v.n_events, np.uint32), event_thr32=_arr(v.event_thr32, v.n_events, np.uint32),
            event_thr64=_arr(v.event_thr64, v.n_events, np.uint64),
            gate_child_off=_arr(v.gate_child_off, v.n_gates + 1, np.uint32), gate_child=_arr(v.gate_child, v.n_edges, np.uint32),
            gate_op=_arr(v.gate_op, v.n_gates, np.uint8), gate_flags=_arr(v.gate_flags, v.n_gates, np.uint8),
            gate_kmin=_arr(v.gate_kmin, v.n_gates, np.uint16), gate_kmax=_arr(v.gate_kmax, v.n_gates, np.uint16),
            layer_off=_arr(v.layer_off, v.n_layers + 1, np.uint32),
            tally_kind=_arr(v.tally_kind, v.n_tallies, np.uint8), tally_gate=_arr(v.tally_gate, v.n_tallies, np.uint32),
            event_k=_arr(v.event_k, v.n_events, np.uint8) if v.event_k else np.zeros(v.n_events, dtype=np.uint8))

    def _opts(self, seed, sample_begin, n_samples, device, batch_slabs, stream, overlap, profile) -> _lib.CMcOpts:
        o = _lib.CMcOpts()
        _lib.lib().canopy_mc_opts_default(C.byref(o))
        o.seed, o.sample_begin, o.n_samples = int(seed), int(sample_begin), int(n_samples)
        o.device, o.batch_slabs = int(device), int(batch_slabs)
        # None: the library's own stream.  A torch handle of 0 is the legacy default stream: the C ABI takes NULL as
        # "library-owned", so the default streams are passed explicitly (cudaStreamLegacy = 1, include/canopy.h).
        o.stream = None if stream is None else (1 if int(stream) == 0 else int(stream))
        o.overlap, o.profile = int(bool(overlap)), int(bool(profile))
        return o

    def run(self, n_samples: int, seed: int = 0, sample_begin: int = 0, device: int = 0, batch_slabs: int = 0,
            stream=None, overlap: bool = True, profile: bool = False):
        """Tallies into a HOST buffer (h2d/d2h inside the call). Returns (uint64[n_tallies], Report)."""
        o = self._opts(seed, sample_begin, n_samples, device, batch_slabs, stream, overlap, profile)
        out = np.zeros(max(self.n_tallies, 1), dtype=np.uint64)
        rep = _lib.CMcReport()
        _lib.check(_lib.lib().canopy_mc_run(self._h, C.byref(o), out.ctypes.data_as(C.POINTER(C.c_uint64)), C.byref(rep)))
        return out[: self.n_tallies], _report(rep)

    def run_device(self, tallies_ptr: int, n_samples: int, seed: int = 0, sample_begin: int = 0, device: int = 0,
                   batch_slabs: int = 0, stream=None, overlap: bool = True, profile: bool = False, want_report: bool = False,
                   profile_walk_shape: int = 0):
        """Tallies left in DEVICE memory at `tallies_ptr` (uint64[n_tallies]), ordered on `stream`.
        With want_report=False nothing synchronises the host."""
        o = self._opts(seed, sample_begin, n_samples, device, batch_slabs, stream, overlap, profile)
        o.profile_walk_shape = int(profile_walk_shape)   # canopy_mc_opts.profile_walk_shape: 1 = the pipeline's one CTA per SM
        rep = _lib.CMcReport()
        _lib.check(_lib.lib().canopy_mc_run_device(self._h, C.byref(o), C.c_void_p(tallies_ptr),
                                                   C.byref(rep) if want_report else None))
        return _report(rep) if want_report else None

    def run_static(self, n_samples: int, seed: int = 0, sample_begin: int = 0, device: int = 0, stream=None,
                   tallies_ptr: int | None = None, host: bool = True):
        """canopy_mc_run_static: the fused engine over the library's built-in scratch array, for hosts that cannot let a
        plug-in allocate device memory (see include/canopy.h for who that is and for the sharing rules: one array per GPU,
        calls must not overlap on the GPU).  Returns uint64[n_tallies] when `host`, else None (tallies are then left at the
        caller-owned device address `tallies_ptr`, ordered on `stream`)."""
        o = self._opts(seed, sample_begin, n_samples, device, 0, stream, False, False)
        out = np.zeros(max(self.n_tallies, 1), dtype=np.uint64) if host else None
        _lib.check(_lib.lib().canopy_mc_run_static(
            self._h, C.byref(o), out.ctypes.data_as(C.POINTER(C.c_uint64)) if host else None,
            C.c_void_p(tallies_ptr) if tallies_ptr else None))
        return out[: self.n_tallies] if host else None

    def run_multi(self, devices, n_samples: int, seed: int = 0, sample_begin: int = 0, batch_slabs: int = 0,
                  overlap: bool = True):
        """Single-process multi-GPU with one NCCL all-reduce of the tallies (canopy_mc_run_multi)."""
        o = self._opts(seed, sample_begin, n_samples, 0, batch_slabs, None, overlap, False)
        devs = (C.c_int32 * len(devices))(*devices)
        out = np.zeros(max(self.n_tallies, 1), dtype=np.uint64)
        rep = _lib.CMcReport()
        _lib.check(_lib.lib().canopy_mc_run_multi(self._h, C.byref(o), devs, len(devices),
                                                  out.ctypes.data_as(C.POINTER(C.c_uint64)), C.byref(rep)))
        return out[: self.n_tallies], _report(rep)

    def fingerprint(self) -> int:
        """64-bit hash of the lowered graph, thresholds and options (canopy_plan_fingerprint)."""
        out = C.c_uint64()
        _lib.check(_lib.lib().canopy_plan_fingerprint(self._h, C.byref(out)))
        return int(out.value)

    def update_p(self, p) -> None:
        """New basic-event probabilities for the same lowered graph (parameter-uncertainty outer loop,
        SURVEY.md 8f-4): thresholds are re-quantised and re-uploaded; raises if the constant folding changes."""
        p = np.ascontiguousarray(p, dtype=np.float64)
        _lib.check(_lib.lib().canopy_plan_update_p(self._h, p.ctypes.data_as(C.POINTER(C.c_double)), p.shape[0]))

    def debug_set_sparse_table(self, live_event: int, cdf) -> None:
        """Spec 4 test hook (include/canopy-testing.h): replace the count table of one sparse event."""
        t = np.ascontiguousarray(cdf, dtype=np.uint64)
        _lib.check(_lib.lib().canopy_plan_debug_set_sparse_table(self._h, live_event, t.ctypes.data_as(C.POINTER(C.c_uint64))))

    def sample_words(self, seed: int, first_word: int, n_words: int, device: int = 0) -> np.ndarray:
        """Raw sampler output [n_events_live, n_words] (parity hook)."""
        o = self._opts(seed, 0, 0, device, 0, None, False, False)
        n_ev = int(self.stats["n_basic_live"])
        out = np.zeros((max(n_ev, 1), n_words), dtype=np.uint64)
        _lib.check(_lib.lib().canopy_mc_sample_words(self._h, C.byref(o), first_word, n_words,
                                                     out.ctypes.data_as(C.POINTER(C.c_uint64))))
        return out[:n_ev]


class Session:
    """A restartable run: state = (seed, next sample index, tallies).  `advance(n)` evaluates the
    next n samples and accumulates; `save`/`load` round-trip the record through an .npz file.  Any chunking gives
    the tallies of one run over the same range (tests/test_gpu_properties.py)."""

    def __init__(self, plan: Plan, seed: int = 0, sample_begin: int = 0):
        self.plan = plan
        self.ck = _lib.CCheckpoint()
        _lib.check(_lib.lib().canopy_checkpoint_init(plan._h, int(seed), int(sample_begin), C.byref(self.ck)))
        self.tallies = np.zeros(max(plan.n_tallies, 1), dtype=np.uint64)

    @property
    def n_done(self) -> int:
        return int(self.ck.next_sample - self.ck.sample_begin)

    def advance(self, n_samples: int, device: int = 0, batch_slabs: int = 0, overlap: bool = True):
        o = self.plan._opts(0, 0, n_samples, device, batch_slabs, None, overlap, False)
        rep = _lib.CMcReport()
        _lib.check(_lib.lib().canopy_mc_run_resume(self.plan._h, C.byref(o), C.byref(self.ck),
                                                   self.tallies.ctypes.data_as(C.POINTER(C.c_uint64)), C.byref(rep)))
        return _report(rep)

    def estimates(self):
        """[(p_hat, sigma, ci_lo, ci_hi)] per output over the samples done so far."""
        return [estimate(int(t), self.n_done) for t in self.tallies[: self.plan.n_tallies]]

    def save(self, path: str) -> None:
        rec = np.array([getattr(self.ck, f) for f, _ in _lib.CCheckpoint._fields_], dtype=np.uint64)
        np.savez(path, record=rec, tallies=self.tallies)

    @classmethod
    def load(cls, path: str, plan: Plan) -> "Session":
        z = np.load(path)
        s = cls.__new__(cls)
        s.plan = plan
        s.ck = _lib.CCheckpoint(*[int(x) for x in z["record"]])
        s.tallies = np.ascontiguousarray(z["tallies"], dtype=np.uint64).copy()
        if s.tallies.shape[0] != max(plan.n_tallies, 1):
            raise _lib.CanopyError(_lib.E_INVALID, "checkpoint has another number of tallies than the plan")
        return s


# canopy_deviate_kind (include/canopy.h)
DEV_CONSTANT, DEV_UNIFORM, DEV_NORMAL, DEV_LOGNORMAL, DEV_LOGNORMAL_EF, DEV_GAMMA, DEV_BETA, DEV_HISTOGRAM = range(8)


class PSource:
    """Probability source of the uncertainty loop (canopy_psource): BasicEvent::p() of every event as the mean
    (Expression::value()) or as the trial-th joint draw of all random deviates (Expression::Sample()/Reset(),
    reference expression.h:98-117, expr/random_deviate.h).  Built from a loaded MEF model (`load_mef` leaves one in
    Graph.meta["psource"]) or from one distribution per event: tuples (kind, a, b[, c]) or
    (DEV_HISTOGRAM, bounds, weights)."""

    def __init__(self, handle):
        self._h = handle
        nb, nd = C.c_uint32(), C.c_uint32()
        _lib.check(_lib.lib().canopy_psource_info(self._h, C.byref(nb), C.byref(nd)))
        self.n_basic, self.n_deviates = int(nb.value), int(nd.value)

    @classmethod
    def from_deviates(cls, dists) -> "PSource":
        arr = (_lib.CDeviate * max(len(dists), 1))()
        keep = []
        for i, d in enumerate(dists):
            arr[i].kind = int(d[0])
            if d[0] == DEV_HISTOGRAM:
                bounds = np.ascontiguousarray(d[1], dtype=np.float64)
                weights = np.ascontiguousarray(d[2], dtype=np.float64)
                if bounds.shape[0] != weights.shape[0] + 1:
                    raise ValueError("histogram: need len(bounds) == len(weights) + 1")
                keep += [bounds, weights]
                arr[i].n_bins = weights.shape[0]
                arr[i].bounds = bounds.ctypes.data_as(C.POINTER(C.c_double))
                arr[i].weights = weights.ctypes.data_as(C.POINTER(C.c_double))
            else:
                vals = list(d[1:]) + [0.0, 0.0, 0.0]
                arr[i].a, arr[i].b, arr[i].c = float(vals[0]), float(vals[1]), float(vals[2])
        h = C.c_void_p()
        _lib.check(_lib.lib().canopy_psource_from_deviates(arr, len(dists), C.byref(h)))
        return cls(h)

    def mean(self) -> np.ndarray:
        out = np.zeros(max(self.n_basic, 1), dtype=np.float64)
        _lib.check(_lib.lib().canopy_psource_mean(self._h, out.ctypes.data_as(C.POINTER(C.c_double))))
        return out[: self.n_basic]

    def sample(self, seed: int, trial: int) -> np.ndarray:
        out = np.zeros(max(self.n_basic, 1), dtype=np.float64)
        _lib.check(_lib.lib().canopy_psource_sample(self._h, int(seed), int(trial), out.ctypes.data_as(C.POINTER(C.c_double))))
        return out[: self.n_basic]

    def close(self):
        if self._h:
            _lib.lib().canopy_psource_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def uncertainty(graph: Graph, settings: Settings, source, n_outer: int, device: int = 0, deviate_seed: int | None = None,
                num_quantiles: int = 20, num_bins: int = 20):
    """Parameter-uncertainty outer loop (reference Settings::uncertainty_analysis, settings.h:281-295).
    `source` is a PSource: every trial draws all deviates (counter-based, canopy_psource_sample), re-quantises the
    thresholds of ONE lowered plan on the device and reruns the inner Monte Carlo of settings.num_trials samples on a
    disjoint counter range, all inside canopy_uncertainty_run.  A callable `source(k) -> float64[n_basic]` is also
    accepted (probabilities drawn by the caller; same loop through canopy_plan_update_p).
    Returns dict(p=[n_outer, n_tallies], mean, std, ci_lo, ci_hi, error_factor, quantiles=[n_tallies, num_quantiles],
    bin_edges, bin_counts, q05, q50, q95)."""
    settings.validate()
    with Plan(graph, settings.tally_all_gates, settings.threshold_bits, settings.sampler) as plan:
        nt = plan.n_tallies
        if isinstance(source, PSource):
            o = plan._opts(settings.seed, settings.sample_begin, settings.num_trials, device, settings.batch_size, None,
                           settings.overlap, False)
            u = _lib.CUncertaintyOpts()
            u.n_trials, u.first_trial = int(n_outer), 0
            u.seed = int(settings.seed if deviate_seed is None else deviate_seed)
            u.num_quantiles, u.num_bins = int(num_quantiles), int(num_bins)
            stats = (_lib.CUncertaintyStats * max(nt, 1))()
            quant = np.zeros((max(nt, 1), num_quantiles)); edges = np.zeros((max(nt, 1), num_bins + 1))
            counts = np.zeros((max(nt, 1), num_bins), dtype=np.uint64); p = np.zeros((n_outer, max(nt, 1)))
            dp = C.POINTER(C.c_double)
            _lib.check(_lib.lib().canopy_uncertainty_run(plan._h, source._h, C.byref(o), C.byref(u), stats,
                                                         quant.ctypes.data_as(dp), edges.ctypes.data_as(dp),
                                                         counts.ctypes.data_as(C.POINTER(C.c_uint64)), p.ctypes.data_as(dp)))
            p = p[:, :nt]
            f = lambda name: np.array([getattr(stats[i], name) for i in range(nt)])
            return dict(p=p, mean=f("mean"), std=f("sigma"), ci_lo=f("ci_lo"), ci_hi=f("ci_hi"), error_factor=f("error_factor"),
                        quantiles=quant[:nt], bin_edges=edges[:nt], bin_counts=counts[:nt],
                        q05=np.quantile(p, 0.05, axis=0), q50=np.quantile(p, 0.5, axis=0), q95=np.quantile(p, 0.95, axis=0))
        out = []
        for k in range(n_outer):
            plan.update_p(source(k))
            t, _ = plan.run(settings.num_trials, settings.seed, settings.sample_begin + k * settings.num_trials, device,
                            settings.batch_size, overlap=settings.overlap)
            out.append(t.astype(np.float64) / settings.num_trials)
    p = np.stack(out)
    return dict(p=p, mean=p.mean(axis=0), std=p.std(axis=0, ddof=1) if n_outer > 1 else np.zeros(p.shape[1]),
                q05=np.quantile(p, 0.05, axis=0), q50=np.quantile(p, 0.5, axis=0), q95=np.quantile(p, 0.95, axis=0))


def quantify(graph: Graph, settings: Settings, device: int = 0):
    """One-call analysis: lower, run, estimate.  Returns dict(tallies, p, sigma, ci_lo, ci_hi, report, stats)."""
    settings.validate()
    with Plan(graph, settings.tally_all_gates, settings.threshold_bits, settings.sampler) as plan:
        t, rep = plan.run(settings.num_trials, settings.seed, settings.sample_begin, device, settings.batch_size,
                          overlap=settings.overlap, profile=settings.profile)
        est = [estimate(int(x), settings.num_trials) for x in t]
        return dict(tallies=t, p=np.array([e[0] for e in est]), sigma=np.array([e[1] for e in est]),
                    ci_lo=np.array([e[2] for e in est]), ci_hi=np.array([e[3] for e in est]), report=rep,
                    stats=dict(plan.stats))


def estimate(tally: int, n_samples: int):
    e = _lib.CEstimate()
    _lib.check(_lib.lib().canopy_estimate_from_tally(tally, n_samples, C.byref(e)))
    return e.p, e.sigma, e.ci_lo, e.ci_hi


def shard_range(sample_begin: int, n_samples: int, rank: int, world: int):
    b, n = C.c_uint64(), C.c_uint64()
    _lib.check(_lib.lib().canopy_shard_range(sample_begin, n_samples, rank, world, C.byref(b), C.byref(n)))
    return int(b.value), int(n.value)


def plan_cache_clear() -> None:
    """Empty the process-wide cache of lowered plans (canopy_plan_cache_clear)."""
    _lib.check(_lib.lib().canopy_plan_cache_clear())


def trim_device_cache(device: int = -1) -> None:
    """Return the value buffers parked by destroyed plans to the driver (canopy_trim_device_cache)."""
    _lib.check(_lib.lib().canopy_trim_device_cache(device))


def microbench_int(kind: int, device: int = 0):
    ops, ms = C.c_double(), C.c_double()
    _lib.check(_lib.lib().canopy_microbench_int(device, kind, C.byref(ops), C.byref(ms)))
    return ops.value, ms.value


def load_mef(paths=None, text: str | None = None, ccf_analysis: bool = False, mission_time: float | None = None) -> Graph:
    """Open-PSA MEF XML -> Graph via the C++ loader in libcanopy (canopy_model_load_mef*_ex).
    `ccf_analysis` / `mission_time` mirror reference Settings (settings.h:190-197,298-308).
    Graph.meta["root_names"] names the outputs: fault-tree top events, then "<initiating event>.<sequence>"."""
    h = C.c_void_p()
    L = _lib.lib()
    o = _lib.CMefOpts()
    L.canopy_mef_opts_default(C.byref(o))
    o.ccf_analysis = int(bool(ccf_analysis))
    if mission_time is not None:
        o.mission_time = float(mission_time)
    if text is not None:
        _lib.check(L.canopy_model_load_mef_string_ex(text.encode(), C.byref(o), C.byref(h)))
    else:
        if isinstance(paths, (str, bytes)):
            paths = [paths]
        arr = (C.c_char_p * len(paths))(*[p.encode() if isinstance(p, str) else p for p in paths])
        _lib.check(L.canopy_model_load_mef_ex(arr, len(paths), C.byref(o), C.byref(h)))
    try:
        d = CGraphDesc()
        _lib.check(L.canopy_model_desc(h, C.byref(d)))
        def cp(ptr, n, dt):
            return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dt, copy=True) if n else np.zeros(0, dtype=dt)
        g = Graph(p=cp(d.p, d.n_basic, np.float64), op=cp(d.op, d.n_gates, np.uint8),
                  kmin=cp(d.kmin, d.n_gates, np.uint16), kmax=cp(d.kmax, d.n_gates, np.uint16),
                  child_off=cp(d.child_off, d.n_gates + 1, np.uint64), child=cp(d.child, d.n_edges, np.uint32),
                  roots=cp(d.roots, d.n_roots, np.uint32),
                  basic_names=[L.canopy_model_basic_name(h, i).decode() for i in range(d.n_basic)],
                  gate_names=[L.canopy_model_gate_name(h, i).decode() for i in range(d.n_gates)])
        g.meta["root_names"] = [L.canopy_model_root_name(h, i).decode() for i in range(d.n_roots)]
        ps = C.c_void_p()
        _lib.check(L.canopy_psource_from_model(h, C.byref(ps)))
        g.meta["psource"] = PSource(ps)   # the model's expressions: mean and sampled probabilities (uncertainty loop)
        return g
    finally:
        L.canopy_model_destroy(h)
