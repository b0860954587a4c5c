"""ctypes loader for the in-tree libcanopy.so (C ABI in include/canopy.h).

There is no Python or CPU fallback: if the shared library is missing this raises, loudly.
"""
from __future__ import annotations

import ctypes as C
import os

from .graph import CGraphDesc

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CANOPY_LIB") or os.path.join(_HERE, "libcanopy.so")  # CANOPY_LIB: kernel-tuning builds only

STATUS_NAMES = {0: "OK", 1: "E_INVALID", 2: "E_VALIDITY", 3: "E_CYCLE", 4: "E_UNSUPPORTED", 5: "E_CUDA",
                6: "E_NCCL", 7: "E_OOM", 8: "E_DOMAIN", 9: "E_IO"}
OK, E_INVALID, E_VALIDITY, E_CYCLE, E_UNSUPPORTED, E_CUDA, E_NCCL, E_OOM, E_DOMAIN, E_IO = range(10)


class CanopyError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"canopy {STATUS_NAMES.get(status, status)}: {message}")
        self.status = status
        self.message = message


class CPlanOpts(C.Structure):
    _fields_ = [("tally_all_gates", C.c_int32), ("threshold_bits", C.c_int32), ("sampler", C.c_int32),
                ("no_plan_cache", C.c_int32), ("engine", C.c_int32), ("reserved", C.c_int32 * 3)]


class CPlanStats(C.Structure):
    _fields_ = [("n_basic_in", C.c_uint32), ("n_gates_in", C.c_uint32), ("n_edges_in", C.c_uint64),
                ("n_basic_live", C.c_uint32), ("n_gates_live", C.c_uint32), ("n_edges_live", C.c_uint64),
                ("n_gates_stored", C.c_uint32), ("n_gates_tallied", C.c_uint32), ("n_layers", C.c_uint32),
                ("max_fanin", C.c_uint32), ("mean_fanin", C.c_double), ("n_tallies", C.c_uint32),
                ("n_const_nodes", C.c_uint32), ("n_alias_nodes", C.c_uint32), ("spec_version", C.c_uint32),
                ("plan_cache_hit", C.c_uint32)]


class CPlanView(C.Structure):
    _fields_ = [("n_events", C.c_uint32), ("n_gates", C.c_uint32), ("n_layers", C.c_uint32), ("n_tallies", C.c_uint32),
                ("n_edges", C.c_uint64),
                ("event_orig", C.POINTER(C.c_uint32)), ("event_thr32", C.POINTER(C.c_uint32)),
                ("event_thr64", C.POINTER(C.c_uint64)),
                ("gate_child_off", C.POINTER(C.c_uint32)), ("gate_child", C.POINTER(C.c_uint32)),
                ("gate_op", C.POINTER(C.c_uint8)), ("gate_flags", C.POINTER(C.c_uint8)),
                ("gate_kmin", C.POINTER(C.c_uint16)), ("gate_kmax", C.POINTER(C.c_uint16)),
                ("layer_off", C.POINTER(C.c_uint32)),
                ("tally_kind", C.POINTER(C.c_uint8)), ("tally_gate", C.POINTER(C.c_uint32)),
                ("event_k", C.POINTER(C.c_uint8))]


class CMcOpts(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("sample_begin", C.c_uint64), ("n_samples", C.c_uint64),
                ("device", C.c_int32), ("batch_slabs", C.c_int32), ("stream", C.c_void_p),
                ("overlap", C.c_int32), ("profile", C.c_int32), ("profile_walk_shape", C.c_int32), ("reserved", C.c_int32 * 3)]


class CMcReport(C.Structure):
    _fields_ = [("total_ms", C.c_double), ("sampler_ms", C.c_double), ("gate_ms", C.c_double),
                ("sampler_launches", C.c_uint64), ("gate_launches", C.c_uint64), ("other_launches", C.c_uint64),
                ("n_batches", C.c_uint64), ("slab_samples", C.c_uint64), ("batch_slabs", C.c_uint64),
                ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64)]


class CCheckpoint(C.Structure):
    _fields_ = [("magic", C.c_uint32), ("spec_version", C.c_uint32), ("threshold_bits", C.c_uint32), ("n_tallies", C.c_uint32),
                ("plan_fingerprint", C.c_uint64), ("seed", C.c_uint64), ("sample_begin", C.c_uint64), ("next_sample", C.c_uint64)]


class CMefOpts(C.Structure):
    _fields_ = [("ccf_analysis", C.c_int32), ("reserved0", C.c_int32), ("mission_time", C.c_double),
                ("reserved", C.c_int32 * 4)]


class CDeviate(C.Structure):
    _fields_ = [("kind", C.c_int32), ("n_bins", C.c_uint32), ("a", C.c_double), ("b", C.c_double), ("c", C.c_double),
                ("bounds", C.POINTER(C.c_double)), ("weights", C.POINTER(C.c_double))]


class CUncertaintyOpts(C.Structure):
    _fields_ = [("n_trials", C.c_uint64), ("first_trial", C.c_uint64), ("seed", C.c_uint64),
                ("num_quantiles", C.c_uint32), ("num_bins", C.c_uint32), ("reserved", C.c_int32 * 4)]


class CUncertaintyStats(C.Structure):
    _fields_ = [("mean", C.c_double), ("sigma", C.c_double), ("ci_lo", C.c_double), ("ci_hi", C.c_double),
                ("error_factor", C.c_double), ("min", C.c_double), ("max", C.c_double)]


class CPlanStreams(C.Structure):
    _fields_ = [("n_dev_child", C.c_uint32), ("warps", C.c_uint32), ("n_layers", C.c_uint32), ("reserved", C.c_uint32),
                ("dev_child", C.POINTER(C.c_uint32)), ("gate_child_off", C.POINTER(C.c_uint32)),
                ("split", C.POINTER(C.c_uint32)), ("split_child", C.POINTER(C.c_uint32))]


class CEstimate(C.Structure):
    _fields_ = [("p", C.c_double), ("sigma", C.c_double), ("ci_lo", C.c_double), ("ci_hi", C.c_double)]


# every symbol include/canopy.h declares: name -> (restype, argtypes)
_u64p = C.POINTER(C.c_uint64)
SYMBOLS = {
    "canopy_plan_create": (C.c_int, [C.POINTER(CGraphDesc), C.POINTER(CPlanOpts), C.POINTER(C.c_void_p)]),
    "canopy_plan_destroy": (None, [C.c_void_p]),
    "canopy_plan_info": (C.c_int, [C.c_void_p, C.POINTER(CPlanStats)]),
    "canopy_plan_view_get": (C.c_int, [C.c_void_p, C.POINTER(CPlanView)]),
    "canopy_mc_opts_default": (None, [C.POINTER(CMcOpts)]),
    "canopy_mc_run": (C.c_int, [C.c_void_p, C.POINTER(CMcOpts), _u64p, C.POINTER(CMcReport)]),
    "canopy_mc_run_device": (C.c_int, [C.c_void_p, C.POINTER(CMcOpts), C.c_void_p, C.POINTER(CMcReport)]),
    "canopy_mc_run_static": (C.c_int, [C.c_void_p, C.POINTER(CMcOpts), _u64p, C.c_void_p]),
    "canopy_mc_run_multi": (C.c_int, [C.c_void_p, C.POINTER(CMcOpts), C.POINTER(C.c_int32), C.c_int32, _u64p, C.POINTER(CMcReport)]),
    "canopy_trim_device_cache": (C.c_int, [C.c_int32]),
    "canopy_plan_cache_clear": (C.c_int, []),
    "canopy_plan_fingerprint": (C.c_int, [C.c_void_p, _u64p]),
    "canopy_checkpoint_init": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint64, C.POINTER(CCheckpoint)]),
    "canopy_mc_run_resume": (C.c_int, [C.c_void_p, C.POINTER(CMcOpts), C.POINTER(CCheckpoint), _u64p, C.POINTER(CMcReport)]),
    "canopy_plan_update_p": (C.c_int, [C.c_void_p, C.POINTER(C.c_double), C.c_uint32]),
    "canopy_mc_sample_words": (C.c_int, [C.c_void_p, C.POINTER(CMcOpts), C.c_uint64, C.c_uint64, _u64p]),
    "canopy_shard_range": (C.c_int, [C.c_uint64, C.c_uint64, C.c_int32, C.c_int32, _u64p, _u64p]),
    "canopy_estimate_from_tally": (C.c_int, [C.c_uint64, C.c_uint64, C.POINTER(CEstimate)]),
    "canopy_microbench_int": (C.c_int, [C.c_int32, C.c_int32, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "canopy_model_load_mef": (C.c_int, [C.POINTER(C.c_char_p), C.c_int32, C.POINTER(C.c_void_p)]),
    "canopy_model_load_mef_string": (C.c_int, [C.c_char_p, C.POINTER(C.c_void_p)]),
    "canopy_mef_opts_default": (None, [C.POINTER(CMefOpts)]),
    "canopy_model_load_mef_ex": (C.c_int, [C.POINTER(C.c_char_p), C.c_int32, C.POINTER(CMefOpts), C.POINTER(C.c_void_p)]),
    "canopy_model_load_mef_string_ex": (C.c_int, [C.c_char_p, C.POINTER(CMefOpts), C.POINTER(C.c_void_p)]),
    "canopy_model_root_name": (C.c_char_p, [C.c_void_p, C.c_uint32]),
    "canopy_model_desc": (C.c_int, [C.c_void_p, C.POINTER(CGraphDesc)]),
    "canopy_model_basic_name": (C.c_char_p, [C.c_void_p, C.c_uint32]),
    "canopy_model_gate_name": (C.c_char_p, [C.c_void_p, C.c_uint32]),
    "canopy_model_destroy": (None, [C.c_void_p]),
    "canopy_psource_from_model": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "canopy_psource_from_deviates": (C.c_int, [C.POINTER(CDeviate), C.c_uint32, C.POINTER(C.c_void_p)]),
    "canopy_psource_info": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "canopy_psource_mean": (C.c_int, [C.c_void_p, C.POINTER(C.c_double)]),
    "canopy_psource_sample": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint64, C.POINTER(C.c_double)]),
    "canopy_psource_destroy": (None, [C.c_void_p]),
    "canopy_uncertainty_run": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(CMcOpts), C.POINTER(CUncertaintyOpts),
                                         C.POINTER(CUncertaintyStats), C.POINTER(C.c_double), C.POINTER(C.c_double),
                                         _u64p, C.POINTER(C.c_double)]),
    "canopy_last_error": (C.c_char_p, []),
    "canopy_status_string": (C.c_char_p, [C.c_int]),
    "canopy_version": (C.c_uint32, []),
    "canopy_spec_version": (C.c_uint32, []),
    "canopy_sparse_class": (C.c_int, [C.c_uint64, C.POINTER(C.c_uint32)]),
    "canopy_sparse_cdf": (C.c_int, [C.c_uint64, C.c_uint32, C.POINTER(C.c_uint64)]),
}
# include/canopy-testing.h: exported by libcanopy_testing.so only (the product library plus csrc/capi_testing.cpp)
TESTING_SYMBOLS = {
    "canopy_plan_debug_streams": (C.c_int, [C.c_void_p, C.POINTER(CPlanStreams)]),
    "canopy_plan_debug_set_sparse_table": (C.c_int, [C.c_void_p, C.c_uint32, C.POINTER(C.c_uint64)]),
    "canopy_debug_force_row_copies": (C.c_int, [C.c_int32]),
}
TESTING_LIB_PATH = os.path.join(_HERE, "libcanopy_testing.so")
_TESTING = False


def use_testing_library():
    """Called by tests/conftest.py before anything is loaded: bind libcanopy_testing.so, which adds the test hooks of
    include/canopy-testing.h to the very same objects libcanopy.so is linked from.  Product code never calls this."""
    global LIB_PATH, _TESTING
    if _LIB is not None and not _TESTING:
        raise RuntimeError("libcanopy.so is already loaded; use_testing_library() must come first")
    if not os.environ.get("CANOPY_LIB"):
        LIB_PATH = TESTING_LIB_PATH
    _TESTING = True

_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "or `make -C canopy_b200/csrc`. canopy_b200 has no CPU or pure-Python fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in list(SYMBOLS.items()) + (list(TESTING_SYMBOLS.items()) if _TESTING else []):
            fn = getattr(L, name)  # AttributeError here means the .so does not match include/canopy.h
            fn.restype = res
            fn.argtypes = args
        _LIB = L
    return _LIB


def check(status: int):
    if status != 0:
        msg = lib().canopy_last_error()
        raise CanopyError(status, msg.decode() if msg else "")
