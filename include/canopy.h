/*
 * canopy.h -- C ABI of the B200-native Monte Carlo quantification engine.
 *
 * Packaging contract kept from the reference (openpra-org/canopy):
 *   - shared library target `canopy`              (reference src/CMakeLists.txt:9-12, CMakeLists.txt:21)
 *   - installed public headers canopy.h / canopy-cpp.h   (reference CMakeLists.txt:60-68)
 *   - export macros keyed on CANOPY_BUILD / CANOPY_SHARED (reference src/CMakeLists.txt:21-22)
 * The reference header itself is an empty include guard (reference include/canopy.h:33-36)
 * and the reference ships no quantification code, so every entry point below is a
 * *designed* boundary; each one cites the reference type it consumes or stands in for.
 *
 * Conventions
 *   - every function returns canopy_status and never throws across the boundary; the
 *     reference signals errors with C++ exceptions (reference src/mef/openpsa/error.h:23-96),
 *     which map to the status codes below; the message is available from
 *     canopy_last_error() (thread-local).
 *   - caller owns every array in a descriptor and every output buffer; the library owns
 *     canopy_plan and all device memory.
 *   - there is NO CPU fallback: canopy_mc_* fail with CANOPY_E_CUDA when no sm_100 device
 *     is usable. Lowering (canopy_plan_*) is host-only and works without a GPU.
 */
#ifndef CANOPY_H
#define CANOPY_H

#include <stddef.h>
#include <stdint.h>

#ifdef CANOPY_SHARED
#  if defined(_WIN32) && !defined(__MINGW32__)
#    ifdef CANOPY_BUILD
#      define CANOPY_API __declspec(dllexport)
#    else
#      define CANOPY_API __declspec(dllimport)
#    endif
#  else
#    define CANOPY_API __attribute__((visibility("default")))
#  endif
#else
#  define CANOPY_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define CANOPY_VERSION_MAJOR 0
#define CANOPY_VERSION_MINOR 1
/* Version of the sampling specification. Tallies are a pure function of
 * (spec version, graph, seed, sample range). Three specs are implemented and selected per plan
 * (canopy_plan_opts.sampler); all draw Philox4x32-10 variates and compare them with thresholds derived from
 * the event probability:
 *   spec 1  "direct": one 32/64-bit variate per (event, sample), compared whole.
 *   spec 2  "sliced": the same comparison evaluated MSB-first; the top CANOPY_SLICED_HEAD_BITS bits of
 *           the variates of 128 consecutive samples come from that many bit-sliced Philox calls, the
 *           remaining bits from tail calls that only the ~2^-H samples still undecided after the head
 *           bits ever consume (4 undecided samples per call; 2 with 64-bit thresholds).
 *   spec 4  "classed sparse": spec 2 for every event with p >= 2^-CANOPY_SPARSE_MIN_ZEROS.  A rarer event works on groups
 *           of 2^k samples, k = min(CANOPY_SPARSE_MAX_LOG2, z + 2) with z the leading zero bits of its 64-bit threshold
 *           (fewer than four expected ones per group): ONE Philox call per group; a 64-bit variate compared with the
 *           event's Binomial(2^k, thr/2^64) cumulative table (CANOPY_SPARSE_CDF_LEN entries, exact integer arithmetic)
 *           gives the number of ones of the group, k-bit fields of the same call place them (duplicates skipped;
 *           further calls only if the fields run out).  Same i.i.d. Bernoulli(thr/2^bits) law per sample up to the
 *           table's rounding (below 2^-50) and the mass beyond CANOPY_SPARSE_CDF_LEN ones (below 1e-17).  Serves 32- and
 *           64-bit thresholds alike (a 32-bit threshold is widened to thr32 << 32).
 *   (spec 3, this spec with 128-sample groups only, was retired in its favour: CANOPY_E_UNSUPPORTED.)
 * CANOPY_SPEC_VERSION is the default spec of this build. */
#define CANOPY_SPEC_VERSION 4
#ifndef CANOPY_SLICED_HEAD_BITS
#define CANOPY_SLICED_HEAD_BITS 7   /* H of spec 2; part of the spec (changing it changes every tally) */
#endif
#define CANOPY_SAMPLER_DEFAULT 0   /* = the build default (spec 4) */
#define CANOPY_SAMPLER_DIRECT  1   /* spec 1 */
#define CANOPY_SAMPLER_SLICED  2   /* spec 2 */
#define CANOPY_SAMPLER_SPARSE  4   /* spec 4 */
#ifndef CANOPY_SPARSE_MIN_ZEROS
#define CANOPY_SPARSE_MIN_ZEROS 5  /* spec 4: events with p < 2^-5 take the sparse path */
#endif
#ifndef CANOPY_SPARSE_MAX_LOG2
#define CANOPY_SPARSE_MAX_LOG2 13  /* spec 4: largest group is 2^13 = 8192 samples (one slab row) */
#endif
#ifndef CANOPY_SPARSE_CDF_LEN
#define CANOPY_SPARSE_CDF_LEN  32  /* spec 4: entries of the cumulative count table = largest count a group can draw */
#endif

typedef enum canopy_status {
    CANOPY_OK = 0,
    CANOPY_E_INVALID = 1,     /* null pointer, bad option, id out of range          (LogicError, error.h:55)      */
    CANOPY_E_VALIDITY = 2,    /* connective arity / min / max rules violated         (ValidityError, error.h:76)   */
    CANOPY_E_CYCLE = 3,       /* gate graph is not acyclic                           (CycleError, error.h:93)      */
    CANOPY_E_UNSUPPORTED = 4, /* valid MEF construct this engine does not evaluate                                 */
    CANOPY_E_CUDA = 5,        /* CUDA failure, no device, or device is not sm_100                                  */
    CANOPY_E_NCCL = 6,        /* NCCL failure or libnccl.so.2 not loadable                                         */
    CANOPY_E_OOM = 7,         /* host or device allocation failed                                                  */
    CANOPY_E_DOMAIN = 8,      /* probability outside [0,1] or NaN                    (DomainError, error.h:98)     */
    CANOPY_E_IO = 9           /* model file unreadable / malformed XML               (IOError, io/xml/error.h:34)  */
} canopy_status;

/* Same numeric values as `enum Connective` in reference src/mef/openpsa/event/event.h:143-157. */
typedef enum canopy_connective {
    CANOPY_AND = 0,
    CANOPY_OR = 1,
    CANOPY_ATLEAST = 2,      /* k-of-n, k = kmin                                   */
    CANOPY_XOR = 3,          /* exactly 2 args                                     */
    CANOPY_NOT = 4,          /* exactly 1 arg                                      */
    CANOPY_NAND = 5,
    CANOPY_NOR = 6,
    CANOPY_NULL = 7,         /* pass-through, exactly 1 arg                        */
    CANOPY_IFF = 8,          /* exactly 2 args                                     */
    CANOPY_IMPLY = 9,        /* exactly 2 args, arg0 -> arg1                       */
    CANOPY_CARDINALITY = 10  /* kmin <= (#true args) <= kmax                       */
} canopy_connective;
#define CANOPY_NUM_CONNECTIVES 11

/* Child reference encoding (one uint32 per formula argument, reference
 * src/mef/openpsa/event/formula.h:9-15 `Arg{complement, event}`):
 *   bit 31     complement flag
 *   bits 0-30  node id: [0, n_basic) basic event, [n_basic, n_basic+n_gates) gate,
 *              or one of the two constants (HouseEvent::kFalse/kTrue, event.h:37-38). */
#define CANOPY_CHILD_COMPLEMENT 0x80000000u
#define CANOPY_CHILD_ID_MASK    0x7FFFFFFFu
#define CANOPY_CONST_FALSE      0x7FFFFFFEu
#define CANOPY_CONST_TRUE       0x7FFFFFFFu

/* Flat, caller-owned description of a MEF gate DAG. This is what a walk over
 * Model -> FaultTree::top_events() -> Gate::formula().args() (reference fault_tree.h:142,
 * event/gate.h:45-48, event/formula.h:100) produces; house events are passed as the two
 * constants. Read during the call only. */
typedef struct canopy_graph_desc {
    uint32_t n_basic;
    uint32_t n_gates;
    uint32_t n_roots;
    uint32_t reserved0;
    uint64_t n_edges;
    const double*   p;          /* [n_basic]   BasicEvent::p() (event.h:93-96), must be in [0,1]           */
    const uint8_t*  op;         /* [n_gates]   canopy_connective                                           */
    const uint16_t* kmin;       /* [n_gates]   Formula::min_number (formula.h:145); ignored unless atleast/cardinality */
    const uint16_t* kmax;       /* [n_gates]   Formula::max_number (formula.h:146); ignored unless cardinality         */
    const uint64_t* child_off;  /* [n_gates+1] CSR offsets into child[]                                    */
    const uint32_t* child;      /* [n_edges]   encoded as above, argument order preserved (matters for imply) */
    const uint32_t* roots;      /* [n_roots]   node ids (gate or basic event) whose tallies are reported first */
} canopy_graph_desc;

typedef struct canopy_plan_opts {
    int32_t tally_all_gates;   /* 0: tallies = [n_roots]; 1: tallies = [n_roots] ++ [n_gates] (descriptor order) */
    int32_t threshold_bits;    /* 32 (default) or 64: width of the variate / threshold comparison                */
    int32_t sampler;           /* CANOPY_SAMPLER_*: which sampling spec the plan's tallies follow (0 = default)   */
    int32_t no_plan_cache;     /* 1: lower from scratch and do not keep the result (see canopy_plan_create)       */
    int32_t engine;            /* CANOPY_ENGINE_*: how runs of this plan are executed (same tallies either way)   */
    int32_t reserved[3];
} canopy_plan_opts;
/* AUTO: the fused kernel when the node rows of one slab fit in shared memory (about 200 live nodes; sampler, gate walk and
 * tallies in one launch, nothing but tallies leaves the SM), else the streamed pair of kernels over the value buffer in HBM.
 * STREAMED forces the latter; FUSED fails at plan creation (CANOPY_E_UNSUPPORTED) when the model is too large or the plan
 * uses the direct sampler (spec 1). */
#define CANOPY_ENGINE_AUTO     0
#define CANOPY_ENGINE_STREAMED 1
#define CANOPY_ENGINE_FUSED    2

/* Lowering statistics, reported before and after dead-code elimination / folding. */
typedef struct canopy_plan_stats {
    uint32_t n_basic_in, n_gates_in;
    uint64_t n_edges_in;
    uint32_t n_basic_live;     /* basic events that are actually sampled                         */
    uint32_t n_gates_live;     /* device gates after alias/constant folding + DCE                */
    uint64_t n_edges_live;     /* E: sum of fan-in over live gates                               */
    uint32_t n_gates_stored;   /* G': live gates whose output word is consumed by a later layer  */
    uint32_t n_gates_tallied;
    uint32_t n_layers;
    uint32_t max_fanin;
    double   mean_fanin;
    uint32_t n_tallies;        /* length of the tallies output array                             */
    uint32_t n_const_nodes;    /* roots/gates folded to constants                                */
    uint32_t n_alias_nodes;    /* gates folded to (possibly complemented) aliases                */
    uint32_t spec_version;     /* the spec this plan samples with (1 direct, 2 sliced, 4 classed sparse) */
    uint32_t plan_cache_hit;   /* 1: the lowered form came from the plan cache                    */
} canopy_plan_stats;

typedef struct canopy_plan canopy_plan; /* opaque; immutable lowered graph + lazily created per-device state */

/* Host lowering: validate (reference event.cpp:140-204 arity/min/max/nesting rules), fold constants
 * and NOT/NULL aliases, eliminate dead nodes, levelise, emit CSR. No GPU needed.
 * The last few lowered forms are kept in a process-wide cache keyed by a hash of the descriptor and verified by
 * comparing the arrays, so creating a plan for a model that was lowered before costs one pass over the descriptor
 * instead of the lowering (c4: 42 ms); opts->no_plan_cache = 1 bypasses it, canopy_plan_cache_clear() empties it. */
CANOPY_API canopy_status canopy_plan_create(const canopy_graph_desc* desc, const canopy_plan_opts* opts,
                                            canopy_plan** out_plan);
CANOPY_API void          canopy_plan_destroy(canopy_plan* plan);
CANOPY_API canopy_status canopy_plan_info(const canopy_plan* plan, canopy_plan_stats* out);
CANOPY_API canopy_status canopy_plan_cache_clear(void);

/* Introspection of the lowered form (used by the host-logic tests; all pointers stay valid until
 * canopy_plan_destroy). Device node ids: [0,n_basic_live) events, then gates in layer order. */
typedef struct canopy_plan_view {
    uint32_t n_events, n_gates, n_layers, n_tallies;
    uint64_t n_edges;
    const uint32_t* event_orig;   /* [n_events] original basic-event index (feeds the Philox counter)      */
    const uint32_t* event_thr32;  /* [n_events] floor(p * 2^32)                                             */
    const uint64_t* event_thr64;  /* [n_events] floor(p * 2^64) (valid when threshold_bits == 64)           */
    const uint32_t* gate_child_off; /* [n_gates+1]                                                          */
    const uint32_t* gate_child;   /* [n_edges] device node id | complement bit 31                           */
    const uint8_t*  gate_op;      /* [n_gates] device op: 0 and, 1 or, 2 xor, 3 count-in-range              */
    const uint8_t*  gate_flags;   /* [n_gates] bit0 invert output, bit1 store, bit2 tally                   */
    const uint16_t* gate_kmin;    /* [n_gates]                                                              */
    const uint16_t* gate_kmax;    /* [n_gates]                                                              */
    const uint32_t* layer_off;    /* [n_layers+1] gate index ranges                                         */
    const uint8_t*  tally_kind;   /* [n_tallies] 0 const-false, 1 const-true, 2 device gate, 3 complement of device gate */
    const uint32_t* tally_gate;   /* [n_tallies] device gate index for kinds 2/3                            */
    const uint8_t*  event_k;      /* [n_events] spec 4: 0 dense (planes), else log2 of the event's group size; NULL under specs 1, 2 */
} canopy_plan_view;
CANOPY_API canopy_status canopy_plan_view_get(const canopy_plan* plan, canopy_plan_view* out);

/* Mirrors the Monte Carlo fields of reference Settings (settings.h:118-151,178-187,324-327):
 * num_trials -> n_samples, seed -> seed (both widened from int to 64 bit); batch_size -> batch_slabs
 * (0 = engine default). sample_begin makes a run restartable / shardable by counter range. */
typedef struct canopy_mc_opts {
    uint64_t seed;
    uint64_t sample_begin;
    uint64_t n_samples;
    int32_t  device;          /* CUDA ordinal                                                       */
    int32_t  batch_slabs;     /* slabs (canopy_mc_report.slab_samples samples each) per batch; 0 = default */
    void*    stream;          /* cudaStream_t to order the run on; NULL = library-owned stream (pass
                                 cudaStreamLegacy / cudaStreamPerThread explicitly for the default streams) */
    int32_t  overlap;         /* 1 (set by canopy_mc_opts_default): the sampler of batch i+1 runs beside the gate walk of
                                 batch i (programmatic dependent launch on the run's stream) */
    int32_t  profile;         /* 1: time sampler / gate kernels separately with CUDA events (serialises the two) */
    int32_t  profile_walk_shape; /* profile = 1 only.  0: each gate walk is launched as a walk that has the GPU to itself
                                 (several CTAs per SM); 1: in the shape the overlapped pipeline gives it (one CTA per SM,
                                 the rest of the SM left to the sampler of the next batch), but still alone -- what the
                                 timed steps actually launch, for the roofline */
    int32_t  reserved[3];
} canopy_mc_opts;
CANOPY_API void canopy_mc_opts_default(canopy_mc_opts* opts);

/* Per-run measurements taken with CUDA events on the run's stream. */
typedef struct canopy_mc_report {
    double   total_ms;          /* device time first launch -> last kernel done                     */
    double   sampler_ms;        /* sum over batches (profile=1 only, else 0)                        */
    double   gate_ms;           /* sum over batches (profile=1 only, else 0)                        */
    uint64_t sampler_launches;
    uint64_t gate_launches;
    uint64_t other_launches;    /* tally clear etc.                                                 */
    uint64_t n_batches;
    uint64_t slab_samples;      /* samples per slab (8192: one 1-KB row per node)                   */
    uint64_t batch_slabs;
    uint64_t h2d_bytes, d2h_bytes;
} canopy_mc_report;

/* Run S = n_samples trials of the whole DAG on one device; tallies[i] = number of samples in
 * [sample_begin, sample_begin+n_samples) for which output i is true. Host output buffer. */
CANOPY_API canopy_status canopy_mc_run(canopy_plan* plan, const canopy_mc_opts* opts,
                                       uint64_t* tallies_host, canopy_mc_report* report);
/* Same, but tallies are left in device memory (uint64[n_tallies], caller-allocated on opts->device),
 * ordered on opts->stream: the buffer a collective (NCCL all-reduce) consumes with no host hop. */
CANOPY_API canopy_status canopy_mc_run_device(canopy_plan* plan, const canopy_mc_opts* opts,
                                              uint64_t* tallies_device, canopy_mc_report* report);
/* The same tallies as canopy_mc_run / canopy_mc_run_device, computed WITHOUT allocating anything on the device.
 * (This entry point has no counterpart in the reference, which has no device code at all.)
 *
 * Who needs it.  This library gets loaded into other people's processes: a monitoring service or a solver whose own
 * allocator took the GPU's memory at start-up, a job inside a MIG slice or container quota that is already spoken for, a
 * control loop that bans cudaMalloc / cudaFree after initialisation because both stall every stream on the device while
 * they run.  For such a host the first canopy_mc_run of a plan is a problem: that is when the plan is made resident
 * (cudaMalloc for the plan arrays, a stream, events).  This call never does any of that -- no cudaMalloc, cudaFree,
 * stream or event creation, on any call, including the first.
 *
 * How.  It runs the fused engine (CANOPY_ENGINE_FUSED: sampler, gate walk and tallies in one kernel over shared-memory
 * rows).  The little global memory that engine needs besides the caller's output -- a device copy of the plan arrays,
 * the spec-4 count tables, one counter per gate -- sits in a fixed array of CANOPY_STATIC_SCRATCH_BYTES compiled into the
 * library.  Every GPU on which the library's device code is loaded carries that array until the process exits.
 *
 * Limits.  Only plans the fused engine can run: at least one live gate, at most about 200 live nodes (the rows of one slab
 * must fit in shared memory), sampler spec 2 or 4; and the plan's device image must fit in the array (every plan within
 * the node limit seen so far does, with room to spare).  Anything larger is refused with CANOPY_E_UNSUPPORTED and a
 * message that names the size and the limit; use canopy_mc_run for it.  The plan's own `engine` option is not consulted:
 * tallies do not depend on the engine.
 *
 * Arguments.  opts->seed / sample_begin / n_samples / device as for canopy_mc_run.  opts->stream orders the work; NULL means
 * the calling thread's cudaStreamPerThread (there is no library-owned stream here, creating one would allocate).
 * batch_slabs / overlap / profile are ignored.  Outputs: tallies_host (uint64[n_tallies], any host memory; the call then
 * returns after the stream has drained) and/or tallies_device (uint64[n_tallies] in memory the CALLER already owns on
 * opts->device; with tallies_host == NULL the call returns with the work still queued on the stream).  At least one of
 * the two must be given unless the plan has no tallies.
 *
 * Sharing.  There is ONE array per GPU, used by every call, every plan and every thread, and it keeps whatever the last
 * call left in it.  The call therefore re-uploads the plan image and zeroes the counters every time (a few tens of KB),
 * and two calls whose GPU work overlaps in time -- different streams, or device-output calls issued back to back from two
 * threads -- overwrite each other's counters and give wrong tallies.  Host-side entry is serialised per device; keeping
 * the GPU work of successive calls in order (same stream, or an event between streams) is up to the caller.
 *
 * First call.  The first call on each device looks up the array's device address and sets kernel attributes for that
 * device.  Make it during initialisation, before any stream capture is begun. */
#define CANOPY_STATIC_SCRATCH_BYTES (128u * 1024u)
CANOPY_API canopy_status canopy_mc_run_static(canopy_plan* plan, const canopy_mc_opts* opts,
                                              uint64_t* tallies_host, uint64_t* tallies_device);
/* Single-process multi-GPU: shards [sample_begin, +n_samples) into n_devices contiguous counter
 * ranges, one host thread per device, then one ncclAllReduce(ncclUint64, ncclSum) over the tallies. */
CANOPY_API canopy_status canopy_mc_run_multi(canopy_plan* plan, const canopy_mc_opts* opts,
                                             const int32_t* devices, int32_t n_devices,
                                             uint64_t* tallies_host, canopy_mc_report* report);

/* Value buffers (GBs) of destroyed plans are parked per device and reused by the next plan, so that a one-shot
 * quantification or a loop that lowers a new plan per trial does not pay cudaMalloc/cudaFree of that size each time
 * (at most two buffers per device are kept).  This returns them to the driver; device < 0 = every device. */
CANOPY_API canopy_status canopy_trim_device_cache(int32_t device);

/* Restart record (SURVEY.md 8f-3, 5 "checkpoint / resume"): with a counter-based generator the whole state of a
 * run is (seed, next sample index, tallies).  The caller keeps `canopy_checkpoint` + the uint64 tally array (plain
 * data, safe to write to disk); canopy_mc_run_resume() evaluates the NEXT opts->n_samples samples, adds their
 * tallies into `tallies_host_accum` and advances the record.  opts->seed / opts->sample_begin are ignored (the
 * record holds them).  Tallies accumulated in any chunking equal a single run over the same range. */
#define CANOPY_CHECKPOINT_MAGIC 0x59504E43u  /* "CNPY" */
typedef struct canopy_checkpoint {
    uint32_t magic;
    uint32_t spec_version;      /* sampling spec of the plan                                      */
    uint32_t threshold_bits;
    uint32_t n_tallies;
    uint64_t plan_fingerprint;  /* canopy_plan_fingerprint() of the plan the tallies belong to    */
    uint64_t seed;
    uint64_t sample_begin;      /* tallies cover samples [sample_begin, next_sample)              */
    uint64_t next_sample;
} canopy_checkpoint;
/* 64-bit FNV-1a over the lowered arrays, thresholds and plan options: equal for equal (graph, options). */
CANOPY_API canopy_status canopy_plan_fingerprint(const canopy_plan* plan, uint64_t* out);
CANOPY_API canopy_status canopy_checkpoint_init(const canopy_plan* plan, uint64_t seed, uint64_t sample_begin,
                                                canopy_checkpoint* out);
CANOPY_API canopy_status canopy_mc_run_resume(canopy_plan* plan, const canopy_mc_opts* opts, canopy_checkpoint* ck,
                                              uint64_t* tallies_host_accum, canopy_mc_report* report);

/* Parameter-uncertainty outer loop: replace the basic-event probabilities of an existing plan
 * (same graph, same lowering) and re-quantise the thresholds; the next run uses them.  p[n_basic] in descriptor
 * order.  The constant folding done at lowering must stay valid: an event that was folded to a constant
 * (p <= 0, p >= 1, or a threshold of 0) must fold to the same constant, and a sampled event must stay sampled;
 * otherwise CANOPY_E_INVALID (create a new plan).  Changes canopy_plan_fingerprint().
 * May be called while other threads run the same plan: it waits for the host side of every run in progress (and for
 * kernels a device-output run left in flight), and a run started meanwhile waits for it, so each run samples with the
 * old probabilities or with the new ones, never with a mixture. */
CANOPY_API canopy_status canopy_plan_update_p(canopy_plan* plan, const double* p, uint32_t n_basic);

/* Parity hook: run only the sampler kernel and return the packed words of the live basic events,
 * words_host[e * n_words + w], e in canopy_plan_view.event_orig order, w = global word index
 * first_word .. first_word+n_words-1 (sample s is bit s&63 of word s>>6). */
CANOPY_API canopy_status canopy_mc_sample_words(canopy_plan* plan, const canopy_mc_opts* opts,
                                                uint64_t first_word, uint64_t n_words,
                                                uint64_t* words_host);

/* Shard helper: contiguous, 4096-sample-aligned split of a sample range over `world`
 * ranks; the union over ranks is exactly the input range for every world size. */
CANOPY_API canopy_status canopy_shard_range(uint64_t sample_begin, uint64_t n_samples,
                                            int32_t rank, int32_t world,
                                            uint64_t* out_begin, uint64_t* out_n);

/* Point estimate from a tally: p_hat = tally/n, sigma = sqrt(p(1-p)/n), 95% normal CI. */
typedef struct canopy_estimate { double p, sigma, ci_lo, ci_hi; } canopy_estimate;
CANOPY_API canopy_status canopy_estimate_from_tally(uint64_t tally, uint64_t n_samples, canopy_estimate* out);

/* Integer-pipe microbenchmarks used for the sampler roofline (lane-instructions per second).
 * kind: 0 LOP3, 1 IMAD.WIDE.U32 (+1 LOP3 each, counted once), 2 Philox4x32-10 rounds (2 IMAD.WIDE + 2 LOP3 per round),
 *       3 POPC, 4 mix of 1 IMAD.WIDE.U32 : 2 LOP3 (all three counted), 5 uniform-datapath pairs, 6 kind 5 beside 2 LOP3 */
CANOPY_API canopy_status canopy_microbench_int(int32_t device, int32_t kind, double* out_lane_ops_per_s,
                                               double* out_ms);

/* Open-PSA MEF XML -> flat graph: fault trees, components, model-data, house events, parameters and
 * expressions (mean value), CCF groups, event trees (one root per initiating event and sequence).
 * The returned model owns the arrays that canopy_model_desc() points to; roots are the fault-tree top events
 * (reference fault_tree.h:151-186) followed by the sequences. Stands in for reference Initializer
 * (initializer.h:65-70); coverage is listed at the top of csrc/mef_model.cpp and in INTEGRATION.md. */
typedef struct canopy_model canopy_model;
typedef struct canopy_mef_opts {
    int32_t ccf_analysis;   /* Settings::ccf_analysis (settings.h:298-308): replace CCF members by their proxy gates */
    int32_t reserved0;
    double  mission_time;   /* Settings::mission_time (settings.h:190-197,330), hours; default 8760          */
    int32_t reserved[4];
} canopy_mef_opts;
CANOPY_API void          canopy_mef_opts_default(canopy_mef_opts* opts);
CANOPY_API canopy_status canopy_model_load_mef(const char* const* xml_paths, int32_t n_paths, canopy_model** out);
CANOPY_API canopy_status canopy_model_load_mef_string(const char* xml_text, canopy_model** out);
CANOPY_API canopy_status canopy_model_load_mef_ex(const char* const* xml_paths, int32_t n_paths,
                                                  const canopy_mef_opts* opts, canopy_model** out);
CANOPY_API canopy_status canopy_model_load_mef_string_ex(const char* xml_text, const canopy_mef_opts* opts,
                                                         canopy_model** out);
CANOPY_API canopy_status canopy_model_desc(const canopy_model* model, canopy_graph_desc* out);
CANOPY_API const char*   canopy_model_basic_name(const canopy_model* model, uint32_t i);
CANOPY_API const char*   canopy_model_gate_name(const canopy_model* model, uint32_t i);
CANOPY_API const char*   canopy_model_root_name(const canopy_model* model, uint32_t i);
CANOPY_API void          canopy_model_destroy(canopy_model* model);

/* ---- Parameter uncertainty ---------------------------------------------------------------------
 * A probability source yields BasicEvent::p() of every event either as the mean (Expression::value(), expression.h:74)
 * or as one joint draw of all random deviates (Expression::Sample()/Reset(), expression.h:98-117: each deviate is drawn
 * once per trial and shared by every expression that refers to it, e.g. one <parameter> used by many events).  The six
 * deviates of the reference (expr/random_deviate.h:25-264; DoSample has no body there) are drawn from counter-based
 * Philox streams: draw j of deviate d in trial t is a pure function of (seed, t, d, j) -- csrc/expr_tape.h states the
 * mapping and the algorithms.  From a loaded model the source carries every expression of the MEF file (parameters,
 * arithmetic, exponential/GLM/Weibull/periodic-test, CCF factor models, collected event-tree expressions); from a flat
 * list it is one independent distribution per basic event. */
typedef enum canopy_deviate_kind {
    CANOPY_DEV_CONSTANT = 0,     /* a = value                                                                    */
    CANOPY_DEV_UNIFORM = 1,      /* a = min, b = max                         (UniformDeviate, random_deviate.h:46) */
    CANOPY_DEV_NORMAL = 2,       /* a = mean, b = sigma                      (NormalDeviate, :70)                  */
    CANOPY_DEV_LOGNORMAL = 3,    /* a = mu, b = sigma of the underlying normal (LognormalDeviate::Normal, :150)    */
    CANOPY_DEV_LOGNORMAL_EF = 4, /* a = mean, b = error factor, c = confidence level (LognormalDeviate::Logarithmic, :133) */
    CANOPY_DEV_GAMMA = 5,        /* a = k (shape), b = theta (scale)         (GammaDeviate, :171)                  */
    CANOPY_DEV_BETA = 6,         /* a = alpha, b = beta                      (BetaDeviate, :195)                   */
    CANOPY_DEV_HISTOGRAM = 7     /* bounds[n_bins + 1] strictly increasing, weights[n_bins] >= 0 (Histogram, :222) */
} canopy_deviate_kind;
typedef struct canopy_deviate {
    int32_t  kind;               /* canopy_deviate_kind */
    uint32_t n_bins;             /* histogram only */
    double   a, b, c;
    const double* bounds;        /* histogram only; read during the call */
    const double* weights;
} canopy_deviate;
typedef struct canopy_psource canopy_psource;
CANOPY_API canopy_status canopy_psource_from_model(const canopy_model* model, canopy_psource** out);
CANOPY_API canopy_status canopy_psource_from_deviates(const canopy_deviate* dist, uint32_t n_basic, canopy_psource** out);
CANOPY_API canopy_status canopy_psource_info(const canopy_psource* src, uint32_t* n_basic, uint32_t* n_deviates);
CANOPY_API canopy_status canopy_psource_mean(const canopy_psource* src, double* p_out /* [n_basic] */);
/* The trial-th joint draw, each probability clamped to [0, 1] (a deviate may leave it). */
CANOPY_API canopy_status canopy_psource_sample(const canopy_psource* src, uint64_t seed, uint64_t trial,
                                               double* p_out /* [n_basic] */);
CANOPY_API void          canopy_psource_destroy(canopy_psource* src);

/* Outer loop (Settings::uncertainty_analysis, settings.h:281-295): for trial t in [first_trial, first_trial + n_trials)
 * draw every p, re-quantise the thresholds of `plan` on its device (canopy_plan_update_p: same lowering, new thresholds
 * and count tables, no new plan), run mc->n_samples samples [mc->sample_begin + (t - first_trial) n_samples, ...) so that
 * the aleatory streams of different trials are disjoint, and record p_hat_t = tally / n_samples of every output.  A drawn
 * probability is clamped into the sampled range (2^-threshold_bits, 1 - 2^-53) when the event is sampled at the mean, and
 * ignored when the event is a constant at the mean (the lowering folded it).  The plan is left at the mean probabilities.
 * Per output: mean and standard deviation over the trials, a 95 % normal interval of the mean, the error factor
 * q95 / q50, num_quantiles quantiles (levels i / num_quantiles, i = 1 .. num_quantiles, as order statistics;
 * Settings::num_quantiles, settings.h:154-163) and a histogram of num_bins equal bins over [min, max]
 * (Settings::num_bins, settings.h:166-175). */
typedef struct canopy_uncertainty_opts {
    uint64_t n_trials;
    uint64_t first_trial;
    uint64_t seed;            /* of the deviates; the sample-level seed is mc->seed */
    uint32_t num_quantiles;   /* 0 = 20, the reference default (settings.h:328) */
    uint32_t num_bins;        /* 0 = 20 (settings.h:329) */
    int32_t  reserved[4];
} canopy_uncertainty_opts;
typedef struct canopy_uncertainty_stats { double mean, sigma, ci_lo, ci_hi, error_factor, min, max; } canopy_uncertainty_stats;
CANOPY_API canopy_status canopy_uncertainty_run(canopy_plan* plan, const canopy_psource* src, const canopy_mc_opts* mc,
                                                const canopy_uncertainty_opts* opts,
                                                canopy_uncertainty_stats* stats /* [n_tallies] */,
                                                double* quantiles    /* [n_tallies][num_quantiles] or NULL */,
                                                double* bin_edges    /* [n_tallies][num_bins + 1] or NULL */,
                                                uint64_t* bin_counts /* [n_tallies][num_bins] or NULL */,
                                                double* p_hat        /* [n_trials][n_tallies] or NULL */);

CANOPY_API const char* canopy_last_error(void);   /* thread-local; valid until the next call on this thread */
CANOPY_API const char* canopy_status_string(canopy_status s);
CANOPY_API uint32_t    canopy_version(void);      /* major<<16 | minor */
CANOPY_API uint32_t    canopy_spec_version(void);
/* Spec 4, host only, parity hooks for the tests: the group exponent k of an event with 64-bit threshold thr64 (0 = dense
 * event; a 32-bit threshold is thr32 << 32), and the cumulative count table (CANOPY_SPARSE_CDF_LEN words, 0.64 fixed point)
 * the engine builds for a sparse event (k = 7 .. CANOPY_SPARSE_MAX_LOG2, thr64 < 2^(64 - CANOPY_SPARSE_MIN_ZEROS)). */
CANOPY_API canopy_status canopy_sparse_class(uint64_t thr64, uint32_t* out_k);
CANOPY_API canopy_status canopy_sparse_cdf(uint64_t thr64, uint32_t k, uint64_t* out_cdf);

#ifdef __cplusplus
} /* extern "C" */
#endif
#endif /* CANOPY_H */
