/*
 * canopy_oracle.c -- CPU oracle for the bit-packed Monte Carlo hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * PARITY UNPINNED: the reference (openpra-org/canopy) contains no implementation of this path, no
 * tests and no golden vectors.  This oracle is therefore a restatement of
 *   (a) the connective semantics of the Open-PSA MEF as constrained by the reference's own
 *       validation rules (reference src/mef/openpsa/event/event.cpp:140-204, event/event.h:143-157,
 *       event/formula.h:9-15), and
 *   (b) this repository's written sampling specs (spec 1 "direct", spec 2 "sliced", spec 4 "classed sparse"),
 * pinned only by the published Philox4x32-10 known-answer vectors (Random123 kat_vectors) and by
 * exact 2^B enumeration on small trees.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * this file's shared object.  Nothing under canopy_b200/ links or calls it.
 *
 * Three independent evaluators live here on purpose:
 *   oracle_mc_scalar   one sample at a time, one `unsigned char` per node, recursion over the
 *                      ORIGINAL descriptor (no lowering, no bit tricks)          -> parity checker
 *   oracle_exact       sum over all 2^B basic-event states with the QUANTISED probabilities
 *   oracle_mc_packed   64 samples per word, pthreads over sample blocks          -> timed CPU baseline
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/canopy.h" /* descriptor struct + connective enum only */

/* ------------------------------------------------------------------------------------------------
 * Philox4x32-10 (Salmon et al., SC'11).  Constants as in Random123 / cuRAND
 * (curand_philox4x32_x.h:88-91).  SPEC v1 section "RNG".
 * ---------------------------------------------------------------------------------------------- */
#define PHILOX_M0 0xD2511F53u
#define PHILOX_M1 0xCD9E8D57u
#define PHILOX_W0 0x9E3779B9u
#define PHILOX_W1 0xBB67AE85u

void oracle_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
    uint32_t k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)PHILOX_M0 * c0;
        uint64_t p1 = (uint64_t)PHILOX_M1 * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += PHILOX_W0; k1 += PHILOX_W1;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* Eight Philox calls that share counter words 0-2 and differ in word 3, lane-wise (structure of arrays) so that the
 * compiler vectorises the rounds (AVX2 where the host has it: the function is compiled twice and dispatched at load time).
 * Only the timed CPU port uses it; the values are those of oracle_philox4x32_10, which the tests compare it with. */
#if defined(__x86_64__) && defined(__GNUC__)
__attribute__((target_clones("avx2", "default")))
#endif
static void philox_batch8(uint32_t a0, uint32_t a1, uint32_t a2, const uint32_t c3[8], const uint32_t key[2], uint32_t out[4][8]) {
    uint32_t c0[8], c1[8], c2[8], cc3[8];
    for (int i = 0; i < 8; ++i) { c0[i] = a0; c1[i] = a1; c2[i] = a2; cc3[i] = c3[i]; }
    uint32_t k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        for (int i = 0; i < 8; ++i) {
            const uint64_t p0 = (uint64_t)PHILOX_M0 * c0[i], p1 = (uint64_t)PHILOX_M1 * c2[i];
            const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1[i] ^ k0, n2 = (uint32_t)(p0 >> 32) ^ cc3[i] ^ k1;
            c1[i] = (uint32_t)p1; cc3[i] = (uint32_t)p0; c0[i] = n0; c2[i] = n2;
        }
        k0 += PHILOX_W0; k1 += PHILOX_W1;
    }
    for (int i = 0; i < 8; ++i) { out[0][i] = c0[i]; out[1][i] = c1[i]; out[2][i] = c2[i]; out[3][i] = cc3[i]; }
}
void oracle_philox_batch8(uint32_t a0, uint32_t a1, uint32_t a2, const uint32_t c3[8], const uint32_t key[2], uint32_t out[32]) {
    philox_batch8(a0, a1, a2, c3, key, (uint32_t(*)[8])out);   /* test entry: out[w * 8 + lane] */
}

/* ------------------------------------------------------------------------------------------------
 * SPEC v1 "Threshold quantisation".  kind: 0 = always false (p <= 0), 1 = always true (p >= 1),
 * 2 = sampled with thr = floor(p * 2^bits).  Returns -1 for NaN / out of [0,1].
 * ---------------------------------------------------------------------------------------------- */
int oracle_threshold(double p, int bits, uint64_t* thr) {
    if (!(p >= 0.0) || !(p <= 1.0)) return -1;
    if (p <= 0.0) { *thr = 0; return 0; }
    if (p >= 1.0) { *thr = 0; return 1; }
    if (bits == 32) *thr = (uint64_t)(p * 4294967296.0);              /* exact: scaling by 2^32 */
    else            *thr = (uint64_t)(p * 18446744073709551616.0);    /* exact: scaling by 2^64, p < 1 */
    return 2;
}

/* Quantised probability actually realised by the sampler for one event. */
double oracle_quantised_p(double p, int bits) {
    uint64_t thr; int k = oracle_threshold(p, bits, &thr);
    if (k <= 0) return 0.0;
    if (k == 1) return 1.0;
    return bits == 32 ? (double)thr / 4294967296.0 : (double)thr / 18446744073709551616.0;
}

/* ------------------------------------------------------------------------------------------------
 * spec 4 "classed sparse" (DESIGN.md section 1).  Everything is a function of the 64-bit threshold t (32-bit
 * thresholds are widened: t = thr32 << 32).  z = leading zero bits of t.  An event with z < Z (= CANOPY_SPARSE_MIN_ZEROS)
 * samples as spec 2.  A rarer event works on groups of 2^k samples, k = min(KMAX, z + 2): ONE Philox call per group,
 * ctr = (g_lo, g_hi, e, 0x400) with g = s >> k, gives a 64-bit variate U (words 1:0) and position fields (words 2, 3);
 * N = #{j : U >= cdf_j} ones are placed at the first N distinct k-bit fields.
 * Count table, 0.64 fixed point, all integer:  q = 2^64 - t;  r = floor(t 2^64 / q);  a_0 = q^(2^k) by k truncating
 * squarings;  a_{j+1} = floor(floor(a_j r / 2^51) (2^k - j) / (j + 1)) >> 13;  cdf_j = min(2^64 - 1, a_0 + ... + a_j).
 * ---------------------------------------------------------------------------------------------- */
unsigned oracle_sparse_class_k(uint64_t t) {
    if (t == 0) return 0;
    unsigned z = 0;
    while (!((t << z) >> 63)) ++z;
    if (z < CANOPY_SPARSE_MIN_ZEROS) return 0;
    return z + 2 < CANOPY_SPARSE_MAX_LOG2 ? z + 2 : CANOPY_SPARSE_MAX_LOG2;
}

void oracle_sparse_cdf(uint64_t t, unsigned k, uint64_t* cdf) {
    const uint64_t q = (uint64_t)0 - t;
    const uint64_t r = (uint64_t)(((unsigned __int128)t << 64) / q);
    uint64_t a = q;
    for (unsigned sq = 0; sq < k; ++sq) a = (uint64_t)(((unsigned __int128)a * a) >> 64);
    unsigned __int128 acc = 0;
    for (unsigned j = 0; j < CANOPY_SPARSE_CDF_LEN; ++j) {
        acc += a;
        cdf[j] = (acc >> 64) ? ~0ull : (uint64_t)acc;
        unsigned __int128 y = ((unsigned __int128)a * r) >> 51;
        y = y * ((1u << k) - j);
        y = y / (j + 1u);
        a = (uint64_t)(y >> 13);
    }
}

static uint64_t thr64_of(uint64_t thr, int bits) { return bits == 32 ? thr << 32 : thr; }

/* group exponent k of a sampled event under `spec`, 0 when the event takes the compare (plane) path */
static unsigned sparse_k(int spec, uint64_t thr, int bits) {
    /* thr = 0 (p below 2^-bits) never fires under any spec; it stays on the compare path, where u < 0 is never true */
    if (spec != 4 || thr == 0) return 0;
    return oracle_sparse_class_k(thr64_of(thr, bits));
}

/* Field i of group g of event e, straight from the text: fpw = floor(32 / k) fields per word; fields 0 .. 2 fpw - 1 are
 * the digits of words 2 and 3 of call 0x400; field 2 fpw + (4 c + w) fpw + d is digit d of word w of call 0x401 + c. */
static unsigned sparse_field(const uint32_t key[2], uint32_t e, uint64_t g, unsigned k, unsigned i) {
    uint32_t ctr[4] = {(uint32_t)g, (uint32_t)(g >> 32), e, 0x400u}, out[4];
    const unsigned fpw = 32u / k;
    unsigned word, digit = i % fpw;
    if (i < 2 * fpw) word = 2 + i / fpw;
    else { unsigned i2 = i - 2 * fpw; ctr[3] = 0x401u + i2 / (4 * fpw); word = (i2 / fpw) & 3u; }
    oracle_philox4x32_10(ctr, key, out);
    return (out[word] >> (k * digit)) & ((1u << k) - 1u);
}

/* Definition form: the 2^k sample values of one group as one flag per sample (flag has room for 2^KMAX). */
static void sparse_group_flags_tab(uint64_t seed, uint32_t e, uint64_t g, unsigned k, const uint64_t* cdf, unsigned char* flag) {
    uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    uint32_t ctr[4] = {(uint32_t)g, (uint32_t)(g >> 32), e, 0x400u}, out[4];
    oracle_philox4x32_10(ctr, key, out);
    uint64_t U = ((uint64_t)out[1] << 32) | out[0];
    unsigned n = 0;
    for (unsigned j = 0; j < CANOPY_SPARSE_CDF_LEN; ++j) n += U >= cdf[j];
    memset(flag, 0, (size_t)1 << k);
    for (unsigned i = 0, placed = 0; placed < n; ++i) {
        unsigned pos = sparse_field(key, e, g, k, i);
        if (!flag[pos]) { flag[pos] = 1; ++placed; }
    }
}

/* The variate u(e, s) of DESIGN.md section 1, straight from its definition, and the basic-event bit
 * x = u < thr.  spec 1 "direct": one Philox call per 4 (2) samples, output word s&3 (pair s&1).
 * spec 2 "sliced": the top H = CANOPY_SLICED_HEAD_BITS bits of u from H bit-sliced calls of the sample's
 * 128-group; a sample whose head bits equal the threshold's takes its low 32-H (64-H) bits from the tail
 * calls of the group, which are shared out in order among the group's undecided samples. */
int oracle_sample_bit(uint64_t seed, uint32_t e, uint64_t s, int kind, uint64_t thr, int bits, int spec) {
    if (kind == 0) return 0;
    if (kind == 1) return 1;
    uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    uint32_t ctr[4], out[4];
    const unsigned k = sparse_k(spec, thr, bits);
    if (k) {
        unsigned char flag[1u << CANOPY_SPARSE_MAX_LOG2];
        uint64_t cdf[CANOPY_SPARSE_CDF_LEN];
        oracle_sparse_cdf(thr64_of(thr, bits), k, cdf);
        sparse_group_flags_tab(seed, e, s >> k, k, cdf, flag);
        return flag[s & ((1u << k) - 1u)];
    }
    /* spec 4 is spec 2 for every other event */
    if (spec == 1) {
        if (bits == 32) {
            uint64_t q = s >> 2;
            ctr[0] = (uint32_t)q; ctr[1] = (uint32_t)(q >> 32); ctr[2] = e; ctr[3] = 0;
            oracle_philox4x32_10(ctr, key, out);
            return out[s & 3] < (uint32_t)thr;
        } else {
            uint64_t q = s >> 1;
            ctr[0] = (uint32_t)q; ctr[1] = (uint32_t)(q >> 32); ctr[2] = e; ctr[3] = 1;
            oracle_philox4x32_10(ctr, key, out);
            unsigned j = (unsigned)(s & 1) * 2;
            uint64_t u = ((uint64_t)out[j + 1] << 32) | out[j];
            return u < thr;
        }
    }
    /* spec 2: head bits of all 128 samples of the group (the tail word of a sample depends on how many
     * samples before it in the group are still undecided after the head planes) */
    const unsigned H = CANOPY_SLICED_HEAD_BITS;
    uint64_t g = s >> 7; unsigned pos = (unsigned)(s & 127);
    uint32_t und[4] = {~0u, ~0u, ~0u, ~0u}, res[4] = {0, 0, 0, 0};
    unsigned head_thr = (unsigned)(thr >> (bits - H));
    for (unsigned k = 0; k < H; ++k) {
        ctr[0] = (uint32_t)g; ctr[1] = (uint32_t)(g >> 32); ctr[2] = e; ctr[3] = 0x100u + k;
        oracle_philox4x32_10(ctr, key, out);
        int tb = (head_thr >> (H - 1 - k)) & 1;
        for (int j = 0; j < 4; ++j) {
            if (tb) { res[j] |= und[j] & ~out[j]; und[j] &= out[j]; }
            else und[j] &= ~out[j];
        }
    }
    if (!((und[pos >> 5] >> (pos & 31)) & 1u)) return (res[pos >> 5] >> (pos & 31)) & 1u;
    unsigned rank = 0;  /* undecided samples of the group before this one */
    for (unsigned q = 0; q < pos; ++q) rank += (und[q >> 5] >> (q & 31)) & 1u;
    if (bits == 32) {
        ctr[0] = (uint32_t)g; ctr[1] = (uint32_t)(g >> 32); ctr[2] = e; ctr[3] = 0x200u + (rank >> 2);
        oracle_philox4x32_10(ctr, key, out);
        return (out[rank & 3] >> H) < ((uint32_t)thr & (0xFFFFFFFFu >> H));
    } else {
        ctr[0] = (uint32_t)g; ctr[1] = (uint32_t)(g >> 32); ctr[2] = e; ctr[3] = 0x300u + (rank >> 1);
        oracle_philox4x32_10(ctr, key, out);
        unsigned j = (rank & 1) * 2;
        uint64_t v = ((uint64_t)out[j + 1] << 32) | out[j];
        return (v >> H) < (thr & (~0ull >> H));
    }
}

static uint64_t packed_sample_word(uint64_t seed, uint32_t e, uint64_t w, int kind, uint64_t thr, int bits) {
    if (kind == 0) return 0;
    if (kind == 1) return ~0ull;
    uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    uint32_t ctr[4], out[4];
    uint64_t word = 0;
    if (bits == 32) {
        uint32_t t = (uint32_t)thr;
        uint64_t q0 = w * 16;
        for (unsigned c = 0; c < 16; ++c) {
            uint64_t q = q0 + c;
            ctr[0] = (uint32_t)q; ctr[1] = (uint32_t)(q >> 32); ctr[2] = e; ctr[3] = 0;
            oracle_philox4x32_10(ctr, key, out);
            uint64_t nib = (uint64_t)(out[0] < t) | ((uint64_t)(out[1] < t) << 1) | ((uint64_t)(out[2] < t) << 2) | ((uint64_t)(out[3] < t) << 3);
            word |= nib << (4 * c);
        }
    } else {
        uint64_t q0 = w * 32;
        for (unsigned c = 0; c < 32; ++c) {
            uint64_t q = q0 + c;
            ctr[0] = (uint32_t)q; ctr[1] = (uint32_t)(q >> 32); ctr[2] = e; ctr[3] = 1;
            oracle_philox4x32_10(ctr, key, out);
            uint64_t u0 = ((uint64_t)out[1] << 32) | out[0], u1 = ((uint64_t)out[3] << 32) | out[2];
            word |= ((uint64_t)(u0 < thr) | ((uint64_t)(u1 < thr) << 1)) << (2 * c);
        }
    }
    return word;
}

/* spec 2: the four 32-sample result words of 128-sample group g, evaluated lazily (MSB-first on bit
 * planes, tail call only for samples still undecided after the H head planes).  Must agree bit for bit
 * with oracle_sample_bit(..., spec = 2), which builds every variate in full; tests/test_oracle.py checks it. */
static void sliced_group_words(uint64_t seed, uint32_t e, uint64_t g, int kind, uint64_t thr, int bits, uint64_t out2[2]) {
    if (kind == 0) { out2[0] = out2[1] = 0; return; }
    if (kind == 1) { out2[0] = out2[1] = ~0ull; return; }
    const unsigned H = CANOPY_SLICED_HEAD_BITS;
    uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    uint32_t ctr[4], x[4];
    uint32_t und[4] = {~0u, ~0u, ~0u, ~0u}, res[4] = {0, 0, 0, 0};
    unsigned head = (unsigned)(thr >> (bits - H));
    /* the H <= 7 head calls and tail call 0 in one batch of eight (they differ in counter word 3 only) */
    uint32_t c3[8], y[4][8];
    for (unsigned k = 0; k < 8; ++k) c3[k] = k < H ? 0x100u + k : (bits == 32 ? 0x200u : 0x300u);
    const int batched = H <= 7;
    if (batched) philox_batch8((uint32_t)g, (uint32_t)(g >> 32), e, c3, key, y);
    for (unsigned k = 0; k < H; ++k) {
        if (batched) { x[0] = y[0][k]; x[1] = y[1][k]; x[2] = y[2][k]; x[3] = y[3][k]; }
        else {
            ctr[0] = (uint32_t)g; ctr[1] = (uint32_t)(g >> 32); ctr[2] = e; ctr[3] = 0x100u + k;
            oracle_philox4x32_10(ctr, key, x);
        }
        int tb = (head >> (H - 1 - k)) & 1;
        for (int j = 0; j < 4; ++j) {
            if (tb) { res[j] |= und[j] & ~x[j]; und[j] &= x[j]; }
            else und[j] &= ~x[j];
        }
    }
    /* tail: the r-th undecided sample of the group (ascending sample index) takes output word r&3 of tail
     * call r>>2 (32-bit thresholds), or words 2(r&1)+1 : 2(r&1) of tail call r>>1 (64-bit thresholds) */
    unsigned rank = 0, have = ~0u;
    if (batched) { x[0] = y[0][7]; x[1] = y[1][7]; x[2] = y[2][7]; x[3] = y[3][7]; have = 0; }
    for (unsigned j = 0; j < 4; ++j) {
        uint32_t u = und[j];
        while (u) {
            const unsigned t = (unsigned)__builtin_ctz(u);
            u &= u - 1;
            unsigned call = bits == 32 ? rank >> 2 : rank >> 1;
            if (call != have) {
                ctr[0] = (uint32_t)g; ctr[1] = (uint32_t)(g >> 32); ctr[2] = e; ctr[3] = (bits == 32 ? 0x200u : 0x300u) + call;
                oracle_philox4x32_10(ctr, key, x);
                have = call;
            }
            int lt;
            if (bits == 32) lt = (x[rank & 3] >> H) < ((uint32_t)thr & (0xFFFFFFFFu >> H));
            else { unsigned w = (rank & 1) * 2; lt = ((((uint64_t)x[w + 1] << 32) | x[w]) >> H) < (thr & (~0ull >> H)); }
            if (lt) res[j] |= 1u << t;
            ++rank;
        }
    }
    out2[0] = ((uint64_t)res[1] << 32) | res[0];
    out2[1] = ((uint64_t)res[3] << 32) | res[2];
}

/* spec 4, sparse event: the ones of group g (2^k samples) in streaming form -- one call, fields consumed word by word,
 * further calls only when the fields run out.  Writes the N distinct positions to pos[] (order of appearance) and returns
 * N.  Must agree with sparse_group_flags_tab; tests/test_oracle.py checks it. */
static unsigned sparse_group_positions(uint64_t seed, uint32_t e, uint64_t g, unsigned k, const uint64_t* cdf, uint16_t* pos) {
    uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    uint32_t ctr[4] = {(uint32_t)g, (uint32_t)(g >> 32), e, 0x400u}, x[4];
    oracle_philox4x32_10(ctr, key, x);
    uint64_t U = ((uint64_t)x[1] << 32) | x[0];
    unsigned n = 0;
    while (n < CANOPY_SPARSE_CDF_LEN && U >= cdf[n]) ++n;   /* cdf is non-decreasing */
    const unsigned fpw = 32u / k, mask = (1u << k) - 1u;
    unsigned word = 2, digit = 0, placed = 0;
    while (placed < n) {
        if (digit == fpw) {
            digit = 0;
            if (++word == 4) { word = 0; ++ctr[3]; oracle_philox4x32_10(ctr, key, x); }
        }
        unsigned p = (x[word] >> (k * digit)) & mask;
        ++digit;
        unsigned dup = 0;
        for (unsigned i = 0; i < placed; ++i) dup |= pos[i] == p;
        if (!dup) pos[placed++] = (uint16_t)p;
    }
    return n;
}

/* The 128-sample window `g128` of a sparse event: bits of the enclosing 2^k-group that fall into it. */
static void sparse_window_words(uint64_t seed, uint32_t e, uint64_t g128, unsigned k, const uint64_t* cdf, uint64_t out2[2]) {
    uint16_t pos[CANOPY_SPARSE_CDF_LEN];
    const unsigned n = sparse_group_positions(seed, e, g128 >> (k - 7), k, cdf, pos);
    const unsigned lo = (unsigned)(g128 & ((1u << (k - 7)) - 1u)) * 128u;   /* first position of the window inside the group */
    out2[0] = out2[1] = 0;
    for (unsigned i = 0; i < n; ++i)
        if (pos[i] >= lo && pos[i] < lo + 128u) out2[(pos[i] - lo) >> 6] |= 1ull << ((pos[i] - lo) & 63u);
}

/* Words 2g and 2g+1 (samples [128g, 128g+128)) of one basic event under any spec. */
static void group_words_tab(int spec, uint64_t seed, uint32_t e, uint64_t g, int kind, uint64_t thr, int bits, const uint64_t* tab, uint64_t out2[2]) {
    unsigned k;
    if (spec == 1) {
        out2[0] = packed_sample_word(seed, e, 2 * g, kind, thr, bits);
        out2[1] = packed_sample_word(seed, e, 2 * g + 1, kind, thr, bits);
    } else if (kind == 2 && (k = sparse_k(spec, thr, bits)) != 0) {
        uint64_t cdf[CANOPY_SPARSE_CDF_LEN];
        if (!tab) { oracle_sparse_cdf(thr64_of(thr, bits), k, cdf); tab = cdf; }   /* callers that loop over groups pass the event's table */
        sparse_window_words(seed, e, g, k, tab, out2);
    } else sliced_group_words(seed, e, g, kind, thr, bits, out2);
}
static void group_words(int spec, uint64_t seed, uint32_t e, uint64_t g, int kind, uint64_t thr, int bits, uint64_t out2[2]) {
    group_words_tab(spec, seed, e, g, kind, thr, bits, NULL, out2);
}

/* Count tables of every sparse event of a descriptor (spec 4), [B][CANOPY_SPARSE_CDF_LEN]; NULL under the other specs. */
static uint64_t* sparse_tables(int spec, int bits, uint32_t B, const int* kind, const uint64_t* thr) {
    if (spec != 4) return NULL;
    uint64_t* tab = (uint64_t*)calloc((size_t)(B ? B : 1) * CANOPY_SPARSE_CDF_LEN, sizeof(uint64_t));
    for (uint32_t e = 0; e < B; ++e) {
        unsigned k = kind[e] == 2 ? sparse_k(spec, thr[e], bits) : 0;
        if (k) oracle_sparse_cdf(thr64_of(thr[e], bits), k, tab + (size_t)e * CANOPY_SPARSE_CDF_LEN);
    }
    return tab;
}

/* ------------------------------------------------------------------------------------------------
 * Connective semantics on plain ints (arity rules event.cpp:140-191).
 * x[] are the child values AFTER the per-edge complement (formula.h:12-15).
 * ---------------------------------------------------------------------------------------------- */
static int eval_connective(int op, const unsigned char* x, uint32_t n, unsigned kmin, unsigned kmax) {
    unsigned c = 0;
    for (uint32_t i = 0; i < n; ++i) c += x[i] ? 1u : 0u;
    switch (op) {
        case CANOPY_AND:         return c == n;
        case CANOPY_OR:          return c > 0;
        case CANOPY_ATLEAST:     return c >= kmin;
        case CANOPY_XOR:         return (x[0] != 0) != (x[1] != 0);
        case CANOPY_NOT:         return !x[0];
        case CANOPY_NAND:        return c != n;
        case CANOPY_NOR:         return c == 0;
        case CANOPY_NULL:        return x[0] != 0;
        case CANOPY_IFF:         return (x[0] != 0) == (x[1] != 0);
        case CANOPY_IMPLY:       return !x[0] || x[1];
        case CANOPY_CARDINALITY: return c >= kmin && c <= kmax;
        default:                 return 0;
    }
}

/* Topological order of gates by DFS over the original descriptor.  Returns 0, or -3 on a cycle,
 * -1 on a bad id.  order[] receives gate indices, children first. */
static int topo_order(const canopy_graph_desc* d, uint32_t* order) {
    uint32_t G = d->n_gates, B = d->n_basic;
    unsigned char* mark = (unsigned char*)calloc(G ? G : 1, 1); /* 0 new, 1 on stack, 2 done */
    uint32_t* stack = (uint32_t*)malloc(sizeof(uint32_t) * (G ? G : 1));
    uint64_t* pos = (uint64_t*)malloc(sizeof(uint64_t) * (G ? G : 1));
    uint32_t n_out = 0; int rc = 0;
    for (uint32_t g0 = 0; g0 < G && rc == 0; ++g0) {
        if (mark[g0]) continue;
        uint32_t sp = 0; stack[sp] = g0; pos[sp] = d->child_off[g0]; mark[g0] = 1; sp = 1;
        while (sp && rc == 0) {
            uint32_t g = stack[sp - 1];
            if (pos[sp - 1] == d->child_off[g + 1]) { mark[g] = 2; order[n_out++] = g; --sp; continue; }
            uint32_t id = d->child[pos[sp - 1]++] & CANOPY_CHILD_ID_MASK;
            if (id == CANOPY_CONST_FALSE || id == CANOPY_CONST_TRUE || id < B) continue;
            if (id >= B + G) { rc = -1; break; }
            uint32_t c = id - B;
            if (mark[c] == 1) { rc = -3; break; }
            if (mark[c] == 0) { mark[c] = 1; stack[sp] = c; pos[sp] = d->child_off[c]; ++sp; }
        }
    }
    free(mark); free(stack); free(pos);
    return rc;
}

static size_t n_tallies_of(const canopy_graph_desc* d, int tally_all) {
    return (size_t)d->n_roots + (tally_all ? d->n_gates : 0);
}

/* ------------------------------------------------------------------------------------------------
 * Scalar Monte Carlo: tallies[i] = #{s in [begin, begin+n) : output_i(s)}.
 * Outputs: roots in descriptor order, then (tally_all) every gate in descriptor order.
 * ---------------------------------------------------------------------------------------------- */
int oracle_mc_scalar(const canopy_graph_desc* d, uint64_t seed, uint64_t begin, uint64_t n,
                     int tally_all, int bits, int spec, uint64_t* tallies) {
    if (spec != 1 && spec != 2 && spec != 4) return -9;
    uint32_t B = d->n_basic, G = d->n_gates;
    uint32_t* order = (uint32_t*)malloc(sizeof(uint32_t) * (G ? G : 1));
    int rc = topo_order(d, order);
    if (rc) { free(order); return rc; }
    int* kind = (int*)malloc(sizeof(int) * (B ? B : 1));
    uint64_t* thr = (uint64_t*)malloc(sizeof(uint64_t) * (B ? B : 1));
    for (uint32_t e = 0; e < B; ++e) {
        kind[e] = oracle_threshold(d->p[e], bits, &thr[e]);
        if (kind[e] < 0) { free(order); free(kind); free(thr); return -8; }
    }
    /* which basic events are referenced at all (skip sampling the rest: they cannot matter) */
    unsigned char* used = (unsigned char*)calloc(B ? B : 1, 1);
    for (uint64_t i = 0; i < d->n_edges; ++i) { uint32_t id = d->child[i] & CANOPY_CHILD_ID_MASK; if (id < B) used[id] = 1; }
    for (uint32_t r = 0; r < d->n_roots; ++r) if (d->roots[r] < B) used[d->roots[r]] = 1;
    size_t nt = n_tallies_of(d, tally_all);
    memset(tallies, 0, sizeof(uint64_t) * nt);
    unsigned char* val = (unsigned char*)malloc((size_t)B + G + 1);
    uint32_t maxf = 1;
    for (uint32_t g = 0; g < G; ++g) { uint64_t f = d->child_off[g + 1] - d->child_off[g]; if (f > maxf) maxf = (uint32_t)f; }
    unsigned char* x = (unsigned char*)malloc(maxf);
    /* basic-event bits are produced 128 samples (one group) at a time; gates are still evaluated one
     * sample at a time on plain bytes */
    uint64_t* evw = (uint64_t*)malloc(sizeof(uint64_t) * 2 * (B ? B : 1));
    uint64_t cur_g = ~0ull;
    for (uint64_t s = begin; s < begin + n; ++s) {
        if ((s >> 7) != cur_g) {
            cur_g = s >> 7;
            for (uint32_t e = 0; e < B; ++e)
                if (used[e]) group_words(spec, seed, e, cur_g, kind[e], thr[e], bits, evw + 2 * (size_t)e);
        }
        for (uint32_t e = 0; e < B; ++e)
            val[e] = used[e] ? (unsigned char)((evw[2 * (size_t)e + ((s >> 6) & 1)] >> (s & 63)) & 1u) : 0;
        for (uint32_t i = 0; i < G; ++i) {
            uint32_t g = order[i];
            uint64_t o = d->child_off[g]; uint32_t f = (uint32_t)(d->child_off[g + 1] - o);
            for (uint32_t j = 0; j < f; ++j) {
                uint32_t c = d->child[o + j]; uint32_t id = c & CANOPY_CHILD_ID_MASK;
                unsigned char v = id == CANOPY_CONST_TRUE ? 1 : id == CANOPY_CONST_FALSE ? 0 : val[id];
                x[j] = (c & CANOPY_CHILD_COMPLEMENT) ? !v : v;
            }
            val[B + g] = (unsigned char)eval_connective(d->op[g], x, f, d->kmin ? d->kmin[g] : 0, d->kmax ? d->kmax[g] : 0);
        }
        for (uint32_t r = 0; r < d->n_roots; ++r) tallies[r] += val[d->roots[r]];
        if (tally_all) for (uint32_t g = 0; g < G; ++g) tallies[d->n_roots + g] += val[B + g];
    }
    free(evw);
    free(order); free(kind); free(thr); free(used); free(val); free(x);
    return 0;
}

/* Raw sampler words, same layout as canopy_mc_sample_words: words[i*n_words + w] for the events
 * listed in event_ids[0..n_events). */
int oracle_sample_words(const canopy_graph_desc* d, uint64_t seed, int bits, int spec, const uint32_t* event_ids,
                        uint32_t n_events, uint64_t first_word, uint64_t n_words, uint64_t* words) {
    if (spec != 1 && spec != 2 && spec != 4) return -9;
    for (uint32_t i = 0; i < n_events; ++i) {
        uint32_t e = event_ids[i]; uint64_t thr; int kind = oracle_threshold(d->p[e], bits, &thr);
        if (kind < 0) return -8;
        for (uint64_t w = 0; w < n_words; ++w) {
            uint64_t word = 0;
            for (unsigned b = 0; b < 64; ++b)
                word |= (uint64_t)oracle_sample_bit(seed, e, (first_word + w) * 64 + b, kind, thr, bits, spec) << b;
            words[(uint64_t)i * n_words + w] = word;
        }
    }
    return 0;
}

/* Same output as oracle_sample_words through the lazily evaluated group routine (fast). */
int oracle_sample_words_fast(const canopy_graph_desc* d, uint64_t seed, int bits, int spec, const uint32_t* event_ids,
                             uint32_t n_events, uint64_t first_word, uint64_t n_words, uint64_t* words) {
    if (spec != 1 && spec != 2 && spec != 4) return -9;
    for (uint32_t i = 0; i < n_events; ++i) {
        uint32_t e = event_ids[i]; uint64_t thr; int kind = oracle_threshold(d->p[e], bits, &thr);
        if (kind < 0) return -8;
        uint64_t pair[2], have = ~0ull, tab[CANOPY_SPARSE_CDF_LEN];
        unsigned k = kind == 2 ? sparse_k(spec, thr, bits) : 0;
        int sparse = k != 0;
        if (sparse) oracle_sparse_cdf(thr64_of(thr, bits), k, tab);
        for (uint64_t w = 0; w < n_words; ++w) {
            uint64_t gw = first_word + w;
            if ((gw >> 1) != have) { have = gw >> 1; group_words_tab(spec, seed, e, have, kind, thr, bits, sparse ? tab : NULL, pair); }
            words[(uint64_t)i * n_words + w] = pair[gw & 1];
        }
    }
    return 0;
}

/* Test hook (spec 4): the words of one sparse event of group exponent k under an explicit count table instead of the one
 * its threshold gives; stream = 1 uses the streaming form, 0 the definition form (one flag per sample, one Philox call per
 * field). */
int oracle_sparse_words_with_table(uint64_t seed, uint32_t e, unsigned k, const uint64_t* tab, int stream, uint64_t first_word,
                                   uint64_t n_words, uint64_t* words) {
    if (k < 7 || k > CANOPY_SPARSE_MAX_LOG2) return -9;
    uint64_t pair[2] = {0, 0}, have = ~0ull;
    unsigned char* flag = (unsigned char*)malloc((size_t)1 << CANOPY_SPARSE_MAX_LOG2);
    for (uint64_t w = 0; w < n_words; ++w) {
        uint64_t gw = first_word + w, g = gw >> 1;   /* 128-sample window */
        if (g != have) {
            have = g;
            if (stream) sparse_window_words(seed, e, g, k, tab, pair);
            else {
                sparse_group_flags_tab(seed, e, g >> (k - 7), k, tab, flag);
                const unsigned lo = (unsigned)(g & ((1u << (k - 7)) - 1u)) * 128u;
                pair[0] = pair[1] = 0;
                for (unsigned q = 0; q < 128; ++q) if (flag[lo + q]) pair[q >> 6] |= 1ull << (q & 63);
            }
        }
        words[w] = pair[gw & 1];
    }
    free(flag);
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * Exact probabilities by enumeration of all 2^B states (B <= 24), with the quantised p.
 * probs[] has the same layout as tallies.
 * ---------------------------------------------------------------------------------------------- */
int oracle_exact(const canopy_graph_desc* d, int tally_all, int bits, double* probs) {
    uint32_t B = d->n_basic, G = d->n_gates;
    if (B > 24) return -4;
    uint32_t* order = (uint32_t*)malloc(sizeof(uint32_t) * (G ? G : 1));
    int rc = topo_order(d, order);
    if (rc) { free(order); return rc; }
    double* q = (double*)malloc(sizeof(double) * (B ? B : 1));
    for (uint32_t e = 0; e < B; ++e) {
        if (!(d->p[e] >= 0.0) || !(d->p[e] <= 1.0)) { free(order); free(q); return -8; }
        q[e] = oracle_quantised_p(d->p[e], bits);
    }
    size_t nt = n_tallies_of(d, tally_all);
    long double* acc = (long double*)calloc(nt ? nt : 1, sizeof(long double));
    unsigned char* val = (unsigned char*)malloc((size_t)B + G + 1);
    uint32_t maxf = 1;
    for (uint32_t g = 0; g < G; ++g) { uint64_t f = d->child_off[g + 1] - d->child_off[g]; if (f > maxf) maxf = (uint32_t)f; }
    unsigned char* x = (unsigned char*)malloc(maxf);
    for (uint64_t st = 0; st < (1ull << B); ++st) {
        long double w = 1.0L;
        for (uint32_t e = 0; e < B; ++e) { val[e] = (st >> e) & 1; w *= val[e] ? (long double)q[e] : 1.0L - (long double)q[e]; }
        if (w == 0.0L) continue;
        for (uint32_t i = 0; i < G; ++i) {
            uint32_t g = order[i];
            uint64_t o = d->child_off[g]; uint32_t f = (uint32_t)(d->child_off[g + 1] - o);
            for (uint32_t j = 0; j < f; ++j) {
                uint32_t c = d->child[o + j]; uint32_t id = c & CANOPY_CHILD_ID_MASK;
                unsigned char v = id == CANOPY_CONST_TRUE ? 1 : id == CANOPY_CONST_FALSE ? 0 : val[id];
                x[j] = (c & CANOPY_CHILD_COMPLEMENT) ? !v : v;
            }
            val[B + g] = (unsigned char)eval_connective(d->op[g], x, f, d->kmin ? d->kmin[g] : 0, d->kmax ? d->kmax[g] : 0);
        }
        for (uint32_t r = 0; r < d->n_roots; ++r) if (val[d->roots[r]]) acc[r] += w;
        if (tally_all) for (uint32_t g = 0; g < G; ++g) if (val[B + g]) acc[d->n_roots + g] += w;
    }
    for (size_t i = 0; i < nt; ++i) probs[i] = (double)acc[i];
    free(order); free(q); free(acc); free(val); free(x);
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * Bit-packed multi-threaded CPU evaluator: the "port" timed as the CPU baseline.
 * Work unit = one 64-sample word column of the whole DAG; threads take blocks of word columns.
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    const canopy_graph_desc* d;
    const uint32_t* order;
    const int* kind; const uint64_t* thr; const unsigned char* used;
    const uint64_t* tab; /* spec 4 count tables, or NULL */
    const uint32_t* edge_idx; const uint64_t* edge_mask; /* per argument: slot in val[] (constants have their own two), complement mask */
    uint64_t seed, begin, n; int tally_all, bits, spec;
    uint64_t first_group, n_groups; /* global 128-sample group range covering [begin, begin+n) */
    int tid, nthreads;
    uint64_t* tallies; /* private */
} packed_job;

/* count-in-range on 64 lanes with a ripple bit-sliced counter */
static uint64_t packed_count_range(const uint64_t* x, uint32_t n, unsigned lo, unsigned hi) {
    uint64_t plane[17]; memset(plane, 0, sizeof plane);
    for (uint32_t i = 0; i < n; ++i) {
        uint64_t carry = x[i];
        for (int j = 0; j < 17 && carry; ++j) { uint64_t t = plane[j] & carry; plane[j] ^= carry; carry = t; }
    }
    /* ge_lo: count >= lo ; le_hi: count <= hi, both by MSB-first comparison against a constant */
    uint64_t gt = 0, eq = ~0ull;
    for (int j = 16; j >= 0; --j) { if ((lo >> j) & 1) eq &= plane[j]; else { gt |= eq & plane[j]; eq &= ~plane[j]; } }
    uint64_t ge_lo = gt | eq;
    uint64_t lt = 0; eq = ~0ull;
    for (int j = 16; j >= 0; --j) { if ((hi >> j) & 1) { lt |= eq & ~plane[j]; eq &= plane[j]; } else eq &= ~plane[j]; }
    uint64_t le_hi = lt | eq;
    return ge_lo & le_hi;
}

static uint64_t word_mask(uint64_t w, uint64_t begin, uint64_t end) {
    uint64_t s0 = w * 64, mask = ~0ull;
    if (s0 + 64 <= begin || s0 >= end) return 0;
    if (s0 < begin) mask &= ~0ull << (begin - s0);
    if (s0 + 64 > end) mask &= (1ull << (end - s0)) - 1;
    return mask;
}

/* Work unit = one 128-sample window (two 64-bit word columns) of the whole DAG; a thread takes whole blocks of 64
 * consecutive windows (8192 samples, the largest sparse group), so that the ones of a sparse event's 2^k-sample group are
 * drawn once (one Philox call) and then dealt out to the 2^(k-7) windows of the group as they come by. */
static void* packed_worker(void* arg) {
    packed_job* J = (packed_job*)arg;
    const canopy_graph_desc* d = J->d;
    uint32_t B = d->n_basic, G = d->n_gates;
    uint64_t* val = (uint64_t*)malloc(sizeof(uint64_t) * 2 * ((size_t)B + G + 2));
    val[2 * ((size_t)B + G)] = val[2 * ((size_t)B + G) + 1] = 0;              /* slot B+G: constant false */
    val[2 * ((size_t)B + G + 1)] = val[2 * ((size_t)B + G + 1) + 1] = ~0ull;  /* slot B+G+1: constant true */
    uint32_t maxf = 1;
    for (uint32_t g = 0; g < G; ++g) { uint64_t f = d->child_off[g + 1] - d->child_off[g]; if (f > maxf) maxf = (uint32_t)f; }
    uint64_t* x = (uint64_t*)malloc(sizeof(uint64_t) * 2 * maxf);
    /* per sparse event: the positions of the 2^k-group the thread is currently inside */
    unsigned char* ev_k = (unsigned char*)calloc(B ? B : 1, 1);
    unsigned char* cache_n = (unsigned char*)calloc(B ? B : 1, 1);
    uint64_t* cache_g = (uint64_t*)malloc(sizeof(uint64_t) * (B ? B : 1));
    uint16_t* cache_pos = (uint16_t*)malloc(sizeof(uint16_t) * CANOPY_SPARSE_CDF_LEN * (size_t)(B ? B : 1));
    for (uint32_t e = 0; e < B; ++e) {
        cache_g[e] = ~0ull;
        if (J->used[e] && J->kind[e] == 2) ev_k[e] = (unsigned char)sparse_k(J->spec, J->thr[e], J->bits);
    }
    uint64_t end = J->begin + J->n;
    const uint64_t first_block = J->first_group >> 6, last_block = (J->first_group + J->n_groups - 1) >> 6;
    for (uint64_t blk = first_block + (uint64_t)J->tid; blk <= last_block; blk += (uint64_t)J->nthreads)
    for (uint64_t grp = blk << 6; grp < (blk + 1) << 6; ++grp) {
        if (grp < J->first_group || grp >= J->first_group + J->n_groups) continue;
        uint64_t mask[2] = {word_mask(2 * grp, J->begin, end), word_mask(2 * grp + 1, J->begin, end)};
        for (uint32_t e = 0; e < B; ++e) {
            uint64_t* v = val + 2 * (size_t)e;
            if (!J->used[e]) { v[0] = v[1] = 0; continue; }
            const unsigned k = ev_k[e];
            if (!k) { group_words_tab(J->spec, J->seed, e, grp, J->kind[e], J->thr[e], J->bits, NULL, v); continue; }
            const uint64_t big = grp >> (k - 7);
            uint16_t* pos = cache_pos + (size_t)e * CANOPY_SPARSE_CDF_LEN;
            if (cache_g[e] != big) {
                cache_g[e] = big;
                cache_n[e] = (unsigned char)sparse_group_positions(J->seed, e, big, k, J->tab + (size_t)e * CANOPY_SPARSE_CDF_LEN, pos);
            }
            const unsigned lo = (unsigned)(grp & ((1u << (k - 7)) - 1u)) * 128u;
            v[0] = v[1] = 0;
            for (unsigned i = 0; i < cache_n[e]; ++i) {
                const unsigned q = (unsigned)pos[i] - lo;      /* wraps for positions below the window */
                if (q < 128u) v[q >> 6] |= 1ull << (q & 63u);
            }
        }
        /* both 64-sample halves of the window at once: every gate is two words wide */
        for (uint32_t i = 0; i < G; ++i) {
            const uint32_t g = J->order[i];
            const uint64_t o = d->child_off[g]; const uint32_t f = (uint32_t)(d->child_off[g + 1] - o);
            const uint32_t* idx = J->edge_idx + o; const uint64_t* cm = J->edge_mask + o;
            uint64_t r0 = 0, r1 = 0;
#define ARG0(j) (val[2 * (size_t)idx[j]] ^ cm[j])
#define ARG1(j) (val[2 * (size_t)idx[j] + 1] ^ cm[j])
            switch (d->op[g]) {
                case CANOPY_AND: case CANOPY_NAND:
                    r0 = r1 = ~0ull; for (uint32_t j = 0; j < f; ++j) { r0 &= ARG0(j); r1 &= ARG1(j); }
                    if (d->op[g] == CANOPY_NAND) { r0 = ~r0; r1 = ~r1; }
                    break;
                case CANOPY_OR: case CANOPY_NOR:
                    for (uint32_t j = 0; j < f; ++j) { r0 |= ARG0(j); r1 |= ARG1(j); }
                    if (d->op[g] == CANOPY_NOR) { r0 = ~r0; r1 = ~r1; }
                    break;
                case CANOPY_XOR:  r0 = ARG0(0) ^ ARG0(1); r1 = ARG1(0) ^ ARG1(1); break;
                case CANOPY_IFF:  r0 = ~(ARG0(0) ^ ARG0(1)); r1 = ~(ARG1(0) ^ ARG1(1)); break;
                case CANOPY_IMPLY: r0 = ~ARG0(0) | ARG0(1); r1 = ~ARG1(0) | ARG1(1); break;
                case CANOPY_NOT:  r0 = ~ARG0(0); r1 = ~ARG1(0); break;
                case CANOPY_NULL: r0 = ARG0(0); r1 = ARG1(0); break;
                case CANOPY_ATLEAST: case CANOPY_CARDINALITY: {
                    for (uint32_t j = 0; j < f; ++j) { x[j] = ARG0(j); x[maxf + j] = ARG1(j); }
                    const unsigned hi = d->op[g] == CANOPY_ATLEAST ? 0xFFFFu : d->kmax[g];
                    r0 = packed_count_range(x, f, d->kmin[g], hi); r1 = packed_count_range(x + maxf, f, d->kmin[g], hi);
                    break;
                }
                default: break;
            }
#undef ARG0
#undef ARG1
            val[2 * ((size_t)B + g)] = r0; val[2 * ((size_t)B + g) + 1] = r1;
        }
        for (int h = 0; h < 2; ++h) {
            if (!mask[h]) continue;
            for (uint32_t r = 0; r < d->n_roots; ++r)
                J->tallies[r] += (uint64_t)__builtin_popcountll(val[2 * (size_t)d->roots[r] + h] & mask[h]);
            if (J->tally_all)
                for (uint32_t g = 0; g < G; ++g)
                    J->tallies[d->n_roots + g] += (uint64_t)__builtin_popcountll(val[2 * ((size_t)B + g) + h] & mask[h]);
        }
    }
    free(val); free(x); free(ev_k); free(cache_n); free(cache_g); free(cache_pos);
    return NULL;
}

int oracle_mc_packed(const canopy_graph_desc* d, uint64_t seed, uint64_t begin, uint64_t n,
                     int tally_all, int bits, int spec, int nthreads, uint64_t* tallies) {
    uint32_t B = d->n_basic, G = d->n_gates;
    if (spec != 1 && spec != 2 && spec != 4) return -9;
    if (nthreads < 1) nthreads = 1;
    uint32_t* order = (uint32_t*)malloc(sizeof(uint32_t) * (G ? G : 1));
    int rc = topo_order(d, order);
    if (rc) { free(order); return rc; }
    int* kind = (int*)malloc(sizeof(int) * (B ? B : 1));
    uint64_t* thr = (uint64_t*)malloc(sizeof(uint64_t) * (B ? B : 1));
    for (uint32_t e = 0; e < B; ++e) {
        kind[e] = oracle_threshold(d->p[e], bits, &thr[e]);
        if (kind[e] < 0) { free(order); free(kind); free(thr); return -8; }
    }
    unsigned char* used = (unsigned char*)calloc(B ? B : 1, 1);
    for (uint64_t i = 0; i < d->n_edges; ++i) { uint32_t id = d->child[i] & CANOPY_CHILD_ID_MASK; if (id < B) used[id] = 1; }
    for (uint32_t r = 0; r < d->n_roots; ++r) if (d->roots[r] < B) used[d->roots[r]] = 1;
    size_t nt = n_tallies_of(d, tally_all);
    memset(tallies, 0, sizeof(uint64_t) * nt);
    if (n == 0) { free(order); free(kind); free(thr); free(used); return 0; }
    uint64_t first_group = begin / 128, last_group = (begin + n - 1) / 128;
    uint64_t* tab = sparse_tables(spec, bits, B, kind, thr);
    uint32_t* edge_idx = (uint32_t*)malloc(sizeof(uint32_t) * (d->n_edges ? d->n_edges : 1));
    uint64_t* edge_mask = (uint64_t*)malloc(sizeof(uint64_t) * (d->n_edges ? d->n_edges : 1));
    for (uint64_t i = 0; i < d->n_edges; ++i) {
        const uint32_t c = d->child[i], id = c & CANOPY_CHILD_ID_MASK;
        edge_idx[i] = id == CANOPY_CONST_TRUE ? B + G + 1 : id == CANOPY_CONST_FALSE ? B + G : id;
        edge_mask[i] = (c & CANOPY_CHILD_COMPLEMENT) ? ~0ull : 0ull;
    }
    packed_job* jobs = (packed_job*)calloc((size_t)nthreads, sizeof(packed_job));
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)nthreads);
    for (int t = 0; t < nthreads; ++t) {
        packed_job* J = &jobs[t];
        J->d = d; J->order = order; J->kind = kind; J->thr = thr; J->used = used; J->tab = tab;
        J->edge_idx = edge_idx; J->edge_mask = edge_mask;
        J->seed = seed; J->begin = begin; J->n = n; J->tally_all = tally_all; J->bits = bits; J->spec = spec;
        J->first_group = first_group; J->n_groups = last_group - first_group + 1;
        J->tid = t; J->nthreads = nthreads;
        J->tallies = (uint64_t*)calloc(nt ? nt : 1, sizeof(uint64_t));
        pthread_create(&th[t], NULL, packed_worker, J);
    }
    for (int t = 0; t < nthreads; ++t) {
        pthread_join(th[t], NULL);
        for (size_t i = 0; i < nt; ++i) tallies[i] += jobs[t].tallies[i];
        free(jobs[t].tallies);
    }
    free(jobs); free(th); free(order); free(kind); free(thr); free(used); free(tab); free(edge_idx); free(edge_mask);
    return 0;
}

int oracle_spec_version(void) { return CANOPY_SPEC_VERSION; }
int oracle_sparse_cdf_len(void) { return CANOPY_SPARSE_CDF_LEN; }
int oracle_sparse_min_zeros(void) { return CANOPY_SPARSE_MIN_ZEROS; }
int oracle_sparse_max_log2(void) { return CANOPY_SPARSE_MAX_LOG2; }
