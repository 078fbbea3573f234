// Tile engine: <psi|H|psi>, per-term <P_k> and H psi with several xmask groups evaluated per HBM read of the state.
// See tile_plan.hpp for the decomposition.  One CTA = 256 threads owns one tile (a coset of the batch's subspace:
// 4096 amplitudes, 64 KB) in shared memory; each thread also keeps its own 16 amplitudes in registers.  The batch's
// groups arrive as RECORDS (TileRec: the even-nY or odd-nY terms of one group) sorted by key = (which, XR), XR being the
// register part of the group's virtual flip mask, and every key has its own fully specialised loop: the partner of
// register r is row r ^ XR of thread v0 ^ xt, an LDS.128 at an immediate offset.  <psi|H|psi> visits each pair
// {v, v ^ xv} once (the 8 registers with bit msb(XR) clear; all 16 when XR = 0) and accumulates
//     E += u s(j) Re w  |  u s(j) Im w,   w = conj(psi[j^x]) psi[j],
// where the sign of a term at register r factorises into a thread constant (one POPC over the coset part xor the
// thread part) and a Walsh function of r: one 8-point signed butterfly (7 DFMA) per term on top of the record's
// Re w / Im w products.
//
// Kernels: k_tile_expect<2, PT> (single buffer; PT = per-term sums instead of the energy), k_tile_expect_db<3>
// (2048-amplitude tiles, cp.async double buffering, opt-in), k_sum_term_partials; H psi: kernels_tile_apply_pipe.cuh.
// HBM traffic of <psi|H|psi>: 16 B per amplitude per BATCH (>= kTileBits - C groups), no writes.
#pragma once
#include "kernels.cuh"
#include "tile_plan.hpp"

namespace b200vqe {

__device__ __forceinline__ double flip_bit(double v, unsigned word, int r) {
    // flips the sign of v when bit r of word is set
    const unsigned m = (r == 31) ? (word & 0x80000000u) : ((word << (31 - r)) & 0x80000000u);
    return __hiloint2double(__double2hiint(v) ^ (int)m, __double2loint(v));
}

struct TileTermRaw { unsigned long long packed; double u; };

// descriptor tables are read through generic pointers: shared memory when the batch's tables were staged
// there (the usual case), global memory otherwise
__device__ __forceinline__ TileTermRaw load_term(const TileTerm *terms, uint32_t k) {
    const double2 raw = *(reinterpret_cast<const double2 *>(terms) + k);
    TileTermRaw t;
    t.packed = (unsigned long long)__double_as_longlong(raw.x);
    t.u = raw.y;
    return t;
}

// Signed sum over the 8 visited registers:  sum_i (-1)^{parity(i & zi)} w[i]  as a 3-stage butterfly whose
// stage signs are the +-1.0 doubles s0..s2 (one DFMA per node, 7 in total).
__device__ __forceinline__ double signed_sum8(const double (&w)[8], double s0, double s1, double s2) {
    const double y0 = fma(s0, w[1], w[0]), y1 = fma(s0, w[3], w[2]), y2 = fma(s0, w[5], w[4]), y3 = fma(s0, w[7], w[6]);
    const double x0 = fma(s1, y1, y0), x1 = fma(s1, y3, y2);
    return fma(s2, x1, x0);
}

__device__ __forceinline__ double pm_one_hi(unsigned hi_word, int bit) {   // bit of hi_word set ? -1.0 : 1.0
    const unsigned m = bit == 31 ? hi_word : (hi_word << (31 - bit));
    return __hiloint2double((int)((m & 0x80000000u) | 0x3ff00000u), 0);
}

// ---- class-ordered group walk over global-memory tables (used by k_sweep_small, kernels_small.cuh) ----
// One xmask group on the 8 registers r whose bit HS equals BV (the other element of each pair has the
// bit flipped when the group's xv has register bit HS; groups without a register bit call this for
// BV = 0 and BV = 1).  WHICH = 0: even-nY terms on Re w, 1: odd-nY terms on Im w.
template <int HS, int BV, int WHICH>
__device__ __forceinline__ void tile_terms(const double2 (&a)[16], const double2 *__restrict__ tile, unsigned pidx,
                                           unsigned v0, unsigned rep, uint32_t k, uint32_t count,
                                           const TileTerm *terms, double &acc) {
    TileTermRaw t = load_term(terms, k);
    double w[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int r = ((i >> HS) << (HS + 1)) | (BV << HS) | (i & ((1 << HS) - 1));
        const double2 b = tile[pidx ^ ((unsigned)r << 8)];
        w[i] = WHICH == 0 ? fma(b.x, a[r].x, b.y * a[r].y) : fma(b.x, a[r].y, -(b.y * a[r].x));
    }
    for (uint32_t e = 0; e < count; ++e) {
        const unsigned zl = (unsigned)t.packed;                   // z on the high index bits
        const unsigned hw = (unsigned)(t.packed >> 32);           // bits 0..7: thread bits of zv, 8..11: register bits
        const double u = t.u;
        if (e + 1 < count) t = load_term(terms, k + e + 1);      // prefetch the next term
        // coset part and thread part share one POPC: parity(a) ^ parity(b) = parity(a ^ b)
        unsigned par = (unsigned)__popc((rep & zl) ^ (v0 & hw & 0xffu));
        if (BV) par += hw >> (8 + HS);
        // free register bits: i-bit k <-> r-bit (k < HS ? k : k + 1)
        const double s0 = pm_one_hi(hw, 8 + (0 < HS ? 0 : 1));
        const double s1 = pm_one_hi(hw, 8 + (1 < HS ? 1 : 2));
        const double s2 = pm_one_hi(hw, 8 + (2 < HS ? 2 : 3));
        const double us = __hiloint2double(__double2hiint(u) ^ (int)(par << 31), __double2loint(u));
        acc = fma(us, signed_sum8(w, s0, s1, s2), acc);
    }
}

template <int HS, int BV>
__device__ __forceinline__ void tile_group(const double2 (&a)[16], const double2 *__restrict__ tile, unsigned v0,
                                           unsigned rep, const int4 g, const TileTerm *terms, double &acc) {
    // g = {kb, n_even, n_odd, xv | hsel << 16}
    const unsigned pidx = ((unsigned)g.w ^ v0) & 0xfffu;      // partner of register 0; register r: pidx ^ (r << 8)
    if (g.y) tile_terms<HS, BV, 0>(a, tile, pidx, v0, rep, (uint32_t)g.x, (uint32_t)g.y, terms, acc);
    if (g.z) tile_terms<HS, BV, 1>(a, tile, pidx, v0, rep, (uint32_t)(g.x + g.y), (uint32_t)g.z, terms, acc);
}

// all groups [g0, g1) of one pair-selector class
template <int HS, bool BOTH>
__device__ __forceinline__ void tile_class(const double2 (&a)[16], const double2 *__restrict__ tile, unsigned v0, unsigned rep,
                                           const int4 *gp, int g0, int g1, const TileTerm *tp, double &acc) {
    if (g0 >= g1) return;
    int4 g = gp[g0];
    for (int gi = g0; gi < g1; ++gi) {
        const int4 gc = g;
        if (gi + 1 < g1) g = gp[gi + 1];        // prefetch the next group record
        tile_group<HS, 0>(a, tile, v0, rep, gc, tp, acc);
        if (BOTH) tile_group<HS, 1>(a, tile, v0, rep, gc, tp, acc);
    }
}

// ---- expectation engine, per record ----------------------------------------------------------------
// A record is the even-nY (WHICH = 0, uses Re w) or odd-nY (WHICH = 1, uses Im w) terms of one xmask group,
// w = conj(psi[j^x]) psi[j].  The record's key fixes WHICH and XR = xv >> 8 at COMPILE time (records of a batch are
// sorted by key, one specialised loop per key), so the partner of register r is row r ^ XR of thread v0 ^ xt:
// LDS.128 at immediate offsets from one per-record address.  RB = register bits of the tile (4: 16 amplitudes per
// thread, 3: 8 amplitudes per thread); NP = 2^(RB-1) pairs per thread per record.
//   XR != 0: HS = top bit of XR; a thread visits its NP registers with bit HS clear (each pair once, weight 2 in u).
//   XR == 0: the partner sits in the same row; all registers are visited (two halves of NP), weight 1.
template <int XR> struct XrInfo { static constexpr int HS = XR >= 8 ? 3 : (XR >= 4 ? 2 : (XR >= 2 ? 1 : 0)); };

__device__ __forceinline__ double signed_sum(const double (&w)[8], double s0, double s1, double s2) { return signed_sum8(w, s0, s1, s2); }
__device__ __forceinline__ double signed_sum(const double (&w)[4], double s0, double s1, double) {
    return fma(s1, fma(s0, w[3], w[2]), fma(s0, w[1], w[0]));
}

// one term on the NP products w[]: sign = thread constant (one POPC) x Walsh function of the visited registers
// PT (per-term output): the term's signed pair sum, without its coefficient, is reduced over the warp and added to the
// warp's own accumulator of that term in shared memory (*slot; lane 0 only) -- what getExpectationValueZ() returns per kernel.
template <int RB, int HS, int BV, bool PT = false>
__device__ __forceinline__ void tile_term(const double (&w)[1 << (RB - 1)], unsigned long long packed, double u, unsigned v0,
                                          unsigned rep, double &acc, double *slot = nullptr, double weight = 1.0) {
    const unsigned zl = (unsigned)packed;                     // z on the high index bits
    const unsigned hw = (unsigned)(packed >> 32);             // bits 0..7: thread bits of zv, 8..8+RB-1: register bits
    // coset part and thread part share one POPC: parity(a) ^ parity(b) = parity(a ^ b)
    unsigned par = (unsigned)__popc((rep & zl) ^ (v0 & hw & 0xffu));
    if (BV) par += hw >> (8 + HS);
    // free register bits: i-bit k <-> r-bit (k < HS ? k : k + 1)
    const double s0 = pm_one_hi(hw, 8 + (0 < HS ? 0 : 1));
    const double s1 = pm_one_hi(hw, 8 + (1 < HS ? 1 : 2));
    const double s2 = RB == 4 ? pm_one_hi(hw, 8 + (2 < HS ? 2 : 3)) : 1.0;
    if (!PT) {
        const double us = __hiloint2double(__double2hiint(u) ^ (int)(par << 31), __double2loint(u));
        acc = fma(us, signed_sum(w, s0, s1, s2), acc);
    } else {
        const double ws = __hiloint2double(__double2hiint(weight) ^ (int)(par << 31), 0);      // +-1.0 or +-2.0
        double v = ws * signed_sum(w, s0, s1, s2);
#pragma unroll
        for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((v0 & 31u) == 0) *slot += v;
    }
}

// PURE: the record's (sheared) flip mask has no thread part -- the partner of register r is register r ^ XR of the SAME
// thread, no shared-memory read at all (tile_plan.hpp, TilePlanOpts::shear)
template <int RB, int WHICH, int XR, int BV, bool PT = false, bool PURE = false>
__device__ __forceinline__ void tile_rec_half(const double2 (&a)[1 << RB], const double2 *__restrict__ P, unsigned v0, unsigned rep,
                                              const int4 ra, const int4 rb, const double2 *__restrict__ s_terms, double &acc,
                                              double *part = nullptr) {
    constexpr int HS = XR ? XrInfo<XR>::HS : RB - 1;
    constexpr int NP = 1 << (RB - 1);
    double w[NP];
#pragma unroll
    for (int i = 0; i < NP; ++i) {
        const int r = ((i >> HS) << (HS + 1)) | (BV << HS) | (i & ((1 << HS) - 1));
        const double2 b = PURE ? a[r ^ XR] : P[(r ^ XR) << 8];
        w[i] = WHICH == 0 ? fma(b.x, a[r].x, b.y * a[r].y) : fma(b.x, a[r].y, -(b.y * a[r].x));
    }
    constexpr double WEIGHT = XR ? 2.0 : 1.0;      // each pair once | every amplitude
    tile_term<RB, HS, BV, PT>(w, ((unsigned long long)(unsigned)rb.y << 32) | (unsigned)rb.x, __hiloint2double(rb.w, rb.z), v0, rep, acc,
                              PT ? part + ra.x : nullptr, WEIGHT);
    const int count = ra.y;
    if (count > 1) {
        const double2 *tp = s_terms + ra.x;
        double2 t = tp[1];
#pragma unroll 1
        for (int e = 1; e < count; ++e) {
            const double2 c = t;
            if (e + 1 < count) t = tp[e + 1];
            tile_term<RB, HS, BV, PT>(w, (unsigned long long)__double_as_longlong(c.x), c.y, v0, rep, acc, PT ? part + ra.x + e : nullptr, WEIGHT);
        }
    }
}

// all records of the batch, key by key: s_recs holds (n_r + 1) records as int4 pairs, the last one a sentinel (key 255)
// NK = number of keys compiled in: 2 * kTileRecKeys (generic + pure-register records) or kTileRecKeys (generic only)
template <int RB, int K, bool PT = false, int NK = 2 * kTileRecKeys>
__device__ __forceinline__ void tile_runs(const double2 (&a)[1 << RB], const double2 *__restrict__ tile, unsigned v0, unsigned rep,
                                          const int4 *__restrict__ s_recs, const double2 *__restrict__ s_terms, int &gi, int4 &ra,
                                          int4 &rb, double &acc, double *part = nullptr) {
    if constexpr (K < NK) {
        constexpr int WHICH = (K >> 4) & 1;
        constexpr int XR = ((K & 15) + 1) & 15;
        constexpr bool PURE = K >= kTileRecKeys;
        // a pure record with XR = 0 is the diagonal group x = 0: even nY only
        if constexpr (XR < (1 << RB) && !(PURE && XR == 0 && WHICH == 1)) {
            while (((unsigned)ra.z >> 16) == (unsigned)K) {
                const int4 ca = ra, cb = rb;
                ++gi;
                ra = s_recs[2 * gi];                                   // next record (or the sentinel)
                rb = s_recs[2 * gi + 1];
                const double2 *P = tile + ((v0 ^ (unsigned)ca.z) & 0xffu);     // row 0 of the partner thread
                tile_rec_half<RB, WHICH, XR, 0, PT, PURE>(a, P, v0, rep, ca, cb, s_terms, acc, part);
                if (XR == 0) tile_rec_half<RB, WHICH, XR, 1, PT, PURE>(a, P, v0, rep, ca, cb, s_terms, acc, part);
            }
        }
        tile_runs<RB, K + 1, PT, NK>(a, tile, v0, rep, s_recs, s_terms, gi, ra, rb, acc, part);
    }
}

// the same walk with one jump-table dispatch per record instead of the chain of key tests (whose branch sites are spread over
// the whole function: 64 instruction-cache lines per tile)
template <int RB, int K, bool PT>
__device__ __forceinline__ void tile_one(const double2 (&a)[1 << RB], const double2 *__restrict__ tile, unsigned v0, unsigned rep,
                                         const int4 ca, const int4 cb, const double2 *__restrict__ s_terms, double &acc, double *part) {
    constexpr int WHICH = (K >> 4) & 1;
    constexpr int XR = ((K & 15) + 1) & 15;
    constexpr bool PURE = K >= kTileRecKeys;
    if constexpr (XR < (1 << RB) && !(PURE && XR == 0 && WHICH == 1)) {
        const double2 *P = tile + ((v0 ^ (unsigned)ca.z) & 0xffu);
        tile_rec_half<RB, WHICH, XR, 0, PT, PURE>(a, P, v0, rep, ca, cb, s_terms, acc, part);
        if (XR == 0) tile_rec_half<RB, WHICH, XR, 1, PT, PURE>(a, P, v0, rep, ca, cb, s_terms, acc, part);
    }
}

template <int RB, bool PT = false>
__device__ __forceinline__ void tile_switch(const double2 (&a)[1 << RB], const double2 *__restrict__ tile, unsigned v0, unsigned rep,
                                            const int4 *__restrict__ s_recs, const double2 *__restrict__ s_terms, int n_r, double &acc,
                                            double *part = nullptr) {
    int4 ra = s_recs[0], rb = s_recs[1];
#pragma unroll 1
    for (int gi = 0; gi < n_r; ++gi) {
        const int4 ca = ra, cb = rb;
        ra = s_recs[2 * gi + 2];                                       // next record (or the sentinel)
        rb = s_recs[2 * gi + 3];
#define B200VQE_CASE(K) case K: tile_one<RB, K, PT>(a, tile, v0, rep, ca, cb, s_terms, acc, part); break;
#define B200VQE_CASE8(K) B200VQE_CASE(K) B200VQE_CASE(K + 1) B200VQE_CASE(K + 2) B200VQE_CASE(K + 3) B200VQE_CASE(K + 4) B200VQE_CASE(K + 5) B200VQE_CASE(K + 6) B200VQE_CASE(K + 7)
        switch ((unsigned)ca.z >> 16) {
            B200VQE_CASE8(0) B200VQE_CASE8(8) B200VQE_CASE8(16) B200VQE_CASE8(24) B200VQE_CASE8(32) B200VQE_CASE8(40) B200VQE_CASE8(48) B200VQE_CASE8(56)
            default: break;
        }
#undef B200VQE_CASE8
#undef B200VQE_CASE
    }
}

__device__ __forceinline__ unsigned tile_rep(const TileBatch &sb, unsigned t) {
    unsigned rep = 0;
#pragma unroll
    for (int i = 0; i <= kTileMaxDim; ++i) rep |= (t & sb.seg[i]) << i;
    return rep;
}

__device__ __forceinline__ void prefetch_l2(const void *p) {
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}

// MINB = resident CTAs per SM the register allocation is held to (2: 128 registers, 3: 80 with spills)
// PT = per-term output (b200vqe_expectation(per_term), RDM): instead of the energy, every CTA leaves the signed pair
// sums of each term over ITS tiles in term_partials[blockIdx.x][plan term] (per-warp accumulators in shared memory,
// flushed at every batch switch; fixed order, no atomics); k_sum_term_partials adds the CTAs in order.
// Records are dispatched to their specialised bodies by one switch per record (tile_switch).  Round 2 measured the alternatives
// on 28 q / 10^4 terms (profiles/r2_expect_kernel_variants_28q_1e4terms.txt): a chain of key tests over all 64 keys +38 %
// (its 64 test sites are spread over 139 KB of code), three CTAs per SM (80 registers, spills) and the unsheared planner-1
// kernel slower as well; they are no longer built.
template <int MINB, bool PT = false>
__global__ void __launch_bounds__(kBlock, MINB)
k_tile_expect(const double2 *__restrict__ psi, int C, int M, unsigned chunk_bytes, const TileBatch *__restrict__ batches, int n_batches,
              unsigned long long tile_begin, unsigned long long tile_count, const TileRec *__restrict__ recs,
              const TileTerm *__restrict__ terms, double *partials, unsigned *counter, double *out,
              double *term_partials = nullptr, int n_plan_terms = 0, const double2 *__restrict__ psi_hi = nullptr, unsigned sel_mask = 0) {
    extern __shared__ __align__(16) unsigned char tile_raw[];
    double2 *tile = reinterpret_cast<double2 *>(tile_raw);
    __shared__ TileBatch sb;
    __shared__ __align__(16) int4 s_recs[2 * (kStageRecs + 1)];
    __shared__ __align__(16) double2 s_terms[kStageTerms];
    __shared__ double s_part[PT ? kBlock / 32 : 1][PT ? kStageTerms : 1];
    double *part = PT ? &s_part[threadIdx.x >> 5][0] : nullptr;
    auto flush_terms = [&]() {      // PT: this CTA's sums of the finished batch -> global, warps added in order
        __syncthreads();
        const int n_t = sb.t_end - sb.t_begin;
        for (int k = threadIdx.x; k < n_t; k += kBlock) {
            double sum = 0.0;
#pragma unroll
            for (int wp = 0; wp < kBlock / 32; ++wp) sum += s_part[PT ? wp : 0][PT ? k : 0];
            term_partials[(size_t)blockIdx.x * n_plan_terms + sb.t_begin + k] = sum;
        }
    };
    const unsigned v0 = threadIdx.x;
    const unsigned lo = v0 & ((1u << C) - 1u);
    double acc = 0.0;
    int cur_b = -1;
    unsigned Tc = 0, rb0 = 0, rb1 = 0, rb2 = 0, rb3 = 0;
    unsigned lm0 = 0, lm1 = 0, lm2 = 0, lm3 = 0;                     // shear of the batch, as byte offsets inside a chunk
    const char *psi_b = reinterpret_cast<const char *>(psi);
    const unsigned lo_b = lo << 4;
    const char *psi_lo = reinterpret_cast<const char *>(psi + lo);   // line of this thread's column in chunk hi: psi_lo + hi * chunk_bytes
    unsigned pcA = 0;                    // combo of the two rows this thread prefetches for the next tile
    unsigned rep_next = 0;
    bool have_next = false;
    const unsigned long long total = (unsigned long long)n_batches * tile_count;

    for (unsigned long long w = blockIdx.x; w < total; w += gridDim.x) {
        const int b = (int)(w / tile_count);
        const unsigned long long t = tile_begin + (w - (unsigned long long)b * tile_count);
        if (b != cur_b) {                       // uniform over the CTA
            if (PT && cur_b >= 0) flush_terms();
            __syncthreads();
            if (threadIdx.x < kTileBatchWords)
                reinterpret_cast<uint32_t *>(&sb)[threadIdx.x] = reinterpret_cast<const uint32_t *>(&batches[b])[threadIdx.x];
            __syncthreads();
            cur_b = b;
            Tc = 0;
            for (int i = 0; i < 8 - C; ++i) if ((v0 >> (C + i)) & 1u) Tc ^= sb.basis[i];
            rb0 = sb.basis[M - 4]; rb1 = sb.basis[M - 3]; rb2 = sb.basis[M - 2]; rb3 = sb.basis[M - 1];
            lm0 = sb.lam[0] << 4; lm1 = sb.lam[1] << 4; lm2 = sb.lam[2] << 4; lm3 = sb.lam[3] << 4;
            // the 8 lanes sharing a 128-byte line split its 16 rows between them: this thread takes rows 2(lane & 7), +1
            pcA = Tc ^ ((v0 & 1u) ? rb1 : 0u) ^ ((v0 & 2u) ? rb2 : 0u) ^ ((v0 & 4u) ? rb3 : 0u);
            have_next = false;
            // stage the batch's records (term indices made relative) + a sentinel, and its terms; the planner keeps every
            // batch within kStageRecs / kStageTerms
            const int n_r = sb.r_end - sb.r_begin;
            const int n_t = sb.t_end - sb.t_begin;
            const int4 *rsrc = reinterpret_cast<const int4 *>(recs + sb.r_begin);
            for (int i = threadIdx.x; i < 2 * n_r; i += kBlock) s_recs[i] = rsrc[i];
            if (threadIdx.x == 0) s_recs[2 * n_r] = make_int4(0, 0, 0xff << 16, 0);
            const double2 *tsrc = reinterpret_cast<const double2 *>(terms) + sb.t_begin;
            for (int i = threadIdx.x; i < n_t; i += kBlock) s_terms[i] = tsrc[i];
            if (PT) {
                for (int i = threadIdx.x; i < n_t; i += kBlock)
#pragma unroll
                    for (int wp = 0; wp < kBlock / 32; ++wp) s_part[PT ? wp : 0][PT ? i : 0] = 0.0;
            }
            __syncthreads();
        }
        const unsigned rep = have_next ? rep_next : tile_rep(sb, (unsigned)t);   // high index (j >> C) has at most 32 bits
        const unsigned base = rep ^ Tc;
        double2 a[16];
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            const unsigned hi = base ^ ((r & 1) ? rb0 : 0u) ^ ((r & 2) ? rb1 : 0u) ^ ((r & 4) ? rb2 : 0u) ^ ((r & 8) ? rb3 : 0u);
            // register r holds the amplitude with low index bits lo ^ Lambda(r): still one 128-byte line per 8 lanes
            const unsigned lor = lo_b ^ ((r & 1) ? lm0 : 0u) ^ ((r & 2) ? lm1 : 0u) ^ ((r & 4) ? lm2 : 0u) ^ ((r & 8) ? lm3 : 0u);
            // pair register of a sharded register (sel_mask != 0): chunks with that high-index bit set live in the peer's slab
            const char *slab = (hi & sel_mask) ? reinterpret_cast<const char *>(psi_hi) : psi_b;
            a[r] = __ldcs(reinterpret_cast<const double2 *>(slab + ((unsigned long long)(hi & ~sel_mask) * chunk_bytes + lor)));   // streamed once per batch
        }
        __syncthreads();                        // readers of the previous tile are done
#pragma unroll
        for (int r = 0; r < 16; ++r) tile[((unsigned)r << 8) | v0] = a[r];
        __syncthreads();
        {   // pull the CTA's next tile into L2 while this one is being worked on: 512 lines, two per thread
            const unsigned long long wn = w + gridDim.x;
            have_next = wn < total && wn / tile_count == (unsigned long long)b;
            if (have_next) {
                rep_next = tile_rep(sb, (unsigned)(tile_begin + (wn - (unsigned long long)b * tile_count)));
                const unsigned h0 = rep_next ^ pcA, h1 = h0 ^ rb0;
                const char *hi_lo = reinterpret_cast<const char *>(psi_hi + lo);
                prefetch_l2(((h0 & sel_mask) ? hi_lo : psi_lo) + (unsigned long long)(h0 & ~sel_mask) * chunk_bytes);
                prefetch_l2(((h1 & sel_mask) ? hi_lo : psi_lo) + (unsigned long long)(h1 & ~sel_mask) * chunk_bytes);
            }
        }
        tile_switch<4, PT>(a, tile, v0, rep, s_recs, s_terms, sb.r_end - sb.r_begin, acc, part);
    }
    if (PT) {
        if (cur_b >= 0) flush_terms();
        return;
    }
    double v[1] = {acc};
    grid_reduce<1>(v, partials, counter, out, 1.0);
}

// out[k] = sum over CTAs (in order) of term_partials[cta][k]
__global__ void __launch_bounds__(kBlock, 4)
k_sum_term_partials(const double *__restrict__ term_partials, int n_cta, int n_plan_terms, double *__restrict__ out) {
    const int k = blockIdx.x * kBlock + threadIdx.x;
    if (k >= n_plan_terms) return;
    double sum = 0.0;
    for (int c = 0; c < n_cta; ++c) sum += term_partials[(size_t)c * n_plan_terms + k];
    out[k] = sum;
}


// ---------------------------------------------------------------------------------------------
// <psi|H|psi>, double-buffered: 2048-amplitude tiles (RB = 3, 8 amplitudes per thread, 32 KB), two tile buffers per CTA,
// three CTAs per SM.  While the records of the current tile are walked, the CTA's next tile streams from HBM into the
// other buffer with cp.async (16 B per thread per row, no registers, no LSU store), so the load latency that the
// single-buffer kernel exposes at every tile is hidden and 24 warps per SM keep the shared-memory pipe busy.
// A thread reads its own 8 amplitudes back from the tile (8 LDS.128: what the register path spends on stores).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

template <int RB>
__global__ void __launch_bounds__(kBlock, 3)
k_tile_expect_db(const double2 *__restrict__ psi, int C, int M, unsigned chunk_bytes, const TileBatch *__restrict__ batches, int n_batches,
                 unsigned long long tile_begin, unsigned long long tile_count, const TileRec *__restrict__ recs,
                 const TileTerm *__restrict__ terms, double *partials, unsigned *counter, double *out) {
    constexpr int REGS = 1 << RB;
    constexpr unsigned TILE = 256u << RB;                 // amplitudes per tile buffer
    extern __shared__ __align__(16) unsigned char tile_raw[];
    double2 *buf = reinterpret_cast<double2 *>(tile_raw);  // [2][TILE]
    __shared__ TileBatch sb;
    __shared__ __align__(16) int4 s_recs[2 * (kStageRecs + 1)];
    __shared__ __align__(16) double2 s_terms[kStageTerms];
    const unsigned v0 = threadIdx.x;
    const unsigned lo = v0 & ((1u << C) - 1u);
    double acc = 0.0;
    int cur_b = -1;
    unsigned Tc = 0, rbv[RB], lmv[RB];
    const char *psi_b = reinterpret_cast<const char *>(psi);
    const unsigned lo_b = lo << 4;
    const unsigned long long total = (unsigned long long)n_batches * tile_count;
    unsigned p = 0;                      // buffer holding the current tile
    bool in_flight = false;              // the current tile was requested during the previous iteration
    unsigned rep = 0;

    auto request = [&](unsigned rep_t, unsigned which) {      // this thread's REGS rows of the tile with representative rep_t
        const unsigned base = rep_t ^ Tc;
        double2 *dst = buf + which * TILE + v0;
#pragma unroll
        for (int r = 0; r < REGS; ++r) {
            unsigned hi = base, lor = lo_b;
#pragma unroll
            for (int k = 0; k < RB; ++k) if (r & (1 << k)) { hi ^= rbv[k]; lor ^= lmv[k]; }
            cp_async16(dst + ((unsigned)r << 8), psi_b + ((unsigned long long)hi * chunk_bytes + lor));
        }
        cp_async_commit();
    };

    for (unsigned long long w = blockIdx.x; w < total; w += gridDim.x) {
        const int b = (int)(w / tile_count);
        if (b != cur_b) {                       // uniform over the CTA; no request is in flight across a batch switch
            __syncthreads();
            if (threadIdx.x < kTileBatchWords)
                reinterpret_cast<uint32_t *>(&sb)[threadIdx.x] = reinterpret_cast<const uint32_t *>(&batches[b])[threadIdx.x];
            __syncthreads();
            cur_b = b;
            Tc = 0;
            for (int i = 0; i < 8 - C; ++i) if ((v0 >> (C + i)) & 1u) Tc ^= sb.basis[i];
#pragma unroll
            for (int k = 0; k < RB; ++k) { rbv[k] = sb.basis[M - RB + k]; lmv[k] = sb.lam[k] << 4; }
            const int n_r = sb.r_end - sb.r_begin;
            const int n_t = sb.t_end - sb.t_begin;
            const int4 *rsrc = reinterpret_cast<const int4 *>(recs + sb.r_begin);
            for (int i = threadIdx.x; i < 2 * n_r; i += kBlock) s_recs[i] = rsrc[i];
            if (threadIdx.x == 0) s_recs[2 * n_r] = make_int4(0, 0, 0xff << 16, 0);
            const double2 *tsrc = reinterpret_cast<const double2 *>(terms) + sb.t_begin;
            for (int i = threadIdx.x; i < n_t; i += kBlock) s_terms[i] = tsrc[i];
            in_flight = false;
        }
        if (!in_flight) {
            rep = tile_rep(sb, (unsigned)(tile_begin + (w - (unsigned long long)b * tile_count)));
            request(rep, p);
        }
        cp_async_wait_all();
        __syncthreads();                        // the current tile is complete; everyone is done with the other buffer
        unsigned rep_next = 0;
        {
            const unsigned long long wn = w + gridDim.x;
            in_flight = wn < total && wn / tile_count == (unsigned long long)b;
            if (in_flight) {
                rep_next = tile_rep(sb, (unsigned)(tile_begin + (wn - (unsigned long long)b * tile_count)));
                request(rep_next, p ^ 1u);
            }
        }
        const double2 *tile = buf + p * TILE;
        double2 a[REGS];
#pragma unroll
        for (int r = 0; r < REGS; ++r) a[r] = tile[((unsigned)r << 8) | v0];
        int gi = 0;
        int4 ra = s_recs[0], rb = s_recs[1];
        tile_runs<RB, 0>(a, tile, v0, rep, s_recs, s_terms, gi, ra, rb, acc);
        rep = rep_next;
        p ^= 1u;
    }
    cp_async_wait_all();
    double v[1] = {acc};
    grid_reduce<1>(v, partials, counter, out, 1.0);
}

}  // namespace b200vqe
