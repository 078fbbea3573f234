// <psi|H|psi> on the tile engine, pipelined:  k_tile_expect_pipe.
//
// k_tile_expect (kernels_tile.cuh) fills a tile and walks its records in turn: ld.global x16 -> barrier -> STS x16 ->
// barrier -> walk.  The ncu source page of that kernel puts 46 % of the warp-stall samples into the per-tile code (the
// first STS waiting for the loads, the two barriers, two integer divisions and the address arithmetic of the 16 loads).
// Here the fill is taken off the critical path:
//
//   * one CTA per SM, 3 tile buffers of 64 KB in shared memory used as a ring;
//   * tiles stream from HBM with cp.async (LDGSTS, 16 B per thread per row: every thread copies exactly the 16 amplitudes
//     it owns in view 0, so the tile lands in the sheared (thread, register) layout the records are planned for);
//     completion is an mbarrier per (stage, walking group): cp.async.mbarrier.arrive.noinc by every copying thread;
//   * TWO groups of 256 threads each own every other tile of the CTA: wait on the barrier, read their 16 amplitudes of
//     the current VIEW from the landed tile into registers, walk the view's records, switch to the next view (tile_plan.hpp:
//     another register subspace of the same tile, in which four more xmask groups pair amplitudes of ONE thread), and
//     finally refill the stage with the CTA's tile i + 3.  While one group walks tile i and the other tile i+1, tile i+2
//     is in flight.
//
// A bulk-copy (cp.async.bulk / UBLKCP) fill was built and measured in round 2 and removed again: UBLKCP is a
// uniform-datapath instruction, one copy per warp-instruction with its operands in uniform registers, so a warp whose
// lanes hold 32 chunk addresses is serialised into an ELECT / R2UR / UBLKCP / BRA.U.ANY loop (~9 issue slots per 128-byte
// chunk, 512 chunks per tile; > 300 s per step against 1.06 s).  The cosets of a GF(2) subspace are XOR-scattered 128-byte
// chunks: no larger contiguous run for a bulk copy, no affine tensor map.  Evidence: profiles/r2_sass_ring_and_pipe_fill.txt,
// profiles/r2_expect_kernel_variants_28q_1e4terms.txt.
//
// Work items (batch, tile) are visited without divisions: every role advances its own (batch, tile) pair, and the coset
// representative of the next tile of the same batch is a masked add:  rep' = ((rep | P) + dstep) & ~P  with P the pivot
// mask of the batch's subspace and dstep the deposited tile stride.
#pragma once
#include "kernels_tile.cuh"

namespace b200vqe {

constexpr int kRingStages = 3;
constexpr int kRingGroups = 2;                                   // consumer groups of kBlock threads
constexpr int kRingThreads = kRingGroups * kBlock;               // 16 warps x 128 registers = the whole register file
constexpr int kRingTileBytes = (int)sizeof(double2) << kTileBits;

struct RingTables {                                              // per consumer group
    TileBatch sb;
    int4 recs[2 * (kStageRecs + kTileMaxViews + 1)];             // records + view markers + the end sentinel
    double2 terms[kStageTerms];
    uint16_t vbase[kTileMaxViews][kBlock];                       // byte offset of (thread, register 0) of every view of the batch
};

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    unsigned done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    } while (!done);
}
__device__ __forceinline__ void group_sync(int g) {               // named barrier of one consumer group
    if (g == 0) asm volatile("bar.sync 1, %0;" ::"n"(kBlock) : "memory");
    else asm volatile("bar.sync 2, %0;" ::"n"(kBlock) : "memory");
}
__device__ __forceinline__ double2 lds_f64x2(unsigned addr) {
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(addr));
    return v;
}

// (batch, tile) of a role, advanced without divisions
struct RingCursor {
    int b;
    unsigned long long t;            // tile of batch b, relative to tile_begin
    __device__ __forceinline__ void advance(unsigned long long step, unsigned long long tile_count) {
        t += step;
        while (t >= tile_count) { t -= tile_count; ++b; }
    }
};

// ---------------------------------------------------------------------------------------------
// k_tile_expect_pipe.  Every thread copies exactly the 16 amplitudes it owns in view 0 with cp.async (LDGSTS, 16 B each), so the
// tile lands in the SHEARED (thread, register) layout the records are planned for; completion is an mbarrier per (stage,
// walking group): every copying thread issues cp.async.mbarrier.arrive.noinc.  The group that walked item i is the one that
// refills its stage with item i + 3 (after a barrier of that group only), so no "empty" barrier is needed: group 0 streams
// the odd items, group 1 the even ones.  (Why not cp.async.bulk: see the header of this file.)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void cp_async_arrive_noinc(unsigned bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void cp_async16_u(unsigned dst, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}

__device__ __forceinline__ unsigned lds_u16(unsigned addr) {
    unsigned short v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint2 lds_u32x2(unsigned addr) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
    return v;
}
__device__ __forceinline__ int4 lds_i4(unsigned addr) {
    int4 v;
    asm volatile("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}

// One half of a record, lean form: every shared-memory operand is a 32-bit shared address (no generic-pointer
// arithmetic to rematerialise), the second half of the record descriptor is fetched inside the body, and the partner
// amplitudes arrive four at a time so that 8 x 4 load registers are never live at once.
template <int WHICH, int XR, int BV, bool PURE>
__device__ __forceinline__ void pipe_rec_half(const double2 (&a)[16], unsigned Pu, unsigned v0, unsigned rep, unsigned rec_u, int kb, int count,
                                              unsigned terms_u, double &acc) {
    constexpr int RB = 4;
    constexpr int HS = XR ? XrInfo<XR>::HS : RB - 1;
    constexpr int NP = 1 << (RB - 1);
    const int4 rb = lds_i4(rec_u + 16);                 // first term inline: packed z parts, coefficient
    double w[NP];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        double2 b[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int i = 4 * h + k;
            const int r = ((i >> HS) << (HS + 1)) | (BV << HS) | (i & ((1 << HS) - 1));
            b[k] = PURE ? a[r ^ XR] : lds_f64x2(Pu + ((unsigned)(r ^ XR) << 12));
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int i = 4 * h + k;
            const int r = ((i >> HS) << (HS + 1)) | (BV << HS) | (i & ((1 << HS) - 1));
            w[i] = WHICH == 0 ? fma(b[k].x, a[r].x, b[k].y * a[r].y) : fma(b[k].x, a[r].y, -(b[k].y * a[r].x));
        }
    }
    double *none = nullptr;
    tile_term<RB, HS, BV, false>(w, ((unsigned long long)(unsigned)rb.y << 32) | (unsigned)rb.x, __hiloint2double(rb.w, rb.z), v0, rep, acc, none);
    if (count > 1) {
        unsigned tu = terms_u + ((unsigned)kb << 4);
#pragma unroll 1
        for (int e = 1; e < count; ++e) {
            tu += 16;
            const double2 c = lds_f64x2(tu);
            tile_term<RB, HS, BV, false>(w, (unsigned long long)__double_as_longlong(c.x), c.y, v0, rep, acc, none);
        }
    }
}

template <int K>
__device__ __forceinline__ void pipe_one(const double2 (&a)[16], unsigned tile_u, unsigned v0, unsigned rep, const int4 ra, unsigned rec_u,
                                         unsigned terms_u, double &acc) {
    constexpr int WHICH = (K >> 4) & 1;
    constexpr int XR = ((K & 15) + 1) & 15;
    constexpr bool PURE = K >= kTileRecKeys;
    if constexpr (!(PURE && XR == 0 && WHICH == 1)) {
        const unsigned Pu = tile_u + (((v0 ^ (unsigned)ra.z) & 0xffu) << 4);      // row 0 of the partner thread
        pipe_rec_half<WHICH, XR, 0, PURE>(a, Pu, v0, rep, rec_u, ra.x, ra.y, terms_u, acc);
        if (XR == 0) pipe_rec_half<WHICH, XR, 1, PURE>(a, Pu, v0, rep, rec_u, ra.x, ra.y, terms_u, acc);
    }
}

// all records of the batch: one dispatch per RUN of equal keys (records are sorted by key)
// up to the next view marker or the end sentinel, which is returned (ru: its address)
__device__ __forceinline__ int4 pipe_walk(const double2 (&a)[16], unsigned tile_u, unsigned v0, unsigned rep, unsigned &ru, unsigned terms_u,
                                          double &acc) {
    int4 ra = lds_i4(ru);
#pragma unroll 1
    for (;;) {
        const unsigned key = (unsigned)ra.z >> 16;
        if (key >= 2u * kTileRecKeys) break;                          // view marker or end sentinel
#define B200VQE_CASE(K)                                                                        \
        case K:                                                                                \
            do {                                                                               \
                pipe_one<K>(a, tile_u, v0, rep, ra, ru, terms_u, acc);                         \
                ru += 32;                                                                      \
                ra = lds_i4(ru);                                                               \
            } while (((unsigned)ra.z >> 16) == (unsigned)(K));                                 \
            break;
#define B200VQE_CASE8(K) B200VQE_CASE(K) B200VQE_CASE(K + 1) B200VQE_CASE(K + 2) B200VQE_CASE(K + 3) B200VQE_CASE(K + 4) B200VQE_CASE(K + 5) B200VQE_CASE(K + 6) B200VQE_CASE(K + 7)
        switch (key) {
            B200VQE_CASE8(0) B200VQE_CASE8(8) B200VQE_CASE8(16) B200VQE_CASE8(24) B200VQE_CASE8(32) B200VQE_CASE8(40) B200VQE_CASE8(48) B200VQE_CASE8(56)
            default: break;
        }
#undef B200VQE_CASE8
#undef B200VQE_CASE
    }
    return ra;
}

__device__ __forceinline__ double lds_f64(unsigned addr) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}

// The fused block of a view (tile_plan.hpp, TileFusedSlot): the four generator groups of the view -- XR = 1, 2, 4, 8, one term
// each, the first NE of them with even nY -- in one straight-line block: four independent chains of 8 pair products, the
// signed butterfly and one POPC, no dispatch, no partner reads.  su: shared address of slot 0.
template <int NE>
__device__ __forceinline__ void fused_view(const double2 (&a)[16], unsigned su, unsigned v0, unsigned rep, double &acc) {
    double part[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int4 q0 = lds_i4(su + 64u * j);                // zl, zt, u
        const double2 s01 = lds_f64x2(su + 64u * j + 16u);
        const double s2 = lds_f64(su + 64u * j + 48u);
        double w[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int r = ((i >> j) << (j + 1)) | (i & ((1 << j) - 1));      // bit j clear
            const double2 b = a[r | (1 << j)];
            w[i] = j < NE ? fma(b.x, a[r].x, b.y * a[r].y) : fma(b.x, a[r].y, -(b.y * a[r].x));
        }
        const unsigned par = (unsigned)__popc((rep & (unsigned)q0.x) ^ (v0 & (unsigned)q0.y));
        const double us = __hiloint2double(q0.w ^ (int)(par << 31), q0.z);
        part[j] = us * signed_sum8(w, s01.x, s01.y, s2);
    }
    acc += (part[0] + part[1]) + (part[2] + part[3]);
}

struct PipeFillBatch {               // what a fill needs of its batch (shared memory, per group)
    uint32_t basis[kTileMaxDim];
    uint32_t seg[kTileMaxDim + 1];
    uint32_t lam[4];
    uint32_t piv;
    uint32_t pad[6];
};
constexpr int kPipeSmemBytes = kRingStages * kRingTileBytes + kRingGroups * (int)(sizeof(RingTables) + sizeof(PipeFillBatch)) + kRingGroups * kRingStages * 8 + 32 * 8;

// FILL = 0: one register of at most 32 qubits (sel_mask = 0): a tile row is addressed by a 32-bit ELEMENT index
// (high index << C | low bits) XORed with one of 16 per-batch row constants -- LOP3 + IMAD.WIDE + LDGSTS per row.
// FILL = 1: pair register of a sharded handle, at most 32 qubits together: the same element index; its top bit picks the slab
// (this rank's or the one received from the peer), the rest indexes it -- two SELs more per row.
// FILL = 2: anything larger: 64-bit chunk arithmetic per row.
template <int FILL>
__global__ void __launch_bounds__(kRingThreads, 1)
k_tile_expect_pipe(const double2 *__restrict__ psi, const double2 *__restrict__ psi_hi, unsigned sel_mask, int C, int M,
                   const TileBatch *__restrict__ batches, int n_batches,
                   unsigned long long tile_begin, unsigned long long tile_count, const TileRec *__restrict__ recs,
                   const TileTerm *__restrict__ terms, double *partials, unsigned *counter, double *out) {
    extern __shared__ __align__(128) unsigned char ring_raw[];
    unsigned char *tiles = ring_raw;                                                     // [kRingStages][64 KB]
    RingTables *tabs = reinterpret_cast<RingTables *>(ring_raw + kRingStages * kRingTileBytes);
    PipeFillBatch *fbs = reinterpret_cast<PipeFillBatch *>(tabs + kRingGroups);
    // full[stage][group that walks the item]: a stage is walked by the two groups in turn, and a group must see EVERY phase of a
    // barrier it waits on (a parity wait cannot tell phase k from phase k + 2: with one barrier per stage a group that ran two
    // items ahead of a slow fill passed the wait on the previous tile's phase), so each (stage, walker) pair has its own
    unsigned long long *bars = reinterpret_cast<unsigned long long *>(fbs + kRingGroups);
    double *s_red = reinterpret_cast<double *>(bars + kRingGroups * kRingStages);         // [16] warp sums
    const unsigned tiles_u = smem_u32(tiles);
    const unsigned full_u = smem_u32(bars);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kRingGroups * kRingStages; ++s) mbar_init(full_u + 8 * s, kBlock);      // one (noinc) arrive per copying thread
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const unsigned long long total = (unsigned long long)n_batches * tile_count;
    const unsigned long long n_items = blockIdx.x < total ? (total - blockIdx.x + gridDim.x - 1) / gridDim.x : 0ull;   // once per kernel
    const int g = warp / (kBlock / 32);
    const unsigned v0 = threadIdx.x & (kBlock - 1);
    const unsigned lo_b = (v0 & ((1u << C) - 1u)) << 4;
    const unsigned chunk_shift = (unsigned)C + 4u;
    RingTables &tb = tabs[g];
    PipeFillBatch &fb = fbs[g];
    const unsigned recs_u = smem_u32(tb.recs), terms_u = smem_u32(tb.terms);
    const unsigned vbase_u = smem_u32(tb.vbase), vcol_u = smem_u32(tb.sb.vcol);
    const unsigned long long step = (unsigned long long)kRingGroups * gridDim.x;
    // pair register of a sharded <psi|H|psi> (backend_ham.inc): chunks whose high index has bit sel_mask set live in the
    // slab received from the peer rank (psi_hi), the others in this rank's own slab; sel_mask = 0: one register
    const char *psi_b = reinterpret_cast<const char *>(psi);
    const char *psi_h = reinterpret_cast<const char *>(psi_hi);
    double acc = 0.0;

    // ---- fill side: this group streams items 1 - g, 3 - g, 5 - g, ... (stage = item % 3) ----
    RingCursor fcur{0, 0};
    fcur.advance(blockIdx.x + (unsigned long long)(1 - g) * gridDim.x, tile_count);
    unsigned long long fi = (unsigned long long)(1 - g);
    int fs = 1 - g;
    int f_b = -1;
    unsigned f_rep = 0, f_dstep = 0, f_tc = 0;
    auto fill_one = [&]() {                              // all 256 threads of the group
        if (fcur.b != f_b) {
            f_b = fcur.b;
            group_sync(g);                               // earlier fills have read the previous descriptor
            const TileBatch *B = batches + f_b;
            if (v0 < kTileMaxDim) fb.basis[v0] = B->basis[v0];
            else if (v0 < 2 * kTileMaxDim + 1) fb.seg[v0 - kTileMaxDim] = B->seg[v0 - kTileMaxDim];
            else if (v0 < 2 * kTileMaxDim + 5) fb.lam[v0 - 2 * kTileMaxDim - 1] = B->lam[v0 - 2 * kTileMaxDim - 1];
            else if (v0 == 2 * kTileMaxDim + 5) fb.piv = B->pad[0];
            group_sync(g);
            f_tc = 0;
            for (int k = 0; k < 8 - C; ++k) if ((v0 >> (C + k)) & 1u) f_tc ^= fb.basis[k];
            unsigned r0 = 0, d0 = 0;
            const unsigned t_abs = (unsigned)(tile_begin + fcur.t), st = (unsigned)step;
#pragma unroll
            for (int k = 0; k <= kTileMaxDim; ++k) { r0 |= (t_abs & fb.seg[k]) << k; d0 |= (st & fb.seg[k]) << k; }
            f_rep = r0; f_dstep = d0;
        }
        const unsigned rb0 = fb.basis[M - 4], rb1 = fb.basis[M - 3], rb2 = fb.basis[M - 2], rb3 = fb.basis[M - 1];
        const unsigned base = f_rep ^ f_tc;
        const unsigned dst = tiles_u + (unsigned)fs * kRingTileBytes + (v0 << 4);
        if constexpr (FILL == 2) {
            const unsigned lm0 = fb.lam[0] << 4, lm1 = fb.lam[1] << 4, lm2 = fb.lam[2] << 4, lm3 = fb.lam[3] << 4;
#pragma unroll
            for (int r = 0; r < 16; ++r) {
                const unsigned hi = base ^ ((r & 1) ? rb0 : 0u) ^ ((r & 2) ? rb1 : 0u) ^ ((r & 4) ? rb2 : 0u) ^ ((r & 8) ? rb3 : 0u);
                const unsigned lor = lo_b ^ ((r & 1) ? lm0 : 0u) ^ ((r & 2) ? lm1 : 0u) ^ ((r & 4) ? lm2 : 0u) ^ ((r & 8) ? lm3 : 0u);
                const char *slab = (hi & sel_mask) ? psi_h : psi_b;
                cp_async16_u(dst + ((unsigned)r << 12), slab + (((unsigned long long)(hi & ~sel_mask) << chunk_shift) + lor));
            }
        } else {
            // element index of (this thread, register r):  e0 ^ (XOR of the row generators of the set bits of r)
            const unsigned e0 = (base << C) | (lo_b >> 4);
            const unsigned g0 = (rb0 << C) | fb.lam[0], g1 = (rb1 << C) | fb.lam[1], g2 = (rb2 << C) | fb.lam[2], g3 = (rb3 << C) | fb.lam[3];
#pragma unroll
            for (int r = 0; r < 16; ++r) {
                const unsigned e = e0 ^ ((r & 1) ? g0 : 0u) ^ ((r & 2) ? g1 : 0u) ^ ((r & 4) ? g2 : 0u) ^ ((r & 8) ? g3 : 0u);
                if constexpr (FILL == 0) cp_async16_u(dst + ((unsigned)r << 12), psi + e);
                else {
                    // sel_mask is the slab bit of the HIGH index (j >> C): element bit sel_mask << C
                    const unsigned top = sel_mask << C;
                    cp_async16_u(dst + ((unsigned)r << 12), ((e & top) ? psi_hi : psi) + (e & (top - 1u)));
                }
            }
        }
        cp_async_arrive_noinc(full_u + 8 * (kRingGroups * fs + (1 - g)));   // the item is walked by the other group
        // next item of this group's fill sequence: same batch -> masked add of the deposited stride
        fcur.advance(step, tile_count);
        f_rep = ((f_rep | fb.piv) + f_dstep) & ~fb.piv;
        fi += kRingGroups;
        fs += kRingGroups;
        if (fs >= kRingStages) fs -= kRingStages;
    };
    // prologue: items 0, 1, 2
    if (fi < n_items) fill_one();
    if (g == 1 && fi < n_items) fill_one();

    // ---- walk side: this group owns items g, g + 2, ... ----
    RingCursor cur{0, 0};
    cur.advance(blockIdx.x + (unsigned long long)g * gridDim.x, tile_count);
    int cur_b = -1;
    unsigned piv = 0, dstep = 0, rep = 0;
    int s = g;                                              // stage of item i = i % kRingStages
    unsigned u = 0;                                         // uses of barrier (s, g) so far: every third item of this group
    int s_wraps = 0;
    for (unsigned long long i = g; i < n_items; i += kRingGroups) {
        if (cur.b != cur_b) {                               // uniform over the group
            cur_b = cur.b;
            group_sync(g);                                  // everyone is done with the previous batch's tables
            if (v0 < kTileBatchWords) reinterpret_cast<uint32_t *>(&tb.sb)[v0] = reinterpret_cast<const uint32_t *>(&batches[cur_b])[v0];
            group_sync(g);
            const int n_r = tb.sb.r_end - tb.sb.r_begin, n_t = tb.sb.t_end - tb.sb.t_begin;
            const int4 *rsrc = reinterpret_cast<const int4 *>(recs + tb.sb.r_begin);
            for (int k = (int)v0; k < 2 * n_r; k += kBlock) tb.recs[k] = rsrc[k];
            if (v0 == 0) tb.recs[2 * n_r] = make_int4(0, 0, kTileRecKeyEnd << 16, 0);
            const double2 *tsrc = reinterpret_cast<const double2 *>(terms) + tb.sb.t_begin;
            for (int k = (int)v0; k < n_t; k += kBlock) tb.terms[k] = tsrc[k];
            piv = tb.sb.pad[0];
            dstep = tile_rep(tb.sb, (unsigned)step);
            {   // where this thread's registers live in every view of the batch (read back by this thread only)
                const int nv = (int)tb.sb.n_views;
                for (int k = 0; k < nv; ++k) {
                    unsigned bs = 0;
#pragma unroll
                    for (int b = 0; b < kTileThreadBits; ++b) if ((v0 >> b) & 1u) bs ^= tb.sb.vcol[k][b];
                    tb.vbase[k][v0] = (uint16_t)bs;
                }
            }
            group_sync(g);
            rep = tile_rep(tb.sb, (unsigned)(tile_begin + cur.t));
        }
        mbar_wait(full_u + 8 * (kRingGroups * s + g), u & 1u);     // the tile has landed (all 256 copying threads' cp.async are complete)
        const unsigned tile_u = tiles_u + (unsigned)s * kRingTileBytes;
        {
            unsigned ru = recs_u, view = 0, pat = 0;
            {   // plans with fused view blocks open every view, view 0 included, with a marker
                const int4 m = lds_i4(ru);
                if (((unsigned)m.z >> 16) == (unsigned)kTileRecKeyView) { view = (unsigned)m.w; pat = (unsigned)m.y; ru += 32; }
            }
#pragma unroll 1
            for (;;) {
                // registers of this view: 16 amplitudes of the tile at  base ^ (XOR of the view's register columns)
                const unsigned base = lds_u16(vbase_u + ((view << 8) + v0) * 2u);
                const uint2 gc = lds_u32x2(vcol_u + view * 24u + 16u);      // four 16-bit register columns
                const unsigned g0 = gc.x & 0xffffu, g1 = gc.x >> 16, g2 = gc.y & 0xffffu, g3 = gc.y >> 16;
                double2 a[16];
#pragma unroll
                for (int r = 0; r < 16; ++r)
                    a[r] = lds_f64x2(tile_u + (base ^ ((r & 1) ? g0 : 0u) ^ ((r & 2) ? g1 : 0u) ^ ((r & 4) ? g2 : 0u) ^ ((r & 8) ? g3 : 0u)));
                if (pat) {                                  // the view's four generator groups, fused (pat - 1 of them with even nY)
                    switch (pat) {
                        case 1: fused_view<0>(a, ru, v0, rep, acc); break;
                        case 2: fused_view<1>(a, ru, v0, rep, acc); break;
                        case 3: fused_view<2>(a, ru, v0, rep, acc); break;
                        case 4: fused_view<3>(a, ru, v0, rep, acc); break;
                        default: fused_view<4>(a, ru, v0, rep, acc); break;
                    }
                    ru += 4u * (unsigned)sizeof(TileFusedSlot);
                }
                const int4 stop = pipe_walk(a, tile_u, v0, rep, ru, terms_u, acc);
                if (((unsigned)stop.z >> 16) != (unsigned)kTileRecKeyView) break;
                view = (unsigned)stop.w;                    // marker: the records that follow want the registers of this view
                pat = (unsigned)stop.y;
                ru += 32;
            }
        }
        // the stage is refilled by this same group: a barrier of the group is all the release it needs
        if (i + kRingStages < n_items) {
            group_sync(g);
            fill_one();                                     // item i + 3 -> stage s
        }
        cur.advance(step, tile_count);
        rep = ((rep | piv) + dstep) & ~piv;
        s += kRingGroups;
        if (s >= kRingStages) s -= kRingStages;
        if (++s_wraps == kRingStages) { s_wraps = 0; ++u; }      // barrier (s, g) comes round every kRingStages items of this group
    }
    const double wsum = warp_sum(acc);
    if (lane == 0) s_red[warp] = wsum;
    __syncthreads();
    // deterministic grid reduction: per-CTA sum in warp order, the last CTA adds the CTAs in order
    __shared__ bool s_last;
    if (threadIdx.x == 0) {
        double sum = 0.0;
#pragma unroll
        for (int w = 0; w < kRingGroups * (kBlock / 32); ++w) sum += s_red[w];
        partials[blockIdx.x] = sum;
        __threadfence();
        s_last = atomicAdd(counter, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last && threadIdx.x < 32) {
        __threadfence();
        double sum = 0.0;
        for (unsigned b = lane; b < gridDim.x; b += 32) sum += __ldcg(&partials[b]);
        sum = warp_sum(sum);
        if (lane == 0) { out[0] = sum; *counter = 0u; }
    }
}

}  // namespace b200vqe
