// K2 on the tile engine:  k_tile_apply_pipe8.   dst (+)= in_scale * sum_{groups of ONE batch} f_g(j) src[j ^ x_g]
//
// One launch per batch (stream order makes batch b's read-modify-write of dst see batch b-1's).  A tile of src (a coset of the
// batch's subspace: 4096 amplitudes, 64 KB) sits in shared memory; a thread keeps complex accumulators for its own outputs,
// walks every record of the batch -- the partner of row r is row r ^ XR of thread v0 ^ xt, LDS.128 at immediate offsets -- and
// adds the tile's contribution to dst once.  HBM traffic per batch: 16 B (src) + 32 B (dst read + write) per amplitude instead
// of 16 B per GROUP.  Signs are evaluated at the thread's own index j (s(j ^ x) = s(j) (-1)^nY is folded into the coefficient
// on the host): u = coef for even nY, -Im(coef) for odd nY.
//   first != 0:  dst = in_scale * acc - prev_scale * dst   (prev_scale = 0 unless fuse_prev)   [Lanczos: w' = H v_j - beta_j v_{j-1}]
//   first == 0:  dst += in_scale * acc
//
// Pipeline (as k_tile_expect_pipe, kernels_tile_ring.cuh): one CTA per SM, a ring of three 64 KB tile buffers filled with
// cp.async (LDGSTS), one mbarrier per (stage, walking group), two groups that own every other tile of the CTA and refill the
// stage they have just walked with the CTA's tile i + 3.  A group is 512 threads: its two halves (h = 0, 1) split the 16 rows of
// every thread column -- thread (v0, h) owns rows 8h .. 8h + 7, 8 outputs, 8 complex accumulators (32 registers) -- so the CTA
// holds 1024 threads at 64 registers; 8 warps per scheduler hide the LDS -> DFMA latency that kept the 256-thread-per-tile
// forms (round 1: k_tile_apply, fill / barrier / walk / update in turn, 3240 ms on 28 q / 10^4 terms; its pipelined 256-thread
// successor: 2796 ms) at one instruction per SM-cycle.  The dst lines of a tile are pulled into L2 when its walk starts.
//
// Records (tile_plan.hpp, build_tile_records(..., apply_keys = true)): key = which * 8 + (XR & 7) -- sixteen specialised
// bodies -- and bit 3 of XR, which only swaps the two halves, is a run-time row offset.  A single-term record updates its
// factor +-u IN PLACE along a Gray code of the rows (one LOP3 per row on the high word: 0, 1, 3, 2 | 6, 7, 5, 4).
//
// Addresses: shards of at most 32 local qubits are addressed by a 32-bit ELEMENT index (high index << C | low bits); row r of a
// thread is  e0 ^ (XOR of the row generators of the set bits of r)  for the source fill, the dst prefetch and the dst update
// alike: LOP3 + IMAD.WIDE per access.  (H psi plans have no shear: the low index bits of a thread are those of its lane.)
//
// ALPHA (last batch of an H psi whose caller wants <src|dst>): after a tile's outputs are final the thread adds
// Re conj(src[j]) dst[j] of its outputs -- the source amplitudes are still in the stage -- and the CTA's sum goes through
// the deterministic grid reduction; the separate k_alpha pass over both vectors (32 B per amplitude) is not needed.
#pragma once
#include "kernels_tile_ring.cuh"

namespace b200vqe {

struct ApplyTables {                 // one batch per launch: shared by both groups
    TileBatch sb;
    int4 recs[2 * (kStageRecs + 1)];
    double2 terms[kStageTerms];
};

constexpr int kApply8GroupThreads = 2 * kBlock;                       // 512 threads walk one tile
constexpr int kApply8Threads = kRingGroups * kApply8GroupThreads;     // 1024
constexpr int kApplyKeys = 16;                                        // (which, XR & 7)

__device__ __forceinline__ void group_sync8(int g) {                   // named barrier of one 512-thread group
    if (g == 0) asm volatile("bar.sync 1, %0;" ::"n"(kApply8GroupThreads) : "memory");
    else asm volatile("bar.sync 2, %0;" ::"n"(kApply8GroupThreads) : "memory");
}

// one record on the 8 outputs of a thread (all operands as 32-bit shared addresses, the record's second half fetched here)
template <int WHICH, int X7>
__device__ __forceinline__ void apply_rec_body8(double2 (&acc)[8], unsigned tile_u, unsigned h8, unsigned v0, unsigned rep,
                                                const int4 ra, unsigned ru, unsigned terms_u) {
    // row 8 (h ^ XR bit 3) of the partner thread v0 ^ xt
    const unsigned Pu = tile_u + (((v0 ^ (unsigned)ra.z) & 0xffu) << 4) + ((h8 ^ (((unsigned)ra.z >> 8) & 8u)) << 12);
    const int count = ra.y;
    if (count == 1) {
        const int4 rb = lds_i4(ru + 16);                       // first term inline: packed z parts, coefficient
        const unsigned zl = (unsigned)rb.x, hw = (unsigned)rb.y;
        // sign at row 8h + r:  parity(rep & z_hi) ^ parity(thread & zt) ^ (h & zr_3)  ^  parity(r & zr_{0..2})
        const unsigned par = (unsigned)__popc((rep & zl) ^ (v0 & hw & 0xffu)) ^ ((hw >> 11) & (h8 >> 3));
        const unsigned g0 = (hw << 23) & 0x80000000u, g1 = (hw << 22) & 0x80000000u, g2 = (hw << 21) & 0x80000000u;
        int fh = rb.w ^ (int)(par << 31);                      // high word of the factor +-u, updated in place along a Gray code
        const int fl = rb.z;
#define B200VQE_ROW(R, B)                                                            \
        {                                                                            \
            const double f = __hiloint2double(fh, fl);                               \
            if (WHICH == 0) {                                                        \
                acc[R].x = fma(f, (B).x, acc[R].x);                                  \
                acc[R].y = fma(f, (B).y, acc[R].y);                                  \
            } else {                                                                 \
                acc[R].x = fma(-f, (B).y, acc[R].x);                                 \
                acc[R].y = fma(f, (B).x, acc[R].y);                                  \
            }                                                                        \
        }
        {
            const double2 b0 = lds_f64x2(Pu + ((unsigned)(0 ^ X7) << 12)), b1 = lds_f64x2(Pu + ((unsigned)(1 ^ X7) << 12));
            const double2 b3 = lds_f64x2(Pu + ((unsigned)(3 ^ X7) << 12)), b2 = lds_f64x2(Pu + ((unsigned)(2 ^ X7) << 12));
            B200VQE_ROW(0, b0) fh ^= (int)g0;
            B200VQE_ROW(1, b1) fh ^= (int)g1;
            B200VQE_ROW(3, b3) fh ^= (int)g0;
            B200VQE_ROW(2, b2) fh ^= (int)g2;
        }
        {
            const double2 b6 = lds_f64x2(Pu + ((unsigned)(6 ^ X7) << 12)), b7 = lds_f64x2(Pu + ((unsigned)(7 ^ X7) << 12));
            const double2 b5 = lds_f64x2(Pu + ((unsigned)(5 ^ X7) << 12)), b4 = lds_f64x2(Pu + ((unsigned)(4 ^ X7) << 12));
            B200VQE_ROW(6, b6) fh ^= (int)g0;
            B200VQE_ROW(7, b7) fh ^= (int)g1;
            B200VQE_ROW(5, b5) fh ^= (int)g0;
            B200VQE_ROW(4, b4)
        }
#undef B200VQE_ROW
        return;
    }
    const unsigned tu0 = terms_u + ((unsigned)ra.x << 4);
#pragma unroll
    for (int part = 0; part < 2; ++part) {       // four rows at a time keeps the factor sums in registers
        double f[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll 1
        for (int e = 0; e < count; ++e) {
            const double2 t = lds_f64x2(tu0 + ((unsigned)e << 4));
            const unsigned long long packed = (unsigned long long)__double_as_longlong(t.x);
            const unsigned zl = (unsigned)packed, hw = (unsigned)(packed >> 32);
            const unsigned par = (unsigned)__popc((rep & zl) ^ (v0 & hw & 0xffu)) & 1u;
            const unsigned S = ((hw >> 16) ^ (0u - par)) >> h8;       // bit r: sign of the term at row 8h + r
#pragma unroll
            for (int i = 0; i < 4; ++i) f[i] += flip_bit(t.y, S, part * 4 + i);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int r = part * 4 + i;
            const double2 b = lds_f64x2(Pu + ((unsigned)(r ^ X7) << 12));
            if (WHICH == 0) {
                acc[r].x = fma(f[i], b.x, acc[r].x);
                acc[r].y = fma(f[i], b.y, acc[r].y);
            } else {
                acc[r].x = fma(-f[i], b.y, acc[r].x);
                acc[r].y = fma(f[i], b.x, acc[r].y);
            }
        }
    }
}

template <int K>
__device__ __forceinline__ void apply_runs8(double2 (&acc)[8], unsigned tile_u, unsigned h8, unsigned v0, unsigned rep, unsigned &ru, int4 &ra,
                                            unsigned terms_u) {
    if constexpr (K < kApplyKeys) {
        while ((((unsigned)ra.z >> 16) & 0xffu) == (unsigned)K) {
            apply_rec_body8<(K >> 3), (K & 7)>(acc, tile_u, h8, v0, rep, ra, ru, terms_u);
            ru += 32;
            ra = lds_i4(ru);
        }
        apply_runs8<K + 1>(acc, tile_u, h8, v0, rep, ru, ra, terms_u);
    }
}

struct Apply8Consts {                 // per launch, read from shared memory where they are used (not kept in registers)
    unsigned g0, g1, g2, piv, dstep, pad[3];
    double in_scale, prev_scale;
};
constexpr int kApply8SmemBytes = kRingStages * kRingTileBytes + (int)sizeof(ApplyTables) + (int)sizeof(Apply8Consts) + kRingGroups * kRingStages * 8 + 64 * 8;

template <bool ALPHA>
__global__ void __launch_bounds__(kApply8Threads, 1)
k_tile_apply_pipe8(const double2 *__restrict__ src, double2 *__restrict__ dst, int C, int M, const TileBatch *__restrict__ batch,
                   unsigned tile_count, const TileRec *__restrict__ recs, const TileTerm *__restrict__ terms, int first,
                   const double *norm_src, const double *norm_prev, int fuse_prev, double *partials, unsigned *counter, double *alpha_out) {
    extern __shared__ __align__(128) unsigned char ring_raw[];
    unsigned char *tiles = ring_raw;                                                     // [kRingStages][64 KB]
    ApplyTables &tb = *reinterpret_cast<ApplyTables *>(ring_raw + kRingStages * kRingTileBytes);
    Apply8Consts &cs = *reinterpret_cast<Apply8Consts *>(&tb + 1);
    unsigned long long *bars = reinterpret_cast<unsigned long long *>(&cs + 1);          // full[stage][walking group]
    double *s_red = reinterpret_cast<double *>(bars + kRingGroups * kRingStages);         // [32] warp sums
    const unsigned tiles_u = smem_u32(tiles);
    const unsigned full_u = smem_u32(bars);
    const unsigned cs_u = smem_u32(&cs);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kRingGroups * kRingStages; ++s) mbar_init(full_u + 8 * s, kApply8GroupThreads);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (threadIdx.x < kTileBatchWords) reinterpret_cast<uint32_t *>(&tb.sb)[threadIdx.x] = reinterpret_cast<const uint32_t *>(batch)[threadIdx.x];
    __syncthreads();
    {
        const int n_r = tb.sb.r_end - tb.sb.r_begin, n_t = tb.sb.t_end - tb.sb.t_begin;
        const int4 *rsrc = reinterpret_cast<const int4 *>(recs + tb.sb.r_begin);
        for (int i = threadIdx.x; i < 2 * n_r; i += kApply8Threads) tb.recs[i] = rsrc[i];
        if (threadIdx.x == 0) tb.recs[2 * n_r] = make_int4(0, 0, kTileRecKeyEnd << 16, 0);
        const double2 *tsrc = reinterpret_cast<const double2 *>(terms) + tb.sb.t_begin;
        for (int i = threadIdx.x; i < n_t; i += kApply8Threads) tb.terms[i] = tsrc[i];
        if (threadIdx.x == 0) {
            cs.g0 = tb.sb.basis[M - 4] << C; cs.g1 = tb.sb.basis[M - 3] << C; cs.g2 = tb.sb.basis[M - 2] << C;
            cs.piv = tb.sb.pad[0];
            cs.dstep = tile_rep(tb.sb, kRingGroups * gridDim.x);       // tile t -> t + step: masked add of the deposited stride
            cs.in_scale = norm_src ? 1.0 / *norm_src : 1.0;
            cs.prev_scale = fuse_prev ? (*norm_src) / (*norm_prev) : 0.0;
        }
    }
    __syncthreads();

    const int g = threadIdx.x >> 9;
    const unsigned v0 = threadIdx.x & (kBlock - 1);
    const unsigned h8 = (threadIdx.x >> 5) & 8u;                        // 8 * h: first row of this thread (warp-uniform)
    const unsigned n_items = blockIdx.x < tile_count ? (tile_count - blockIdx.x + gridDim.x - 1) / gridDim.x : 0u;
    const unsigned recs_u = smem_u32(tb.recs), terms_u = smem_u32(tb.terms);
    // element index of (this thread, its first row) of the tile with representative rep:  (rep << C) ^ e_thread
    unsigned e_thread = (v0 & ((1u << C) - 1u)) ^ (h8 ? tb.sb.basis[M - 1] << C : 0u);
    for (int i = 0; i < 8 - C; ++i) if ((v0 >> (C + i)) & 1u) e_thread ^= tb.sb.basis[i] << C;
    const bool reads_dst = !first || fuse_prev;

    unsigned fi = (unsigned)(1 - g);
    int fs = 1 - g;
    unsigned f_rep = tile_rep(tb.sb, blockIdx.x + fi * gridDim.x);
    auto fill_one = [&]() {
        const uint2 gg = lds_u32x2(cs_u);                // g0, g1
        const uint2 gp = lds_u32x2(cs_u + 8);            // g2, piv
        const unsigned e0 = (f_rep << C) ^ e_thread;
        const unsigned d0 = tiles_u + (unsigned)fs * kRingTileBytes + (v0 << 4) + (h8 << 12);
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            const unsigned e = e0 ^ ((r & 1) ? gg.x : 0u) ^ ((r & 2) ? gg.y : 0u) ^ ((r & 4) ? gp.x : 0u);
            cp_async16_u(d0 + ((unsigned)r << 12), src + e);
        }
        cp_async_arrive_noinc(full_u + 8 * (kRingGroups * fs + (1 - g)));
        f_rep = ((f_rep | gp.y) + lds_u32x2(cs_u + 16).x) & ~gp.y;
        fi += kRingGroups;
        fs += kRingGroups;
        if (fs >= kRingStages) fs -= kRingStages;
    };
    if (fi < n_items) fill_one();
    if (g == 1 && fi < n_items) fill_one();

    unsigned rep = tile_rep(tb.sb, blockIdx.x + (unsigned)g * gridDim.x);
    int s = g, s_wraps = 0;
    unsigned u = 0;
    double alpha = 0.0;
    for (unsigned i = (unsigned)g; i < n_items; i += kRingGroups) {
        if (reads_dst) {     // this tile's dst lines -> L2 while its records are walked: the 8 lanes sharing a 128-byte line and the two
                             // halves split its 16 rows, this thread takes row 8h + (lane & 7)
            const uint2 gg = lds_u32x2(cs_u);
            const unsigned g2 = lds_u32x2(cs_u + 8).x;
            prefetch_l2(dst + ((rep << C) ^ e_thread ^ ((v0 & 1u) ? gg.x : 0u) ^ ((v0 & 2u) ? gg.y : 0u) ^ ((v0 & 4u) ? g2 : 0u)));
        }
        mbar_wait(full_u + 8 * (kRingGroups * s + g), u & 1u);
        const unsigned tile_u = tiles_u + (unsigned)s * kRingTileBytes;
        double2 acc[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) acc[r] = make_double2(0.0, 0.0);
        {
            unsigned ru = recs_u;
            int4 ra = lds_i4(ru);
            apply_runs8<0>(acc, tile_u, h8, v0, rep, ru, ra, terms_u);
        }
        {
            const uint2 gg = lds_u32x2(cs_u);
            const unsigned g2 = lds_u32x2(cs_u + 8).x;
            const double2 sc = lds_f64x2(cs_u + 32);       // in_scale, prev_scale
            const unsigned e0 = (rep << C) ^ e_thread;
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const unsigned e = e0 ^ ((r & 1) ? gg.x : 0u) ^ ((r & 2) ? gg.y : 0u) ^ ((r & 4) ? g2 : 0u);
                double2 *p = dst + e;
                double2 o = make_double2(acc[r].x * sc.x, acc[r].y * sc.x);
                if (first) {
                    if (fuse_prev) {
                        const double2 q = *p;
                        o.x -= sc.y * q.x;
                        o.y -= sc.y * q.y;
                    }
                } else {
                    const double2 q = *p;
                    o.x += q.x;
                    o.y += q.y;
                }
                __stcs(p, o);
                if (ALPHA) {
                    const double2 a = lds_f64x2(tile_u + (v0 << 4) + ((h8 + (unsigned)r) << 12));       // src[j] of this output
                    alpha = fma(a.x, o.x, fma(a.y, o.y, alpha));
                }
            }
        }
        if (i + kRingStages < n_items) {
            group_sync8(g);
            fill_one();
        }
        {
            const unsigned piv = lds_u32x2(cs_u + 8).y, dstep = lds_u32x2(cs_u + 16).x;
            rep = ((rep | piv) + dstep) & ~piv;
        }
        s += kRingGroups;
        if (s >= kRingStages) s -= kRingStages;
        if (++s_wraps == kRingStages) { s_wraps = 0; ++u; }
    }
    if (!ALPHA) return;
    const double wsum = warp_sum(alpha);
    if (lane == 0) s_red[warp] = wsum;
    __syncthreads();
    __shared__ bool s_last;
    if (threadIdx.x == 0) {
        double sum = 0.0;
#pragma unroll
        for (int w = 0; w < kApply8Threads / 32; ++w) sum += s_red[w];
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
        if (lane == 0) { alpha_out[0] = lds_f64x2(cs_u + 32).x * sum; *counter = 0u; }
    }
}

}  // namespace b200vqe
