// K2 on the tile engine, pipelined:  k_tile_apply_pipe  (dst (+)= in_scale * sum over the groups of ONE batch, as k_tile_apply).
//
// k_tile_apply (kernels_tile.cuh) fills a tile, walks the batch's records and updates dst in turn:
// ld.global x16 -> barrier -> STS x16 -> barrier -> walk -> dst read-modify-write, two CTAs of 8 warps per SM.  Here the
// fill is taken off the critical path exactly as in k_tile_expect_pipe (kernels_tile_ring.cuh): one CTA per SM, a ring of
// three 64 KB tile buffers filled with cp.async (LDGSTS, every thread copies the 16 source amplitudes of its own 16 outputs),
// one mbarrier per (stage, walking group), two groups of 256 threads that own every other tile of the CTA and refill the stage
// they have just walked with the CTA's tile i + 3.  The dst lines of a tile are pulled into L2 when its walk starts, so the
// read-modify-write at the end of the walk finds them there.
//
// Addresses: a register of at most 32 qubits is addressed by a 32-bit ELEMENT index (high index << C | low bits); row r of a
// thread is  e0 ^ (XOR of the four row generators of the set bits of r)  for the source fill, the dst prefetch and the dst
// update alike: LOP3 + IMAD.WIDE per access.  (H psi plans have no shear: the low index bits of a thread are those of its lane.)
//
// ALPHA (last batch of an H psi whose caller wants <src|dst>): after a tile's outputs are final the thread adds
// Re conj(src[j]) dst[j] of its 16 outputs -- the source amplitudes are still in the stage -- and the CTA's sum goes through
// the deterministic grid reduction; the separate k_alpha pass over both vectors (32 B per amplitude) is not needed.
#pragma once
#include "kernels_tile_ring.cuh"

namespace b200vqe {

struct ApplyTables {                 // one batch per launch: shared by both groups
    TileBatch sb;
    int4 recs[2 * (kStageRecs + 1)];
    double2 terms[kStageTerms];
};
constexpr int kApplyPipeSmemBytes = kRingStages * kRingTileBytes + (int)sizeof(ApplyTables) + kRingGroups * kRingStages * 8 + 32 * 8;

template <bool ALPHA>
__global__ void __launch_bounds__(kRingThreads, 1)
k_tile_apply_pipe(const double2 *__restrict__ src, double2 *__restrict__ dst, int C, int M, const TileBatch *__restrict__ batch,
                  unsigned tile_count, const TileRec *__restrict__ recs, const TileTerm *__restrict__ terms, int first,
                  const double *norm_src, const double *norm_prev, int fuse_prev, double *partials, unsigned *counter, double *alpha_out) {
    extern __shared__ __align__(128) unsigned char ring_raw[];
    unsigned char *tiles = ring_raw;                                                     // [kRingStages][64 KB]
    ApplyTables &tb = *reinterpret_cast<ApplyTables *>(ring_raw + kRingStages * kRingTileBytes);
    unsigned long long *bars = reinterpret_cast<unsigned long long *>(&tb + 1);          // full[stage][walking group]
    double *s_red = reinterpret_cast<double *>(bars + kRingGroups * kRingStages);         // [16] warp sums
    const unsigned tiles_u = smem_u32(tiles);
    const unsigned full_u = smem_u32(bars);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kRingGroups * kRingStages; ++s) mbar_init(full_u + 8 * s, kBlock);      // one (noinc) arrive per copying thread
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (threadIdx.x < kTileBatchWords) reinterpret_cast<uint32_t *>(&tb.sb)[threadIdx.x] = reinterpret_cast<const uint32_t *>(batch)[threadIdx.x];
    __syncthreads();
    {   // the batch's records (+ sentinel) and terms; the planner keeps every batch within kStageRecs / kStageTerms
        const int n_r = tb.sb.r_end - tb.sb.r_begin, n_t = tb.sb.t_end - tb.sb.t_begin;
        const int4 *rsrc = reinterpret_cast<const int4 *>(recs + tb.sb.r_begin);
        for (int i = threadIdx.x; i < 2 * n_r; i += kRingThreads) tb.recs[i] = rsrc[i];
        if (threadIdx.x == 0) tb.recs[2 * n_r] = make_int4(0, 0, kTileRecKeyEnd << 16, 0);
        const double2 *tsrc = reinterpret_cast<const double2 *>(terms) + tb.sb.t_begin;
        for (int i = threadIdx.x; i < n_t; i += kRingThreads) tb.terms[i] = tsrc[i];
    }
    __syncthreads();

    const int g = warp / (kBlock / 32);
    const unsigned v0 = threadIdx.x & (kBlock - 1);
    const unsigned n_items = blockIdx.x < tile_count ? (tile_count - blockIdx.x + gridDim.x - 1) / gridDim.x : 0u;
    // element index of (this thread, register 0) of the tile with representative rep:  ((rep ^ Tc) << C) | lo
    unsigned e_thread = v0 & ((1u << C) - 1u);
    for (int i = 0; i < 8 - C; ++i) if ((v0 >> (C + i)) & 1u) e_thread ^= tb.sb.basis[i] << C;
    const unsigned g0 = tb.sb.basis[M - 4] << C, g1 = tb.sb.basis[M - 3] << C, g2 = tb.sb.basis[M - 2] << C, g3 = tb.sb.basis[M - 1] << C;
    // tile t -> t + step: masked add of the deposited stride (tile_plan.hpp)
    const unsigned piv = tb.sb.pad[0];
    const unsigned dstep = tile_rep(tb.sb, kRingGroups * gridDim.x);
    const double in_scale = norm_src ? 1.0 / *norm_src : 1.0;
    const double prev_scale = fuse_prev ? (*norm_src) / (*norm_prev) : 0.0;
    const bool reads_dst = !first || fuse_prev;
    // the 8 lanes sharing a 128-byte line split its 16 rows between them: this thread prefetches rows 2 (lane & 7) and 2 (lane & 7) + 1
    const unsigned e_pf = e_thread ^ ((v0 & 1u) ? g1 : 0u) ^ ((v0 & 2u) ? g2 : 0u) ^ ((v0 & 4u) ? g3 : 0u);

    // ---- fill side: this group streams items 1 - g, 3 - g, ... (stage = item % 3); the item is walked by the other group ----
    unsigned fi = (unsigned)(1 - g);
    int fs = 1 - g;
    unsigned f_rep = tile_rep(tb.sb, blockIdx.x + fi * gridDim.x);
    auto fill_one = [&]() {
        const unsigned e0 = (f_rep << C) ^ e_thread;
        const unsigned d0 = tiles_u + (unsigned)fs * kRingTileBytes + (v0 << 4);
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            const unsigned e = e0 ^ ((r & 1) ? g0 : 0u) ^ ((r & 2) ? g1 : 0u) ^ ((r & 4) ? g2 : 0u) ^ ((r & 8) ? g3 : 0u);
            cp_async16_u(d0 + ((unsigned)r << 12), src + e);
        }
        cp_async_arrive_noinc(full_u + 8 * (kRingGroups * fs + (1 - g)));
        f_rep = ((f_rep | piv) + dstep) & ~piv;
        fi += kRingGroups;
        fs += kRingGroups;
        if (fs >= kRingStages) fs -= kRingStages;
    };
    if (fi < n_items) fill_one();
    if (g == 1 && fi < n_items) fill_one();

    // ---- walk side: this group owns items g, g + 2, ... ----
    unsigned rep = tile_rep(tb.sb, blockIdx.x + (unsigned)g * gridDim.x);
    int s = g, s_wraps = 0;
    unsigned u = 0;
    double alpha = 0.0;
    for (unsigned i = (unsigned)g; i < n_items; i += kRingGroups) {
        if (reads_dst) {     // this tile's dst lines -> L2 while its records are walked (one tile ahead: three tiles ahead they were evicted again)
            const unsigned ep = (rep << C) ^ e_pf;
            prefetch_l2(dst + ep);
            prefetch_l2(dst + (ep ^ g0));
        }
        mbar_wait(full_u + 8 * (kRingGroups * s + g), u & 1u);
        const unsigned tile_u = tiles_u + (unsigned)s * kRingTileBytes;
        double2 acc[16];
#pragma unroll
        for (int r = 0; r < 16; ++r) acc[r] = make_double2(0.0, 0.0);
        {
            const double2 *tile = reinterpret_cast<const double2 *>(tiles + (size_t)s * kRingTileBytes);
            int gi = 0;
            int4 ra = tb.recs[0], rb = tb.recs[1];
            apply_runs<0>(acc, tile, v0, rep, tb.recs, tb.terms, gi, ra, rb);
        }
        const unsigned e0 = (rep << C) ^ e_thread;
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            const unsigned e = e0 ^ ((r & 1) ? g0 : 0u) ^ ((r & 2) ? g1 : 0u) ^ ((r & 4) ? g2 : 0u) ^ ((r & 8) ? g3 : 0u);
            double2 *p = dst + e;
            double2 o = make_double2(acc[r].x * in_scale, acc[r].y * in_scale);
            if (first) {
                if (fuse_prev) {
                    const double2 q = *p;
                    o.x -= prev_scale * q.x;
                    o.y -= prev_scale * q.y;
                }
            } else {
                const double2 q = *p;
                o.x += q.x;
                o.y += q.y;
            }
            __stcs(p, o);
            if (ALPHA) {
                const double2 a = lds_f64x2(tile_u + (v0 << 4) + ((unsigned)r << 12));       // src[j] of this output
                alpha = fma(a.x, o.x, fma(a.y, o.y, alpha));
            }
        }
        if (i + kRingStages < n_items) {
            group_sync(g);                                  // the stage is refilled by this same group
            fill_one();
        }
        rep = ((rep | piv) + dstep) & ~piv;
        s += kRingGroups;
        if (s >= kRingStages) s -= kRingStages;
        if (++s_wraps == kRingStages) { s_wraps = 0; ++u; }
    }
    if (!ALPHA) return;
    // alpha = in_scale * Re <src|dst>: CTA sum in warp order, the last CTA adds the CTAs in order (deterministic)
    const double wsum = warp_sum(alpha);
    if (lane == 0) s_red[warp] = wsum;
    __syncthreads();
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
        if (lane == 0) { alpha_out[0] = in_scale * sum; *counter = 0u; }
    }
}


// ---------------------------------------------------------------------------------------------
// k_tile_apply_pipe8: the same pipeline with 512 threads per tile.  The two halves (h = 0, 1) of a walking group split the
// 16 rows of every thread column between them -- thread (v0, h) owns rows 8h .. 8h + 7, i.e. 8 outputs and 8 complex
// accumulators (32 registers instead of 64) -- so the CTA holds 1024 threads at 64 registers: 8 warps per scheduler instead
// of 4 hide the LDS -> DFMA latency that keeps k_tile_apply_pipe at ~1 instruction per SM-cycle.  Same plans, same records:
// the partner of row r is row r ^ XR of thread v0 ^ xt; the half only moves the first partner row (XR bit 3 swaps the halves).
// ---------------------------------------------------------------------------------------------
constexpr int kApply8GroupThreads = 2 * kBlock;                       // 512 threads walk one tile
constexpr int kApply8Threads = kRingGroups * kApply8GroupThreads;     // 1024

__device__ __forceinline__ void group_sync8(int g) {                   // named barrier of one 512-thread group
    if (g == 0) asm volatile("bar.sync 1, %0;" ::"n"(kApply8GroupThreads) : "memory");
    else asm volatile("bar.sync 2, %0;" ::"n"(kApply8GroupThreads) : "memory");
}

// one record on the 8 outputs of a thread (all operands as 32-bit shared addresses, the record's second half fetched here)
template <int WHICH, int XR>
__device__ __forceinline__ void apply_rec_body8(double2 (&acc)[8], unsigned tile_u, unsigned h8, unsigned v0, unsigned rep,
                                                const int4 ra, unsigned ru, unsigned terms_u) {
    constexpr int X7 = XR & 7;
    // row (8h ^ (XR & 8)) of the partner thread v0 ^ xt
    const unsigned Pu = tile_u + (((v0 ^ (unsigned)ra.z) & 0xffu) << 4) + ((h8 ^ (unsigned)(XR & 8)) << 12);
    const int count = ra.y;
    if (count == 1) {
        const int4 rb = lds_i4(ru + 16);                       // first term inline: packed z parts, coefficient
        const unsigned zl = (unsigned)rb.x, hw = (unsigned)rb.y;
        const unsigned par = (unsigned)__popc((rep & zl) ^ (v0 & hw & 0xffu)) & 1u;
        const unsigned S = ((hw >> 16) ^ (0u - par)) >> h8;   // bit r: sign of the term at row 8h + r
        const double u = __hiloint2double(rb.w, rb.z);
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {                       // four partner rows in flight at a time
            double2 b[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) b[k] = lds_f64x2(Pu + ((unsigned)((4 * hh + k) ^ X7) << 12));
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int r = 4 * hh + k;
                const double bx = flip_bit(b[k].x, S, r), by = flip_bit(b[k].y, S, r);      // the sign rides on the partner amplitude
                if (WHICH == 0) {
                    acc[r].x = fma(u, bx, acc[r].x);
                    acc[r].y = fma(u, by, acc[r].y);
                } else {
                    acc[r].x = fma(-u, by, acc[r].x);
                    acc[r].y = fma(u, bx, acc[r].y);
                }
            }
        }
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
            const unsigned S = ((hw >> 16) ^ (0u - par)) >> h8;
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
    if constexpr (K < kTileRecKeys) {
        while (((unsigned)ra.z >> 16) == (unsigned)K) {
            apply_rec_body8<(K >> 4), (((K & 15) + 1) & 15)>(acc, tile_u, h8, v0, rep, ra, ru, terms_u);
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
