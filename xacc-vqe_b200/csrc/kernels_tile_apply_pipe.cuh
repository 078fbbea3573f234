// K2 on the tile engine, pipelined:  k_tile_apply_pipe  (dst (+)= in_scale * sum over the groups of ONE batch, as k_tile_apply).
//
// k_tile_apply (kernels_tile.cuh) fills a tile, walks the batch's records and updates dst in turn:
// ld.global x16 -> barrier -> STS x16 -> barrier -> walk -> dst read-modify-write, two CTAs of 8 warps per SM.  Here the
// fill is taken off the critical path exactly as in k_tile_expect_pipe (kernels_tile_ring.cuh): one CTA per SM, a ring of
// three 64 KB tile buffers filled with cp.async (LDGSTS, every thread copies the 16 source amplitudes of its own 16 outputs),
// one mbarrier per (stage, walking group), two groups of 256 threads that own every other tile of the CTA and refill the stage
// they have just walked with the CTA's tile i + 3.  The dst lines of a tile are pulled into L2 when its source tile is
// requested (three tiles ahead), so the read-modify-write at the end of the walk finds them there.
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
        if (reads_dst) {
            const unsigned ep = (f_rep << C) ^ e_pf;
            prefetch_l2(dst + ep);
            prefetch_l2(dst + (ep ^ g0));
        }
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

}  // namespace b200vqe
