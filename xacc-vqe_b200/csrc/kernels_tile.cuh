// Tile engine: <psi|H|psi> with several xmask groups evaluated per HBM read of the state.
// See tile_plan.hpp for the decomposition.  One CTA = 256 threads owns one tile (a coset of the
// batch's subspace: 4096 amplitudes, 64 KB) in shared memory; each thread also keeps its own 16
// amplitudes in registers.  For every group of the batch a thread visits the 8 of its amplitudes
// that are the "bit h clear" element of their pair {v, v^xv} (all 16 when the group has no register
// bit), reads the partner from shared memory (one conflict-free LDS.128) and accumulates
//     E += fe(j) Re w + go(j) Im w,   w = conj(psi[j^x]) psi[j],
// fe/go being the signed coefficient sums of the group's even-/odd-nY terms at j.  Signs come from
// one 64-bit POPC per term (coset part), one 8-bit POPC (thread part) and a host-made 16-bit sign
// word over the register index, so the per-amplitude cost of a term is a shift, a LOP3 and a DADD.
//
// HBM traffic: 16 B per amplitude per BATCH (>= kTileBits - C groups), no writes.
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

__device__ __forceinline__ TileTermRaw load_term(const TileTerm *__restrict__ terms, uint32_t k) {
    const double2 raw = __ldg(reinterpret_cast<const double2 *>(terms) + k);
    TileTermRaw t;
    t.packed = (unsigned long long)__double_as_longlong(raw.x);
    t.u = raw.y;
    return t;
}

// visits the 8 registers r whose bit HS equals BV
template <int HS, int BV>
__device__ __forceinline__ void tile_group(const double2 (&a)[16], const double2 *__restrict__ tile, unsigned v0,
                                           unsigned long long rep, const TileGroup g,
                                           const TileTerm *__restrict__ terms, double (&acc)[4]) {
    constexpr int NA = 8;
    double fe[NA], go[NA];
#pragma unroll
    for (int i = 0; i < NA; ++i) { fe[i] = 0.0; go[i] = 0.0; }
    uint32_t k = g.kb;
    for (uint32_t e = 0; e < g.n_even; ++e, ++k) {
        const TileTermRaw t = load_term(terms, k);
        const unsigned p0 = (unsigned)(__popcll(rep & t.packed) + __popc(v0 & (unsigned)(t.packed >> 40) & 0xffu)) & 1u;
        const unsigned S = (unsigned)(t.packed >> 48) ^ (0u - p0);
#pragma unroll
        for (int i = 0; i < NA; ++i) {
            const int r = ((i >> HS) << (HS + 1)) | (BV << HS) | (i & ((1 << HS) - 1));
            fe[i] += flip_bit(t.u, S, r);
        }
    }
    for (uint32_t o = 0; o < g.n_odd; ++o, ++k) {
        const TileTermRaw t = load_term(terms, k);
        const unsigned p0 = (unsigned)(__popcll(rep & t.packed) + __popc(v0 & (unsigned)(t.packed >> 40) & 0xffu)) & 1u;
        const unsigned S = (unsigned)(t.packed >> 48) ^ (0u - p0);
#pragma unroll
        for (int i = 0; i < NA; ++i) {
            const int r = ((i >> HS) << (HS + 1)) | (BV << HS) | (i & ((1 << HS) - 1));
            go[i] += flip_bit(t.u, S, r);
        }
    }
    const unsigned xr = (unsigned)g.xv >> 8;
    const unsigned plo = (v0 ^ (unsigned)g.xv) & 0xffu;
#pragma unroll
    for (int i = 0; i < NA; ++i) {
        const int r = ((i >> HS) << (HS + 1)) | (BV << HS) | (i & ((1 << HS) - 1));
        const double2 b = tile[(((unsigned)r ^ xr) << 8) | plo];
        const double wre = b.x * a[r].x + b.y * a[r].y;
        const double wim = b.x * a[r].y - b.y * a[r].x;
        acc[i & 3] += fe[i] * wre + go[i] * wim;
    }
}

__global__ void __launch_bounds__(kBlock, 2)
k_tile_expect(const double2 *__restrict__ psi, int C, int M, const TileBatch *__restrict__ batches, int n_batches,
              unsigned long long tile_begin, unsigned long long tile_count, const TileGroup *__restrict__ groups,
              const TileTerm *__restrict__ terms, double *partials, unsigned *counter, double *out) {
    extern __shared__ __align__(16) unsigned char tile_raw[];
    double2 *tile = reinterpret_cast<double2 *>(tile_raw);
    __shared__ TileBatch sb;
    const unsigned v0 = threadIdx.x;
    const unsigned lo = v0 & ((1u << C) - 1u);
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    int cur_b = -1;
    unsigned long long Tc = 0, rb0 = 0, rb1 = 0, rb2 = 0, rb3 = 0;
    const unsigned long long total = (unsigned long long)n_batches * tile_count;

    for (unsigned long long w = blockIdx.x; w < total; w += gridDim.x) {
        const int b = (int)(w / tile_count);
        const unsigned long long t = tile_begin + (w - (unsigned long long)b * tile_count);
        if (b != cur_b) {                       // uniform over the CTA
            __syncthreads();
            if (threadIdx.x < 32)
                reinterpret_cast<uint32_t *>(&sb)[threadIdx.x] = reinterpret_cast<const uint32_t *>(&batches[b])[threadIdx.x];
            __syncthreads();
            cur_b = b;
            Tc = 0;
            for (int i = 0; i < 8 - C; ++i) if ((v0 >> (C + i)) & 1u) Tc ^= sb.basis[i];
            rb0 = sb.basis[M - 4]; rb1 = sb.basis[M - 3]; rb2 = sb.basis[M - 2]; rb3 = sb.basis[M - 1];
        }
        unsigned long long rep = t;
        for (int i = 0; i < M; ++i) {
            const int p = sb.pivot[i];
            rep = ((rep >> p) << (p + 1)) | (rep & ((1ull << p) - 1ull));
        }
        const unsigned long long base = rep ^ Tc;
        double2 a[16];
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            const unsigned long long hi = base ^ ((r & 1) ? rb0 : 0ull) ^ ((r & 2) ? rb1 : 0ull) ^ ((r & 4) ? rb2 : 0ull) ^
                                          ((r & 8) ? rb3 : 0ull);
            a[r] = psi[(hi << C) | lo];
        }
        __syncthreads();                        // readers of the previous tile are done
#pragma unroll
        for (int r = 0; r < 16; ++r) tile[((unsigned)r << 8) | v0] = a[r];
        __syncthreads();
        for (int g = sb.g_begin; g < sb.g_end; ++g) {
            const int4 raw = __ldg(reinterpret_cast<const int4 *>(groups) + g);
            TileGroup tg;
            tg.kb = (uint32_t)raw.x; tg.n_even = (uint32_t)raw.y; tg.n_odd = (uint32_t)raw.z;
            tg.xv = (uint16_t)((unsigned)raw.w & 0xffffu);
            tg.hsel = (int16_t)((unsigned)raw.w >> 16);
            switch (tg.hsel) {
            case 0: tile_group<0, 0>(a, tile, v0, rep, tg, terms, acc); break;
            case 1: tile_group<1, 0>(a, tile, v0, rep, tg, terms, acc); break;
            case 2: tile_group<2, 0>(a, tile, v0, rep, tg, terms, acc); break;
            case 3: tile_group<3, 0>(a, tile, v0, rep, tg, terms, acc); break;
            default:    // no register bit in xv: visit every amplitude (weight 1), in two halves
                tile_group<3, 0>(a, tile, v0, rep, tg, terms, acc);
                tile_group<3, 1>(a, tile, v0, rep, tg, terms, acc);
                break;
            }
        }
    }
    double v[1] = {(acc[0] + acc[1]) + (acc[2] + acc[3])};
    grid_reduce<1>(v, partials, counter, out, 1.0);
}

}  // namespace b200vqe
