// Hand-written sm_100a kernels for the Pauli-sum x statevector path.  No tensor cores: this is
// bit-indexed HBM streaming (coalesced 16-byte double2 loads, parities by POPC, warp-shuffle +
// block reductions in fp64).  See DESIGN.md for the byte model of each kernel.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200vqe {

constexpr int kBlock = 256;        // threads per CTA for every streaming kernel
constexpr int kMaxWorld = 8;

// One xmask group: terms [kb, kb+n_even) have even nY, [kb+n_even, kb+n_even+n_odd) odd nY.
struct GroupDesc {
    uint64_t x;
    uint32_t kb;
    uint32_t n_even;
    uint32_t n_odd;
    uint32_t pad;
};

// Slab table: slab t holds amplitudes with (index >> nl) == t.  Unsharded: p[0] only.
struct Slabs {
    const double2 *p[kMaxWorld];
};

__device__ __forceinline__ double flip_sign(double v, unsigned par) {
    return __hiloint2double(__double2hiint(v) ^ (int)(par << 31), __double2loint(v));
}
__device__ __forceinline__ unsigned parity64(uint64_t v) {
    return (unsigned)__popcll(v) & 1u;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Deterministic grid reduction of NV doubles per CTA: every CTA stores its block sums, the last
// CTA to arrive (atomic ticket) adds them in CTA order and writes out[0..NV).  *counter is reset.
template <int NV>
__device__ __forceinline__ void grid_reduce(double (&v)[NV], double *partials, unsigned *counter,
                                            double *out, double scale, bool take_sqrt = false) {
    __shared__ double s_red[NV][kBlock / 32];
    __shared__ bool s_last;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        const double w = warp_sum(v[i]);
        if (lane == 0) s_red[i][wid] = w;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            double s = 0.0;
#pragma unroll
            for (int w = 0; w < kBlock / 32; ++w) s += s_red[i][w];
            partials[(size_t)i * gridDim.x + blockIdx.x] = s;
        }
        __threadfence();
        const unsigned ticket = atomicAdd(counter, 1u);
        s_last = (ticket == gridDim.x - 1);
    }
    __syncthreads();
    if (s_last) {
        __threadfence();
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            double s = 0.0;
            for (unsigned b = threadIdx.x; b < gridDim.x; b += kBlock)
                s += __ldcg(&partials[(size_t)i * gridDim.x + b]);
            s = warp_sum(s);
            __syncthreads();
            if (lane == 0) s_red[i][wid] = s;
            __syncthreads();
            if (threadIdx.x == 0) {
                double t = 0.0;
#pragma unroll
                for (int w = 0; w < kBlock / 32; ++w) t += s_red[i][w];
                out[i] = take_sqrt ? sqrt(t * scale) : t * scale;
            }
        }
        if (threadIdx.x == 0) *counter = 0u;
    }
}

// ---------------------------------------------------------------------------------------------
// K1  <psi|H|psi>  for one bucket of xmask groups.
//
// Tile-stationary: a CTA keeps R amplitudes per thread (a tile of 256*R) in registers and walks
// every group of the bucket, gathering only the partner amplitudes psi[j ^ x] (512-byte coalesced
// segments, lanes permuted by the low bits of x).  Buckets with h >= log2(tile) enumerate only
// tiles whose bit h is 0, so each pair {j, j^x} is visited once (weight 2): HBM traffic per group
// is 8 bytes per amplitude, half of the 16 B/amplitude model.  The bucket h = -1 (top set bit of x
// inside the tile, incl. the diagonal group x = 0) visits every j (weight 1); its partners sit in
// the tile just loaded, i.e. in L1/L2.
//   per element: w = conj(psi[j^x]) psi[j];  E += fe(j) Re w + go(j) Im w,
//   fe = sum_{k even} u_k (-1)^popc(j&z_k),  go = sum_{k odd} u_k (-1)^popc(j&z_k),
//   u_k = coef_k (even nY) or i*coef_k (odd nY)  -- real for a Hermitian H (REALC).
// ---------------------------------------------------------------------------------------------
template <int R, bool REALC>
__global__ void __launch_bounds__(kBlock, 2)
k_expect_energy(const __grid_constant__ Slabs slabs, const double2 *__restrict__ psi, int nl, int rank,
                const GroupDesc *__restrict__ gd, int g_begin, int g_end,
                const uint64_t *__restrict__ tz, const double2 *__restrict__ tu, int h,
                uint64_t n_tiles, double weight, double *partials, unsigned *counter, double *out) {
    constexpr int TS = kBlock * R;
    constexpr int LB = 8;   // log2(kBlock)
    const uint64_t nloc = 1ull << nl;
    const uint64_t lmask = nloc - 1;
    const uint64_t rank_bits = (uint64_t)rank << nl;
    double acc_re = 0.0, acc_im = 0.0;

    for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const uint64_t p0 = tile * TS;
        const uint64_t base = (h < 0) ? p0 : (((p0 >> h) << (h + 1)) | (p0 & ((1ull << h) - 1)));
        const uint64_t j0 = base + threadIdx.x;
        double2 a[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const uint64_t j = j0 + ((uint64_t)r << LB);
            a[r] = (j < nloc) ? psi[j] : make_double2(0.0, 0.0);
        }
        const uint64_t jg0 = rank_bits | j0;
        for (int g = g_begin; g < g_end; ++g) {
            const GroupDesc d = gd[g];
            const uint64_t xl = d.x & lmask;
            const double2 *__restrict__ pb = slabs.p[rank ^ (int)(d.x >> nl)];
            double wre[R], wim[R];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const uint64_t j = j0 + ((uint64_t)r << LB);
                const double2 b = (j < nloc) ? pb[j ^ xl] : make_double2(0.0, 0.0);
                wre[r] = b.x * a[r].x + b.y * a[r].y;
                wim[r] = b.x * a[r].y - b.y * a[r].x;
            }
            double fe[R], go[R], fei[R], goi[R];
#pragma unroll
            for (int r = 0; r < R; ++r) { fe[r] = 0.0; go[r] = 0.0; fei[r] = 0.0; goi[r] = 0.0; }
            uint32_t k = d.kb;
            for (uint32_t e = 0; e < d.n_even; ++e, ++k) {
                const uint64_t z = tz[k];
                const double2 u = tu[k];
                const unsigned p0par = parity64(jg0 & z);
                const unsigned zr = (unsigned)(z >> LB);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const unsigned par = p0par ^ (__popc(zr & (unsigned)r) & 1u);
                    fe[r] += flip_sign(u.x, par);
                    if (!REALC) fei[r] += flip_sign(u.y, par);
                }
            }
            for (uint32_t o = 0; o < d.n_odd; ++o, ++k) {
                const uint64_t z = tz[k];
                const double2 u = tu[k];
                const unsigned p0par = parity64(jg0 & z);
                const unsigned zr = (unsigned)(z >> LB);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const unsigned par = p0par ^ (__popc(zr & (unsigned)r) & 1u);
                    go[r] += flip_sign(u.x, par);
                    if (!REALC) goi[r] += flip_sign(u.y, par);
                }
            }
#pragma unroll
            for (int r = 0; r < R; ++r) {
                acc_re += fe[r] * wre[r] + go[r] * wim[r];
                if (!REALC) acc_im += fei[r] * wre[r] + goi[r] * wim[r];
            }
        }
    }
    double v[2] = {acc_re, acc_im};
    grid_reduce<2>(v, partials, counter, out, weight);
}

// ---------------------------------------------------------------------------------------------
// K1t  per-term <psi|P_k|psi> for KC consecutive terms (what getExpectationValueZ returns per
// kernel; RDMGenerator needs every term separately).  All-j form, one accumulator per term per
// thread kept in registers over the whole sweep, one grid reduction at the end.
//   <P_k> = sum_j s_k(j) * (Re w or Im w),  w = conj(psi[j^x_k]) psi[j], sign/phase r_k on host.
// ---------------------------------------------------------------------------------------------
struct TermChunk {
    uint64_t x[16];
    uint64_t z[16];
    uint32_t odd;   // bit k set: nY odd -> use Im w
    uint32_t n;
};

template <int KC>
__global__ void __launch_bounds__(kBlock, 2)
k_expect_terms(const __grid_constant__ Slabs slabs, const double2 *__restrict__ psi, int nl, int rank, TermChunk tc,
               double *partials, unsigned *counter, double *out) {
    const uint64_t nloc = 1ull << nl;
    const uint64_t lmask = nloc - 1;
    const uint64_t rank_bits = (uint64_t)rank << nl;
    double acc[KC];
#pragma unroll
    for (int k = 0; k < KC; ++k) acc[k] = 0.0;
    for (uint64_t j = (uint64_t)blockIdx.x * kBlock + threadIdx.x; j < nloc; j += (uint64_t)gridDim.x * kBlock) {
        const double2 a = psi[j];
        const uint64_t jg = rank_bits | j;
#pragma unroll
        for (int k = 0; k < KC; ++k) {
            if (k < (int)tc.n) {
                const uint64_t x = tc.x[k];
                const double2 b = slabs.p[rank ^ (int)(x >> nl)][(j ^ x) & lmask];
                const double wre = b.x * a.x + b.y * a.y;
                const double wim = b.x * a.y - b.y * a.x;
                const double v = ((tc.odd >> k) & 1u) ? wim : wre;
                acc[k] += flip_sign(v, parity64(jg & tc.z[k]));
            }
        }
    }
    grid_reduce<KC>(acc, partials, counter, out, 1.0);
}

}  // namespace b200vqe
