// K3, fused:  k_rot_tile -- a RUN of consecutive Pauli rotations exp(-i phi_r P_r) applied to the resident state in ONE HBM pass.
//
// One rotation pairs amplitude j with j ^ x_r; applied on its own (k_pauli_rotation, kernels_apply.cuh) it is one read and one
// write of the whole register (32 B per amplitude), and a UCCSD ansatz is thousands of them (ir/algorithms/uccsd/UCCSD.cpp:259-359
// emits them term by term; 2080 at 12 qubits).  Rotations do not commute, but a run of CONSECUTIVE rotations whose flip masks
// span a subspace of dimension <= M of the high index bits acts inside the cosets of that subspace -- the tiles of the tile
// engine (tile_plan.hpp) -- so the host cuts the rotation list into such runs (the eight strings of one double excitation share
// one flip mask: a run holds tens of rotations) and this kernel keeps a tile of 4096 amplitudes in shared memory, applies
// the run's rotations to it in order (each thread 8 pairs {v, v ^ xv}; one barrier between rotations) and writes it back.
// HBM traffic: 32 B per amplitude per RUN instead of per rotation; a rotation costs two LDS.128 + two STS.128 per pair.
//
// Signs:  s(j) = (-1)^{parity(j & z)} splits into parity(rep & z_hi) of the coset and parity(v & zv) of the virtual index, as in
// the expectation engine; the rank bits of z (sharded registers) are a constant sign the host folds into m.
//   pair {v, t = v ^ xv}, bit hb (top bit of xv) of v clear:   a' = c a + m s(t) b,   b' = c b + m s(v) a,   m = -i sin(phi) i^nY
//   xv = 0 (diagonal string):                                 a' = (c + m s(v)) a
#pragma once
#include "kernels_tile.cuh"

namespace b200vqe {

struct RotRec {                       // 48 bytes
    uint32_t zl;                      // z on the high index bits (j >> C)
    uint16_t xv, zv;                  // virtual flip mask / virtual z mask (kTileBits bits)
    uint32_t hb;                      // top bit of xv (pair selector); unused when xv = 0
    uint32_t flip;                    // parity(x & z): s(t) = s(v) (-1)^flip
    double c, mre, mim;               // cos(phi), m = -i sin(phi) i^nY (times the rank-bit sign)
    uint64_t pad;
};
static_assert(sizeof(RotRec) == 48, "RotRec layout");
constexpr int kRotStage = 160;        // rotations of one run staged in shared memory

__global__ void __launch_bounds__(kBlock, 2)
k_rot_tile(double2 *__restrict__ psi, int C, int M, const TileBatch *__restrict__ batch, unsigned tile_count, const RotRec *__restrict__ rots) {
    extern __shared__ __align__(16) unsigned char tile_raw[];
    double2 *tile = reinterpret_cast<double2 *>(tile_raw);
    __shared__ TileBatch sb;
    __shared__ __align__(16) RotRec s_rot[kRotStage];
    const unsigned v0 = threadIdx.x;
    if (threadIdx.x < kTileBatchWords) reinterpret_cast<uint32_t *>(&sb)[threadIdx.x] = reinterpret_cast<const uint32_t *>(batch)[threadIdx.x];
    __syncthreads();
    const int n_rot = sb.r_end - sb.r_begin;
    {
        const int4 *src = reinterpret_cast<const int4 *>(rots + sb.r_begin);
        int4 *dst = reinterpret_cast<int4 *>(s_rot);
        for (int i = threadIdx.x; i < 3 * n_rot; i += kBlock) dst[i] = src[i];
    }
    // element index of (this thread, row 0) relative to the tile's representative; rows = the top four basis vectors
    unsigned e_thread = v0 & ((1u << C) - 1u);
    for (int i = 0; i < 8 - C; ++i) if ((v0 >> (C + i)) & 1u) e_thread ^= sb.basis[i] << C;
    const unsigned g0 = sb.basis[M - 4] << C, g1 = sb.basis[M - 3] << C, g2 = sb.basis[M - 2] << C, g3 = sb.basis[M - 1] << C;
    __syncthreads();
    for (unsigned t = blockIdx.x; t < tile_count; t += gridDim.x) {
        const unsigned rep = tile_rep(sb, t);
        const unsigned e0 = (rep << C) ^ e_thread;
        __syncthreads();                         // the previous tile has been written back by everyone
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            const unsigned e = e0 ^ ((r & 1) ? g0 : 0u) ^ ((r & 2) ? g1 : 0u) ^ ((r & 4) ? g2 : 0u) ^ ((r & 8) ? g3 : 0u);
            tile[((unsigned)r << 8) | v0] = __ldcs(psi + e);
        }
        for (int k = 0; k < n_rot; ++k) {
            __syncthreads();                     // rotation k - 1 (or the fill) is complete
            const RotRec &R = s_rot[k];
            const unsigned xv = R.xv, zv = R.zv;
            const unsigned cpar = (unsigned)__popc(rep & R.zl);
            const double c = R.c, mre = R.mre, mim = R.mim;
            if (xv == 0) {
#pragma unroll 4
                for (int i = 0; i < 16; ++i) {
                    const unsigned v = ((unsigned)i << 8) | v0;
                    const unsigned par = (cpar + (unsigned)__popc(v & zv)) & 1u;
                    const double fr = c + flip_sign(mre, par), fi = flip_sign(mim, par);
                    const double2 a = tile[v];
                    tile[v] = make_double2(fr * a.x - fi * a.y, fr * a.y + fi * a.x);
                }
                continue;
            }
            const unsigned hb = R.hb, lowm = (1u << hb) - 1u, fl = R.flip;
#pragma unroll 4
            for (int i = 0; i < 8; ++i) {
                const unsigned p = ((unsigned)i << 8) | v0;                       // pair index 0 .. 2047
                const unsigned v = ((p >> hb) << (hb + 1)) | (p & lowm);          // bit hb clear
                const unsigned w = v ^ xv;
                const double2 a = tile[v], b = tile[w];
                const unsigned pv = (cpar + (unsigned)__popc(v & zv)) & 1u, pw = pv ^ fl;
                const double tr = flip_sign(mre, pw), ti = flip_sign(mim, pw);
                const double jr = flip_sign(mre, pv), ji = flip_sign(mim, pv);
                tile[v] = make_double2(c * a.x + tr * b.x - ti * b.y, c * a.y + tr * b.y + ti * b.x);
                tile[w] = make_double2(c * b.x + jr * a.x - ji * a.y, c * b.y + jr * a.y + ji * a.x);
            }
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            const unsigned e = e0 ^ ((r & 1) ? g0 : 0u) ^ ((r & 2) ? g1 : 0u) ^ ((r & 4) ? g2 : 0u) ^ ((r & 8) ? g3 : 0u);
            psi[e] = tile[((unsigned)r << 8) | v0];
        }
    }
}

}  // namespace b200vqe
