// Batched small-register path: the optimizer-driven VQE loop (task/tasks/VQEMinimizeTask.hpp:50-53 calling
// ComputeEnergyVQETask::execute per parameter vector) for registers of up to 12 qubits.
//
// k_sweep_small: one CTA per parameter vector.  The whole statevector (padded to 2^12 amplitudes = 64 KB) lives in
// shared memory in its natural order for the entire ansatz: every fused Pauli rotation exp(-i phi P) is one pass over
// the 2048 pairs {j, j^x} (8 per thread) between two __syncthreads, with no kernel launch and no HBM traffic; the
// energy is then evaluated by the tile engine's group loops (kernels_tile.cuh) on the same shared-memory tile through a
// "natural" plan (virtual index == amplitude index).  HBM traffic per energy: the rotation/term tables only.
#pragma once
#include "kernels.cuh"
#include "kernels_tile.cuh"

namespace b200vqe {

struct SweepRot { uint32_t x, z; };    // masks of one rotation (n <= 12)

__global__ void __launch_bounds__(kBlock, 2)
k_sweep_small(int n_rot, const SweepRot *__restrict__ rots, const double2 *__restrict__ cs /* [n_vec][n_rot] (cos phi, sin phi) */,
              unsigned basis_index, const TileBatch *__restrict__ batch, const TileGroup *__restrict__ groups,
              const TileTerm *__restrict__ terms, double ident, double *__restrict__ energies) {
    extern __shared__ __align__(16) unsigned char tile_raw[];
    double2 *tile = reinterpret_cast<double2 *>(tile_raw);
    __shared__ double s_red[kBlock / 32];
    const unsigned v0 = threadIdx.x;
    const double2 *my_cs = cs + (size_t)blockIdx.x * n_rot;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const unsigned j = v0 + ((unsigned)k << 8);
        tile[j] = make_double2(j == basis_index ? 1.0 : 0.0, 0.0);
    }
    __syncthreads();
    for (int r = 0; r < n_rot; ++r) {
        const double2 c_s = __ldg(&my_cs[r]);
        if (c_s.y == 0.0 && c_s.x == 1.0) continue;           // phi = 0 (uniform over the CTA)
        const SweepRot rot = rots[r];
        // psi <- cos(phi) psi - i sin(phi) P psi,  m = -i sin(phi) i^nY
        const int ny = __popc(rot.x & rot.z) & 3;
        const double mre = (ny == 1) ? c_s.y : (ny == 3 ? -c_s.y : 0.0);
        const double mim = (ny == 0) ? -c_s.y : (ny == 2 ? c_s.y : 0.0);
        if (rot.x == 0) {
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                const unsigned j = v0 + ((unsigned)k << 8);
                const unsigned par = __popc(j & rot.z) & 1u;
                const double fr = c_s.x + flip_sign(mre, par), fi = flip_sign(mim, par);
                const double2 a = tile[j];
                tile[j] = make_double2(fr * a.x - fi * a.y, fr * a.y + fi * a.x);
            }
        } else {
            const int h = 31 - __clz(rot.x);
            const unsigned lowm = (1u << h) - 1u;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const unsigned p = v0 + ((unsigned)k << 8);
                const unsigned j = ((p >> h) << (h + 1)) | (p & lowm);
                const unsigned t = j ^ rot.x;
                const double2 a = tile[j], b = tile[t];
                const unsigned pj = __popc(j & rot.z) & 1u, pt = __popc(t & rot.z) & 1u;
                const double tr = flip_sign(mre, pt), ti = flip_sign(mim, pt);
                const double jr = flip_sign(mre, pj), ji = flip_sign(mim, pj);
                tile[j] = make_double2(c_s.x * a.x + tr * b.x - ti * b.y, c_s.x * a.y + tr * b.y + ti * b.x);
                tile[t] = make_double2(c_s.x * b.x + jr * a.x - ji * a.y, c_s.x * b.y + jr * a.y + ji * a.x);
            }
        }
        __syncthreads();
    }
    // <psi|H|psi> on the resident tile (natural plan: one batch, coset representative 0)
    double2 a[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) a[k] = tile[((unsigned)k << 8) | v0];
    double acc = 0.0;
    const int g0 = batch->g_begin;
    const int4 *gp = reinterpret_cast<const int4 *>(groups) + g0;
    const int n_g = batch->g_end - g0;
    const int c0 = batch->cls_end[0], c1 = batch->cls_end[1], c2 = batch->cls_end[2], c3 = batch->cls_end[3];
    tile_class<0, false>(a, tile, v0, 0u, gp, 0, c0, terms, acc);
    tile_class<1, false>(a, tile, v0, 0u, gp, c0, c1, terms, acc);
    tile_class<2, false>(a, tile, v0, 0u, gp, c1, c2, terms, acc);
    tile_class<3, false>(a, tile, v0, 0u, gp, c2, c3, terms, acc);
    tile_class<3, true>(a, tile, v0, 0u, gp, c3, n_g, terms, acc);
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < kBlock / 32; ++w) s += s_red[w];
        energies[blockIdx.x] = ident + s;
    }
}

}  // namespace b200vqe
