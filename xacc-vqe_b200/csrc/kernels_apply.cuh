// K2 (H psi), K4 (Lanczos vector updates), K3 (gates / fused Pauli rotations), state initialisers.
#pragma once
#include "kernels.cuh"

namespace b200vqe {

// ---------------------------------------------------------------------------------------------
// K2  dst = in_scale * H src  [- beta_prev * dst_old]   and  alpha = Re <in_scale*src | dst>
//
// Tile-stationary over OUTPUTS: a thread owns R output amplitudes and walks every xmask group,
// gathering src[j ^ x] (one 16-byte load per amplitude per group = the 16*N*G byte model) and
// multiplying by f(t) = sum_k coef_k (-1)^popc(t & z_k), t = j ^ x.  One write of dst at the end.
// REALC (Hermitian H): coef_k is real for even nY and purely imaginary for odd nY, so
// f = A + iB with A from the even block and B from the odd block; tv[k].x holds that real number.
// Scalars come from device memory so a Lanczos step needs no host round trip:
//   in_scale = 1 / *inv_scale_src (if non-null),  beta_prev = *beta_ptr (if non-null).
// ---------------------------------------------------------------------------------------------
template <int R, bool REALC>
__global__ void __launch_bounds__(kBlock, 2)
k_apply(const __grid_constant__ Slabs slabs, const double2 *__restrict__ src_local, double2 *__restrict__ dst, int nl, int rank,
        const GroupDesc *__restrict__ gd, int n_groups, const uint64_t *__restrict__ tz,
        const double2 *__restrict__ tv, uint64_t n_tiles, const double *norm_src, const double *norm_prev,
        int fuse_prev, double *partials, unsigned *counter, double *alpha_out) {
    constexpr int TS = kBlock * R;
    constexpr int LB = 8;
    const uint64_t nloc = 1ull << nl;
    const uint64_t lmask = nloc - 1;
    const uint64_t rank_bits = (uint64_t)rank << nl;
    const double in_scale = norm_src ? 1.0 / *norm_src : 1.0;
    // w' = H v_j - beta_j v_{j-1}, v_{j-1} = W_{j-1}/beta_{j-1}, beta_j = *norm_src, beta_{j-1} = *norm_prev
    const double prev_scale = fuse_prev ? (*norm_src) / (*norm_prev) : 0.0;
    double alpha = 0.0;

    for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const uint64_t j0 = tile * TS + threadIdx.x;
        const uint64_t jg0 = rank_bits | j0;
        double2 acc[R];
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = make_double2(0.0, 0.0);
        for (int g = 0; g < n_groups; ++g) {
            const GroupDesc d = gd[g];
            const uint64_t xl = d.x & lmask;
            const double2 *__restrict__ pb = slabs.p[rank ^ (int)(d.x >> nl)];
            double2 b[R];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const uint64_t j = j0 + ((uint64_t)r << LB);
                b[r] = (j < nloc) ? pb[j ^ xl] : make_double2(0.0, 0.0);
            }
            const uint64_t tg0 = jg0 ^ d.x;   // partner (global) index of r = 0
            double fa[R], fb[R];
#pragma unroll
            for (int r = 0; r < R; ++r) { fa[r] = 0.0; fb[r] = 0.0; }
            uint32_t k = d.kb;
            for (uint32_t e = 0; e < d.n_even; ++e, ++k) {
                const uint64_t z = tz[k];
                const double2 c = tv[k];
                const unsigned p0 = parity64(tg0 & z);
                const unsigned zr = (unsigned)(z >> LB);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const unsigned par = p0 ^ (__popc(zr & (unsigned)r) & 1u);
                    fa[r] += flip_sign(c.x, par);
                    if (!REALC) fb[r] += flip_sign(c.y, par);
                }
            }
            for (uint32_t o = 0; o < d.n_odd; ++o, ++k) {
                const uint64_t z = tz[k];
                const double2 c = tv[k];
                const unsigned p0 = parity64(tg0 & z);
                const unsigned zr = (unsigned)(z >> LB);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const unsigned par = p0 ^ (__popc(zr & (unsigned)r) & 1u);
                    if (REALC) {
                        fb[r] += flip_sign(c.x, par);
                    } else {
                        fa[r] += flip_sign(c.x, par);
                        fb[r] += flip_sign(c.y, par);
                    }
                }
            }
#pragma unroll
            for (int r = 0; r < R; ++r) {
                acc[r].x += fa[r] * b[r].x - fb[r] * b[r].y;
                acc[r].y += fa[r] * b[r].y + fb[r] * b[r].x;
            }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const uint64_t j = j0 + ((uint64_t)r << LB);
            if (j < nloc) {
                double2 o = make_double2(acc[r].x * in_scale, acc[r].y * in_scale);
                if (fuse_prev) {
                    const double2 p = dst[j];
                    o.x -= prev_scale * p.x;
                    o.y -= prev_scale * p.y;
                }
                dst[j] = o;
                const double2 s = src_local[j];
                alpha += in_scale * (s.x * o.x + s.y * o.y);
            }
        }
    }
    if (alpha_out) {
        double v[1] = {alpha};
        grid_reduce<1>(v, partials, counter, alpha_out, 1.0);
    }
}

// <bra|ket> (complex) -- used for alpha of the public apply_h and for state_inner / norm2.
__global__ void __launch_bounds__(kBlock, 4)
k_inner(const double2 *__restrict__ bra, const double2 *__restrict__ ket, uint64_t n,
        double *partials, unsigned *counter, double *out) {
    double re = 0.0, im = 0.0;
    for (uint64_t j = (uint64_t)blockIdx.x * kBlock + threadIdx.x; j < n; j += (uint64_t)gridDim.x * kBlock) {
        const double2 a = bra[j], b = ket[j];
        re += a.x * b.x + a.y * b.y;
        im += a.x * b.y - a.y * b.x;
    }
    double v[2] = {re, im};
    grid_reduce<2>(v, partials, counter, out, 1.0);
}

// alpha = in_scale * Re <src|dst>  (the tile-engine H psi leaves alpha to this pass; in_scale = 1 / *norm_src)
__global__ void __launch_bounds__(kBlock, 4)
k_alpha(const double2 *__restrict__ src, const double2 *__restrict__ dst, uint64_t n, const double *norm_src,
        double *partials, unsigned *counter, double *out) {
    const double in_scale = norm_src ? 1.0 / *norm_src : 1.0;
    double re = 0.0;
    for (uint64_t j = (uint64_t)blockIdx.x * kBlock + threadIdx.x; j < n; j += (uint64_t)gridDim.x * kBlock) {
        const double2 a = src[j], b = dst[j];
        re += a.x * b.x + a.y * b.y;
    }
    double v[1] = {re};
    grid_reduce<1>(v, partials, counter, out, in_scale);
}

// K4  W_{j+1} = w' - (alpha_j / beta_j) W_j ;  out[0] = ||W_{j+1}||^2 (sqrt taken by k_sqrt)
//     optional Ritz accumulation  y += (s_j / beta_j) W_j
__global__ void __launch_bounds__(kBlock, 4)
k_lanczos_update(double2 *__restrict__ w, const double2 *__restrict__ v, double2 *__restrict__ y, uint64_t n,
                 const double *alpha, const double *norm_v, const double *ritz_coef,
                 double *partials, unsigned *counter, double *out, bool take_sqrt) {
    const double inv = 1.0 / *norm_v;
    const double c = *alpha * inv;
    const double yc = ritz_coef ? *ritz_coef * inv : 0.0;
    double nrm = 0.0;
    for (uint64_t j = (uint64_t)blockIdx.x * kBlock + threadIdx.x; j < n; j += (uint64_t)gridDim.x * kBlock) {
        const double2 vv = v[j];
        double2 ww = w[j];
        ww.x -= c * vv.x;
        ww.y -= c * vv.y;
        w[j] = ww;
        nrm += ww.x * ww.x + ww.y * ww.y;
        if (y) {
            double2 yy = y[j];
            yy.x += yc * vv.x;
            yy.y += yc * vv.y;
            y[j] = yy;
        }
    }
    double r[1] = {nrm};
    grid_reduce<1>(r, partials, counter, out, 1.0, take_sqrt);
}

__global__ void k_sqrt_inplace(double *v) { *v = sqrt(*v); }

// out[k] = full[bit-reversal over n bits of (first + k)]: the register in the reference's index order (qubit q <-> bit n-1-q,
// task/tasks/DiagonalizeTask.cpp:189-193); `full` = the whole register in device order (sharded: the gathered slabs)
__global__ void __launch_bounds__(kBlock, 4)
k_extract_reference(const double2 *__restrict__ full, double2 *__restrict__ out, int n, uint64_t first, uint64_t count) {
    for (uint64_t k = (uint64_t)blockIdx.x * kBlock + threadIdx.x; k < count; k += (uint64_t)gridDim.x * kBlock)
        out[k] = full[__brevll(first + k) >> (64 - n)];
}

__global__ void __launch_bounds__(kBlock, 4)
k_scale(double2 *__restrict__ v, uint64_t n, double s) {
    for (uint64_t j = (uint64_t)blockIdx.x * kBlock + threadIdx.x; j < n; j += (uint64_t)gridDim.x * kBlock) {
        double2 a = v[j];
        a.x *= s;
        a.y *= s;
        v[j] = a;
    }
}

__global__ void __launch_bounds__(kBlock, 4)
k_axpy(double2 *__restrict__ y, const double2 *__restrict__ x, uint64_t n, double s) {
    for (uint64_t j = (uint64_t)blockIdx.x * kBlock + threadIdx.x; j < n; j += (uint64_t)gridDim.x * kBlock) {
        double2 a = y[j];
        const double2 b = x[j];
        a.x += s * b.x;
        a.y += s * b.y;
        y[j] = a;
    }
}

// ---------------------------------------------------------------------------------------------
// K3  gates on the resident state (32 bytes per amplitude: one read + one write)
// ---------------------------------------------------------------------------------------------
// fused Pauli rotation  psi <- cos(phi) psi - i sin(phi) P psi.   m = -i sin(phi) i^nY.
// pair {j, t = j^x} with bit h of j clear:  a' = c a + m s(t) b ,  b' = c b + m s(j) a
__global__ void __launch_bounds__(kBlock, 4)
k_pauli_rotation(double2 *__restrict__ psi, int nl, int rank, uint64_t x, uint64_t z, int h,
                 double c, double mre, double mim) {
    const uint64_t nloc = 1ull << nl;
    const uint64_t rank_bits = (uint64_t)rank << nl;
    if (x == 0) {
        for (uint64_t j = (uint64_t)blockIdx.x * kBlock + threadIdx.x; j < nloc; j += (uint64_t)gridDim.x * kBlock) {
            const unsigned par = parity64((rank_bits | j) & z);
            const double fr = c + flip_sign(mre, par), fi = flip_sign(mim, par);
            const double2 a = psi[j];
            psi[j] = make_double2(fr * a.x - fi * a.y, fr * a.y + fi * a.x);
        }
        return;
    }
    const uint64_t half = nloc >> 1;
    const uint64_t lowm = (1ull << h) - 1;
    for (uint64_t p = (uint64_t)blockIdx.x * kBlock + threadIdx.x; p < half; p += (uint64_t)gridDim.x * kBlock) {
        const uint64_t j = ((p >> h) << (h + 1)) | (p & lowm);
        const uint64_t t = j ^ x;
        const double2 a = psi[j], b = psi[t];
        const unsigned pj = parity64((rank_bits | j) & z), pt = parity64((rank_bits | t) & z);
        const double tr = flip_sign(mre, pt), ti = flip_sign(mim, pt);
        const double jr = flip_sign(mre, pj), ji = flip_sign(mim, pj);
        psi[j] = make_double2(c * a.x + tr * b.x - ti * b.y, c * a.y + tr * b.y + ti * b.x);
        psi[t] = make_double2(c * b.x + jr * a.x - ji * a.y, c * b.y + jr * a.y + ji * a.x);
    }
}

struct Mat2 { double m[8]; };   // m00 m01 m10 m11, (re, im) each

__global__ void __launch_bounds__(kBlock, 4)
k_gate1(double2 *__restrict__ psi, uint64_t nloc, int q, Mat2 g) {
    const uint64_t half = nloc >> 1;
    const uint64_t qb = 1ull << q;
    for (uint64_t p = (uint64_t)blockIdx.x * kBlock + threadIdx.x; p < half; p += (uint64_t)gridDim.x * kBlock) {
        const uint64_t j = ((p >> q) << (q + 1)) | (p & (qb - 1));
        const uint64_t t = j | qb;
        const double2 a = psi[j], b = psi[t];
        psi[j] = make_double2(g.m[0] * a.x - g.m[1] * a.y + g.m[2] * b.x - g.m[3] * b.y,
                              g.m[0] * a.y + g.m[1] * a.x + g.m[2] * b.y + g.m[3] * b.x);
        psi[t] = make_double2(g.m[4] * a.x - g.m[5] * a.y + g.m[6] * b.x - g.m[7] * b.y,
                              g.m[4] * a.y + g.m[5] * a.x + g.m[6] * b.y + g.m[7] * b.x);
    }
}

__global__ void __launch_bounds__(kBlock, 4)
k_cnot(double2 *__restrict__ psi, uint64_t nloc, int ctrl, int tgt) {
    const uint64_t quarter = nloc >> 2;
    const int lo = ctrl < tgt ? ctrl : tgt, hi = ctrl < tgt ? tgt : ctrl;
    for (uint64_t p = (uint64_t)blockIdx.x * kBlock + threadIdx.x; p < quarter; p += (uint64_t)gridDim.x * kBlock) {
        uint64_t j = ((p >> lo) << (lo + 1)) | (p & ((1ull << lo) - 1));
        j = ((j >> hi) << (hi + 1)) | (j & ((1ull << hi) - 1));
        j |= 1ull << ctrl;
        const uint64_t t = j | (1ull << tgt);
        const double2 a = psi[j];
        psi[j] = psi[t];
        psi[t] = a;
    }
}

// ---------------------------------------------------------------------------------------------
// K3 on sharded qubits.  The flip reaches the shard bits, so every partner of this rank's slab lives in ONE
// peer's slab (rank ^ x_rank); the host exchanges the two slabs (ncclSend/ncclRecv, NVLink) into `peer`
// and each rank updates its own amplitudes: 48 bytes per amplitude (own read + write, partner read).
// ---------------------------------------------------------------------------------------------
// psi[j] <- c psi[j] + m s(t) peer[j ^ xl],  t = (peer_rank << nl) | (j ^ xl)
__global__ void __launch_bounds__(kBlock, 4)
k_pauli_rotation_remote(double2 *__restrict__ psi, const double2 *__restrict__ peer, int nl, int peer_rank, uint64_t xl,
                        uint64_t z, double c, double mre, double mim) {
    const uint64_t nloc = 1ull << nl;
    const uint64_t peer_bits = (uint64_t)peer_rank << nl;
    for (uint64_t j = (uint64_t)blockIdx.x * kBlock + threadIdx.x; j < nloc; j += (uint64_t)gridDim.x * kBlock) {
        const uint64_t t = j ^ xl;
        const double2 a = psi[j], b = __ldcs(peer + t);
        const unsigned pt = parity64((peer_bits | t) & z);
        const double tr = flip_sign(mre, pt), ti = flip_sign(mim, pt);
        psi[j] = make_double2(c * a.x + tr * b.x - ti * b.y, c * a.y + tr * b.y + ti * b.x);
    }
}

// one-qubit gate on a shard bit: psi[j] <- m_own psi[j] + m_peer peer[j]   (row of the 2x2 matrix chosen by this rank's bit)
__global__ void __launch_bounds__(kBlock, 4)
k_gate1_remote(double2 *__restrict__ psi, const double2 *__restrict__ peer, uint64_t nloc, double ore, double oim, double pre, double pim) {
    for (uint64_t j = (uint64_t)blockIdx.x * kBlock + threadIdx.x; j < nloc; j += (uint64_t)gridDim.x * kBlock) {
        const double2 a = psi[j], b = __ldcs(peer + j);
        psi[j] = make_double2(ore * a.x - oim * a.y + pre * b.x - pim * b.y, ore * a.y + oim * a.x + pre * b.y + pim * b.x);
    }
}

// diagonal one-qubit gate on a shard bit: the whole slab takes one phase
__global__ void __launch_bounds__(kBlock, 4)
k_scale_complex(double2 *__restrict__ v, uint64_t n, double re, double im) {
    for (uint64_t j = (uint64_t)blockIdx.x * kBlock + threadIdx.x; j < n; j += (uint64_t)gridDim.x * kBlock) {
        const double2 a = v[j];
        v[j] = make_double2(re * a.x - im * a.y, re * a.y + im * a.x);
    }
}

// CNOT whose target is a shard bit and whose control is local: amplitudes with the control bit set come from the peer
__global__ void __launch_bounds__(kBlock, 4)
k_take_where_bit(double2 *__restrict__ psi, const double2 *__restrict__ peer, uint64_t nloc, int ctrl) {
    const uint64_t half = nloc >> 1;
    const uint64_t cb = 1ull << ctrl;
    for (uint64_t p = (uint64_t)blockIdx.x * kBlock + threadIdx.x; p < half; p += (uint64_t)gridDim.x * kBlock) {
        const uint64_t j = ((p >> ctrl) << (ctrl + 1)) | (p & (cb - 1)) | cb;
        psi[j] = __ldcs(peer + j);
    }
}

// ---------------------------------------------------------------------------------------------
// Particle-number sector (diag-number-symmetry, task/tasks/DiagonalizeTask.cpp:54-183): zero every amplitude whose
// occupation vector x_i = parity(rows[i] & j) does not hold exactly ne ones.  JW: rows are unit vectors, x = j.
// ---------------------------------------------------------------------------------------------
struct SectorRows { uint64_t r[40]; };

template <bool UNIT>
__global__ void __launch_bounds__(kBlock, 4)
k_project_sector(double2 *__restrict__ psi, uint64_t nloc, uint64_t first, int n, int ne, const __grid_constant__ SectorRows rows) {
    for (uint64_t j = (uint64_t)blockIdx.x * kBlock + threadIdx.x; j < nloc; j += (uint64_t)gridDim.x * kBlock) {
        const uint64_t g = first + j;
        int occ;
        if (UNIT) {
            occ = __popcll(g);
        } else {
            occ = 0;
            for (int i = 0; i < n; ++i) occ += (int)parity64(rows.r[i] & g);
        }
        if (occ != ne) psi[j] = make_double2(0.0, 0.0);
    }
}

// ---------------------------------------------------------------------------------------------
// state initialisers
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock, 4)
k_set_basis(double2 *__restrict__ psi, uint64_t nloc, uint64_t first, uint64_t basis) {
    for (uint64_t j = (uint64_t)blockIdx.x * kBlock + threadIdx.x; j < nloc; j += (uint64_t)gridDim.x * kBlock)
        psi[j] = make_double2((first + j) == basis ? 1.0 : 0.0, 0.0);
}

__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

// complex Gaussian by Box-Muller on a counter-based hash of (seed, global index); not normalised
__global__ void __launch_bounds__(kBlock, 4)
k_random(double2 *__restrict__ psi, uint64_t nloc, uint64_t first, uint64_t seed) {
    for (uint64_t j = (uint64_t)blockIdx.x * kBlock + threadIdx.x; j < nloc; j += (uint64_t)gridDim.x * kBlock) {
        const uint64_t h1 = splitmix64(seed ^ splitmix64(2 * (first + j)));
        const uint64_t h2 = splitmix64(seed ^ splitmix64(2 * (first + j) + 1));
        const double u1 = ((double)(h1 >> 11) + 1.0) * (1.0 / 9007199254740993.0);
        const double u2 = (double)(h2 >> 11) * (1.0 / 9007199254740992.0);
        const double r = sqrt(-2.0 * log(u1));
        double s, c;
        sincospi(2.0 * u2, &s, &c);
        psi[j] = make_double2(r * c, r * s);
    }
}

struct ProductAngles { double c[40]; double sre[40]; double sim[40]; };   // amp0 = c, amp1 = (sre, sim)

__global__ void __launch_bounds__(kBlock, 4)
k_product(double2 *__restrict__ psi, uint64_t nloc, uint64_t first, int n, ProductAngles pa) {
    for (uint64_t j = (uint64_t)blockIdx.x * kBlock + threadIdx.x; j < nloc; j += (uint64_t)gridDim.x * kBlock) {
        const uint64_t g = first + j;
        double re = 1.0, im = 0.0;
        for (int q = 0; q < n; ++q) {
            if ((g >> q) & 1) {
                const double nr = re * pa.sre[q] - im * pa.sim[q];
                im = re * pa.sim[q] + im * pa.sre[q];
                re = nr;
            } else {
                re *= pa.c[q];
                im *= pa.c[q];
            }
        }
        psi[j] = make_double2(re, im);
    }
}

}  // namespace b200vqe
