/* CPU ORACLE -- TEST INFRASTRUCTURE ONLY (see oracle/pauli_oracle.py header).
 *
 * Plain-C restatement of the reference's Hamiltonian x statevector path for sizes the
 * numpy oracle cannot reach, and the timed CPU baseline of bench.py.  Two flavours:
 *
 *  "faithful"  string-based, single thread, the way the reference does it:
 *              row bit string MSB first, one string walk per term, strtol back
 *              (task/tasks/DiagonalizeTask.cpp:189-203, XACC Term::computeActionOnBra)
 *  "fast"      the same arithmetic on (xmask, zmask, coef*i^nY) tables with OpenMP over
 *              amplitudes -- the strongest CPU form of the path (BASELINE.md section 4, B-fast)
 *
 * Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline, --impl reference) may
 * load this library.  Parity status: pinned through pauli_oracle.py (same goldens); the
 * functions here are cross-checked against it in tests/test_oracle_golden.py.
 *
 * Index convention of the "fast" functions: qubit q <-> bit q of the amplitude index.
 * P|j> = i^nY (-1)^popc(j & z) |j ^ x>,  coef = c * i^nY.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#ifdef _OPENMP
#include <omp.h>
#endif

int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

static inline double sgn_of(uint64_t v) { return (__builtin_popcountll(v) & 1) ? -1.0 : 1.0; }

/* out = H psi ; psi/out interleaved (re,im), N = 2^n */
void oracle_apply(int n, int64_t T, const uint64_t *xm, const uint64_t *zm,
                  const double *cre, const double *cim, const double *psi, double *out) {
    const uint64_t N = 1ull << n;
#pragma omp parallel for schedule(static)
    for (uint64_t i = 0; i < N; ++i) {
        double ar = 0.0, ai = 0.0;
        for (int64_t k = 0; k < T; ++k) {
            const uint64_t j = i ^ xm[k];
            const double s = sgn_of(j & zm[k]);
            const double pr = psi[2 * j], pi = psi[2 * j + 1];
            const double fr = s * cre[k], fi = s * cim[k];
            ar += fr * pr - fi * pi;
            ai += fr * pi + fi * pr;
        }
        out[2 * i] = ar;
        out[2 * i + 1] = ai;
    }
}

/* per[k] = <psi|P_k|psi> (complex, interleaved; includes i^nY so it is real for a
 * Hermitian string); energy[0..1] = sum_k c_k <P_k> with c_k = coef_k / i^nY. */
void oracle_expect(int n, int64_t T, const uint64_t *xm, const uint64_t *zm,
                   const double *cre, const double *cim, const double *psi,
                   double *energy, double *per) {
    const uint64_t N = 1ull << n;
    double er = 0.0, ei = 0.0;
    for (int64_t k = 0; k < T; ++k) {
        const uint64_t x = xm[k], z = zm[k];
        double sr = 0.0, si = 0.0;
#pragma omp parallel for schedule(static) reduction(+ : sr, si)
        for (uint64_t j = 0; j < N; ++j) {
            /* conj(psi_{j^x}) * sgn(j) * psi_j */
            const uint64_t t = j ^ x;
            const double s = sgn_of(j & z);
            const double ar = psi[2 * j], ai = psi[2 * j + 1];
            const double br = psi[2 * t], bi = psi[2 * t + 1];
            sr += s * (br * ar + bi * ai);
            si += s * (br * ai - bi * ar);
        }
        /* i^nY */
        const int ny = __builtin_popcountll(x & z) & 3;
        double pr, pi;
        switch (ny) {
        case 0: pr = sr; pi = si; break;
        case 1: pr = -si; pi = sr; break;
        case 2: pr = -sr; pi = -si; break;
        default: pr = si; pi = -sr; break;
        }
        if (per) { per[2 * k] = pr; per[2 * k + 1] = pi; }
        /* coef * (sr + i si) : coef already carries i^nY */
        er += cre[k] * sr - cim[k] * si;
        ei += cre[k] * si + cim[k] * sr;
    }
    energy[0] = er;
    energy[1] = ei;
}

/* psi <- exp(-i phi P) psi, in place */
void oracle_pauli_rotation(int n, uint64_t x, uint64_t z, double phi, double *psi) {
    const uint64_t N = 1ull << n;
    const double c = cos(phi), s = sin(phi);
    const int ny = __builtin_popcountll(x & z) & 3;
    static const double PR[4] = {1, 0, -1, 0}, PI[4] = {0, 1, 0, -1};
    const double phr = PR[ny], phi_ = PI[ny];
    if (x == 0) {
#pragma omp parallel for schedule(static)
        for (uint64_t j = 0; j < N; ++j) {
            const double sg = sgn_of(j & z);
            /* P psi_j = ph*sg*psi_j ; psi' = c psi - i s P psi */
            const double ar = psi[2 * j], ai = psi[2 * j + 1];
            const double qr = sg * (phr * ar - phi_ * ai), qi = sg * (phr * ai + phi_ * ar);
            psi[2 * j] = c * ar + s * qi;
            psi[2 * j + 1] = c * ai - s * qr;
        }
        return;
    }
    const int hb = 63 - __builtin_clzll(x);
#pragma omp parallel for schedule(static)
    for (uint64_t p = 0; p < N / 2; ++p) {
        const uint64_t lo = p & ((1ull << hb) - 1);
        const uint64_t j = ((p >> hb) << (hb + 1)) | lo;
        const uint64_t t = j ^ x;
        const double ar = psi[2 * j], ai = psi[2 * j + 1];
        const double br = psi[2 * t], bi = psi[2 * t + 1];
        /* (P psi)_j = ph * sgn(t & z) * psi_t ; (P psi)_t = ph * sgn(j & z) * psi_j */
        const double sj = sgn_of(t & z), st = sgn_of(j & z);
        const double qjr = sj * (phr * br - phi_ * bi), qji = sj * (phr * bi + phi_ * br);
        const double qtr = st * (phr * ar - phi_ * ai), qti = st * (phr * ai + phi_ * ar);
        psi[2 * j] = c * ar + s * qji;
        psi[2 * j + 1] = c * ai - s * qjr;
        psi[2 * t] = c * br + s * qti;
        psi[2 * t + 1] = c * bi - s * qtr;
    }
}

/* gate ids shared with include/b200vqe.h: 0 X, 1 Y, 2 Z, 3 H, 4 Rx, 5 Ry, 6 Rz, 7 CNOT(q0=ctrl,q1=tgt) */
void oracle_apply_gate(int n, int gate, int q0, int q1, double theta, double *psi) {
    const uint64_t N = 1ull << n;
    if (gate == 7) {
        const uint64_t cb = 1ull << q0, tb = 1ull << q1;
#pragma omp parallel for schedule(static)
        for (uint64_t j = 0; j < N; ++j) {
            if ((j & cb) && !(j & tb)) {
                const uint64_t t = j | tb;
                const double r = psi[2 * j], i = psi[2 * j + 1];
                psi[2 * j] = psi[2 * t]; psi[2 * j + 1] = psi[2 * t + 1];
                psi[2 * t] = r; psi[2 * t + 1] = i;
            }
        }
        return;
    }
    double m[8] = {0}; /* m00 m01 m10 m11, (re,im) each */
    const double c = cos(theta / 2), s = sin(theta / 2), h = 1.0 / sqrt(2.0);
    switch (gate) {
    case 0: m[2] = 1; m[4] = 1; break;
    case 1: m[3] = -1; m[5] = 1; break;
    case 2: m[0] = 1; m[6] = -1; break;
    case 3: m[0] = h; m[2] = h; m[4] = h; m[6] = -h; break;
    case 4: m[0] = c; m[3] = -s; m[5] = -s; m[6] = c; break;
    case 5: m[0] = c; m[2] = -s; m[4] = s; m[6] = c; break;
    case 6: m[0] = c; m[1] = -s; m[6] = c; m[7] = s; break;
    default: return;
    }
    const uint64_t qb = 1ull << q0;
#pragma omp parallel for schedule(static)
    for (uint64_t p = 0; p < N / 2; ++p) {
        const uint64_t j = ((p >> q0) << (q0 + 1)) | (p & (qb - 1));
        const uint64_t t = j | qb;
        const double ar = psi[2 * j], ai = psi[2 * j + 1], br = psi[2 * t], bi = psi[2 * t + 1];
        psi[2 * j] = m[0] * ar - m[1] * ai + m[2] * br - m[3] * bi;
        psi[2 * j + 1] = m[0] * ai + m[1] * ar + m[2] * bi + m[3] * br;
        psi[2 * t] = m[4] * ar - m[5] * ai + m[6] * br - m[7] * bi;
        psi[2 * t + 1] = m[4] * ai + m[5] * ar + m[6] * bi + m[7] * br;
    }
}

/* ---- faithful: string-based row builder (DiagonalizeTask.cpp:189-203) ----
 * ops: per term a string of n chars from {I,X,Y,Z}, ops[k*n + q] = Pauli on qubit q;
 * c[k] = plain coefficient (no i^nY).  A is dim x dim complex row-major, zeroed here.
 * Single thread on purpose: the reference is single-threaded. */
void oracle_faithful_dense(int n, int64_t T, const char *ops, const double *cre,
                           const double *cim, double *A) {
    const uint64_t dim = 1ull << n;
    memset(A, 0, sizeof(double) * 2 * dim * dim);
    char *row = (char *)malloc((size_t)n + 1), *nb = (char *)malloc((size_t)n + 1);
    row[n] = nb[n] = 0;
    for (uint64_t r = 0; r < dim; ++r) {
        for (int k = n - 1; k >= 0; --k) row[n - 1 - k] = ((r >> k) & 1) ? '1' : '0';
        for (int64_t t = 0; t < T; ++t) {
            memcpy(nb, row, (size_t)n);
            double fr = cre[t], fi = cim[t];
            for (int q = 0; q < n; ++q) {
                const char p = ops[t * n + q];
                if (p == 'Z') {
                    if (nb[q] == '1') { fr = -fr; fi = -fi; }
                } else if (p == 'X') {
                    nb[q] = nb[q] == '1' ? '0' : '1';
                } else if (p == 'Y') {
                    double tr;
                    if (nb[q] == '1') { tr = -fi; fi = fr; fr = tr; nb[q] = '0'; }   /* * i  */
                    else              { tr = fi; fi = -fr; fr = tr; nb[q] = '1'; }   /* * -i */
                }
            }
            const uint64_t col = (uint64_t)strtol(nb, NULL, 2);
            A[2 * (r * dim + col)] += fr;
            A[2 * (r * dim + col) + 1] += fi;
        }
    }
    free(row);
    free(nb);
}

/* faithful H psi through the same string walk, without storing A (for timing at n > 14):
 * out_r = sum_t coeff_t(r) psi_{col_t(r)} */
void oracle_faithful_apply(int n, int64_t T, const char *ops, const double *cre,
                           const double *cim, const double *psi, double *out,
                           uint64_t row_begin, uint64_t row_end) {
    char *row = (char *)malloc((size_t)n + 1), *nb = (char *)malloc((size_t)n + 1);
    row[n] = nb[n] = 0;
    for (uint64_t r = row_begin; r < row_end; ++r) {
        double ar = 0, ai = 0;
        for (int k = n - 1; k >= 0; --k) row[n - 1 - k] = ((r >> k) & 1) ? '1' : '0';
        for (int64_t t = 0; t < T; ++t) {
            memcpy(nb, row, (size_t)n);
            double fr = cre[t], fi = cim[t];
            for (int q = 0; q < n; ++q) {
                const char p = ops[t * n + q];
                if (p == 'Z') {
                    if (nb[q] == '1') { fr = -fr; fi = -fi; }
                } else if (p == 'X') {
                    nb[q] = nb[q] == '1' ? '0' : '1';
                } else if (p == 'Y') {
                    double tr;
                    if (nb[q] == '1') { tr = -fi; fi = fr; fr = tr; nb[q] = '0'; }
                    else              { tr = fi; fi = -fr; fr = tr; nb[q] = '1'; }
                }
            }
            const uint64_t col = (uint64_t)strtol(nb, NULL, 2);
            ar += fr * psi[2 * col] - fi * psi[2 * col + 1];
            ai += fr * psi[2 * col + 1] + fi * psi[2 * col];
        }
        out[2 * (r - row_begin)] = ar;
        out[2 * (r - row_begin) + 1] = ai;
    }
    free(row);
    free(nb);
}
