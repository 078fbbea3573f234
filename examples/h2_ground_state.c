/* Plain-C use of the drop-in boundary (include/b200vqe.h): FermionCompiler text -> JW PauliOperator -> packed
 * (xmask, zmask, coef) -> device Lanczos.  What a reference-side DiagonalizeBackend does through the C ABI
 * (INTEGRATION.md section 2), without XACC.
 *   gcc -std=c99 -Iinclude examples/h2_ground_state.c -Lxacc-vqe_b200 -lb200vqe -Wl,-rpath,$PWD/xacc-vqe_b200 -o h2
 * Without a GPU it stops after the host part with the library's "no CUDA device" error (there is no CPU fallback). */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "b200vqe.h"

/* H2 STO-3G in the reference's kernel-source format (task/tests/DiagonalizeTaskTester.cpp:43-73) */
static const char *kH2 =
    "__qpu__ kernel() {\n"
    "0.7137758743754461\n"
    "-1.252477303982147 0 1 0 0\n"
    "0.337246551663004 0 1 1 1 1 0 0 0\n"
    "0.0906437679061661 0 1 1 1 3 0 2 0\n"
    "0.0906437679061661 0 1 2 1 0 0 2 0\n"
    "0.3317360224302783 0 1 2 1 2 0 0 0\n"
    "0.0906437679061661 0 1 3 1 1 0 2 0\n"
    "0.3317360224302783 0 1 3 1 3 0 0 0\n"
    "0.337246551663004 1 1 0 1 0 0 1 0\n"
    "0.0906437679061661 1 1 0 1 2 0 3 0\n"
    "-1.252477303982147 1 1 1 0\n"
    "0.0906437679061661 1 1 2 1 0 0 3 0\n"
    "0.3317360224302783 1 1 2 1 2 0 1 0\n"
    "0.0906437679061661 1 1 3 1 1 0 3 0\n"
    "0.3317360224302783 1 1 3 1 3 0 1 0\n"
    "0.3317360224302783 2 1 0 1 0 0 2 0\n"
    "0.0906437679061661 2 1 0 1 2 0 0 0\n"
    "0.3317360224302783 2 1 1 1 1 0 2 0\n"
    "0.0906437679061661 2 1 1 1 3 0 0 0\n"
    "-0.4759344611440753 2 1 2 0\n"
    "0.0906437679061661 2 1 3 1 1 0 0 0\n"
    "0.3486989747346679 2 1 3 1 3 0 2 0\n"
    "0.3317360224302783 3 1 0 1 0 0 3 0\n"
    "0.0906437679061661 3 1 0 1 2 0 1 0\n"
    "0.3317360224302783 3 1 1 1 1 0 3 0\n"
    "0.0906437679061661 3 1 1 1 3 0 1 0\n"
    "0.0906437679061661 3 1 2 1 0 0 1 0\n"
    "0.3486989747346679 3 1 2 1 2 0 3 0\n"
    "-0.4759344611440753 3 1 3 0\n"
    "}\n";

int main(void) {
    b200vqe_op op = NULL;
    if (b200vqe_op_from_fermion_source(kH2, "jw", 0, &op)) {
        fprintf(stderr, "front end: %s\n", b200vqe_front_error());
        return 2;
    }
    int64_t T = 0;
    int n = 0;
    b200vqe_op_n_terms(op, &T);
    b200vqe_op_n_qubits(op, &n);
    uint64_t *x = malloc(sizeof(uint64_t) * T), *z = malloc(sizeof(uint64_t) * T);
    double *re = malloc(sizeof(double) * T), *im = malloc(sizeof(double) * T);
    b200vqe_op_pack(op, x, z, re, im);
    printf("H2: %d qubits, %lld Pauli terms\n", n, (long long)T);

    b200vqe_handle h = NULL;
    if (b200vqe_create(n, 0, &h)) {
        printf("device: %s\n", b200vqe_global_error());
        b200vqe_op_destroy(op);
        free(x); free(z); free(re); free(im);
        return 1;
    }
    int rc = b200vqe_set_hamiltonian(h, T, x, z, re, im);
    b200vqe_lanczos_opts o;
    memset(&o, 0, sizeof(o));
    o.tol = 1e-12; o.max_iter = 64; o.seed = 1;
    double e0 = 0.0, res = 0.0;
    int32_t it = 0;
    if (!rc) rc = b200vqe_lanczos_ground(h, &o, &e0, &it, &res);
    if (rc) fprintf(stderr, "backend: %s\n", b200vqe_last_error(h));
    else printf("E0 = %.15f after %d Lanczos steps (reference golden -1.137270422069755)\n", e0, (int)it);
    b200vqe_destroy(h);
    b200vqe_op_destroy(op);
    free(x); free(z); free(re); free(im);
    return rc ? 3 : 0;
}
