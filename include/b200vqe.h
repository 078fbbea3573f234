/* b200vqe.h -- C ABI of the B200-native Pauli-Hamiltonian x statevector backend.
 *
 * Drop-in boundary for the hot path of ORNL-QCI/xacc-vqe.  Every entry
 * point takes plain pointers and sizes; the caller owns all host arrays; the handle owns device
 * memory.  All functions return 0 on success, a negative B200VQE_E* code otherwise;
 * b200vqe_last_error() gives the message (the reference's convention is xacc::error(msg),
 * e.g. task/VQEProgram.hpp:39-41 -- a wrapper turns the code into that call).
 *
 * Conventions
 *   - qubit q  <->  bit q of masks and of the amplitude index ("device order").  The
 *     reference's dense-matrix order (qubit q <-> bit n-1-q, task/tasks/DiagonalizeTask.cpp:189-193)
 *     is available on download with B200VQE_ORDER_REFERENCE.
 *   - a packed term is (xmask, zmask, coef) with xmask = X|Y sites, zmask = Z|Y sites,
 *     coef = c * i^nY, so that  P|j> = i^nY (-1)^popc(j & z) |j ^ x>.
 *   - amplitudes are interleaved (re, im) IEEE doubles.
 *   - one handle = one host thread = one GPU (one shard when created sharded).
 */
#ifndef B200VQE_H_INCLUDED
#define B200VQE_H_INCLUDED

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200VQE_OK 0
#define B200VQE_EINVAL (-1)
#define B200VQE_ECUDA (-2)
#define B200VQE_ENOMEM (-3)
#define B200VQE_ECOMM (-4)
#define B200VQE_EPARSE (-5)
#define B200VQE_ENOCONV (-6)

#define B200VQE_ORDER_DEVICE 0
#define B200VQE_ORDER_REFERENCE 1

/* gate ids: names as in ir/algorithms/uccsd/UCCSD.cpp:282-365 and ir/algorithms/hwe/HWE.cpp:91-123;
 * R_a(t) = exp(-i t sigma_a / 2); CNOT: q0 = control, q1 = target. */
enum { B200VQE_X = 0, B200VQE_Y = 1, B200VQE_Z = 2, B200VQE_H = 3, B200VQE_RX = 4,
       B200VQE_RY = 5, B200VQE_RZ = 6, B200VQE_CNOT = 7 };

typedef struct { int32_t gate; int32_t q0; int32_t q1; int32_t pad; double theta; } b200vqe_gate;

typedef struct b200vqe_ctx *b200vqe_handle;   /* device backend (one GPU / one shard) */
typedef struct b200vqe_opr *b200vqe_op;       /* host-side PauliOperator */

/* ------------------------------------------------------------------ host front end
 * Replaces what the reference gets from XACC's PauliOperator and its FermionCompiler
 * (compiler/FermionCompiler.cpp:62-147, transformations/jw/JordanWignerIRTransformation.cpp:8-60,
 *  transformations/bk/BravyiKitaevIRTransformation.cpp:9-79). */
const char *b200vqe_global_error(void);   /* message of a failed b200vqe_create* / comm call */
const char *b200vqe_front_error(void);    /* message of a failed front-end (b200vqe_op_*, b200vqe_uccsd_*) call */
int b200vqe_op_from_fermion_source(const char *src, const char *transformation /* "jw" | "bk" */,
                                   int n_qubits /* 0 = max site + 1 */, b200vqe_op *out);
int b200vqe_op_from_string(const char *text /* "(re,im) X0 Z1 + ..." */, b200vqe_op *out);
int b200vqe_op_from_terms(int64_t n_terms, const uint64_t *xmask, const uint64_t *zmask,
                          const double *c_re, const double *c_im /* plain c, not c*i^nY */,
                          b200vqe_op *out);
int b200vqe_op_destroy(b200vqe_op op);
int b200vqe_op_n_terms(b200vqe_op op, int64_t *n_terms);
int b200vqe_op_n_qubits(b200vqe_op op, int *n_qubits);
/* Packing: terms sorted by (xmask, zmask); coef = c * i^nY. */
int b200vqe_op_pack(b200vqe_op op, uint64_t *xmask, uint64_t *zmask, double *coef_re, double *coef_im);
/* Two-call string getters: returns needed size (incl. NUL) in *needed when cap is too small. */
int b200vqe_op_to_string(b200vqe_op op, char *buf, int64_t cap, int64_t *needed);
int b200vqe_op_term_id(b200vqe_op op, int64_t k /* packed order */, char *buf, int64_t cap, int64_t *needed);
int b200vqe_op_add(b200vqe_op a, b200vqe_op b, b200vqe_op *out);
int b200vqe_op_mul(b200vqe_op a, b200vqe_op b, b200vqe_op *out);
int b200vqe_op_scale(b200vqe_op a, double re, double im, b200vqe_op *out);

/* UCCSD ansatz of ir/algorithms/uccsd/UCCSD.cpp:96-365 as fused Pauli rotations:
 * rotation r is exp(-i * imag_coeff[r] * theta[param[r]] * P_r), applied in order r = 0..count-1
 * to |1..1 0..0> (X on qubits 0..n_electrons-1, UCCSD.cpp:361-365).  Call with NULL arrays to
 * get the counts. */
int b200vqe_uccsd_rotations(int n_qubits, int n_electrons, const char *transformation,
                            int64_t *count, int *n_params, uint64_t *xmask, uint64_t *zmask,
                            double *imag_coeff, int32_t *param);
/* The same circuit as the reference's literal gate list evaluated at theta (UCCSD.cpp:259-365). */
int b200vqe_uccsd_gates(int n_qubits, int n_electrons, const char *transformation,
                        const double *theta, int n_theta, int64_t *count, b200vqe_gate *gates);

/* Occupation read-out of the diag-number-symmetry branch (task/tasks/DiagonalizeTask.cpp:54-144): rows[i] (n_qubits
 * entries) such that occupation bit i of basis index j is parity(rows[i] & j).  "jw": unit vectors; "bk": the inverse
 * over GF(2) of the matrix the reference builds at DiagonalizeTask.cpp:76-121. */
int b200vqe_number_sector_rows(int n_qubits, const char *transformation, uint64_t *rows);

/* ------------------------------------------------------------------ device backend */
int b200vqe_create(int n_qubits, int device, b200vqe_handle *out);
/* Sharded: this process holds amplitudes whose top log2(world) index bits equal rank.
 * Collectives go through NCCL; the caller distributes the 128-byte id from rank 0. */
int b200vqe_comm_unique_id(char id[128]);
int b200vqe_create_sharded(int n_qubits, int device, int rank, int world, const char id[128],
                           b200vqe_handle *out);
int b200vqe_destroy(b200vqe_handle h);
const char *b200vqe_last_error(b200vqe_handle h);
/* String options in the manner of xacc::setOption(key, value) (e.g. task/tasks/DiagonalizeTask.cpp:15-30).
 *   "expect-engine"   = "auto" | "tile" | "stream"   which <psi|H|psi> kernel set_hamiltonian prepares
 *   "apply-engine"    = "auto" | "tile" | "stream"   the same for H psi (b200vqe_apply_h, b200vqe_lanczos_ground)
 *   "shard-exchange"  = "auto" | "allgather" | "needed" | "pairwise"   how the tile H psi of a sharded register fetches peer
 *                                                    slabs: one ncclAllGather into `world` staging slabs (auto; fastest measured) |
 *                                                    only the slabs the Hamiltonian's masks reach, one group of ncclSend/ncclRecv |
 *                                                    per-x_rank send/recv pairs pipelined through two staging slabs (least memory)
 *   "expect-kernel"   = "auto" | "pipe" | "classic"  <psi|H|psi> on 4096-amplitude tiles: the pipelined kernel (3-stage cp.async ring,
 *                                                    views, fused view blocks; default) | the single-buffer kernel of round 1
 *   "tile-bits"       = "12" | "11"                  tile of the <psi|H|psi> engine: 4096 amplitudes | 2048 amplitudes double-buffered
 *                                                    with cp.async (three CTAs per SM; k_tile_expect_db)
 *   "tile-chunk-bits" = "2".."8"                     contiguous amplitudes per HBM chunk of the tile engine (default 3: 128 bytes);
 *   "apply-chunk-bits"                               the same for the H psi plans alone (default 4: 256 bytes)
 *   "tile-planner"    = "3" | "2" | "1"              tile planner: 2 + several register subspaces (views) per batch and fused view
 *                                                    blocks (default; walked by the pipelined kernel only) | collision-greedy
 *                                                    subspaces + register subspace spanned by member masks | the first version
 *   "fuse-rotations"  = "on" | "off"                 b200vqe_apply_pauli_rotations: consecutive rotations with local flip masks run as
 *                                                    fused tile passes (one HBM pass per run) | one launch per rotation
 *   "peer-copy"       = "on" | "off"                 sharded <psi|H|psi>: slab swaps as peer-memory copies (CUDA IPC + copy engine,
 *                                                    no SM) | grouped ncclSend/ncclRecv; set before the first sharded expectation
 * Options take effect at the next b200vqe_set_hamiltonian. */
int b200vqe_set_option(b200vqe_handle h, const char *key, const char *value);
int b200vqe_n_slots(b200vqe_handle h, int n_slots);   /* number of resident statevectors (default 1) */
int b200vqe_local_amplitudes(b200vqe_handle h, uint64_t *count, uint64_t *first_index);

/* Hamiltonian tables (replaces the per-term kernels built at task/VQEProgram.hpp:204-220).
 * The identity term, if present, is handled as the reference does (added directly,
 * task/tasks/ComputeEnergyVQETask.cpp:90-106). */
int b200vqe_set_hamiltonian(b200vqe_handle h, int64_t n_terms, const uint64_t *xmask,
                            const uint64_t *zmask, const double *coef_re, const double *coef_im);
int b200vqe_hamiltonian_stats(b200vqe_handle h, int64_t *n_terms, int64_t *n_groups);
/* Tile-engine plan of the handle's Hamiltonian: out[0] batches (= HBM passes over this rank's amplitudes) of <psi|H|psi>,
 * out[1] xmask groups, out[2] batches of H psi (0 until the first H psi plans them), out[3] x_rank classes of a sharded
 * <psi|H|psi>, out[4] chunk bits, out[5] tile bits. */
int b200vqe_plan_stats(b200vqe_handle h, int64_t out[6]);
/* Per-tile cost of that <psi|H|psi> plan, summed over its batches: out[0] register loads (one per view of a batch), out[1] fused view
 * blocks (four xmask groups evaluated from registers alone), out[2] plain records, out[3] plain records that read partner rows from
 * shared memory. */
int b200vqe_plan_costs(b200vqe_handle h, int64_t out[4]);
/* NVLink exchange of the last sharded b200vqe_expectation / b200vqe_apply_h / Lanczos step: CUDA-event time [ms] of the transfers on
 * the communication stream (summed) and the bytes this rank received.  The counterpart of the reference's only collective is the final all-reduce
 * (task/tasks/ComputeEnergyVQETask.cpp:114-126). */
int b200vqe_last_exchange(b200vqe_handle h, double *ms, double *bytes);
/* ... and the time [ms] the compute stream WAITED for those transfers (what the kernels queued before each wait did not hide). */
int b200vqe_last_exchange_exposed(b200vqe_handle h, double *ms);
/* ... and which transport the slab swaps of the sharded <psi|H|psi> use on this handle: 1 peer-memory copies (CUDA IPC + copy engine),
 * -1 grouped ncclSend/ncclRecv, 0 not decided yet (decided collectively at the first sharded <psi|H|psi>; always 0 on one GPU). */
int b200vqe_exchange_mode(b200vqe_handle h, int *mode);

/* The host-only planner exports used by the CPU replay tests (b200vqe_debug_*) are declared in b200vqe_debug.h; a binding of the
 * drop-in ABI does not need them. */

/* State */
int b200vqe_state_reset(b200vqe_handle h, int slot, uint64_t basis_index);
int b200vqe_state_upload(b200vqe_handle h, int slot, const double *amps, uint64_t first, uint64_t count);
int b200vqe_state_download(b200vqe_handle h, int slot, double *amps, uint64_t first, uint64_t count, int order);
        /* order = B200VQE_ORDER_REFERENCE: whole local range only; on a sharded handle the call is COLLECTIVE (every rank must make it,
         * or the others wait in the gather for ever) and rank r receives the reference indices [r 2^nl, (r + 1) 2^nl).  It gathers
         * the whole register on every GPU: 16 B x 2^n for the staging slabs + 16 B x 2^nl for the extraction, i.e. O(2^n) per GPU,
         * not O(2^n / world) -- 72 GiB at 32 qubits on 8 GPUs.  The ranks check the room together first: B200VQE_ENOMEM on all of
         * them if any is short.  Registers that only fit sharded: download in device order and permute on the host. */
int b200vqe_state_random(b200vqe_handle h, int slot, uint64_t seed);            /* normalised complex Gaussian */
int b200vqe_state_product(b200vqe_handle h, int slot, const double *a, const double *b);
                        /* (x)_q (cos a_q |0> + e^{i b_q} sin a_q |1>) */
int b200vqe_state_norm2(b200vqe_handle h, int slot, double *norm2);
int b200vqe_state_inner(b200vqe_handle h, int slot_bra, int slot_ket, double *re_im);

/* Ansatz (replaces handing the gate list to the Accelerator, task/Program.hpp:250-261) */
int b200vqe_apply_gates(b200vqe_handle h, int slot, int64_t count, const b200vqe_gate *gates);
int b200vqe_apply_pauli_rotations(b200vqe_handle h, int slot, int64_t count, const uint64_t *xmask,
                                  const uint64_t *zmask, const double *phi /* exp(-i phi P) */);

/* <psi|H|psi> (replaces task/tasks/ComputeEnergyVQETask.cpp:133-166 and decorators/RDMGenerator.cpp:96-122).
 * energy[0..1] = sum_k c_k <P_k> (re, im); per_term (nullable, length n_terms, order of
 * set_hamiltonian) = <P_k> -- what getExpectationValueZ() returns for term k. */
int b200vqe_expectation(b200vqe_handle h, int slot, double *energy, double *per_term);

/* dst = H src (replaces the computeActionOnBra row loop + matrix, task/tasks/DiagonalizeTask.cpp:196-203,
 * task/tasks/petsc/slepc/SlepcDiagonalizeBackend.cpp:62-69); shape of ARPACK's MultMv(in,out),
 * tpls/arpackpp/include/arseig.h:208-214.  alpha (nullable) = <src|dst> (re, im). */
int b200vqe_apply_h(b200vqe_handle h, int src_slot, int dst_slot, double *alpha);

/* Lowest eigenpair by Lanczos (replaces Eigen dense solve DiagonalizeTask.cpp:205 and
 * EPSLANCZOS/EPS_SMALLEST_REAL, SlepcDiagonalizeBackend.cpp:79-89). */
typedef struct {
    double tol;          /* stop when |beta_m * s_m| <= tol * max(1,|theta|) */
    int32_t max_iter;
    int32_t want_vector; /* leave the Ritz vector in slot 0 */
    uint64_t seed;       /* start vector = normalised complex Gaussian(seed) unless use_slot0 */
    int32_t use_slot0;   /* start from the (non-zero) state already in slot 0 */
    int32_t reserved;    /* bit 0: resume the unfinished run kept in the handle (stopped at max_iter, or b200vqe_lanczos_load) */
} b200vqe_lanczos_opts;
int b200vqe_lanczos_ground(b200vqe_handle h, const b200vqe_lanczos_opts *opts, double *eigenvalue,
                           int32_t *iterations, double *residual);

/* Checkpoint / resume of an unfinished Lanczos run (one file per rank: coefficients + the two live vectors of this
 * rank's slab).  Counterpart, for this backend's long solve, of the reference's restart file
 * (decorators/VQERestartDecorator.cpp:17-75).  A run that returns B200VQE_ENOCONV at max_iter keeps its recurrence in
 * the handle (slots 1 and 2 must be left alone); save it, later: create + set_hamiltonian + load + lanczos_ground with
 * reserved = 1 and a larger max_iter continues exactly where it stopped. */
int b200vqe_lanczos_save(b200vqe_handle h, const char *path);
int b200vqe_lanczos_load(b200vqe_handle h, const char *path);

/* Particle-number sector (replaces the diag-number-symmetry branch, task/tasks/DiagonalizeTask.cpp:54-183, options
 * "diag-number-symmetry" + "n-electrons" + "fermion-transformation"): while set, b200vqe_lanczos_ground diagonalises
 * P H P, P = projector on the basis states whose occupation vector holds n_electrons ones -- the matrix the reference
 * assembles at :153-178 whenever H maps the sector into itself.  n_electrons < 0 clears the sector.
 * b200vqe_state_project_sector applies P to a slot (no renormalisation). */
int b200vqe_set_number_sector(b200vqe_handle h, int n_electrons, const char *transformation /* "jw" | "bk" */);
int b200vqe_state_project_sector(b200vqe_handle h, int slot);

/* Batched small-n path: n_vec parameter vectors of a rotation-form ansatz evaluated in one launch,
 * one statevector per CTA resident in shared memory (n <= 13).  energies has n_vec entries.
 * theta is row-major [n_vec][n_params]. */
int b200vqe_sweep_energies(b200vqe_handle h, uint64_t basis_index, int64_t n_rot, const uint64_t *xmask,
                           const uint64_t *zmask, const double *imag_coeff, const int32_t *param,
                           int n_params, int64_t n_vec, const double *theta, double *energies);

/* ------------------------------------------------------------------ allocation-free entry point
 * No counterpart in the reference (its vectors are host Eigen/PETSc objects; nothing there has a device allocator to avoid).
 *
 * out_re_im[0..1] = <bra|ket> = sum_j conj(bra_j) ket_j of two device vectors the CALLER owns, for hosts that must not
 * (or cannot) let this library allocate device memory: a simulator that has already filled the HBM with its own
 * register, a process whose device memory is managed by a single pool allocator, a loop that cannot take the
 * device-wide stall of cudaMalloc/cudaFree.  It needs no handle and calls no cudaMalloc, cudaFree or other allocating
 * routine.  The block sums of the reduction live in a static device array inside the library image instead
 * (B200VQE_NOALLOC_SCRATCH_BYTES on every GPU that loads the library, held until the process exits).
 *   bra, ket     device pointers on the CURRENT device, 16-byte aligned, n_amplitudes interleaved (re, im) doubles each
 *   out_re_im    device pointer (8-byte aligned), two doubles; written by the kernel in stream order -- the call does not synchronise
 *   stream       a cudaStream_t, or NULL for the default stream
 *   errbuf       nullable; receives the message of a failed call (errbuf_len bytes, NUL-terminated)
 * Limits: 1 <= n_amplitudes <= B200VQE_NOALLOC_MAX_AMPLITUDES, and a device of at most B200VQE_NOALLOC_MAX_CTAS / 4 SMs
 * (a B200 has 148); anything larger is refused with B200VQE_EINVAL and a message, never truncated.
 * The first call on each device looks up the array's address on that device (and may load the library's GPU code):
 * make that first call outside any stream capture.  All calls on one GPU share the one array: the call clears what has to
 * start at zero itself, but two calls whose kernels can overlap on the GPU (different streams, or different host threads
 * without ordering) race on it and return garbage -- order them on one stream or with events.
 * Same summation order as b200vqe_state_inner on a one-GPU handle of the same length. */
#define B200VQE_NOALLOC_MAX_AMPLITUDES ((uint64_t)1 << 33)   /* 2^33 amplitudes = 128 GiB per vector */
#define B200VQE_NOALLOC_MAX_CTAS 1024
#define B200VQE_NOALLOC_SCRATCH_BYTES 16640                  /* 2 x 1024 block sums + one 256-byte line for the ticket */
int b200vqe_inner_noalloc(const double *bra, const double *ket, uint64_t n_amplitudes, double *out_re_im, void *stream,
                          char *errbuf, int64_t errbuf_len);

/* Timing / accounting helpers used by bench.py */
int b200vqe_synchronize(b200vqe_handle h);
int b200vqe_kernel_launches(b200vqe_handle h, int64_t *count);    /* kernels launched by this handle so far */
int b200vqe_last_kernel_ms(b200vqe_handle h, double *ms);          /* CUDA-event time of the last expectation/apply */
int b200vqe_timer_start(b200vqe_handle h);                         /* records a CUDA event on the handle's stream */
int b200vqe_timer_stop(b200vqe_handle h, double *ms);              /* records, synchronises, returns elapsed ms */

#ifdef __cplusplus
}
#endif
#endif /* B200VQE_H_INCLUDED */
