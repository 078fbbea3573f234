/* b200vqe_debug.h -- host-only planner exports of libb200vqe.so.
 *
 * Not part of the drop-in ABI (include/b200vqe.h): nothing here touches a device and no reference interface corresponds to it.  The
 * functions return the tables the tile-engine planners build (csrc/tile_plan.hpp, csrc/kernels_rot_tile.cuh) so that the CPU tests
 * (tests/test_tile_plan_cpu.py, tests/test_sharded_host_gloo.py) can replay the kernels' index arithmetic in numpy against the oracle.
 */
#ifndef B200VQE_DEBUG_H_INCLUDED
#define B200VQE_DEBUG_H_INCLUDED

#include "b200vqe.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Host-only view of the tile-engine plan (see csrc/tile_plan.hpp) for a Hermitian term list, exactly as
 * b200vqe_set_hamiltonian builds it for <psi|H|psi> (batch capacities included): raw TileBatch[160 B],
 * TileGroup[16 B], TileTerm[16 B] records.  tile_bits = 12 (4096-amplitude tiles) | 11.  NULL arrays: counts only. */
int b200vqe_debug_tile_plan(int n_qubits, int chunk_bits, int tile_bits /* 12 | 11 */, int64_t n_terms, const uint64_t *xmask, const uint64_t *zmask,
                            const double *coef_re, const double *coef_im, int64_t *n_batches, int64_t *n_groups,
                            int64_t *n_plan_terms, void *batches, void *groups, void *terms);

/* Host-only: the fused runs of b200vqe_apply_pauli_rotations for `count` rotations with local flip masks on one GPU: TileBatch[320 B]
 * per run and RotRec[48 B] per rotation with phi != 0 (csrc/kernels_rot_tile.cuh); two-call pattern (NULL arrays: sizes only). */
int b200vqe_debug_rotation_runs(int n_qubits, int64_t count, const uint64_t *xmask, const uint64_t *zmask, const double *phi,
                                int64_t *n_runs, int64_t *n_records, void *batches, void *records);

/* Host-only: planner version (the "tile-planner" option) the b200vqe_debug_* plan functions use; default 3. */
int b200vqe_debug_set_planner(int planner);

/* Host-only: the TileRec[32 B] records (csrc/tile_plan.hpp) the <psi|H|psi> kernel walks for that plan; the r_begin / r_end
 * fields of a TileBatch index them.  Call with NULL for the count. */
int b200vqe_debug_tile_records(int n_qubits, int chunk_bits, int tile_bits, int64_t n_terms, const uint64_t *xmask, const uint64_t *zmask,
                               const double *coef_re, const double *coef_im, int64_t *n_records, void *records);

/* Host-only: the tile-engine plan of H psi for shard `rank` of `world` (csrc/tile_plan.hpp records, one plan per x_rank
 * class = which peer slab the partners are read from; classes[3k..3k+2] = x_rank, first batch, end batch).  NULL arrays: counts. */
int b200vqe_debug_apply_plan(int n_qubits, int rank, int world, int chunk_bits, int64_t n_terms, const uint64_t *xmask,
                             const uint64_t *zmask, const double *coef_re, const double *coef_im, int64_t *n_batches,
                             int64_t *n_records, int64_t *n_plan_terms, int64_t *n_classes, void *batches, void *records,
                             void *terms, int32_t *classes);

/* Host-only: the <psi|H|psi> plans of shard `rank` of `world` >= 2: one per x_rank class, built over the pair register {own slab,
 * slab of rank ^ x_rank} (class 0: the own slab); classes[4k..4k+3] = x_rank, qubits of that register, first batch, end batch. */
int b200vqe_debug_expect_class_plan(int n_qubits, int rank, int world, int chunk_bits, int64_t n_terms, const uint64_t *xmask,
                                    const uint64_t *zmask, const double *coef_re, const double *coef_im, int64_t *n_batches,
                                    int64_t *n_records, int64_t *n_plan_terms, int64_t *n_classes, void *batches, void *records,
                                    void *terms, int32_t *classes);

/* Host-only: slab and tile-engine share of `rank` in a register sharded over `world` ranks (what
 * b200vqe_create_sharded / b200vqe_expectation use).  No device needed. */
int b200vqe_debug_shard_layout(int n_qubits, int rank, int world, int tile_bits, uint64_t *n_local, uint64_t *first_index,
                               uint64_t *tile_begin, uint64_t *tile_count);

#ifdef __cplusplus
}
#endif
#endif /* B200VQE_DEBUG_H_INCLUDED */
