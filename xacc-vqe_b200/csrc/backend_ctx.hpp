// Context of one device backend instance (one GPU / one shard).
#pragma once
#include <cuda_runtime.h>
#include <nccl.h>

#include <cstdint>
#include <string>
#include <vector>

#include "kernels.cuh"
#include "tile_plan.hpp"

namespace b200vqe {

constexpr unsigned kFlagRing = 1024;  // pinned ring of flag values (sources of the 4-byte peer flag copies)
constexpr int kOutDoubles = 256;   // device/host scalar scratch (bucket energies, norms, ...)

struct Bucket { int h; int g_begin; int g_end; };   // h = -1: all-j form (weight 1), else pair form (weight 2)

// Device tables of one Hamiltonian.  Two layouts of the same terms:
//   expect: identity stripped, groups ordered by bucket, u_k = coef (even nY) | i*coef (odd nY)
//   apply : identity kept,     groups ordered by xmask,  v_k = coef
struct DeviceHam {
    int64_t n_terms = 0;            // as given by the caller
    int64_t tile_view_loads = 0, tile_fused_blocks = 0, tile_plain_recs = 0, tile_partner_recs = 0;   // of the <psi|H|psi> plan (tile_plan.hpp)
    uint64_t fingerprint = 0;       // FNV-1a of (n, sorted (xmask, zmask, coef) table): Lanczos checkpoints are bound to it
    int64_t n_terms_nonid = 0;
    bool realc = true;
    double ident_re = 0.0, ident_im = 0.0;
    std::vector<int64_t> ident_idx; // caller indices of identity terms
    int tile_R = 1;                 // R the tables were bucketed for

    // expectation
    std::vector<Bucket> buckets;
    int n_groups_e = 0;
    GroupDesc *gd_e = nullptr; uint64_t *tz_e = nullptr; double2 *tu_e = nullptr;
    // apply
    int n_groups_a = 0;
    GroupDesc *gd_a = nullptr; uint64_t *tz_a = nullptr; double2 *tv_a = nullptr;
    // per-term path (host copies, sorted by xmask): caller index, masks, r_k
    std::vector<int64_t> pt_idx; std::vector<uint64_t> pt_x, pt_z; std::vector<int> pt_ny;
    // tile engine (kernels_tile.cuh): batches of xmask groups evaluated per HBM read of the state
    int tile_C = 0, tile_M = 0, tile_TB = kTileBits, n_tile_batches = 0, n_tile_groups = 0, tile_dispatch = 0, tile_views = 1;
    TileBatch *tl_batches = nullptr; TileRec *tl_recs = nullptr; TileTerm *tl_terms = nullptr;
    std::vector<int64_t> tl_term_idx;   // plan term -> caller index (per-term output of the tile engine)
    std::vector<double> tl_term_u;      // unweighted coefficient of the plan term: coef (even nY) | i*coef (odd nY)
    std::vector<double> tl_term_r;      // <P_k> = r_k * (signed pair sum): Re i^nY (even nY) | -Im i^nY (odd nY)
    // ... on a sharded register the full-register plan above serves only the per-term output and is built on its first use
    // (tl_src keeps the terms until then); <psi|H|psi> itself uses one plan per x_rank class (which peer slab the partners live
    // in), built over the PAIR REGISTER {own slab, slab of rank ^ x_rank} of nl + 1 qubits (class 0: the own slab, nl qubits)
    struct TlSrc { uint64_t x, z; double u; int ny; int64_t idx; };
    std::vector<TlSrc> tl_src; bool tl_wanted = false;
    // the per-term output has its own full-register plan (one register subspace per batch), built on its first use
    bool tp_built = false; int tp_C = 0, tp_M = 0, n_tp_batches = 0;
    TileBatch *tp_batches = nullptr; TileRec *tp_recs = nullptr; TileTerm *tp_terms = nullptr;
    struct ExpectClass { int x_rank; int n_reg; int b_begin, b_end; };
    std::vector<ExpectClass> tc_classes;
    TileBatch *tc_batches = nullptr; TileRec *tc_recs = nullptr; TileTerm *tc_terms = nullptr; int n_tc_batches = 0;
    // tile engine for H psi: one plan per x_rank class (which peer slab the partners live in), concatenated
    struct ApplyClass { int x_rank; int b_begin, b_end; };
    struct ApplySrc { uint64_t x, z; double cre, cim; int ny; int64_t idx; };
    bool ta_wanted = false, ta_built = false; int ta_chunk_bits = 3;
    std::vector<ApplySrc> ta_src;   // terms kept on the host until the first H psi plans them
    std::vector<ApplyClass> ta_classes;
    int ta_C = 0, ta_M = 0, n_ta_batches = 0;
    TileBatch *ta_batches = nullptr; TileRec *ta_recs = nullptr; TileTerm *ta_terms = nullptr;
    // natural-layout plan for the shared-memory sweep kernel (registers of up to kTileBits qubits)
    bool has_sweep_plan = false;
    TileBatch *sw_batch = nullptr; TileGroup *sw_groups = nullptr; TileTerm *sw_terms = nullptr;
    // host copies for the small-n kernels / stats
    std::vector<GroupDesc> h_gd_a; std::vector<uint64_t> h_tz_a; std::vector<double2> h_tv_a;

    void free_device() {
        cudaFree(gd_e); cudaFree(tz_e); cudaFree(tu_e); cudaFree(gd_a); cudaFree(tz_a); cudaFree(tv_a);
        cudaFree(tl_batches); cudaFree(tl_recs); cudaFree(tl_terms);
        cudaFree(tc_batches); cudaFree(tc_recs); cudaFree(tc_terms);
        tc_batches = nullptr; tc_recs = nullptr; tc_terms = nullptr;
        cudaFree(tp_batches); cudaFree(tp_recs); cudaFree(tp_terms);
        tp_batches = nullptr; tp_recs = nullptr; tp_terms = nullptr; tp_built = false;
        cudaFree(ta_batches); cudaFree(ta_recs); cudaFree(ta_terms);
        cudaFree(sw_batch); cudaFree(sw_groups); cudaFree(sw_terms);
        sw_batch = nullptr; sw_groups = nullptr; sw_terms = nullptr;
        ta_batches = nullptr; ta_recs = nullptr; ta_terms = nullptr;
        gd_e = gd_a = nullptr; tz_e = tz_a = nullptr; tu_e = tv_a = nullptr;
        tl_batches = nullptr; tl_recs = nullptr; tl_terms = nullptr;
    }
};

}  // namespace b200vqe

struct b200vqe_ctx {
    int n = 0, nl = 0, device = 0, rank = 0, world = 1, n_sm = 148, smem_optin = 0;
    uint64_t nloc = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t comm_stream = nullptr;             // sharded only: the all-gather runs here, overlapped with local groups
    cudaEvent_t ev_src_ready = nullptr, ev_gathered = nullptr; bool gather_pending = false;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_t0 = nullptr, ev_t1 = nullptr;
    std::vector<double2 *> slots;
    double2 *gather = nullptr;          // [world][nloc] staging of the peers' shards
    double2 *xbuf = nullptr;            // [2][nloc] staging of peer shards (pairwise exchanges: ansatz gates, tile H psi) when gather is absent
    // Peer staging (sharded <psi|H|psi>): two slabs + flag words in ONE allocation that every peer maps through CUDA IPC; a peer PUSHES
    // its slab into it with a device-to-device copy (copy engine over NVLink) and raises arrived[b] with a 4-byte copy; this rank tells
    // every peer with 4-byte copies when the class that read buffer b is done.  All waits are stream memory operations on LOCAL flag
    // words (cuStreamWaitValue32): nothing here needs an SM -- the persistent tile kernels hold every register of every SM, so NCCL's
    // send/recv kernels (and any flag kernel) only run when a tile kernel ends.
    double2 *pbuf = nullptr; unsigned *pflags = nullptr;              // flags: [0,1] arrived[b]; [4 + 2 p + b] free_from[p][b]: peer p is done with ITS buffer b
    unsigned *h_seq = nullptr; unsigned h_seq_pos = 0, h_seq_live = 0;                 // pinned ring of sequence numbers: sources of the 4-byte flag copies
    double2 *peer_pbuf[b200vqe::kMaxWorld] = {}; unsigned *peer_pflags[b200vqe::kMaxWorld] = {};
    int peer_state = 0;                                                // 0 not tried, 1 ready, -1 unavailable (NCCL send/recv instead)
    unsigned xseq = 0, buf_last_seq[2] = {0, 0};                       // exchange sequence numbers (identical on every rank)
    cudaEvent_t ev_recv[2] = {nullptr, nullptr}, ev_free[2] = {nullptr, nullptr};   // pipelined slab exchange of the tile H psi
    b200vqe::DeviceHam ham;
    double *partials = nullptr; unsigned *counter = nullptr;
    double *d_out = nullptr, *h_out = nullptr;
    double *d_alpha = nullptr, *d_norm = nullptr, *d_ritz = nullptr; int lanczos_cap = 0;
    // Lanczos recurrence kept between calls (resume / checkpoint): alpha_0..alpha_{m-1}, beta_1..beta_m, W_m and W_{m-1} in slots 1, 2
    std::vector<double> lz_a, lz_b; int lz_m = -1;
    unsigned char *d_rot = nullptr; size_t rot_cap = 0; int64_t rot_runs = 0;   // tables of the fused rotation runs (kernels_rot_tile.cuh); runs launched so far
    double *d_term_partials = nullptr, *d_term_out = nullptr; size_t term_partials_cap = 0, term_out_cap = 0;
    ncclComm_t comm = nullptr;
    std::string err;
    int64_t launches = 0;
    double last_ms = 0.0;
    // NVLink exchange of the last sharded <psi|H|psi> / H psi: CUDA-event time of the transfers on the communication stream
    // (summed over the exchanges of the call) and the bytes this rank received
    cudaEvent_t ev_x0[b200vqe::kMaxWorld] = {}, ev_x1[b200vqe::kMaxWorld] = {}; int n_xev = 0; double last_comm_bytes = 0.0;
    // ... and the time the compute stream spent WAITING for those transfers (the un-hidden part): events around every wait
    cudaEvent_t ev_w0[b200vqe::kMaxWorld] = {}, ev_w1[b200vqe::kMaxWorld] = {}; int n_wev = 0;
    // particle-number sector (b200vqe_set_number_sector); sector_ne < 0: none
    int sector_ne = -1; bool sector_unit = true; uint64_t sector_rows[40] = {0};
    // options (string keys as in the reference's xacc::setOption map)
    int opt_expect_engine = 0;          // 0 auto, 1 tile engine, 2 streaming kernel
    int opt_apply_engine = 0;           // same, for H psi
    int opt_tile_bits = 12;             // tile of the <psi|H|psi> engine: 12 = 4096 amplitudes single-buffered, 11 = 2048 double-buffered
    int opt_shard_exchange = 0;         // tile H psi on a sharded register: 0 all-gather (auto), 1 pairwise pipeline, 2 needed slabs only (one send/recv group)
    int opt_expect_kernel = 0;          // <psi|H|psi> on 4096-amplitude tiles: 0 pipelined kernel (k_tile_expect_pipe: 3-stage cp.async ring), 1 k_tile_expect
    int opt_peer_copy = 1;              // sharded <psi|H|psi>: 1 slab swaps as peer-memory copies (CUDA IPC + copy engine), 0 NCCL send/recv
    int opt_fuse_rotations = 1;         // consecutive Pauli rotations with local flip masks run as fused tile passes (0: one launch per rotation)
    int opt_tile_planner = 3;           // tile planner version (tile_plan.hpp: 3 = 2 + several register subspaces (views) per batch, 2 = collision-greedy subspaces + sheared register subspace, 1 = first version)
    int opt_tile_chunk_bits = 3;        // C: contiguous amplitudes per chunk = 2^C (128 B at C = 3), <psi|H|psi> plans
    int opt_apply_chunk_bits = 4;       // the same for H psi plans: 256-byte chunks make the dst read-modify-write stream measurably faster
                                        // (28 q / 10^4 terms: 2336 ms against 2595 ms at C = 3) although a batch then holds fewer groups
};
