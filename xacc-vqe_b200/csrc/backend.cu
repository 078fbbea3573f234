// Device backend behind the C ABI of include/b200vqe.h.  One context = one GPU (one shard).
#include "../../include/b200vqe.h"
#include "../../include/b200vqe_debug.h"
#include "backend_ctx.hpp"
#include "kernels_apply.cuh"
#include "kernels_small.cuh"
#include "kernels_tile.cuh"
#include "kernels_tile_ring.cuh"
#include "kernels_tile_apply_pipe.cuh"
#include "kernels_rot_tile.cuh"
#include "pauli_host.hpp"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <new>
#include <stdexcept>
#include <dlfcn.h>
#include <nvtx3/nvToolsExt.h>   // header-only: ranges cost nothing unless a profiler injects itself (tracing)
#include <numeric>

using namespace b200vqe;

// ------------------------------------------------------------------------------------------
// NCCL through dlopen: the sharded path needs it, the single-GPU path must not depend on it.
// ------------------------------------------------------------------------------------------
namespace {
struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    bool load(std::string &err) {
        if (lib) return true;
        const char *names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char *n : names) {
            lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (lib) break;
        }
        if (!lib) { err = std::string("cannot dlopen libnccl.so.2: ") + dlerror(); return false; }
#define LOADSYM(field, name) \
        field = reinterpret_cast<decltype(field)>(dlsym(lib, name)); \
        if (!field) { err = std::string("libnccl lacks ") + name; lib = nullptr; return false; }
        LOADSYM(GetUniqueId, "ncclGetUniqueId")
        LOADSYM(CommInitRank, "ncclCommInitRank")
        LOADSYM(CommDestroy, "ncclCommDestroy")
        LOADSYM(AllGather, "ncclAllGather")
        LOADSYM(AllReduce, "ncclAllReduce")
        LOADSYM(Send, "ncclSend")
        LOADSYM(Recv, "ncclRecv")
        LOADSYM(GroupStart, "ncclGroupStart")
        LOADSYM(GroupEnd, "ncclGroupEnd")
        LOADSYM(GetErrorString, "ncclGetErrorString")
#undef LOADSYM
        return true;
    }
};
NcclApi g_nccl;

// stream memory operations of the driver API (no runtime-API counterpart): the front end waits for / writes a 32-bit word, no SM
struct CuApi {
    void *lib = nullptr;
    int (*WaitValue32)(cudaStream_t, unsigned long long, unsigned, unsigned) = nullptr;
    bool tried = false;
    bool load() {
        if (tried) return WaitValue32 != nullptr;
        tried = true;
        lib = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) return false;
        void *f = dlsym(lib, "cuStreamWaitValue32_v2");
        if (!f) f = dlsym(lib, "cuStreamWaitValue32");
        WaitValue32 = reinterpret_cast<int (*)(cudaStream_t, unsigned long long, unsigned, unsigned)>(f);
        return WaitValue32 != nullptr;
    }
};
CuApi g_cu;
thread_local std::string g_err;
}  // namespace

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            h->err = std::string(#call) + ": " + cudaGetErrorString(e_);                           \
            return e_ == cudaErrorMemoryAllocation ? B200VQE_ENOMEM : B200VQE_ECUDA;               \
        }                                                                                          \
    } while (0)
#define NK(call)                                                                                   \
    do {                                                                                           \
        ncclResult_t r_ = (call);                                                                  \
        if (r_ != ncclSuccess) {                                                                   \
            h->err = std::string(#call) + ": " + g_nccl.GetErrorString(r_);                        \
            return B200VQE_ECOMM;                                                                  \
        }                                                                                          \
    } while (0)
#define FAIL(code, msg) do { h->err = (msg); return (code); } while (0)

// Every extern "C" body runs inside this guard: no C++ exception crosses the C ABI (std::bad_alloc of a host vector sized from
// caller- or file-supplied counts becomes B200VQE_ENOMEM, anything else B200VQE_EINVAL with the message kept for last_error).
template <class F>
static int guarded(b200vqe_ctx *h, F &&body) noexcept {
    try {
        return body();
    } catch (const std::bad_alloc &) {
        (h ? h->err : g_err) = "out of host memory";
        return B200VQE_ENOMEM;
    } catch (const std::exception &e) {
        (h ? h->err : g_err) = std::string("internal error: ") + e.what();
        return B200VQE_EINVAL;
    } catch (...) {
        (h ? h->err : g_err) = "internal error";
        return B200VQE_EINVAL;
    }
}
// NVTX range around the hot entry points (set_hamiltonian, expectation, apply_h, lanczos_ground, the ansatz calls, sweep_energies): the
// counterpart of the reference's per-task timing prints; visible in nsys / ncu --nvtx, a no-op otherwise.
struct NvtxRange {
    explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
    NvtxRange(const NvtxRange &) = delete;
    NvtxRange &operator=(const NvtxRange &) = delete;
};

#define LAUNCHED(h) do { ++(h)->launches; CK(cudaGetLastError()); } while (0)

// A public call that writes slot 1 or 2 destroys the two live vectors of an unfinished Lanczos run kept for resume / checkpoint.
static inline void writes_slot(b200vqe_ctx *h, int slot) { if (slot == 1 || slot == 2) h->lz_m = -1; }

static int grid_for(const b200vqe_ctx *h, uint64_t work_items, int per_sm) {
    const uint64_t cap = (uint64_t)h->n_sm * per_sm;
    const uint64_t need = (work_items + kBlock - 1) / kBlock;
    return (int)std::max<uint64_t>(1, std::min(cap, need));
}

static int ensure_slot(b200vqe_ctx *h, int slot) {
    if (slot < 0 || slot >= (int)h->slots.size())
        FAIL(B200VQE_EINVAL, "slot " + std::to_string(slot) + " out of range (call b200vqe_n_slots first; have " + std::to_string(h->slots.size()) + ")");
    if (!h->slots[slot]) {
        CK(cudaSetDevice(h->device));
        CK(cudaMalloc(&h->slots[slot], h->nloc * sizeof(double2)));
    }
    return 0;
}

static int ensure_gather(b200vqe_ctx *h) {
    if (h->world == 1 || h->gather) return 0;
    CK(cudaMalloc(&h->gather, (size_t)h->world * h->nloc * sizeof(double2)));
    if (h->xbuf) { cudaStreamSynchronize(h->stream); if (h->comm_stream) cudaStreamSynchronize(h->comm_stream); cudaFree(h->xbuf); h->xbuf = nullptr; }   // gather now serves the pair exchanges too
    return 0;
}

// Make the slab table for reading `src`: unsharded -> the slot itself; sharded -> all-gather the
// shards into the staging buffer once (each peer shard crosses NVLink once per H application,
// independent of the number of terms), then every group reads HBM-resident slabs.
static int make_slabs(b200vqe_ctx *h, const double2 *src, Slabs &s) {
    for (int i = 0; i < kMaxWorld; ++i) s.p[i] = nullptr;
    if (h->world == 1) { s.p[0] = src; return 0; }
    int rc = ensure_gather(h);
    if (rc) return rc;
    // the gather runs on its own stream: it starts once everything queued so far (producers of src, earlier readers of
    // the staging buffer) is done, and kernels that only read the local slab may be queued behind it on the main stream
    // without waiting; wait_gather() orders the main stream after it before the first remote read
    CK(cudaEventRecord(h->ev_src_ready, h->stream));
    CK(cudaStreamWaitEvent(h->comm_stream, h->ev_src_ready, 0));
    if (h->n_xev < kMaxWorld) CK(cudaEventRecord(h->ev_x0[h->n_xev], h->comm_stream));
    NK(g_nccl.AllGather(src, h->gather, h->nloc * 2, ncclDouble, h->comm, h->comm_stream));
    if (h->n_xev < kMaxWorld) { CK(cudaEventRecord(h->ev_x1[h->n_xev], h->comm_stream)); ++h->n_xev; }
    h->last_comm_bytes += (double)(h->world - 1) * (double)h->nloc * sizeof(double2);
    CK(cudaEventRecord(h->ev_gathered, h->comm_stream));
    h->gather_pending = true;
    for (int t = 0; t < h->world; ++t) s.p[t] = h->gather + (size_t)t * h->nloc;
    s.p[h->rank] = src;
    return 0;
}

// The same, restricted to the peers whose slabs will be read (x_rank offsets present in the Hamiltonian): one NCCL group
// of send/recv pairs, all running concurrently; NVLink bytes = 16 * 2^nl per offset instead of per peer.
static int make_slabs_needed(b200vqe_ctx *h, const double2 *src, const std::vector<int> &x_ranks, Slabs &s) {
    for (int i = 0; i < kMaxWorld; ++i) s.p[i] = nullptr;
    s.p[h->rank] = src;
    int rc = ensure_gather(h);
    if (rc) return rc;
    CK(cudaEventRecord(h->ev_src_ready, h->stream));
    CK(cudaStreamWaitEvent(h->comm_stream, h->ev_src_ready, 0));
    if (h->n_xev < kMaxWorld) CK(cudaEventRecord(h->ev_x0[h->n_xev], h->comm_stream));
    h->last_comm_bytes += (double)x_ranks.size() * (double)h->nloc * sizeof(double2);
    NK(g_nccl.GroupStart());
    for (int c : x_ranks) {
        const int peer = h->rank ^ c;
        NK(g_nccl.Send(src, h->nloc * 2, ncclDouble, peer, h->comm, h->comm_stream));
        NK(g_nccl.Recv(h->gather + (size_t)peer * h->nloc, h->nloc * 2, ncclDouble, peer, h->comm, h->comm_stream));
        s.p[peer] = h->gather + (size_t)peer * h->nloc;
    }
    NK(g_nccl.GroupEnd());
    if (h->n_xev < kMaxWorld) { CK(cudaEventRecord(h->ev_x1[h->n_xev], h->comm_stream)); ++h->n_xev; }
    CK(cudaEventRecord(h->ev_gathered, h->comm_stream));
    h->gather_pending = true;
    return 0;
}

// the compute stream waits for a transfer recorded in `ev`; the events around the wait time its un-hidden part
static int wait_transfer(b200vqe_ctx *h, cudaEvent_t ev) {
    const bool timed = h->n_wev < kMaxWorld;
    if (timed) CK(cudaEventRecord(h->ev_w0[h->n_wev], h->stream));
    CK(cudaStreamWaitEvent(h->stream, ev, 0));
    if (timed) { CK(cudaEventRecord(h->ev_w1[h->n_wev], h->stream)); ++h->n_wev; }
    return 0;
}

static int wait_gather(b200vqe_ctx *h) {
    if (!h->gather_pending) return 0;
    int rc = wait_transfer(h, h->ev_gathered);
    if (rc) return rc;
    h->gather_pending = false;
    return 0;
}

// Pairwise slab exchange for ansatz gates that flip a sharded qubit: this rank's slab goes to `peer` and the peer's
// arrives in a staging buffer (grouped ncclSend/ncclRecv: the two directions of one NVLink pair run concurrently).
// Every rank of a pair must call it; pairs are independent, so ranks outside the exchange simply do not call.
static int exchange_bufs(b200vqe_ctx *h, double2 *bufs[2]) {
    if (h->gather) { bufs[0] = h->gather; bufs[1] = h->gather + h->nloc; return 0; }
    if (!h->xbuf) CK(cudaMalloc(&h->xbuf, 2 * h->nloc * sizeof(double2)));
    bufs[0] = h->xbuf; bufs[1] = h->xbuf + h->nloc;
    return 0;
}

static int exchange_with_peer(b200vqe_ctx *h, const double2 *src, int peer, const double2 **received) {
    double2 *bufs[2];
    int rc = exchange_bufs(h, bufs);
    if (rc) return rc;
    NK(g_nccl.GroupStart());
    NK(g_nccl.Send(src, h->nloc * 2, ncclDouble, peer, h->comm, h->stream));
    NK(g_nccl.Recv(bufs[0], h->nloc * 2, ncclDouble, peer, h->comm, h->stream));
    NK(g_nccl.GroupEnd());
    *received = bufs[0];
    return 0;
}

// The same exchange on the communication stream into staging buffer `which` (tile H psi: the slab of the next x_rank
// class travels while the current class is computed).  The caller orders the streams with ev_free / ev_recv.
static int exchange_async(b200vqe_ctx *h, const double2 *src, int peer, double2 *buf) {
    NK(g_nccl.GroupStart());
    NK(g_nccl.Send(src, h->nloc * 2, ncclDouble, peer, h->comm, h->comm_stream));
    NK(g_nccl.Recv(buf, h->nloc * 2, ncclDouble, peer, h->comm, h->comm_stream));
    NK(g_nccl.GroupEnd());
    return 0;
}

// Peer staging for the sharded <psi|H|psi>: allocate, export through CUDA IPC, map every peer's.  COLLECTIVE (every rank of the
// communicator must call it at the same point); the ranks agree through an all-reduce whether all of them succeeded, else all of
// them stay on NCCL send/recv.  Returns 0 also when unavailable (h->peer_state tells).
static int ensure_peer_staging(b200vqe_ctx *h) {
    if (h->peer_state != 0 || h->world == 1) return 0;
    h->peer_state = -1;
    const size_t slab = h->nloc * sizeof(double2), bytes = 2 * slab + 4096;
    bool ok = h->opt_peer_copy != 0 && g_cu.load();
    if (ok && !h->h_seq && cudaMallocHost(&h->h_seq, kFlagRing * sizeof(unsigned)) != cudaSuccess) { cudaGetLastError(); ok = false; }
    void *block = nullptr;
    struct BlockGuard {          // a CK / NK below returns early: the staging block must not outlive a failed set-up
        void *&p;
        ~BlockGuard() { if (p) cudaFree(p); }
    } block_guard{block};
    cudaIpcMemHandle_t mine{};
    if (ok && cudaMalloc(&block, bytes) != cudaSuccess) { cudaGetLastError(); ok = false; block = nullptr; }
    if (ok && cudaMemsetAsync(static_cast<char *>(block) + 2 * slab, 0, 4096, h->stream) != cudaSuccess) ok = false;
    if (ok && cudaIpcGetMemHandle(&mine, block) != cudaSuccess) { cudaGetLastError(); ok = false; }
    // all handles to every rank (64 bytes each), and a vote
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t size");
    char *d_h = reinterpret_cast<char *>(h->d_out);                     // kOutDoubles * 8 = 2048 bytes >= 64 * kMaxWorld + 8
    double vote = ok ? 1.0 : 0.0;
    CK(cudaMemcpyAsync(d_h + 64 * h->rank, &mine, 64, cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(d_h + 64 * kMaxWorld, &vote, 8, cudaMemcpyHostToDevice, h->stream));
    NK(g_nccl.AllGather(d_h + 64 * h->rank, d_h, 64, ncclChar, h->comm, h->stream));
    NK(g_nccl.AllReduce(d_h + 64 * kMaxWorld, d_h + 64 * kMaxWorld, 1, ncclDouble, ncclSum, h->comm, h->stream));
    std::vector<cudaIpcMemHandle_t> all(h->world);
    CK(cudaMemcpyAsync(all.data(), d_h, 64 * h->world, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(&vote, d_h + 64 * kMaxWorld, 8, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    bool all_ok = vote > h->world - 0.5;
    if (all_ok) {
        for (int p = 0; p < h->world && all_ok; ++p) {
            if (p == h->rank) continue;
            void *q = nullptr;
            if (cudaIpcOpenMemHandle(&q, all[p], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); all_ok = false; break; }
            h->peer_pbuf[p] = static_cast<double2 *>(q);
            h->peer_pflags[p] = reinterpret_cast<unsigned *>(static_cast<char *>(q) + 2 * slab);
        }
    }
    // second vote: every rank could map every peer
    vote = all_ok ? 1.0 : 0.0;
    CK(cudaMemcpyAsync(d_h, &vote, 8, cudaMemcpyHostToDevice, h->stream));
    NK(g_nccl.AllReduce(d_h, d_h, 1, ncclDouble, ncclSum, h->comm, h->stream));
    CK(cudaMemcpyAsync(&vote, d_h, 8, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (vote > h->world - 0.5) {
        h->pbuf = static_cast<double2 *>(block);
        h->pflags = reinterpret_cast<unsigned *>(static_cast<char *>(block) + 2 * slab);
        h->peer_state = 1;
        block = nullptr;         // owned by the handle from here on (b200vqe_destroy frees h->pbuf)
        return 0;
    }
    for (int p = 0; p < h->world; ++p) if (h->peer_pbuf[p]) { cudaIpcCloseMemHandle(h->peer_pbuf[p]); h->peer_pbuf[p] = nullptr; h->peer_pflags[p] = nullptr; }
    return 0;                    // block_guard frees the block
}

// stream `st` waits until the LOCAL flag word has reached `value` (wrap-safe >=); no SM involved
static int flag_wait(b200vqe_ctx *h, cudaStream_t st, const unsigned *flag, unsigned value) {
    const int r = g_cu.WaitValue32(st, (unsigned long long)(uintptr_t)flag, value, 0u /* CU_STREAM_WAIT_VALUE_GEQ */);
    if (r != 0) FAIL(B200VQE_ECUDA, "cuStreamWaitValue32 failed (" + std::to_string(r) + ")");
    return 0;
}
// stream `st` sets a flag word in a PEER's memory: a 4-byte copy from a pinned ring (copy engine, no SM)
static int flag_push(b200vqe_ctx *h, cudaStream_t st, unsigned *remote_flag, unsigned value) {
    // A ring slot is read by its 4-byte copy some time after this function returns, so it may be reused only once that copy is
    // known to be done: the caller resets h_seq_live after the synchronisation that ends its call (every sharded
    // b200vqe_expectation does).  (P - 1) (P + 1) pushes per call at world P = 63 at 8.
    if (h->h_seq_live >= kFlagRing) FAIL(B200VQE_ECOMM, "more than " + std::to_string(kFlagRing) + " flag pushes queued without a synchronisation");
    ++h->h_seq_live;
    unsigned *src = h->h_seq + (h->h_seq_pos++ % kFlagRing);
    *src = value;
    CK(cudaMemcpyAsync(remote_flag, src, sizeof(unsigned), cudaMemcpyDefault, st));
    return 0;
}

static int allreduce_doubles(b200vqe_ctx *h, double *dptr, size_t count) {
    if (h->world == 1) return 0;
    NK(g_nccl.AllReduce(dptr, dptr, count, ncclDouble, ncclSum, h->comm, h->stream));
    return 0;
}

extern "C" {

const char *b200vqe_global_error(void) { return g_err.c_str(); }

int b200vqe_comm_unique_id(char id[128]) {
    return guarded(nullptr, [&]() -> int {
    if (!g_nccl.load(g_err)) return B200VQE_ECOMM;
    ncclUniqueId u;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
    ncclResult_t r = g_nccl.GetUniqueId(&u);
    if (r != ncclSuccess) { g_err = g_nccl.GetErrorString(r); return B200VQE_ECOMM; }
    memcpy(id, &u, 128);
    return 0;
    });
}

static int create_common(int n_qubits, int device, int rank, int world, const char *id, b200vqe_handle *out) {
    if (!out) { g_err = "out is NULL"; return B200VQE_EINVAL; }
    *out = nullptr;
    if (world < 1 || world > kMaxWorld || (world & (world - 1))) { g_err = "world must be 1, 2, 4 or 8"; return B200VQE_EINVAL; }
    int wl = 0;
    while ((1 << wl) < world) ++wl;
    if (n_qubits < 1 || n_qubits > 40 || n_qubits - wl < 1) { g_err = "n_qubits must be in [1, 40] and exceed log2(world)"; return B200VQE_EINVAL; }
    if (rank < 0 || rank >= world) { g_err = "rank out of range"; return B200VQE_EINVAL; }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        g_err = std::string("no CUDA device available (") + cudaGetErrorString(e) + "); this backend has no CPU fallback";
        return B200VQE_ECUDA;
    }
    if (device < 0 || device >= ndev) { g_err = "device index out of range"; return B200VQE_EINVAL; }
    auto *h = new b200vqe_ctx();
    h->n = n_qubits; h->nl = n_qubits - wl; h->nloc = 1ull << h->nl;
    h->device = device; h->rank = rank; h->world = world;
    h->slots.assign(1, nullptr);
    auto fail = [&](const std::string &m, int code) { g_err = m; delete h; return code; };
    if ((e = cudaSetDevice(device)) != cudaSuccess) return fail(cudaGetErrorString(e), B200VQE_ECUDA);
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return fail(cudaGetErrorString(e), B200VQE_ECUDA);
    h->n_sm = prop.multiProcessorCount;
    h->smem_optin = (int)prop.sharedMemPerBlockOptin;
    {   // opt-in shared memory of the tile kernels, per device (a process may hold handles on several)
        const int tile = (int)(sizeof(double2) << kTileBits), two_small = (int)(2 * (sizeof(double2) << kTileBitsSmall));
        cudaError_t ea = cudaSuccess;
        auto set = [&](const void *f, int bytes) { if (ea == cudaSuccess) ea = cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes); };
        set((const void *)k_tile_expect<2, false>, tile);
        set((const void *)k_tile_expect<2, true>, tile);
        set((const void *)k_tile_expect_db<3>, two_small);
        set((const void *)k_tile_apply_pipe8<false>, kApply8SmemBytes);
        set((const void *)k_tile_apply_pipe8<true>, kApply8SmemBytes);
        set((const void *)k_tile_expect_pipe<0>, kPipeSmemBytes);
        set((const void *)k_tile_expect_pipe<1>, kPipeSmemBytes);
        set((const void *)k_tile_expect_pipe<2>, kPipeSmemBytes);
        set((const void *)k_sweep_small, tile);
        set((const void *)k_rot_tile, tile);
        if (ea != cudaSuccess) return fail(std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(ea), B200VQE_ECUDA);
    }
    if ((e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)) != cudaSuccess) return fail(cudaGetErrorString(e), B200VQE_ECUDA);
    cudaEventCreate(&h->ev0);
    cudaEventCreate(&h->ev1);
    cudaEventCreate(&h->ev_t0);
    cudaEventCreate(&h->ev_t1);
    if ((e = cudaMalloc(&h->partials, sizeof(double) * 16 * 4096)) != cudaSuccess) return fail(cudaGetErrorString(e), B200VQE_ENOMEM);
    if ((e = cudaMalloc(&h->counter, sizeof(unsigned))) != cudaSuccess) return fail(cudaGetErrorString(e), B200VQE_ENOMEM);
    cudaMemset(h->counter, 0, sizeof(unsigned));
    if ((e = cudaMalloc(&h->d_out, sizeof(double) * kOutDoubles)) != cudaSuccess) return fail(cudaGetErrorString(e), B200VQE_ENOMEM);
    if ((e = cudaMallocHost(&h->h_out, sizeof(double) * kOutDoubles)) != cudaSuccess) return fail(cudaGetErrorString(e), B200VQE_ENOMEM);
    if (world > 1) {
        int prio_lo = 0, prio_hi = 0;      // the exchange kernels go first whenever CTAs of a compute launch retire
        cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
        if ((e = cudaStreamCreateWithPriority(&h->comm_stream, cudaStreamNonBlocking, prio_hi)) != cudaSuccess) return fail(cudaGetErrorString(e), B200VQE_ECUDA);
        cudaEventCreateWithFlags(&h->ev_src_ready, cudaEventDisableTiming);
        cudaEventCreateWithFlags(&h->ev_gathered, cudaEventDisableTiming);
        for (int i = 0; i < 2; ++i) { cudaEventCreateWithFlags(&h->ev_recv[i], cudaEventDisableTiming); cudaEventCreateWithFlags(&h->ev_free[i], cudaEventDisableTiming); }
        for (int i = 0; i < kMaxWorld; ++i) { cudaEventCreate(&h->ev_x0[i]); cudaEventCreate(&h->ev_x1[i]); cudaEventCreate(&h->ev_w0[i]); cudaEventCreate(&h->ev_w1[i]); }
        if (!g_nccl.load(g_err)) { delete h; return B200VQE_ECOMM; }
        ncclUniqueId u;
        memcpy(&u, id, 128);
        ncclResult_t r = g_nccl.CommInitRank(&h->comm, world, u, rank);
        if (r != ncclSuccess) return fail(std::string("ncclCommInitRank: ") + g_nccl.GetErrorString(r), B200VQE_ECOMM);
    }
    *out = h;
    return 0;
}

int b200vqe_create(int n_qubits, int device, b200vqe_handle *out) {
    return guarded(nullptr, [&]() -> int {
    return create_common(n_qubits, device, 0, 1, nullptr, out);
    });
}
int b200vqe_create_sharded(int n_qubits, int device, int rank, int world, const char id[128], b200vqe_handle *out) {
    return guarded(nullptr, [&]() -> int {
    if (world > 1 && !id) { g_err = "id is NULL"; return B200VQE_EINVAL; }
    return create_common(n_qubits, device, rank, world, id, out);
    });
}

int b200vqe_destroy(b200vqe_handle h) {
    return guarded(h, [&]() -> int {
    if (!h) return 0;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    if (h->comm_stream) { cudaStreamSynchronize(h->comm_stream); cudaStreamDestroy(h->comm_stream); cudaEventDestroy(h->ev_src_ready); cudaEventDestroy(h->ev_gathered);
        for (int i = 0; i < 2; ++i) { cudaEventDestroy(h->ev_recv[i]); cudaEventDestroy(h->ev_free[i]); }
        for (int i = 0; i < kMaxWorld; ++i) { cudaEventDestroy(h->ev_x0[i]); cudaEventDestroy(h->ev_x1[i]); cudaEventDestroy(h->ev_w0[i]); cudaEventDestroy(h->ev_w1[i]); } }
    for (auto p : h->slots) if (p) cudaFree(p);
    if (h->gather) cudaFree(h->gather);
    if (h->xbuf) cudaFree(h->xbuf);
    for (int p = 0; p < kMaxWorld; ++p) if (h->peer_pbuf[p]) cudaIpcCloseMemHandle(h->peer_pbuf[p]);
    if (h->pbuf) cudaFree(h->pbuf);
    if (h->h_seq) cudaFreeHost(h->h_seq);
    h->ham.free_device();
    cudaFree(h->partials); cudaFree(h->counter); cudaFree(h->d_out); cudaFreeHost(h->h_out);
    if (h->d_rot) cudaFree(h->d_rot);
    if (h->d_term_partials) cudaFree(h->d_term_partials);
    if (h->d_term_out) cudaFree(h->d_term_out);
    if (h->d_alpha) cudaFree(h->d_alpha);
    if (h->d_norm) cudaFree(h->d_norm);
    if (h->d_ritz) cudaFree(h->d_ritz);
    if (h->comm) g_nccl.CommDestroy(h->comm);
    cudaEventDestroy(h->ev0); cudaEventDestroy(h->ev1); cudaEventDestroy(h->ev_t0); cudaEventDestroy(h->ev_t1);
    cudaStreamDestroy(h->stream);
    delete h;
    return 0;
    });
}

const char *b200vqe_last_error(b200vqe_handle h) { return h ? h->err.c_str() : g_err.c_str(); }

int b200vqe_n_slots(b200vqe_handle h, int n_slots) {
    return guarded(h, [&]() -> int {
    if (!h) return B200VQE_EINVAL;
    h->lz_m = -1;               // slots may be released below: an unfinished Lanczos run does not survive
    if (n_slots < 1 || n_slots > 16) FAIL(B200VQE_EINVAL, "n_slots must be in [1, 16]");
    CK(cudaSetDevice(h->device));
    for (int s = n_slots; s < (int)h->slots.size(); ++s) if (h->slots[s]) { cudaFree(h->slots[s]); }
    h->slots.resize(n_slots, nullptr);
    return 0;
    });
}

int b200vqe_local_amplitudes(b200vqe_handle h, uint64_t *count, uint64_t *first_index) {
    return guarded(h, [&]() -> int {
    if (!h) return B200VQE_EINVAL;
    if (count) *count = h->nloc;
    if (first_index) *first_index = (uint64_t)h->rank << h->nl;
    return 0;
    });
}

// string options in the manner of xacc::setOption(key, value)
int b200vqe_set_option(b200vqe_handle h, const char *key, const char *value) {
    return guarded(h, [&]() -> int {
    if (!h) return B200VQE_EINVAL;
    if (!key || !value) FAIL(B200VQE_EINVAL, "null option key / value");
    const std::string k(key), v(value);
    if (k == "expect-engine") {
        if (v == "auto") h->opt_expect_engine = 0;
        else if (v == "tile") h->opt_expect_engine = 1;
        else if (v == "stream") h->opt_expect_engine = 2;
        else FAIL(B200VQE_EINVAL, "expect-engine must be auto, tile or stream");
        return 0;
    }
    if (k == "peer-copy") {
        if (v != "on" && v != "1" && v != "off" && v != "0") FAIL(B200VQE_EINVAL, "peer-copy must be on or off");
        if (h->peer_state != 0) FAIL(B200VQE_EINVAL, "peer-copy must be set before the first sharded <psi|H|psi>");   // option left as it was
        h->opt_peer_copy = (v == "on" || v == "1") ? 1 : 0;
        return 0;
    }
    if (k == "fuse-rotations") {
        if (v == "on" || v == "1") h->opt_fuse_rotations = 1;
        else if (v == "off" || v == "0") h->opt_fuse_rotations = 0;
        else FAIL(B200VQE_EINVAL, "fuse-rotations must be on or off");
        return 0;
    }
    if (k == "apply-engine") {
        if (v == "auto") h->opt_apply_engine = 0;
        else if (v == "tile") h->opt_apply_engine = 1;
        else if (v == "stream") h->opt_apply_engine = 2;
        else FAIL(B200VQE_EINVAL, "apply-engine must be auto, tile or stream");
        return 0;
    }
    if (k == "tile-chunk-bits") {
        const int c = atoi(value);
        if (c < 2 || c > 8) FAIL(B200VQE_EINVAL, "tile-chunk-bits must be in [2, 8]");
        h->opt_tile_chunk_bits = c;
        h->opt_apply_chunk_bits = c;        // an explicit choice applies to both engines
        return 0;
    }
    if (k == "apply-chunk-bits") {
        const int c = atoi(value);
        if (c < 2 || c > 8) FAIL(B200VQE_EINVAL, "apply-chunk-bits must be in [2, 8]");
        h->opt_apply_chunk_bits = c;
        return 0;
    }
    if (k == "shard-exchange") {
        if (v == "auto" || v == "allgather") h->opt_shard_exchange = 0;
        else if (v == "needed") h->opt_shard_exchange = 2;
        else if (v == "pairwise") h->opt_shard_exchange = 1;
        else FAIL(B200VQE_EINVAL, "shard-exchange must be auto, allgather, needed or pairwise");
        return 0;
    }
    if (k == "tile-bits") {
        const int c = atoi(value);
        if (c != kTileBitsSmall && c != kTileBits) FAIL(B200VQE_EINVAL, "tile-bits must be 11 or 12");
        h->opt_tile_bits = c;
        return 0;
    }
    if (k == "expect-kernel") {
        if (v == "auto" || v == "pipe") h->opt_expect_kernel = 0;
        else if (v == "classic") h->opt_expect_kernel = 1;
        else FAIL(B200VQE_EINVAL, "expect-kernel must be auto, pipe or classic");
        return 0;
    }
    if (k == "tile-planner") {
        const int c = atoi(value);
        if (c < 1 || c > 3) FAIL(B200VQE_EINVAL, "tile-planner must be 1, 2 or 3");
        h->opt_tile_planner = c;
        return 0;
    }
    FAIL(B200VQE_EINVAL, "unknown option '" + k + "'");
    });
}

int b200vqe_synchronize(b200vqe_handle h) {
    return guarded(h, [&]() -> int {
    if (!h) return B200VQE_EINVAL;
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    return 0;
    });
}
int b200vqe_kernel_launches(b200vqe_handle h, int64_t *count) {
    return guarded(h, [&]() -> int {
    if (!h || !count) return B200VQE_EINVAL;
    *count = h->launches;
    return 0;
    });
}
int b200vqe_last_kernel_ms(b200vqe_handle h, double *ms) {
    return guarded(h, [&]() -> int {
    if (!h || !ms) return B200VQE_EINVAL;
    *ms = h->last_ms;
    return 0;
    });
}

int b200vqe_timer_start(b200vqe_handle h) {
    return guarded(h, [&]() -> int {
    if (!h) return B200VQE_EINVAL;
    CK(cudaSetDevice(h->device));
    CK(cudaEventRecord(h->ev_t0, h->stream));
    return 0;
    });
}
int b200vqe_timer_stop(b200vqe_handle h, double *ms) {
    return guarded(h, [&]() -> int {
    if (!h || !ms) return B200VQE_EINVAL;
    CK(cudaSetDevice(h->device));
    CK(cudaEventRecord(h->ev_t1, h->stream));
    CK(cudaEventSynchronize(h->ev_t1));
    float f = 0.f;
    CK(cudaEventElapsedTime(&f, h->ev_t0, h->ev_t1));
    *ms = f;
    return 0;
    });
}

}  // extern "C"

#include "backend_ham.inc"
#include "backend_state.inc"
#include "backend_small.inc"
#include "backend_lanczos.inc"
