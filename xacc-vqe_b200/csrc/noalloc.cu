// b200vqe_inner_noalloc: <bra|ket> of two caller-owned device vectors without touching the device allocator.
//
// Who this is for.  Every other entry point works on a handle, and creating a handle means cudaMalloc (slots, block-sum
// scratch, result words).  Some hosts of this library cannot afford that:
//   - a simulator plugin that has already claimed nearly all of the HBM for its own register (a 33-qubit state is 128 GiB
//     of the 180 GB) and wants an overlap or a norm of vectors it owns -- a further cudaMalloc there can fail, or succeed
//     and push the owner's next allocation over the edge;
//   - a host whose device memory belongs to one pool allocator (a framework's caching allocator, RMM, a job scheduler's
//     quota): memory taken behind the pool's back is invisible to its accounting and is the classic source of
//     "out of memory with 10 GB free" reports;
//   - an optimiser loop that keeps other streams busy: cudaMalloc and cudaFree synchronise the whole device, so one
//     allocation inside the loop stalls every kernel in flight on every stream.
// For them the block sums and the arrival ticket of the grid reduction live in a fixed array that is part of the library
// image itself.  It costs B200VQE_NOALLOC_SCRATCH_BYTES on every GPU that loads the library, from load until exit.
//
// The kernel is k_inner's loop and grid (kernels_apply.cuh) on top of the same grid_reduce, so the summation order -- and
// therefore the result, bit for bit -- is that of b200vqe_state_inner on a one-GPU handle of the same size.
// Nothing in the handle-based code paths includes or calls this file.
#include "../../include/b200vqe.h"
#include "kernels.cuh"

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <mutex>

namespace {
using namespace b200vqe;

constexpr int kMaxCtas = B200VQE_NOALLOC_MAX_CTAS;
constexpr int kTicketOffset = 2 * kMaxCtas;             // in doubles: the ticket word has the last 256-byte line to itself
constexpr int kScratchDoubles = kTicketOffset + 32;
static_assert(kScratchDoubles * sizeof(double) == B200VQE_NOALLOC_SCRATCH_BYTES, "header and array disagree");

// [0, kMaxCtas)            block sums of Re <bra|ket>, one per CTA
// [kMaxCtas, 2 kMaxCtas)   block sums of Im <bra|ket>   (grid_reduce<2> indexes row i at i * gridDim.x, so a launch with
//                          fewer CTAs uses a prefix [0, 2 gridDim.x) -- always inside the array)
// [kTicketOffset]          arrival ticket (unsigned)
// In host code this name is only a handle for cudaGetSymbolAddress; it is never passed to a kernel as it stands.
static __device__ __align__(256) double g_inner_scratch[kScratchDoubles];

constexpr int kMaxDevices = 64;
struct PerDevice {
    double *scratch = nullptr;   // device address of g_inner_scratch in THIS device's copy of the module
    int n_sm = 0;
};
PerDevice g_dev[kMaxDevices];
std::mutex g_setup_mutex;

__global__ void __launch_bounds__(kBlock, 4)
k_inner_static(const double2 *__restrict__ bra, const double2 *__restrict__ ket, uint64_t n, double *scratch, double *out) {
    double re = 0.0, im = 0.0;
    for (uint64_t j = (uint64_t)blockIdx.x * kBlock + threadIdx.x; j < n; j += (uint64_t)gridDim.x * kBlock) {
        const double2 a = bra[j], b = ket[j];
        re += a.x * b.x + a.y * b.y;
        im += a.x * b.y - a.y * b.x;
    }
    double v[2] = {re, im};
    grid_reduce<2>(v, scratch, reinterpret_cast<unsigned *>(scratch + kTicketOffset), out, 1.0);
}

int fail(char *errbuf, int64_t cap, int code, const char *fmt, ...) {
    if (errbuf && cap > 0) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(errbuf, (size_t)cap, fmt, ap);
        va_end(ap);
    }
    return code;
}

// Resolves, once per device and with that device current, where this device keeps its copy of the array.
int setup_current_device(int dev, PerDevice **out, char *errbuf, int64_t cap) {
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    PerDevice &d = g_dev[dev];
    if (!d.scratch) {
        double *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void **)&p, g_inner_scratch);
        if (e != cudaSuccess || !p)
            return fail(errbuf, cap, B200VQE_ECUDA, "cudaGetSymbolAddress(g_inner_scratch) on device %d: %s", dev, cudaGetErrorString(e));
        int n_sm = 0;
        e = cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess || n_sm <= 0)
            return fail(errbuf, cap, B200VQE_ECUDA, "cudaDeviceGetAttribute(SM count) on device %d: %s", dev, cudaGetErrorString(e));
        d.n_sm = n_sm;
        d.scratch = p;
    }
    *out = &d;
    return 0;
}

int check_device_pointer(const void *p, const char *what, unsigned align, int dev, char *errbuf, int64_t cap) {
    if (!p) return fail(errbuf, cap, B200VQE_EINVAL, "%s is NULL", what);
    if (reinterpret_cast<uintptr_t>(p) & (align - 1)) return fail(errbuf, cap, B200VQE_EINVAL, "%s is not %u-byte aligned", what, align);
    cudaPointerAttributes a;
    cudaError_t e = cudaPointerGetAttributes(&a, p);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(errbuf, cap, B200VQE_EINVAL, "%s: %s", what, cudaGetErrorString(e));
    }
    if (a.type != cudaMemoryTypeDevice && a.type != cudaMemoryTypeManaged)
        return fail(errbuf, cap, B200VQE_EINVAL, "%s must point to device memory (the caller's own allocation); got a host pointer", what);
    if (a.type == cudaMemoryTypeDevice && a.device != dev)
        return fail(errbuf, cap, B200VQE_EINVAL, "%s lives on device %d but device %d is current", what, a.device, dev);
    return 0;
}
}  // namespace

extern "C" int b200vqe_inner_noalloc(const double *bra, const double *ket, uint64_t n_amplitudes, double *out_re_im, void *stream,
                                     char *errbuf, int64_t errbuf_len) {
    try {
        if (errbuf && errbuf_len > 0) errbuf[0] = '\0';
        if (n_amplitudes == 0) return fail(errbuf, errbuf_len, B200VQE_EINVAL, "n_amplitudes is 0");
        if (n_amplitudes > B200VQE_NOALLOC_MAX_AMPLITUDES)
            return fail(errbuf, errbuf_len, B200VQE_EINVAL,
                        "n_amplitudes = %llu exceeds the maximum of this entry point, B200VQE_NOALLOC_MAX_AMPLITUDES = %llu",
                        (unsigned long long)n_amplitudes, (unsigned long long)B200VQE_NOALLOC_MAX_AMPLITUDES);
        int dev = -1;
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return fail(errbuf, errbuf_len, B200VQE_ECUDA, "cudaGetDevice: %s (no CPU fallback)", cudaGetErrorString(e));
        if (dev < 0 || dev >= kMaxDevices) return fail(errbuf, errbuf_len, B200VQE_EINVAL, "device ordinal %d outside [0, %d)", dev, kMaxDevices);
        PerDevice *d = nullptr;
        int rc = setup_current_device(dev, &d, errbuf, errbuf_len);
        if (rc) return rc;
        if ((rc = check_device_pointer(bra, "bra", 16, dev, errbuf, errbuf_len))) return rc;
        if ((rc = check_device_pointer(ket, "ket", 16, dev, errbuf, errbuf_len))) return rc;
        if ((rc = check_device_pointer(out_re_im, "out_re_im", 8, dev, errbuf, errbuf_len))) return rc;

        // the grid of k_inner on a handle (grid_for(h, n, 4)): four CTAs per SM, fewer when the vector is short
        const uint64_t want = (n_amplitudes + kBlock - 1) / kBlock;
        const uint64_t cap_ctas = (uint64_t)d->n_sm * 4;
        const uint64_t grid = want < cap_ctas ? want : cap_ctas;
        if (grid > (uint64_t)kMaxCtas)
            return fail(errbuf, errbuf_len, B200VQE_EINVAL,
                        "this device (%d SMs) would launch %llu CTAs; the static array holds block sums for at most B200VQE_NOALLOC_MAX_CTAS = %d",
                        d->n_sm, (unsigned long long)grid, kMaxCtas);

        cudaStream_t st = (cudaStream_t)stream;
        // The array keeps whatever the previous call (or a call that died half-way) left in it.  Only the ticket has to start
        // at zero: every block-sum slot the last CTA reads has been rewritten by this launch before the ticket reaches it.
        e = cudaMemsetAsync(d->scratch + kTicketOffset, 0, sizeof(unsigned), st);
        if (e != cudaSuccess) return fail(errbuf, errbuf_len, B200VQE_ECUDA, "cudaMemsetAsync(ticket): %s", cudaGetErrorString(e));
        k_inner_static<<<(unsigned)grid, kBlock, 0, st>>>(reinterpret_cast<const double2 *>(bra), reinterpret_cast<const double2 *>(ket),
                                                          n_amplitudes, d->scratch, out_re_im);
        e = cudaGetLastError();
        if (e != cudaSuccess) return fail(errbuf, errbuf_len, B200VQE_ECUDA, "k_inner_static launch: %s", cudaGetErrorString(e));
        return B200VQE_OK;
    } catch (...) {
        return fail(errbuf, errbuf_len, B200VQE_ECUDA, "unexpected C++ exception in b200vqe_inner_noalloc");
    }
}
