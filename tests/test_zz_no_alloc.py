"""b200vqe_inner_noalloc -- the one entry point of the library that allocates no device memory -- against the
handle-based b200vqe_state_inner, both called through the C ABI (ctypes on libb200vqe.so).

Why the entry point exists: a host that has handed its whole HBM to its own statevector, or whose device memory
is owned by one pool allocator, or that cannot take the device-wide stall of cudaMalloc inside its loop, still
wants an overlap / a norm of vectors it already holds on the GPU.  Here torch plays that host: it owns the
vectors and the two result doubles, the library only launches a kernel whose block sums sit in a static array
inside the library image.

The file is named to be collected last: `pytest -x` stops at the first failure, and a fault in here should not
hide the parity tests.  It does not depend on anything the earlier tests leave behind (own handles, own
buffers), and the static array is used by nothing else, so it runs the same alone and after the rest.

Tolerance: 1e-10 relative, the figure of tests/test_gpu_parity.py (the two kernels add in the same order, so the
values are in fact expected to be identical; the test does not insist on it)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

REL = 1e-10


@pytest.fixture(scope="module")
def xb():
    import xaccvqe_b200
    return xaccvqe_b200


def _close(got: complex, want: complex) -> bool:
    return abs(got - want) <= REL * max(abs(want), 1e-300)


def _host_vectors(xb, n_qubits, seed):
    """Two random states made and read back by the existing entry points, plus what b200vqe_state_inner says
    about them: <0|1> and <0|0>."""
    be = xb.Backend(n_qubits)
    be.n_slots(2)
    be.random_state(0, seed)
    be.random_state(1, seed + 1)
    return be.download(0), be.download(1), be.inner(0, 1), be.inner(0, 0)


def _to_device(torch, psi):
    return torch.from_numpy(np.ascontiguousarray(psi).view(np.float64)).to("cuda:0")


def test_inner_noalloc_matches_state_inner_on_consecutive_calls(xb):
    import torch
    torch.cuda.set_device(0)
    big = _host_vectors(xb, 20, seed=11)      # 4096 CTAs wanted -> the full grid of 4 per SM
    small = _host_vectors(xb, 10, seed=23)    # 4 CTAs: reads a short prefix of an array full of the big call's sums
    d_big = (_to_device(torch, big[0]), _to_device(torch, big[1]))
    d_small = (_to_device(torch, small[0]), _to_device(torch, small[1]))
    out = torch.full((4, 2), float("nan"), dtype=torch.float64, device="cuda:0")
    stream = torch.cuda.current_stream().cuda_stream

    def call(bra, ket, row):
        xb.capi.inner_noalloc(bra.data_ptr(), ket.data_ptr(), bra.numel() // 2, out[row].data_ptr(), stream)

    # first call on this device: resolves the address of the static array (kept outside any capture on purpose)
    call(d_big[0], d_big[1], 0)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    # back to back on one stream, no synchronisation in between: different sizes, then the first size again
    call(d_small[0], d_small[1], 1)
    call(d_big[0], d_big[0], 2)
    call(d_big[0], d_big[1], 3)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before, "the allocation-free entry point changed the free device memory"
    got = [complex(r[0], r[1]) for r in out.cpu().numpy()]
    assert _close(got[0], big[2]), (got[0], big[2])
    assert _close(got[1], small[2]), (got[1], small[2])
    assert _close(got[2], big[3]), (got[2], big[3])
    assert _close(got[3], big[2]), (got[3], big[2])
    # and against plain numpy on the downloaded vectors, so that a common mistake of both entry points would show
    assert _close(got[3], np.vdot(big[0], big[1]))
    assert abs(got[2].real - 1.0) < 1e-12 and abs(got[2].imag) < 1e-12


def test_inner_noalloc_refuses_what_it_cannot_do(xb):
    import torch
    torch.cuda.set_device(0)
    v = torch.zeros(64, dtype=torch.float64, device="cuda:0")
    out = torch.zeros(2, dtype=torch.float64, device="cuda:0")
    with pytest.raises(xb.B200Error) as ei:       # beyond the stated maximum: refused before anything is launched
        xb.capi.inner_noalloc(v.data_ptr(), v.data_ptr(), xb.capi.NOALLOC_MAX_AMPLITUDES + 1, out.data_ptr())
    assert ei.value.code == -1 and "B200VQE_NOALLOC_MAX_AMPLITUDES" in str(ei.value)
    with pytest.raises(xb.B200Error) as ei:
        xb.capi.inner_noalloc(v.data_ptr(), v.data_ptr(), 0, out.data_ptr())
    assert ei.value.code == -1
    host = np.zeros(64)
    with pytest.raises(xb.B200Error) as ei:       # a host array where a device vector is expected
        xb.capi.inner_noalloc(host.ctypes.data, v.data_ptr(), 32, out.data_ptr())
    assert ei.value.code == -1 and "bra" in str(ei.value)
    with pytest.raises(xb.B200Error) as ei:
        xb.capi.inner_noalloc(v.data_ptr() + 8, v.data_ptr(), 16, out.data_ptr())
    assert "aligned" in str(ei.value)
    # a refused call leaves the entry point usable
    v.fill_(0.5)
    xb.capi.inner_noalloc(v.data_ptr(), v.data_ptr(), 32, out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert out.cpu().tolist() == [16.0, 0.0]
