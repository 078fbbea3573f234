"""ctypes binding of libb200vqe.so (include/b200vqe.h).  No CPU fallback: if the library is
missing or no CUDA device is present, the device entry points raise."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200vqe.so")
_lib = None

GATE_ID = {"X": 0, "Y": 1, "Z": 2, "H": 3, "Rx": 4, "Ry": 5, "Rz": 6, "CNOT": 7}
ORDER_DEVICE, ORDER_REFERENCE = 0, 1


class Gate(C.Structure):
    _fields_ = [("gate", C.c_int32), ("q0", C.c_int32), ("q1", C.c_int32), ("pad", C.c_int32),
                ("theta", C.c_double)]


class LanczosOpts(C.Structure):
    _fields_ = [("tol", C.c_double), ("max_iter", C.c_int32), ("want_vector", C.c_int32),
                ("seed", C.c_uint64), ("use_slot0", C.c_int32), ("reserved", C.c_int32)]


class B200Error(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"b200vqe error {code}: {msg}")
        self.code = code


def build(force: bool = False) -> str:
    """Compile the library in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    csrc = os.path.join(_HERE, "csrc")
    if force:
        subprocess.check_call(["make", "-C", csrc, "clean"], stdout=subprocess.DEVNULL)
    subprocess.check_call(["make", "-C", csrc], stdout=subprocess.DEVNULL)
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200Error(-2, f"{LIB_PATH} is not built; run __graft_entry__.build() (there is no CPU fallback)")
        _lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        for name in ("b200vqe_last_error", "b200vqe_global_error", "b200vqe_front_error"):
            getattr(_lib, name).restype = C.c_char_p
        _lib.b200vqe_last_error.argtypes = [C.c_void_p]
    return _lib


def _vp(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _u64(a):
    return np.ascontiguousarray(a, dtype=np.uint64)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


NOALLOC_MAX_AMPLITUDES = 1 << 33   # B200VQE_NOALLOC_MAX_AMPLITUDES


def inner_noalloc(bra_ptr: int, ket_ptr: int, n_amplitudes: int, out_ptr: int, stream: int = 0) -> None:
    """b200vqe_inner_noalloc: <bra|ket> of two device vectors the caller owns, written to the caller's two device
    doubles at out_ptr in stream order (no synchronisation, no handle, no device allocation -- for hosts whose HBM is
    already spoken for or belongs to a pool allocator).  The arguments are raw device addresses (e.g.
    tensor.data_ptr()) on the current device and a cudaStream_t as an integer.  Calls share one static scratch array
    per GPU, so keep them ordered on the GPU; see include/b200vqe.h for the limits."""
    L = lib()
    L.b200vqe_inner_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int64]
    L.b200vqe_inner_noalloc.restype = C.c_int
    err = C.create_string_buffer(512)
    rc = L.b200vqe_inner_noalloc(C.c_void_p(bra_ptr), C.c_void_p(ket_ptr), C.c_uint64(n_amplitudes), C.c_void_p(out_ptr),
                                 C.c_void_p(stream), err, C.c_int64(len(err)))
    if rc:
        raise B200Error(rc, err.value.decode())


# --------------------------------------------------------------------------- front end
class OpHandle:
    """Owns a b200vqe_op."""

    def __init__(self, ptr):
        self.ptr = C.c_void_p(ptr) if not isinstance(ptr, C.c_void_p) else ptr

    def __del__(self):
        try:
            if self.ptr:
                lib().b200vqe_op_destroy(self.ptr)
                self.ptr = None
        except Exception:
            pass

    @staticmethod
    def _check(rc):
        if rc:
            raise B200Error(rc, lib().b200vqe_front_error().decode())

    @classmethod
    def from_fermion_source(cls, src: str, transformation: str = "jw", n_qubits: int = 0):
        out = C.c_void_p()
        cls._check(lib().b200vqe_op_from_fermion_source(src.encode(), transformation.encode(), C.c_int(n_qubits), C.byref(out)))
        return cls(out)

    @classmethod
    def from_string(cls, text: str):
        out = C.c_void_p()
        cls._check(lib().b200vqe_op_from_string(text.encode(), C.byref(out)))
        return cls(out)

    @classmethod
    def from_terms(cls, xmask, zmask, coeff):
        xm, zm = _u64(xmask), _u64(zmask)
        c = np.asarray(coeff, dtype=np.complex128)
        cr, ci = _f64(c.real), _f64(c.imag)
        out = C.c_void_p()
        cls._check(lib().b200vqe_op_from_terms(C.c_int64(len(xm)), _vp(xm), _vp(zm), _vp(cr), _vp(ci), C.byref(out)))
        return cls(out)

    def n_terms(self) -> int:
        n = C.c_int64()
        self._check(lib().b200vqe_op_n_terms(self.ptr, C.byref(n)))
        return n.value

    def n_qubits(self) -> int:
        n = C.c_int()
        self._check(lib().b200vqe_op_n_qubits(self.ptr, C.byref(n)))
        return n.value

    def pack(self):
        """-> (xmask u64[T], zmask u64[T], coef c128[T] = c*i^nY) sorted by (xmask, zmask)."""
        T = self.n_terms()
        xm, zm = np.zeros(T, np.uint64), np.zeros(T, np.uint64)
        cr, ci = np.zeros(T), np.zeros(T)
        if T:
            self._check(lib().b200vqe_op_pack(self.ptr, _vp(xm), _vp(zm), _vp(cr), _vp(ci)))
        return xm, zm, cr + 1j * ci

    def _string(self, fn, *args) -> str:
        need = C.c_int64()
        self._check(fn(self.ptr, *args, None, C.c_int64(0), C.byref(need)))
        buf = C.create_string_buffer(need.value)
        self._check(fn(self.ptr, *args, buf, C.c_int64(need.value), C.byref(need)))
        return buf.value.decode()

    def to_string(self) -> str:
        return self._string(lib().b200vqe_op_to_string)

    def term_ids(self):
        return [self._string(lib().b200vqe_op_term_id, C.c_int64(k)) for k in range(self.n_terms())]

    def _binary(self, fn, other):
        out = C.c_void_p()
        self._check(fn(self.ptr, other.ptr, C.byref(out)))
        return OpHandle(out)

    def add(self, other):
        return self._binary(lib().b200vqe_op_add, other)

    def mul(self, other):
        return self._binary(lib().b200vqe_op_mul, other)

    def scale(self, s: complex):
        out = C.c_void_p()
        s = complex(s)
        self._check(lib().b200vqe_op_scale(self.ptr, C.c_double(s.real), C.c_double(s.imag), C.byref(out)))
        return OpHandle(out)


def uccsd_rotations(n_qubits: int, n_electrons: int, transformation: str = "jw"):
    """-> (xmask, zmask, imag_coeff, param, n_params)"""
    cnt, npar = C.c_int64(), C.c_int()
    rc = lib().b200vqe_uccsd_rotations(C.c_int(n_qubits), C.c_int(n_electrons), transformation.encode(),
                                       C.byref(cnt), C.byref(npar), None, None, None, None)
    OpHandle._check(rc)
    n = cnt.value
    xm, zm, ic, pa = np.zeros(n, np.uint64), np.zeros(n, np.uint64), np.zeros(n), np.zeros(n, np.int32)
    rc = lib().b200vqe_uccsd_rotations(C.c_int(n_qubits), C.c_int(n_electrons), transformation.encode(),
                                       C.byref(cnt), C.byref(npar), _vp(xm), _vp(zm), _vp(ic), _vp(pa))
    OpHandle._check(rc)
    return xm, zm, ic, pa, npar.value


def number_sector_rows(n_qubits: int, transformation: str = "jw"):
    """rows[i]: occupation bit i of basis index j = parity(rows[i] & j) (task/tasks/DiagonalizeTask.cpp:54-144)."""
    rows = np.zeros(n_qubits, dtype=np.uint64)
    rc = lib().b200vqe_number_sector_rows(C.c_int(n_qubits), transformation.encode(), _vp(rows))
    if rc:
        raise B200Error(rc, lib().b200vqe_front_error().decode())
    return rows


def uccsd_gates(n_qubits: int, n_electrons: int, theta, transformation: str = "jw"):
    th = _f64(theta)
    cnt = C.c_int64()
    rc = lib().b200vqe_uccsd_gates(C.c_int(n_qubits), C.c_int(n_electrons), transformation.encode(), _vp(th),
                                   C.c_int(len(th)), C.byref(cnt), None)
    OpHandle._check(rc)
    arr = (Gate * cnt.value)()
    rc = lib().b200vqe_uccsd_gates(C.c_int(n_qubits), C.c_int(n_electrons), transformation.encode(), _vp(th),
                                   C.c_int(len(th)), C.byref(cnt), arr)
    OpHandle._check(rc)
    return arr


# --------------------------------------------------------------------------- device backend
class Backend:
    """One GPU (or one shard of a sharded register)."""

    def __init__(self, n_qubits: int, device: int = 0, rank: int = 0, world: int = 1, comm_id: bytes | None = None):
        L = lib()
        self.h = C.c_void_p()
        if world == 1:
            rc = L.b200vqe_create(C.c_int(n_qubits), C.c_int(device), C.byref(self.h))
        else:
            rc = L.b200vqe_create_sharded(C.c_int(n_qubits), C.c_int(device), C.c_int(rank), C.c_int(world),
                                          comm_id, C.byref(self.h))
        if rc:
            raise B200Error(rc, L.b200vqe_global_error().decode())
        self.n_qubits, self.rank, self.world = n_qubits, rank, world
        cnt, first = C.c_uint64(), C.c_uint64()
        L.b200vqe_local_amplitudes(self.h, C.byref(cnt), C.byref(first))
        self.n_local, self.first = cnt.value, first.value
        self.n_terms = 0

    @staticmethod
    def comm_unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        rc = lib().b200vqe_comm_unique_id(buf)
        if rc:
            raise B200Error(rc, lib().b200vqe_global_error().decode())
        return buf.raw

    def close(self):
        if getattr(self, "h", None):
            lib().b200vqe_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc:
            raise B200Error(rc, lib().b200vqe_last_error(self.h).decode())

    def set_option(self, key: str, value):
        """String options (xacc::setOption style); take effect at the next set_hamiltonian."""
        self._ck(lib().b200vqe_set_option(self.h, key.encode(), str(value).encode()))

    def n_slots(self, n: int):
        self._ck(lib().b200vqe_n_slots(self.h, C.c_int(n)))

    def set_hamiltonian(self, xmask, zmask, coef):
        xm, zm = _u64(xmask), _u64(zmask)
        c = np.asarray(coef, dtype=np.complex128)
        cr, ci = _f64(c.real), _f64(c.imag)
        self._ck(lib().b200vqe_set_hamiltonian(self.h, C.c_int64(len(xm)), _vp(xm), _vp(zm), _vp(cr), _vp(ci)))
        self.n_terms = len(xm)

    def hamiltonian_stats(self):
        t, g = C.c_int64(), C.c_int64()
        self._ck(lib().b200vqe_hamiltonian_stats(self.h, C.byref(t), C.byref(g)))
        return t.value, g.value

    def plan_stats(self):
        """Tile-engine plan: batches (= HBM passes over this rank's amplitudes) of <psi|H|psi>, xmask groups, batches of H psi
        (0 until the first H psi), x_rank classes, chunk bits, tile bits."""
        out = (C.c_int64 * 6)()
        self._ck(lib().b200vqe_plan_stats(self.h, out))
        return dict(zip(("expect_batches", "groups", "apply_batches", "expect_classes", "chunk_bits", "tile_bits"), [int(v) for v in out]))

    def plan_costs(self):
        """Per-tile cost of the <psi|H|psi> plan, summed over its batches (b200vqe_plan_costs)."""
        out = (C.c_int64 * 4)()
        self._ck(lib().b200vqe_plan_costs(self.h, out))
        return dict(zip(("view_loads", "fused_blocks", "plain_records", "partner_records"), [int(v) for v in out]))

    def last_exchange(self):
        """(ms, bytes received by this rank) of the NVLink transfers of the last sharded <psi|H|psi> / H psi."""
        ms, by = C.c_double(), C.c_double()
        self._ck(lib().b200vqe_last_exchange(self.h, C.byref(ms), C.byref(by)))
        return ms.value, by.value

    def last_exchange_exposed(self) -> float:
        """ms the compute stream waited for the NVLink transfers of the last sharded <psi|H|psi> / H psi (the un-hidden part)."""
        ms = C.c_double()
        self._ck(lib().b200vqe_last_exchange_exposed(self.h, C.byref(ms)))
        return ms.value

    def exchange_mode(self) -> str:
        """Transport of the slab swaps of the sharded <psi|H|psi>: 'peer-copy' | 'nccl' | 'undecided' (b200vqe_exchange_mode)."""
        m = C.c_int()
        self._ck(lib().b200vqe_exchange_mode(self.h, C.byref(m)))
        return {1: "peer-copy", -1: "nccl"}.get(m.value, "undecided")

    def reset(self, slot: int = 0, basis_index: int = 0):
        self._ck(lib().b200vqe_state_reset(self.h, C.c_int(slot), C.c_uint64(basis_index)))

    def upload(self, psi, slot: int = 0, first: int = 0):
        a = np.ascontiguousarray(psi, dtype=np.complex128)
        self._ck(lib().b200vqe_state_upload(self.h, C.c_int(slot), _vp(a), C.c_uint64(first), C.c_uint64(a.shape[0])))

    def download(self, slot: int = 0, first: int = 0, count: int | None = None, order: int = ORDER_DEVICE, out=None):
        count = self.n_local - first if count is None else count
        a = np.empty(count, dtype=np.complex128) if out is None else out
        self._ck(lib().b200vqe_state_download(self.h, C.c_int(slot), _vp(a), C.c_uint64(first), C.c_uint64(count), C.c_int(order)))
        return a

    def random_state(self, slot: int = 0, seed: int = 0):
        self._ck(lib().b200vqe_state_random(self.h, C.c_int(slot), C.c_uint64(seed)))

    def product_state(self, a, b, slot: int = 0):
        a, b = _f64(a), _f64(b)
        assert len(a) == self.n_qubits and len(b) == self.n_qubits
        self._ck(lib().b200vqe_state_product(self.h, C.c_int(slot), _vp(a), _vp(b)))

    def norm2(self, slot: int = 0) -> float:
        v = C.c_double()
        self._ck(lib().b200vqe_state_norm2(self.h, C.c_int(slot), C.byref(v)))
        return v.value

    def inner(self, bra: int, ket: int) -> complex:
        v = (C.c_double * 2)()
        self._ck(lib().b200vqe_state_inner(self.h, C.c_int(bra), C.c_int(ket), v))
        return complex(v[0], v[1])

    def apply_gates(self, gates, slot: int = 0):
        """gates: iterable of (name, qubits, theta) or a ctypes Gate array."""
        if not isinstance(gates, C.Array):
            lst = list(gates)
            arr = (Gate * len(lst))()
            for i, (name, qs, th) in enumerate(lst):
                arr[i].gate = GATE_ID[name]
                arr[i].q0 = int(qs[0])
                arr[i].q1 = int(qs[1]) if len(qs) > 1 else 0
                arr[i].theta = float(th)
            gates = arr
        self._ck(lib().b200vqe_apply_gates(self.h, C.c_int(slot), C.c_int64(len(gates)), gates))

    def apply_pauli_rotations(self, xmask, zmask, phi, slot: int = 0):
        xm, zm, ph = _u64(xmask), _u64(zmask), _f64(phi)
        self._ck(lib().b200vqe_apply_pauli_rotations(self.h, C.c_int(slot), C.c_int64(len(xm)), _vp(xm), _vp(zm), _vp(ph)))

    def expectation(self, slot: int = 0, per_term: bool = False):
        e = (C.c_double * 2)()
        pt = np.zeros(self.n_terms) if per_term else None
        self._ck(lib().b200vqe_expectation(self.h, C.c_int(slot), e, _vp(pt)))
        return (complex(e[0], e[1]), pt) if per_term else complex(e[0], e[1])

    def apply_h(self, src: int, dst: int, want_alpha: bool = True):
        a = (C.c_double * 2)()
        self._ck(lib().b200vqe_apply_h(self.h, C.c_int(src), C.c_int(dst), a if want_alpha else None))
        return complex(a[0], a[1]) if want_alpha else None

    def lanczos_ground(self, tol: float = 1e-10, max_iter: int = 1000, want_vector: bool = False, seed: int = 1,
                       use_slot0: bool = False, resume: bool = False):
        """resume=True continues the unfinished run kept in the handle (stopped at max_iter, or lanczos_load)."""
        o = LanczosOpts(tol, max_iter, int(want_vector), seed, int(use_slot0), int(resume))
        ev, it, res = C.c_double(), C.c_int32(), C.c_double()
        self._ck(lib().b200vqe_lanczos_ground(self.h, C.byref(o), C.byref(ev), C.byref(it), C.byref(res)))
        return ev.value, it.value, res.value

    def lanczos_save(self, path: str):
        self._ck(lib().b200vqe_lanczos_save(self.h, os.fsencode(path)))

    def lanczos_load(self, path: str):
        self._ck(lib().b200vqe_lanczos_load(self.h, os.fsencode(path)))

    def set_number_sector(self, n_electrons: int, transformation: str = "jw"):
        """diag-number-symmetry (task/tasks/DiagonalizeTask.cpp:54-183): Lanczos runs on P H P; n_electrons < 0 clears."""
        self._ck(lib().b200vqe_set_number_sector(self.h, C.c_int(n_electrons), transformation.encode()))

    def project_sector(self, slot: int = 0):
        self._ck(lib().b200vqe_state_project_sector(self.h, C.c_int(slot)))

    def sweep_energies(self, basis_index, xmask, zmask, imag_coeff, param, n_params, thetas):
        xm, zm, ic = _u64(xmask), _u64(zmask), _f64(imag_coeff)
        pa = np.ascontiguousarray(param, dtype=np.int32)
        th = _f64(thetas).reshape(-1, n_params)
        out = np.zeros(th.shape[0])
        self._ck(lib().b200vqe_sweep_energies(self.h, C.c_uint64(basis_index), C.c_int64(len(xm)), _vp(xm), _vp(zm), _vp(ic),
                                              _vp(pa), C.c_int(n_params), C.c_int64(th.shape[0]), _vp(th), _vp(out)))
        return out

    def synchronize(self):
        self._ck(lib().b200vqe_synchronize(self.h))

    def kernel_launches(self) -> int:
        v = C.c_int64()
        self._ck(lib().b200vqe_kernel_launches(self.h, C.byref(v)))
        return v.value

    def timer_start(self):
        self._ck(lib().b200vqe_timer_start(self.h))

    def timer_stop(self) -> float:
        v = C.c_double()
        self._ck(lib().b200vqe_timer_stop(self.h, C.byref(v)))
        return v.value

    def last_kernel_ms(self) -> float:
        v = C.c_double()
        self._ck(lib().b200vqe_last_kernel_ms(self.h, C.byref(v)))
        return v.value
