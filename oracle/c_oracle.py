"""ctypes loader for oracle/hpsi_oracle.c -- TEST INFRASTRUCTURE ONLY (checker / CPU baseline)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build() -> str:
    so = os.path.join(_HERE, "libhpsi_oracle.so")
    src = os.path.join(_HERE, "hpsi_oracle.c")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "libhpsi_oracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.oracle_num_threads.restype = C.c_int
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _tables(xm, zm, cf):
    xm = np.ascontiguousarray(xm, dtype=np.uint64)
    zm = np.ascontiguousarray(zm, dtype=np.uint64)
    cf = np.asarray(cf, dtype=np.complex128)
    return xm, zm, np.ascontiguousarray(cf.real), np.ascontiguousarray(cf.imag)


def num_threads() -> int:
    return int(lib().oracle_num_threads())


def apply(n, xm, zm, cf, psi):
    xm, zm, cr, ci = _tables(xm, zm, cf)
    psi = np.ascontiguousarray(psi, dtype=np.complex128)
    out = np.empty_like(psi)
    lib().oracle_apply(C.c_int(n), C.c_int64(len(xm)), _p(xm), _p(zm), _p(cr), _p(ci), _p(psi), _p(out))
    return out


def expect(n, xm, zm, cf, psi):
    xm, zm, cr, ci = _tables(xm, zm, cf)
    psi = np.ascontiguousarray(psi, dtype=np.complex128)
    e = np.zeros(2)
    per = np.zeros(len(xm), dtype=np.complex128)
    lib().oracle_expect(C.c_int(n), C.c_int64(len(xm)), _p(xm), _p(zm), _p(cr), _p(ci), _p(psi), _p(e), _p(per))
    return complex(e[0], e[1]), per


def pauli_rotation(n, x, z, phi, psi):
    assert psi.dtype == np.complex128 and psi.flags.c_contiguous
    lib().oracle_pauli_rotation(C.c_int(n), C.c_uint64(int(x)), C.c_uint64(int(z)), C.c_double(phi), _p(psi))
    return psi


GATE_ID = {"X": 0, "Y": 1, "Z": 2, "H": 3, "Rx": 4, "Ry": 5, "Rz": 6, "CNOT": 7}


def apply_gate(n, name, qubits, theta, psi):
    assert psi.dtype == np.complex128 and psi.flags.c_contiguous
    q0 = int(qubits[0])
    q1 = int(qubits[1]) if len(qubits) > 1 else 0
    lib().oracle_apply_gate(C.c_int(n), C.c_int(GATE_ID[name]), C.c_int(q0), C.c_int(q1), C.c_double(theta), _p(psi))
    return psi


def _ops_table(op, n):
    keys = list(op.keys())
    buf = bytearray(b"I" * (len(keys) * n))
    for t, key in enumerate(keys):
        for q, p in key[1]:
            buf[t * n + q] = ord(p)
    c = np.array([op[k] for k in keys], dtype=np.complex128)
    return bytes(buf), np.ascontiguousarray(c.real), np.ascontiguousarray(c.imag), len(keys)


def faithful_dense(op, n):
    ops, cr, ci, T = _ops_table(op, n)
    dim = 1 << n
    A = np.empty((dim, dim), dtype=np.complex128)
    lib().oracle_faithful_dense(C.c_int(n), C.c_int64(T), C.c_char_p(ops), _p(cr), _p(ci), _p(A))
    return A


def faithful_apply_rows(op, n, psi_ref_order, row_begin, row_end):
    ops, cr, ci, T = _ops_table(op, n)
    psi = np.ascontiguousarray(psi_ref_order, dtype=np.complex128)
    out = np.empty(row_end - row_begin, dtype=np.complex128)
    lib().oracle_faithful_apply(C.c_int(n), C.c_int64(T), C.c_char_p(ops), _p(cr), _p(ci), _p(psi), _p(out),
                                C.c_uint64(row_begin), C.c_uint64(row_end))
    return out
