"""CPU ORACLE -- TEST INFRASTRUCTURE ONLY.

A plain numpy / pure-Python restatement of the algorithm the reference
(ORNL-QCI/xacc-vqe, /root/reference) runs on its Hamiltonian x statevector path.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module, and only as the checker. The product
(xacc-vqe_b200/) never imports anything under oracle/.

Parity status: PINNED for the JW/BK Pauli term sets, the H2 ground energy
(exact double of python/tests/eigen_diagonalize_test.py:43), the UCCSD energy
(task/tests/ComputeEnergyVQETaskTester.cpp:99-106, 1e-4) and the 2-RDM energy
(decorators/tests/RDMGeneratorTester.cpp:132,162) -- see tests/test_oracle_golden.py.
UNPINNED (no reference test holds a value): per-row computeActionOnBra outputs,
eigenvector element order, anything above 6 qubits. The Pauli algebra itself lives
in the external XACC framework (eclipse/xacc >= 0.1.15, setup.py:88), which is not
vendored under /root/reference; its semantics are restated here from the
reference's call sites and goldens:
  * term id = concatenated <P><q> in ascending qubit order, "I" for identity
    (python/xaccvqe.py:149-152, ComputeEnergyVQETask.cpp:98-105)
  * text form "(re,im) X0 Z1 + ..." (compiler/optimizers/tests/QubitTaperingTester.cpp:72)
  * action on a bit string indexes the string by qubit number
    (task/tasks/DiagonalizeTask.cpp:189-203 builds the string MSB first)

Conventions used by every function below
  * an operator is a dict  {(var, ((q,'X'),(q','Z'),...)) : complex}  with the
    inner tuple sorted by qubit; var is the symbolic variable name ("" if none)
  * statevectors are numpy complex128 arrays indexed with qubit q <-> bit q of the
    index ("device order"); the reference's dense matrix order (qubit q <-> bit
    n-1-q, DiagonalizeTask.cpp:189-193) is produced by dense_matrix_reference().
"""
from __future__ import annotations

import math
import re
from typing import Dict, Iterable, List, Sequence, Tuple

import numpy as np

Key = Tuple[str, Tuple[Tuple[int, str], ...]]
Op = Dict[Key, complex]

DROP_TOL = 1e-12  # XACC PauliOperator drops terms with |c| below this on +=

# single-qubit Pauli product table: (a, b) -> (phase, result)
_PROD = {
    ("X", "Y"): (1j, "Z"), ("Y", "X"): (-1j, "Z"),
    ("Y", "Z"): (1j, "X"), ("Z", "Y"): (-1j, "X"),
    ("Z", "X"): (1j, "Y"), ("X", "Z"): (-1j, "Y"),
}


# --------------------------------------------------------------------------
# Pauli algebra (restates xacc::quantum::PauliOperator as used by
# transformations/jw/JordanWignerIRTransformation.cpp:35-54)
# --------------------------------------------------------------------------
def pauli(ops: Dict[int, str] | None = None, coeff: complex = 1.0, var: str = "") -> Op:
    ops = ops or {}
    key = tuple(sorted((int(q), p) for q, p in ops.items() if p != "I"))
    return {(var, key): complex(coeff)}


def op_add(a: Op, b: Op) -> Op:
    out = dict(a)
    for k, c in b.items():
        s = out.get(k, 0.0) + c
        if abs(s) < DROP_TOL:
            out.pop(k, None)
        else:
            out[k] = s
    return out


def _term_mul(k1: Key, c1: complex, k2: Key, c2: complex) -> Tuple[Key, complex]:
    var = k1[0] + k2[0]
    ops = dict(k1[1])
    c = c1 * c2
    for q, p in k2[1]:
        if q not in ops:
            ops[q] = p
        elif ops[q] == p:
            del ops[q]
        else:
            ph, r = _PROD[(ops[q], p)]
            c *= ph
            ops[q] = r
    return (var, tuple(sorted(ops.items()))), c


def op_mul(a: Op, b: Op) -> Op:
    out: Op = {}
    for k1, c1 in a.items():
        for k2, c2 in b.items():
            k, c = _term_mul(k1, c1, k2, c2)
            out[k] = out.get(k, 0.0) + c
    return {k: c for k, c in out.items() if abs(c) >= DROP_TOL}


def op_scale(a: Op, s: complex) -> Op:
    return {k: c * s for k, c in a.items()}


def term_id(key: Key) -> str:
    """var + 'X0Z1...' ; 'I' for the identity (python/xaccvqe.py:149-152)."""
    s = key[0] + "".join(f"{p}{q}" for q, p in key[1])
    return s if s else "I"


def op_n_qubits(op: Op) -> int:
    m = -1
    for (_, ops) in op:
        for q, _p in ops:
            m = max(m, q)
    return m + 1


def op_to_string(op: Op) -> str:
    """'(re,im) X0 Z1 + (re,im) I' -- the XACC text form (QubitTaperingTester.cpp:72)."""
    parts = []
    for k in sorted(op, key=term_id):
        c = op[k]
        body = " ".join(f"{p}{q}" for q, p in k[1]) or "I"
        v = f" {k[0]}" if k[0] else ""
        parts.append(f"({c.real:.12g},{c.imag:.12g}){v} {body}")
    return " + ".join(parts)


_TERM_RE = re.compile(r"\(\s*([-+0-9.eE]+)\s*,\s*([-+0-9.eE]+)\s*\)\s*([^+]*)")


def op_from_string(s: str) -> Op:
    """Parse '(re,im) X0 Y1 + ...' (PauliOperator::fromString as used at
    compiler/optimizers/tests/QubitTaperingTester.cpp:72,149,156)."""
    out: Op = {}
    for m in _TERM_RE.finditer(s):
        c = complex(float(m.group(1)), float(m.group(2)))
        ops = {}
        for tok in m.group(3).split():
            if tok == "I":
                continue
            ops[int(tok[1:])] = tok[0]
        out = op_add(out, pauli(ops, c))
    return out


# --------------------------------------------------------------------------
# FermionCompiler text format (compiler/FermionCompiler.cpp:62-103)
# --------------------------------------------------------------------------
def parse_fermion_source(src: str) -> Tuple[List[Tuple[complex, List[Tuple[int, int]]]], int]:
    """Returns ([(coeff, [(site, isCreation), ...]), ...], nQubits)."""
    lines = src.split("\n")
    body = lines[1:-1]
    terms = []
    nq = 0
    for line in body:
        t = line.strip()
        if not t or not re.search(r"[0-9]", t):
            continue
        toks = t.split()
        coeff = float(toks[0])
        ops = []
        # for (i = 1; i < size-1; i += 2)  -- FermionCompiler.cpp:84
        i = 1
        while i < len(toks) - 1:
            site = int(toks[i])
            nq = max(nq, site)
            ops.append((site, int(toks[i + 1])))
            i += 2
        terms.append((complex(coeff), ops))
    return terms, nq + 1


# --------------------------------------------------------------------------
# Jordan-Wigner (transformations/jw/JordanWignerIRTransformation.cpp:22-55)
# --------------------------------------------------------------------------
def jordan_wigner(terms: Iterable[Tuple[complex, Sequence[Tuple[int, int]]]],
                  variables: Sequence[str] | None = None) -> Op:
    result: Op = {}
    for t_idx, (coeff, sites) in enumerate(terms):
        var = variables[t_idx] if variables is not None else ""
        current = pauli({}, coeff, var)
        for (index, is_creation) in sites:
            ycoeff = -0.5j if is_creation else 0.5j
            zs = pauli({j: "Z" for j in range(index)})
            ladder = op_add(pauli({index: "X"}, 0.5), pauli({index: "Y"}, ycoeff))
            current = op_mul(current, op_mul(zs, ladder))
        result = op_add(result, current)
    return result


# --------------------------------------------------------------------------
# Bravyi-Kitaev (transformations/bk/BravyiKitaevIRTransformation.cpp:25-75,
#                transformations/bk/Fenwick.hpp:83-176)
# --------------------------------------------------------------------------
class FenwickTree:
    def __init__(self, n: int):
        self.parent = [-1] * n
        self.children: List[set] = [set() for _ in range(n)]
        if n > 0:
            root = n - 1

            def construct(i1: int, i2: int, par: int) -> None:
                if i1 >= i2:
                    return
                pvt = (i1 + i2) >> 1
                self.children[par].add(pvt)
                self.parent[pvt] = par
                construct(i1, pvt, pvt)
                construct(pvt + 1, i2, par)

            construct(0, n - 1, root)

    def update_set(self, i: int) -> set:          # ancestors, Fenwick.hpp:41-49
        out = set()
        while self.parent[i] >= 0:
            out.add(self.parent[i])
            i = self.parent[i]
        return out

    def children_set(self, i: int) -> set:
        return set(self.children[i])

    def remainder_set(self, i: int) -> set:       # Fenwick.hpp:147-160
        out = set()
        for a in self.update_set(i):
            for c in self.children[a]:
                if c < i:
                    out.add(c)
        return out

    def parity_set(self, i: int) -> set:          # Fenwick.hpp:168-176
        return self.remainder_set(i) | self.children_set(i)


def bravyi_kitaev(terms: Iterable[Tuple[complex, Sequence[Tuple[int, int]]]], n_qubits: int,
                  variables: Sequence[str] | None = None) -> Op:
    tree = FenwickTree(n_qubits)
    result: Op = {}
    for t_idx, (coeff, sites) in enumerate(terms):
        var = variables[t_idx] if variables is not None else ""
        prod = pauli({}, coeff, var)
        for (index, is_creation) in sites:
            dcoeff = -0.5j if is_creation else 0.5j
            d_terms = {index: "Y"}
            c_terms = {index: "X"}
            for ac in tree.remainder_set(index):
                d_terms[ac] = "Z"
            for p in tree.parity_set(index):
                c_terms[p] = "Z"
            for a in tree.update_set(index):
                d_terms[a] = "X"
                c_terms[a] = "X"
            prod = op_mul(prod, op_add(pauli(c_terms, 0.5), pauli(d_terms, dcoeff)))
        result = op_add(result, prod)
    return result


def compile_fermion_source(src: str, transformation: str = "jw", n_qubits: int | None = None) -> Tuple[Op, int]:
    """FermionCompiler::compile (compiler/FermionCompiler.cpp:62-147)."""
    terms, nq = parse_fermion_source(src)
    if n_qubits is not None:
        nq = n_qubits
    if transformation == "jw":
        return jordan_wigner(terms), nq
    if transformation == "bk":
        return bravyi_kitaev(terms, nq), nq
    raise ValueError(transformation)


# --------------------------------------------------------------------------
# String-based term action and the dense build
# (task/tasks/DiagonalizeTask.cpp:184-206; XACC Term::computeActionOnBra/Ket)
# --------------------------------------------------------------------------
def action_on_bra(op: Op, bits: str) -> List[Tuple[str, complex]]:
    out = []
    for (_var, ops), coeff in op.items():
        nb = list(bits)
        c = coeff
        for q, p in ops:
            if p == "Z":
                c *= -1 if nb[q] == "1" else 1
            elif p == "X":
                nb[q] = "0" if nb[q] == "1" else "1"
            elif p == "Y":
                if nb[q] == "1":
                    c *= 1j
                    nb[q] = "0"
                else:
                    c *= -1j
                    nb[q] = "1"
        out.append(("".join(nb), c))
    return out


def action_on_ket(op: Op, bits: str) -> List[Tuple[str, complex]]:
    out = []
    for (_var, ops), coeff in op.items():
        nb = list(bits)
        c = coeff
        for q, p in ops:
            if p == "Z":
                c *= -1 if nb[q] == "1" else 1
            elif p == "X":
                nb[q] = "0" if nb[q] == "1" else "1"
            elif p == "Y":
                if nb[q] == "1":
                    c *= -1j
                    nb[q] = "0"
                else:
                    c *= 1j
                    nb[q] = "1"
        out.append(("".join(nb), c))
    return out


def dense_matrix_reference(op: Op, n: int) -> np.ndarray:
    """A(row, col) built row by row exactly like DiagonalizeTask.cpp:189-203:
    row bit string MSB first, one (string, coeff) per term from computeActionOnBra,
    column = stol(string, 2)."""
    dim = 1 << n
    A = np.zeros((dim, dim), dtype=np.complex128)
    for row in range(dim):
        bits = "".join(str((row >> k) & 1) for k in range(n - 1, -1, -1))
        for s, c in action_on_bra(op, bits):
            A[row, int(s, 2)] += c
    return A


def diagonalize_reference(op: Op, n: int) -> Tuple[float, np.ndarray]:
    """EigenDiagonalizeBackend::diagonalize dense branch (DiagonalizeTask.cpp:184-214).
    Returns (eigenvalues(0), eigenvector column 0 in the reference's index order)."""
    A = dense_matrix_reference(op, n)
    w, v = np.linalg.eigh(A)
    return float(w[0]), v[:, 0]


def bk_sector_matrix(n: int) -> np.ndarray:
    """The GF(2) matrix of DiagonalizeTask.cpp:76-121 ("TranterBK"): Seeley's recursion (:80-104: old block in the two
    diagonal quadrants, ones in the first row of the upper-right quadrant), bottom-right n x n block when the doubling
    overshoots (:98-103), then rows and columns reversed (:108-121; the reference's swap loops only terminate for even n)."""
    B = np.array([[1]], dtype=np.int64)
    while True:
        old = B
        d = old.shape[0]
        B = np.zeros((2 * d, 2 * d), dtype=np.int64)
        B[:d, :d] = old
        B[d:, d:] = old
        B[0, d:] = 1
        if B.shape[0] == n:
            break
        if B.shape[0] > n:
            B = B[B.shape[0] - n:, B.shape[1] - n:]
            break
    return B[::-1, ::-1].copy()


def number_sector_strings(n: int, n_electrons: int, transformation: str = "jw") -> List[str]:
    """Basis strings of the diag-number-symmetry branch in the reference's enumeration order (DiagonalizeTask.cpp:58-72:
    std::next_permutation from 0..01..1 = ascending lexicographic), re-encoded y = B x mod 2 for "bk" (:123-141)."""
    import itertools
    strings = []
    for ones in itertools.combinations(range(n), n_electrons):
        b = ["0"] * n
        for i in ones:
            b[i] = "1"
        strings.append("".join(b))
    strings.sort()
    if transformation == "bk":
        B = bk_sector_matrix(n)
        out = []
        for s in strings:
            x = np.array([1 if ch == "1" else 0 for ch in s], dtype=np.int64)
            y = (B @ x) % 2
            out.append("".join("1" if v else "0" for v in y))
        strings = out
    return strings


def number_subspace_matrix(op: Op, n: int, n_electrons: int, transformation: str = "jw", literal: bool = True) -> np.ndarray:
    """diag-number-symmetry branch (DiagonalizeTask.cpp:146-178).  literal=True keeps the reference's std::map behaviour:
    an image OUTSIDE the sector makes bitsToIdx[...] default-insert index 0 (:172), so its coefficient lands in column 0
    (rows i != 0).  literal=False is the compression P H P.  The two agree whenever H maps the sector into itself."""
    strings = number_sector_strings(n, n_electrons, transformation)
    idx = {s: i for i, s in enumerate(strings)}
    M = np.zeros((len(strings), len(strings)), dtype=np.complex128)
    for i, s in enumerate(strings):
        res = action_on_ket(op, s)
        for r, c in res:
            if r == s:
                M[i, i] += c
        for r, c in res:
            if r in idx:
                k = idx[r]
            elif literal:
                k = 0
            else:
                continue
            if i != k:
                M[i, k] += c
    return M


def diagonalize_number_symmetry(op: Op, n: int, n_electrons: int, transformation: str = "jw", literal: bool = True) -> float:
    """eigenvalues(0) of Eigen::SelfAdjointEigenSolver on that matrix (DiagonalizeTask.cpp:179-181); Eigen reads the
    lower triangle only."""
    M = number_subspace_matrix(op, n, n_electrons, transformation, literal)
    return float(np.linalg.eigvalsh(M, UPLO="L")[0])


# --------------------------------------------------------------------------
# Packing  (semantics of Term::toBinaryVector used at
# compiler/optimizers/QubitTapering.cpp:16-28 and CommutingSetGenerator.hpp:24-36)
# --------------------------------------------------------------------------
def pack_terms(op: Op) -> Tuple[np.ndarray, np.ndarray, np.ndarray, List[str]]:
    """-> (xmask u64[T], zmask u64[T], coef c128[T] = c * i^nY, ids) sorted by
    (xmask, zmask).  P|j> = i^nY (-1)^popc(j & z) |j ^ x>, qubit q <-> bit q."""
    rows = []
    for key, c in op.items():
        x = z = 0
        ny = 0
        for q, p in key[1]:
            if p in ("X", "Y"):
                x |= 1 << q
            if p in ("Z", "Y"):
                z |= 1 << q
            if p == "Y":
                ny += 1
        rows.append((x, z, complex(c) * (1j ** (ny % 4)), term_id(key)))
    rows.sort(key=lambda r: (r[0], r[1]))
    xm = np.array([r[0] for r in rows], dtype=np.uint64)
    zm = np.array([r[1] for r in rows], dtype=np.uint64)
    cf = np.array([r[2] for r in rows], dtype=np.complex128)
    return xm, zm, cf, [r[3] for r in rows]


def _parity(a: np.ndarray) -> np.ndarray:
    a = a.copy()
    for s in (32, 16, 8, 4, 2, 1):
        a ^= a >> np.uint64(s)
    return (a & np.uint64(1)).astype(np.int64)


def apply_masks(xm, zm, cf, psi: np.ndarray) -> np.ndarray:
    """(H psi)_i = sum_k coef_k (-1)^popc((i^x_k) & z_k) psi_{i ^ x_k}."""
    N = psi.shape[0]
    idx = np.arange(N, dtype=np.uint64)
    out = np.zeros(N, dtype=np.complex128)
    for x, z, c in zip(xm, zm, cf):
        j = idx ^ np.uint64(x)
        sgn = 1.0 - 2.0 * _parity(j & np.uint64(z))
        out += c * sgn * psi[j.astype(np.int64)]
    return out


def expect_masks(xm, zm, cf, psi: np.ndarray) -> Tuple[complex, np.ndarray]:
    """-> (sum_k c_k <P_k>, per-term <psi|P_k|psi> (complex; real for Hermitian P))."""
    N = psi.shape[0]
    idx = np.arange(N, dtype=np.uint64)
    per = np.zeros(len(xm), dtype=np.complex128)
    tot = 0.0 + 0.0j
    for k, (x, z, c) in enumerate(zip(xm, zm, cf)):
        ny = bin(int(x) & int(z)).count("1")
        ph = 1j ** (ny % 4)
        j = idx ^ np.uint64(x)
        sgn = 1.0 - 2.0 * _parity(j & np.uint64(z))
        v = np.vdot(psi, ph * sgn * psi[j.astype(np.int64)])
        per[k] = v
        tot += (c / ph) * v
    return tot, per


def dense_matrix_device_order(op: Op, n: int) -> np.ndarray:
    xm, zm, cf, _ = pack_terms(op)
    dim = 1 << n
    A = np.zeros((dim, dim), dtype=np.complex128)
    eye = np.eye(dim, dtype=np.complex128)
    for col in range(dim):
        A[:, col] = apply_masks(xm, zm, cf, eye[:, col])
    return A


def bit_reverse_permutation(n: int) -> np.ndarray:
    """perm[i_ref] = i_dev : reference index (qubit q <-> bit n-1-q) -> device index."""
    out = np.zeros(1 << n, dtype=np.int64)
    for i in range(1 << n):
        r = 0
        for b in range(n):
            if (i >> b) & 1:
                r |= 1 << (n - 1 - b)
        out[i] = r
    return out


# --------------------------------------------------------------------------
# Gate-level statevector simulation (qubit q <-> bit q). Gate conventions:
# R_a(t) = exp(-i t sigma_a / 2) (XACC), CNOT(ctrl, tgt)  (UCCSD.cpp:282-352)
# --------------------------------------------------------------------------
def _apply_1q(psi: np.ndarray, q: int, m: np.ndarray) -> np.ndarray:
    n = int(math.log2(psi.shape[0]))
    t = psi.reshape([2] * n)          # axis a <-> bit n-1-a
    ax = n - 1 - q
    t = np.moveaxis(t, ax, 0)
    t = np.tensordot(m, t, axes=([1], [0]))
    t = np.moveaxis(t, 0, ax)
    return np.ascontiguousarray(t).reshape(-1)


def gate_matrix(name: str, theta: float = 0.0) -> np.ndarray:
    c, s = math.cos(theta / 2), math.sin(theta / 2)
    if name == "X":
        return np.array([[0, 1], [1, 0]], dtype=np.complex128)
    if name == "Y":
        return np.array([[0, -1j], [1j, 0]], dtype=np.complex128)
    if name == "Z":
        return np.array([[1, 0], [0, -1]], dtype=np.complex128)
    if name == "H":
        return np.array([[1, 1], [1, -1]], dtype=np.complex128) / math.sqrt(2.0)
    if name == "Rx":
        return np.array([[c, -1j * s], [-1j * s, c]], dtype=np.complex128)
    if name == "Ry":
        return np.array([[c, -s], [s, c]], dtype=np.complex128)
    if name == "Rz":
        return np.array([[c - 1j * s, 0], [0, c + 1j * s]], dtype=np.complex128)
    raise ValueError(name)


def apply_gate(psi: np.ndarray, name: str, qubits: Sequence[int], theta: float = 0.0) -> np.ndarray:
    if name == "CNOT":
        ctrl, tgt = qubits
        idx = np.arange(psi.shape[0])
        src = np.where((idx >> ctrl) & 1, idx ^ (1 << tgt), idx)
        return psi[src]
    return _apply_1q(psi, qubits[0], gate_matrix(name, theta))


def run_circuit(n: int, gates: Sequence[Tuple[str, Sequence[int], float]],
                psi0: np.ndarray | None = None) -> np.ndarray:
    psi = np.zeros(1 << n, dtype=np.complex128) if psi0 is None else psi0.copy()
    if psi0 is None:
        psi[0] = 1.0
    for name, qs, th in gates:
        psi = apply_gate(psi, name, qs, th)
    return psi


def pauli_rotation(psi: np.ndarray, x: int, z: int, phi: float) -> np.ndarray:
    """exp(-i phi P) psi for the Pauli string P = (x, z) (incl. the i^nY phase)."""
    ny = bin(x & z).count("1")
    idx = np.arange(psi.shape[0], dtype=np.uint64)
    j = idx ^ np.uint64(x)
    sgn = 1.0 - 2.0 * _parity(j & np.uint64(z))
    ppsi = (1j ** (ny % 4)) * sgn * psi[j.astype(np.int64)]
    return math.cos(phi) * psi - 1j * math.sin(phi) * ppsi


# --------------------------------------------------------------------------
# UCCSD (ir/algorithms/uccsd/UCCSD.cpp:96-365, CommutingSetGenerator.hpp:51-93)
# --------------------------------------------------------------------------
def uccsd_fermion_terms(n_qubits: int, n_electrons: int):
    """The cluster operator of UCCSD.cpp:96-219 as [(coeff, sites, var)], plus nParams."""
    n_occ = int(math.ceil(n_electrons / 2.0))
    n_virt = n_qubits // 2 - n_occ
    n_single = n_occ * n_virt
    n_double = n_single * (n_single + 1) // 2
    n_params = n_single + n_double
    params = [f"theta{i}" for i in range(n_params)]
    single = params[:n_single]
    double1 = params[n_single:2 * n_single]
    double2 = params[2 * n_single:]
    fs = [lambda i: 2 * i, lambda i: 2 * i + 1]
    out = []
    count = 0
    for i in range(n_virt):
        for j in range(n_occ):
            vs, os_ = n_occ + i, j
            for s in range(2):
                ti, oi = fs[s], fs[1 - s]
                vt, vo, ot, oo = ti(vs), oi(vs), ti(os_), oi(os_)
                out.append((1.0, [(vt, 1), (ot, 0)], single[count]))
                out.append((-1.0, [(ot, 1), (vt, 0)], single[count]))
                out.append((-1.0, [(vt, 1), (ot, 0), (vo, 1), (oo, 0)], double1[count]))
                out.append((1.0, [(oo, 1), (vo, 0), (ot, 1), (vt, 0)], double1[count]))
            count += 1
    tuples = [(i, j) for i in range(n_virt) for j in range(n_occ)]
    combos = [tuples[i] + tuples[j] for i in range(len(tuples)) for j in range(i + 1, len(tuples))]
    count = 0
    for (p, q, r, s) in combos:
        vs1, os1, vs2, os2 = n_occ + p, q, n_occ + r, s
        for sa in range(2):
            for sb in range(2):
                ia, ib = fs[sa], fs[sb]
                v1a, o1a, v2b, o2b = ia(vs1), ia(os1), ib(vs2), ib(os2)
                out.append((-1.0, [(v1a, 1), (o1a, 0), (v2b, 1), (o2b, 0)], double2[count]))
                # literal restatement of UCCSD.cpp:210 (op6 is NOT the h.c. of op5)
                out.append((1.0, [(o2b, 1), (o1a, 0), (v2b, 1), (o2b, 0)], double2[count]))
        count += 1
    return out, n_params


def _commutes_bv(k1: Key, k2: Key) -> int:
    def bv(k):
        x = z = 0
        for q, p in k[1]:
            if p in ("X", "Y"):
                x |= 1 << q
            if p in ("Z", "Y"):
                z |= 1 << q
        return x, z
    x1, z1 = bv(k1)
    x2, z2 = bv(k2)
    return (bin(x1 & z2).count("1") + bin(z1 & x2).count("1")) % 2


def commuting_sets(keys: List[Key]) -> List[List[Key]]:
    """CommutingSetGenerator::getCommutingSet, including its quirk of comparing
    against allTerms[innerCounter] (CommutingSetGenerator.hpp:72-77)."""
    sets: List[List[Key]] = []
    for i, t in enumerate(keys):
        if i == 0:
            sets.append([t])
            continue
        placed = False
        for s in sets:
            tot = 0
            for inner in range(len(s)):
                tot += _commutes_bv(t, keys[inner])
            if tot == 0:
                s.append(t)
                placed = True
                break
        if not placed:
            sets.append([t])
    return sets


def uccsd_generator(n_qubits: int, n_electrons: int, transformation: str = "jw") -> Tuple[Op, int]:
    ft, n_params = uccsd_fermion_terms(n_qubits, n_electrons)
    terms = [(complex(c), sites) for c, sites, _ in ft]
    variables = [v for _, _, v in ft]
    if transformation == "jw":
        gen = jordan_wigner(terms, variables)
    else:
        gen = bravyi_kitaev(terms, n_qubits, variables)
    return gen, n_params


def uccsd_ordered_terms(gen: Op) -> List[Key]:
    """The reference walks an unordered_map (UCCSD.cpp:247-261); this build fixes the
    order: ascending term id (variable name + Pauli string), then the reference's
    commuting-set grouping."""
    keys = sorted(gen.keys(), key=term_id)
    out = []
    for s in commuting_sets(keys):
        out.extend(s)
    return out


def uccsd_gates(n_qubits: int, n_electrons: int, theta: Sequence[float],
                transformation: str = "jw") -> List[Tuple[str, Sequence[int], float]]:
    """Gate list of UCCSD.cpp:259-365 evaluated at theta."""
    gen, n_params = uccsd_generator(n_qubits, n_electrons, transformation)
    assert len(theta) == n_params, (len(theta), n_params)
    pi = 3.14159265358979323
    gates: List[Tuple[str, Sequence[int], float]] = []
    for key in uccsd_ordered_terms(gen):
        coeff = gen[key]
        th = theta[int(key[0][5:])]
        terms = list(key[1])
        if not terms:
            continue
        pre, mid, post = [], [], []
        for i, (q, p) in enumerate(terms):
            if i < len(terms) - 1:
                mid.append(("CNOT", (q, terms[i + 1][0]), 0.0))
            if p == "X":
                pre.insert(0, ("H", (q,), 0.0))
            elif p == "Y":
                pre.insert(0, ("Rx", (q,), pi / 2.0))
        mid.append(("Rz", (terms[-1][0],), 2 * coeff.imag * th))
        back = []
        for i in range(len(terms) - 1, -1, -1):
            q, p = terms[i]
            if i < len(terms) - 1:
                back.append(("CNOT", (q, terms[i + 1][0]), 0.0))
            if p == "X":
                post.append(("H", (q,), 0.0))
            elif p == "Y":
                post.append(("Rx", (q,), -pi / 2.0))
        gates.extend(pre + mid + back + post)
    xs = [("X", (i,), 0.0) for i in range(n_electrons)]
    return xs + gates


def uccsd_rotations(n_qubits: int, n_electrons: int, transformation: str = "jw"):
    """The same circuit as fused Pauli rotations: [(xmask, zmask, imag_coeff, param_idx)];
    each is exp(-i * imag_coeff * theta[param] * P)."""
    gen, n_params = uccsd_generator(n_qubits, n_electrons, transformation)
    out = []
    for key in uccsd_ordered_terms(gen):
        if not key[1]:
            continue
        x = z = 0
        for q, p in key[1]:
            if p in ("X", "Y"):
                x |= 1 << q
            if p in ("Z", "Y"):
                z |= 1 << q
        out.append((x, z, gen[key].imag, int(key[0][5:])))
    return out, n_params


# --------------------------------------------------------------------------
# Energy (task/tasks/ComputeEnergyVQETask.cpp:47-50,85-106,137-166)
# --------------------------------------------------------------------------
def measured_expectation(psi: np.ndarray, key: Key) -> float:
    """The reference's route: basis change H / Rx(pi/2) on X / Y sites, then the
    Z-parity of the measured qubits (getExpectationValueZ)."""
    phi = psi
    mask = 0
    for q, p in key[1]:
        if p == "X":
            phi = apply_gate(phi, "H", (q,))
        elif p == "Y":
            phi = apply_gate(phi, "Rx", (q,), 3.14159265358979323 / 2.0)
        mask |= 1 << q
    idx = np.arange(psi.shape[0], dtype=np.uint64)
    sgn = 1.0 - 2.0 * _parity(idx & np.uint64(mask))
    return float(np.sum(sgn * (phi.real ** 2 + phi.imag ** 2)))


def energy_reference(op: Op, psi: np.ndarray) -> Tuple[float, Dict[str, float]]:
    """E = c_I + sum_k Re(c_k) <P_k>, per-term values keyed by term id."""
    e = 0.0
    vals = {}
    for key in sorted(op, key=term_id):
        c = op[key].real                       # getCoeff takes the real part only (:47-50)
        if not key[1]:
            e += c
            vals["I"] = 1.0
            continue
        v = measured_expectation(psi, key)
        vals[term_id(key)] = v
        e += c * v
    return e, vals


def uccsd_energy(op: Op, n_qubits: int, n_electrons: int, theta: Sequence[float],
                 transformation: str = "jw") -> float:
    psi = run_circuit(n_qubits, uccsd_gates(n_qubits, n_electrons, theta, transformation))
    return energy_reference(op, psi)[0]


# --------------------------------------------------------------------------
# 2-RDM (decorators/RDMGenerator.cpp:43-149) and the energy contraction of
# decorators/tests/RDMGeneratorTester.cpp:134-158
# --------------------------------------------------------------------------
def hpq_hpqrs(terms, n: int, ) -> Tuple[float, np.ndarray, np.ndarray]:
    """FermionKernel::E_nuc / hpq / hpqrs (ir/FermionKernel.hpp:192-267)."""
    e_nuc = 0.0
    hpq = np.zeros((n, n), dtype=np.complex128)
    hpqrs = np.zeros((n, n, n, n), dtype=np.complex128)
    for c, sites in terms:
        if len(sites) == 0:
            e_nuc = c.real
        elif len(sites) == 2:
            hpq[sites[0][0], sites[1][0]] = c
        elif len(sites) == 4:
            hpqrs[sites[0][0], sites[1][0], sites[2][0], sites[3][0]] = c
    return e_nuc, hpq, hpqrs


def rdm2(psi: np.ndarray, n: int) -> Tuple[np.ndarray, int]:
    """rho_pqrs as RDMGenerator::generate builds it; returns (rho, #unique circuits)."""
    rho = np.zeros((n, n, n, n), dtype=np.complex128)
    functions: Dict[str, Tuple[Key, List[Tuple[Tuple[int, int, int, int], float]]]] = {}
    ident: Dict[Tuple[int, int, int, int], float] = {}
    for m in range(n):
        for nn in range(m + 1, n):
            for v in range(m, n):
                for w in range(v + 1, n):
                    src = (f"__qpu__ k(){{\n0.5 {m} 1 {nn} 1 {w} 0 {v} 0\n"
                           f"0.5 {w} 1 {v} 1 {m} 0 {nn} 0\n}}")
                    terms, _ = parse_fermion_source(src)
                    op = jordan_wigner(terms)
                    for key, c in op.items():
                        t = c.real
                        if key[1]:
                            name = term_id(key)
                            if name in functions:
                                functions[name][1].append(((m, nn, v, w), t))
                            else:
                                functions[name] = (key, [((m, nn, v, w), t)])
                        else:
                            ident[(m, nn, v, w)] = t
    sum_map: Dict[Tuple[int, int, int, int], float] = {}
    for name in sorted(functions):
        key, lst = functions[name]
        ev = measured_expectation(psi, key)
        for el, t in lst:
            sum_map[el] = sum_map.get(el, 0.0) + t * ev
    for el, t in ident.items():
        sum_map[el] = sum_map.get(el, 0.0) + t

    def set_antisym(a, b, c, d, val):
        rho[a, b, c, d] = val
        rho[b, a, c, d] = -val
        rho[a, b, d, c] = -val
        rho[b, a, d, c] = val

    for (a, b, c, d), val in sum_map.items():
        set_antisym(a, b, c, d, val)
        set_antisym(c, d, a, b, val)
    return rho, len(functions)


def rdm_energy(rho: np.ndarray, e_nuc: float, hpq: np.ndarray, hpqrs: np.ndarray) -> float:
    n = hpq.shape[0]
    rho_pq = np.einsum("pqrq->pr", rho)          # trace over dims (1,3)
    e = e_nuc
    for p in range(n):
        for q in range(n):
            for r in range(n):
                for s in range(n):
                    e += 0.5 * (hpqrs[p, q, r, s] * (rho[p, q, s, r] + rho[r, s, q, p])).real
    for p in range(n):
        for q in range(n):
            e += 0.5 * (hpq[p, q] * (rho_pq[p, q] + rho_pq[q, p])).real
    return float(e)


# --------------------------------------------------------------------------
# Nelder-Mead as the reference drives it
# (tpls/cppnumericalsolvers/cppoptlib/solver/neldermeadsolver.h,
#  criteria task/tasks/VQEMinimizeTask.hpp:74-104)
# --------------------------------------------------------------------------
def nelder_mead(fun, x0: Sequence[float], f_delta: float = 1e-6, iterations: int = 10000):
    """tpls/cppnumericalsolvers/cppoptlib/solver/neldermeadsolver.h:47-194, literally:
      * initial simplex (:23-41); loop bound iterations * DIM (:71) but the solver's OWN checkConvergence (:208-219, no gradient
        test) ends the walk once iter > iterations (:209) or max_i |f_i - f_best| < fDelta (:215);
      * second stop (:109-115): spread of f <= tt1 AND spread of x <= tt2, both thresholds max(1e-4, 10 ulp-ish) (:88-90);
      * reflect / expand / outside contraction (f_c <= f_r) / inside contraction (f_c < f_worst, STRICT) / shrink (:130-177);
      * shrink re-evaluates the best vertex as well (:199);
    criteria from task/tasks/VQEMinimizeTask.hpp:74-104 (fDelta 1e-6, iterations 10000, xDelta 0).
    Returns (best vertex, LAST evaluated objective value, iterations): CppOptVQEBackend::minimize reports currentEnergy, the
    value of the most recent call of value() (VQEMinimizeTask.hpp:44,50-53), not f at the best vertex."""
    x0 = np.asarray(x0, dtype=np.float64)
    dim = x0.shape[0]
    rho, xi, gam = 1.0, 2.0, 0.5
    last = [None]

    def ev(x):
        last[0] = float(fun(np.array(x, dtype=np.float64)))
        return last[0]

    S = np.tile(x0[:, None], (1, dim + 1))
    for c in range(1, dim + 1):
        r = c - 1
        S[r, c] = 0.00025 if x0[r] == 0 else 1.05 * x0[r]
    f = [ev(S[:, i]) for i in range(dim + 1)]
    index = sorted(range(dim + 1), key=lambda a: f[a])
    eps = np.finfo(np.float64).eps
    it = 0
    max_iter = iterations * dim
    while it < max_iter:
        best, worst = index[0], index[dim]
        max1 = max(abs(f[index[i]] - f[best]) for i in range(1, dim + 1))
        max2 = max(float(np.max(np.abs(S[:, index[i]] - S[:, best]))) for i in range(1, dim + 1))
        xb = float(np.max(S[:, best]))
        tt1 = max(1e-4, 10 * np.nextafter(f[best], eps) - f[best])
        tt2 = max(1e-4, 10 * (np.nextafter(xb, eps) - xb))
        if iterations != 0 and (it > iterations or (f_delta > 0 and max1 < f_delta)):
            break
        if max1 <= tt1 and max2 <= tt2:
            break
        x_bar = np.sum(S[:, index[:dim]], axis=1) / dim
        x_r = (1 + rho) * x_bar - rho * S[:, worst]
        f_r = ev(x_r)
        shrink = False
        if f_r < f[best]:
            x_e = (1 + rho * xi) * x_bar - rho * xi * S[:, worst]
            f_e = ev(x_e)
            if f_e < f_r:
                S[:, worst], f[worst] = x_e, f_e
            else:
                S[:, worst], f[worst] = x_r, f_r
        elif f_r < f[index[dim - 1]]:
            S[:, worst], f[worst] = x_r, f_r
        elif f_r < f[worst]:
            x_c = (1 + rho * gam) * x_bar - rho * gam * S[:, worst]
            f_c = ev(x_c)
            if f_c <= f_r:
                S[:, worst], f[worst] = x_c, f_c
            else:
                shrink = True
        else:
            x_c = (1 - gam) * x_bar + gam * S[:, worst]
            f_c = ev(x_c)
            if f_c < f[worst]:
                S[:, worst], f[worst] = x_c, f_c
            else:
                shrink = True
        if shrink:
            f[best] = ev(S[:, best])
            for i in range(1, dim + 1):
                S[:, index[i]] = 0.5 * S[:, index[i]] + 0.5 * S[:, best]
                f[index[i]] = ev(S[:, index[i]])
        index = sorted(range(dim + 1), key=lambda a: f[a])
        it += 1
    return S[:, index[0]].copy(), last[0], it


# --------------------------------------------------------------------------
# Matrix-free ground state on the CPU (independent check for n > 14):
# scipy ARPACK on the mask form (SURVEY 8c "Lanczos reference semantics")
# --------------------------------------------------------------------------
def ground_state_eigsh(xm, zm, cf, n: int, tol: float = 1e-12, apply=None):
    from scipy.sparse.linalg import LinearOperator, eigsh
    N = 1 << n
    fn = apply or (lambda v: apply_masks(xm, zm, cf, v.astype(np.complex128)))
    A = LinearOperator((N, N), matvec=fn, dtype=np.complex128)
    w, v = eigsh(A, k=1, which="SA", tol=tol)
    return float(w[0]), v[:, 0]
