"""Host-side mirror of the reference's task layer for the hot path, on top of the C ABI (capi.py).

Keeps the call shape of the reference's Python module (python/xacc-vqe-py.cpp:140-262, :429-477):

    op = compile(src)                                  # FermionCompiler text -> PauliOperator
    r = execute(op, buffer, task="vqe-diagonalize")    # DiagonalizeTask   (task/tasks/DiagonalizeTask.cpp:13-44)
    r = execute(op, buffer, task="compute-energy", ansatz="uccsd", **{"n-electrons": 2, "vqe-params": "0,-0.057"})
    r = execute(op, buffer, task="vqe", ...)           # VQEMinimizeTask   (task/tasks/VQEMinimizeTask.hpp:37-53)
    r.energy, r.angles, r.nQpuCalls, r.vqeIterations   # VQETaskResult     (task/VQETask.hpp:53-65)

Everything numerical runs on the GPU through libb200vqe.so; there is no CPU fallback (Backend() raises
without a CUDA device).  The optimiser itself stays on the host, as in the reference.
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence

import numpy as np

from . import capi
from .capi import B200Error, Backend, OpHandle


# --------------------------------------------------------------------------- operator / buffer / result objects
class PauliOperator:
    """What xaccvqe.compile() returns: the JW/BK image of the fermionic source (or a parsed Pauli string)."""

    def __init__(self, handle: OpHandle):
        self.handle = handle

    @classmethod
    def fromString(cls, text: str) -> "PauliOperator":
        return cls(OpHandle.from_string(text))

    def nTerms(self) -> int:
        return self.handle.n_terms()

    def nQubits(self) -> int:
        return self.handle.n_qubits()

    def toString(self) -> str:
        return self.handle.to_string()

    def pack(self):
        return self.handle.pack()

    def term_ids(self) -> List[str]:
        return self.handle.term_ids()

    def __add__(self, other: "PauliOperator") -> "PauliOperator":
        return PauliOperator(self.handle.add(other.handle))

    def __mul__(self, other) -> "PauliOperator":
        if isinstance(other, PauliOperator):
            return PauliOperator(self.handle.mul(other.handle))
        return PauliOperator(self.handle.scale(complex(other)))

    __rmul__ = __mul__


class AcceleratorBuffer:
    """The slice of xacc::AcceleratorBuffer the tasks write to: extra-info keys and per-term children
    (task/tasks/ComputeEnergyVQETask.cpp:150-160, :185-193)."""

    def __init__(self, name: str = "q", n_qubits: int = 0):
        self.name, self.n_qubits = name, n_qubits
        self.info: Dict[str, object] = {}
        self.children: List[Dict[str, object]] = []

    def addExtraInfo(self, key: str, value) -> None:
        self.info[key] = value

    def getInformation(self, key: Optional[str] = None):
        return self.info if key is None else self.info[key]

    def appendChild(self, record: Dict[str, object]) -> None:
        self.children.append(record)

    def getChildren(self, key: Optional[str] = None, value=None):
        if key is None:
            return self.children
        return [c for c in self.children if c.get(key) == value]


@dataclass
class VQETaskResult:
    """task/VQETask.hpp:17-65, including the CSV persist() of :26-51."""
    energy: float = 0.0
    angles: np.ndarray = field(default_factory=lambda: np.zeros(0))
    nQpuCalls: int = 0
    vqeIterations: int = 0
    ansatzQASM: str = ""
    expVals: Dict[str, float] = field(default_factory=dict)
    fileName: str = ""

    def persist(self) -> None:
        if not self.fileName:
            return
        new = not os.path.exists(self.fileName) or os.path.getsize(self.fileName) == 0
        with open(self.fileName, "a") as f:
            if new:
                f.write("".join(f"t{i}," for i in range(len(self.angles))) + "".join(f"{k}," for k in sorted(self.expVals)) + "E\n")
            f.write("".join(f"{a:g}," for a in self.angles) + "".join(f"{self.expVals[k]:g}," for k in sorted(self.expVals))
                    + f"{self.energy:g}\n")


def compile(src: str, transformation: str = "jw", n_qubits: int = 0) -> PauliOperator:   # noqa: A001 (reference name)
    """xaccvqe.compile(src) (python/xacc-vqe-py.cpp:110-131): FermionCompiler + fermion-transformation."""
    return PauliOperator(OpHandle.from_fermion_source(src, transformation, n_qubits))


# --------------------------------------------------------------------------- ansatz
class Ansatz:
    """A parameterised state preparation given as fused Pauli rotations exp(-i c_r theta_{p_r} P_r) on a basis state."""

    def __init__(self, n_qubits: int, basis_index: int, xmask, zmask, imag_coeff, param, n_params: int, qasm: str = ""):
        self.n_qubits, self.basis_index = n_qubits, basis_index
        self.xmask, self.zmask = np.asarray(xmask, np.uint64), np.asarray(zmask, np.uint64)
        self.imag_coeff, self.param = np.asarray(imag_coeff, float), np.asarray(param, np.int32)
        self.n_params, self.qasm = n_params, qasm

    @classmethod
    def uccsd(cls, n_qubits: int, n_electrons: int, transformation: str = "jw") -> "Ansatz":
        """ir/algorithms/uccsd/UCCSD.cpp:58-368: X on qubits 0..ne-1 (:361-365), then one rotation per generator term."""
        x, z, ic, pa, npar = capi.uccsd_rotations(n_qubits, n_electrons, transformation)
        return cls(n_qubits, (1 << n_electrons) - 1, x, z, ic, pa, npar, qasm=f"uccsd(n-electrons={n_electrons},{transformation})")

    def prepare(self, be: Backend, theta: Sequence[float], slot: int = 0) -> None:
        th = np.asarray(theta, float)
        if len(th) != self.n_params:
            raise B200Error(-1, f"ansatz takes {self.n_params} parameters, got {len(th)}")
        be.reset(slot, self.basis_index)
        if len(self.xmask):
            be.apply_pauli_rotations(self.xmask, self.zmask, self.imag_coeff * th[self.param], slot=slot)


class GateAnsatz:
    """An explicit gate list (name, qubits, theta | parameter index) as the reference's GateFunction ansatz."""

    def __init__(self, n_qubits: int, gates: Sequence, n_params: int):
        self.n_qubits, self.gates, self.n_params, self.qasm = n_qubits, list(gates), n_params, ""

    def prepare(self, be: Backend, theta: Sequence[float], slot: int = 0) -> None:
        th = np.asarray(theta, float)
        be.reset(slot, 0)
        lst = []
        for name, qs, t in self.gates:
            lst.append((name, qs, t(th) if callable(t) else float(t)))
        be.apply_gates(lst, slot=slot)


# --------------------------------------------------------------------------- program = Hamiltonian + ansatz on one backend
class VQEProgram:
    """task/VQEProgram.hpp:106-246: owns the PauliOperator, the ansatz and the (device) accelerator."""

    def __init__(self, op: PauliOperator, n_qubits: Optional[int] = None, ansatz=None, device: int = 0, backend: Optional[Backend] = None):
        self.op = op
        self.n_qubits = int(n_qubits or op.nQubits())
        self.ansatz = ansatz
        self.xmask, self.zmask, self.coef = op.pack()
        self.term_ids = op.term_ids()
        self.backend = backend or Backend(self.n_qubits, device=device)
        self.backend.set_hamiltonian(self.xmask, self.zmask, self.coef)
        self.n_nonidentity = int(np.count_nonzero(self.xmask | self.zmask))

    def getNParameters(self) -> int:
        return 0 if self.ansatz is None else self.ansatz.n_params


def generate_parameters(n_params: int, spec: Optional[str], seed: Optional[int] = None) -> np.ndarray:
    """task/VQEParameterGenerator.hpp:14-73: "a,b,c" string, else uniform random in [-pi, pi)."""
    if spec is not None and str(spec) != "":
        vals = [float(s) for s in str(spec).split(",")]
        out = np.zeros(n_params)
        out[:len(vals)] = vals[:n_params]
        return out
    rng = np.random.default_rng(seed)
    return rng.uniform(-np.pi, np.pi, n_params)


# --------------------------------------------------------------------------- tasks
class ComputeEnergyVQETask:
    """task/tasks/ComputeEnergyVQETask.cpp:12-214 on the device: ansatz -> <psi|H|psi>, per-term children, counters."""
    name = "compute-energy"

    def __init__(self, program: VQEProgram, buffer: Optional[AcceleratorBuffer] = None, persist_file: str = "", per_term: bool = True):
        self.program, self.buffer, self.persist_file, self.per_term = program, buffer, persist_file, per_term
        self.vqeIteration = 0
        self.totalQpuCalls = 0

    def execute(self, parameters: Sequence[float]) -> VQETaskResult:
        p, be = self.program, self.program.backend
        theta = np.asarray(parameters, float)
        if p.ansatz is None:
            raise B200Error(-1, "compute-energy needs an ansatz (kwarg 'ansatz')")
        p.ansatz.prepare(be, theta)
        res = VQETaskResult(angles=theta.copy(), fileName=self.persist_file, ansatzQASM=getattr(p.ansatz, "qasm", ""))
        if self.per_term:
            e, per = be.expectation(per_term=True)
            # energy uses Re(c_k) only (:47-50); identity contributes its coefficient with exp-val 1 (:90-106)
            for k, tid in enumerate(p.term_ids):
                res.expVals[tid] = float(per[k])
                if self.buffer is not None:
                    self.buffer.appendChild({"kernel": tid, "exp-val-z": float(per[k]), "coefficient": float(_plain_coeff(p, k).real),
                                             "parameters": theta.tolist()})
        else:
            e = be.expectation()
        res.energy = float(e.real)
        self.totalQpuCalls += p.n_nonidentity          # nQpuCalls += kernels.size() per energy (:134)
        self.vqeIteration += 1
        res.nQpuCalls, res.vqeIterations = self.totalQpuCalls, self.vqeIteration
        if self.buffer is not None:
            cur = self.buffer.info.get("vqe-energy")
            if cur is None or res.energy < cur:          # keep the lowest energy seen (:185-191)
                self.buffer.addExtraInfo("vqe-energy", res.energy)
                self.buffer.addExtraInfo("vqe-angles", theta.tolist())
            self.buffer.addExtraInfo("vqe-nQPU-calls", self.totalQpuCalls)
        res.persist()
        return res


def _plain_coeff(p: VQEProgram, k: int) -> complex:
    ny = bin(int(p.xmask[k]) & int(p.zmask[k])).count("1") & 3
    return complex(p.coef[k]) / (1j ** ny)


class DiagonalizeTask:
    """task/tasks/DiagonalizeTask.cpp:13-44 with the 'diagonalize-b200' backend: matrix-free Lanczos on the device
    instead of the dense Eigen solve (:184-214) / SLEPc Lanczos (petsc/slepc/SlepcDiagonalizeBackend.cpp:79-89)."""
    name = "vqe-diagonalize"

    def __init__(self, program: VQEProgram, buffer: Optional[AcceleratorBuffer] = None, tol: float = 1e-12, max_iter: int = 4000,
                 seed: int = 1, want_vector: bool = False, n_electrons: Optional[int] = None, transformation: str = "jw"):
        self.program, self.buffer = program, buffer
        # options "diag-number-symmetry" + "n-electrons" + "fermion-transformation" (DiagonalizeTask.cpp:48-57)
        self.n_electrons, self.transformation = n_electrons, transformation
        self.tol, self.max_iter, self.seed, self.want_vector = tol, max_iter, seed, want_vector
        self.ground_state: Optional[np.ndarray] = None
        self.iterations, self.residual = 0, 0.0

    def execute(self, parameters: Sequence[float] = ()) -> VQETaskResult:
        be = self.program.backend
        if self.n_electrons is not None:
            be.set_number_sector(int(self.n_electrons), self.transformation)
        try:
            e0, it, res = be.lanczos_ground(tol=self.tol, max_iter=self.max_iter, want_vector=self.want_vector, seed=self.seed)
        finally:
            if self.n_electrons is not None:
                be.set_number_sector(-1)
        self.iterations, self.residual = it, res
        if self.want_vector:
            # element order of the reference's eigenvector (DiagonalizeTask.cpp:189-193, :222-225)
            self.ground_state = be.download(0, order=capi.ORDER_REFERENCE)      # sharded: this rank's range of the reference order
        if self.buffer is not None:
            self.buffer.addExtraInfo("vqe-energy", e0)
        return VQETaskResult(energy=e0, angles=np.asarray(parameters, float), vqeIterations=it)


def nelder_mead(fun: Callable[[np.ndarray], float], x0: Sequence[float], f_delta: float = 1e-6, iterations: int = 10000):
    """The simplex walk of the reference's optimiser (tpls/cppnumericalsolvers/cppoptlib/solver/neldermeadsolver.h:47-194) with
    its arithmetic and its stop rules, so that iteration counts and nQpuCalls match:
      vertex c = x0 with coordinate c-1 scaled by 1.05 (0.00025 when zero); x_bar = mean of the best DIM vertices;
      reflect (1+rho) x_bar - rho x_w, expand (1+rho xi), contract outside (accepted when f_c <= f_r) / inside (accepted only
      when f_c < f_worst), else shrink by 0.5 towards the best vertex, which is re-evaluated too (:196-204); rho=1, xi=2, gam=0.5.
      Stop (checked before every step): iter > iterations, or spread of f below fDelta (the solver's own checkConvergence,
      :208-219: it has no gradient test, so VQECriteria's gradNorm = 1e-4 is never consulted), or spread of f <= tt1 and
      spread of x <= tt2 (:88-115; tt1 = max(1e-4, 10 * nextafter(f_best, eps) - f_best) exactly as written there,
      tt2 = max(1e-4, 10 ulp of the largest coordinate)); criteria defaults from task/tasks/VQEMinimizeTask.hpp:74-104.
    Returns (best vertex, last evaluated objective value, iterations): CppOptVQEBackend::minimize reports currentEnergy, the
    value of its most recent value() call (VQEMinimizeTask.hpp:44), not f at the best vertex."""
    rho, xi, gam = 1.0, 2.0, 0.5
    x0 = np.asarray(x0, float)
    d = len(x0)
    calls = {"last": 0.0}

    def value(x):
        calls["last"] = float(fun(np.array(x, float)))
        return calls["last"]

    if d == 0:
        return x0, value(x0), 0
    S = np.repeat(x0[:, None], d + 1, axis=1)              # columns are vertices, as in the reference
    for c in range(1, d + 1):
        S[c - 1, c] = 0.00025 if x0[c - 1] == 0 else (1 + 0.05) * x0[c - 1]
    f = [value(S[:, i]) for i in range(d + 1)]
    order = sorted(range(d + 1), key=lambda a: f[a])
    eps = float(np.finfo(float).eps)
    it = 0
    while it < iterations * d:
        lo, hi = order[0], order[d]
        spread_f = max(abs(f[k] - f[lo]) for k in order[1:])
        spread_x = max(float(np.abs(S[:, k] - S[:, lo]).max()) for k in order[1:])
        if iterations != 0 and (it > iterations or (f_delta > 0 and spread_f < f_delta)):
            break
        top = float(S[:, lo].max())
        if spread_f <= max(1e-4, 10 * float(np.nextafter(f[lo], eps)) - f[lo]) and spread_x <= max(1e-4, 10 * (float(np.nextafter(top, eps)) - top)):
            break
        x_bar = sum(S[:, k] for k in order[:d]) / d
        x_r = (1 + rho) * x_bar - rho * S[:, hi]
        f_r = value(x_r)
        accepted = None
        if f_r < f[lo]:
            x_e = (1 + rho * xi) * x_bar - rho * xi * S[:, hi]
            f_e = value(x_e)
            accepted = (x_e, f_e) if f_e < f_r else (x_r, f_r)
        elif f_r < f[order[d - 1]]:
            accepted = (x_r, f_r)
        elif f_r < f[hi]:
            x_c = (1 + rho * gam) * x_bar - rho * gam * S[:, hi]          # outside contraction
            f_c = value(x_c)
            if f_c <= f_r:
                accepted = (x_c, f_c)
        else:
            x_c = (1 - gam) * x_bar + gam * S[:, hi]                     # inside contraction
            f_c = value(x_c)
            if f_c < f[hi]:
                accepted = (x_c, f_c)
        if accepted is not None:
            S[:, hi], f[hi] = accepted
        else:                                                            # shrink towards the best vertex
            f[lo] = value(S[:, lo])
            for k in order[1:]:
                S[:, k] = 0.5 * S[:, k] + 0.5 * S[:, lo]
                f[k] = value(S[:, k])
        order = sorted(range(d + 1), key=lambda a: f[a])
        it += 1
    return S[:, order[0]].copy(), calls["last"], it


class VQEMinimizeTask:
    """task/tasks/VQEMinimizeTask.cpp:10-29 + CppOptVQEBackend (VQEMinimizeTask.hpp:37-53): Nelder-Mead over theta, objective =
    ComputeEnergyVQETask::execute."""
    name = "vqe"

    def __init__(self, program: VQEProgram, buffer: Optional[AcceleratorBuffer] = None, energy_delta: float = 1e-6, iterations: int = 10000):
        self.energy_task = ComputeEnergyVQETask(program, buffer, per_term=False)
        self.energy_delta, self.iterations = energy_delta, iterations

    def execute(self, parameters: Sequence[float]) -> VQETaskResult:
        x, f, it = nelder_mead(lambda th: self.energy_task.execute(th).energy, parameters, self.energy_delta, self.iterations)
        return VQETaskResult(energy=f, angles=x, nQpuCalls=self.energy_task.totalQpuCalls, vqeIterations=self.energy_task.vqeIteration,
                             ansatzQASM=getattr(self.energy_task.program.ansatz, "qasm", ""))


TASKS = {"vqe-diagonalize": DiagonalizeTask, "compute-energy": ComputeEnergyVQETask, "vqe": VQEMinimizeTask}


def execute(op, buffer: Optional[AcceleratorBuffer] = None, **kwargs) -> VQETaskResult:
    """xaccvqe.execute(op, buffer, **kwargs) (python/xacc-vqe-py.cpp:140-262).  Keys honoured: task, ansatz ('uccsd' or an
    Ansatz/GateAnsatz object), n-qubits, n-electrons, vqe-params, transformation, diagonalize-backend, diag-number-symmetry, fermion-transformation, vqe-energy-delta,
    vqe-iterations, vqe-persist-data, device.  'accelerator' and 'mpi-provider' are accepted and ignored: the accelerator is
    the B200 statevector backend, the communicator is NCCL inside a sharded handle."""
    kw = {k.replace("_", "-"): v for k, v in kwargs.items()}
    # ONE fermion transformation drives the compiler, the UCCSD generator and the number-symmetry rows, as the reference's single
    # option 'fermion-transformation' does (set from the kwarg 'transformation', python/xacc-vqe-py.cpp:406-408)
    tr = str(kw.get("transformation", kw.get("fermion-transformation", "jw")))
    if "transformation" in kw and "fermion-transformation" in kw and str(kw["transformation"]) != str(kw["fermion-transformation"]):
        raise B200Error(-1, "'transformation' and 'fermion-transformation' disagree")
    if isinstance(op, str):
        op = compile(op, tr)
    task = kw.get("task", "vqe-diagonalize")           # default task (:178)
    if task not in TASKS:
        raise B200Error(-1, f"unknown task '{task}' (have {sorted(TASKS)})")
    backend_name = kw.get("diagonalize-backend", "diagonalize-b200")
    if backend_name not in ("diagonalize-b200", "diagonalize-eigen", "diagonalize-slepc"):
        raise B200Error(-1, f"unknown diagonalize-backend '{backend_name}'")
    n_qubits = int(kw.get("n-qubits", op.nQubits()))    # max qubit index + 1 unless given (:182-190, :216)
    ansatz = kw.get("ansatz")
    if isinstance(ansatz, str):
        if ansatz != "uccsd":
            raise B200Error(-1, f"unknown ansatz '{ansatz}'")
        if "n-electrons" not in kw:
            raise B200Error(-1, "the uccsd ansatz needs 'n-electrons'")     # UCCSD.cpp:16-22
        ansatz = Ansatz.uccsd(n_qubits, int(kw["n-electrons"]), tr)
    program = VQEProgram(op, n_qubits, ansatz, device=int(kw.get("device", 0)))
    if buffer is not None:
        buffer.n_qubits = n_qubits
    params = generate_parameters(program.getNParameters(), kw.get("vqe-params"), kw.get("seed"))
    if task == "vqe-diagonalize":
        sym = bool(kw.get("diag-number-symmetry", False)) and "n-electrons" in kw        # both options, as DiagonalizeTask.cpp:54-55
        t = DiagonalizeTask(program, buffer, want_vector=bool(kw.get("print-ground-state", False)),
                            n_electrons=int(kw["n-electrons"]) if sym else None,
                            transformation=tr)
    elif task == "compute-energy":
        t = ComputeEnergyVQETask(program, buffer, persist_file=str(kw.get("vqe-persist-data", "")))
    else:
        t = VQEMinimizeTask(program, buffer, float(kw.get("vqe-energy-delta", 1e-6)), int(kw.get("vqe-iterations", 10000)))
    try:
        return t.execute(params)
    finally:
        program.backend.close()


# --------------------------------------------------------------------------- 2-RDM
class RDMGenerator:
    """decorators/RDMGenerator.cpp:8-152 on the device: every 0.5 (a+_m a+_n a_w a_v + h.c.), m<n, v>=m, w>v, is JW-mapped by
    the C++ front end, the Pauli strings are de-duplicated by term id (:58-80), ONE per-term expectation call measures all
    of them against the fixed ansatz state (:96-97), and rho_pqrs is filled with the antisymmetry of :106-149."""

    def __init__(self, n_qubits: int, transformation: str = "jw"):
        self.n = n_qubits
        xs: Dict[str, int] = {}
        self.xmask: List[int] = []
        self.zmask: List[int] = []
        self.contrib: List[List] = []                 # per unique string: [((m,n,v,w), weight), ...]
        self.identity: Dict[tuple, float] = {}
        n = n_qubits
        for m in range(n):
            for nn in range(m + 1, n):
                for v in range(m, n):
                    for w in range(v + 1, n):
                        src = f"__qpu__ k(){{\n0.5 {m} 1 {nn} 1 {w} 0 {v} 0\n0.5 {w} 1 {v} 1 {m} 0 {nn} 0\n}}"
                        op = OpHandle.from_fermion_source(src, transformation, n)
                        xm, zm, cf = op.pack()
                        ids = op.term_ids()
                        for k, tid in enumerate(ids):
                            ny = bin(int(xm[k]) & int(zm[k])).count("1") & 3
                            t = (complex(cf[k]) / (1j ** ny)).real          # std::real(kernel coefficient), :60
                            if not (int(xm[k]) | int(zm[k])):
                                self.identity[(m, nn, v, w)] = t
                                continue
                            if tid not in xs:
                                xs[tid] = len(self.xmask)
                                self.xmask.append(int(xm[k]))
                                self.zmask.append(int(zm[k]))
                                self.contrib.append([])
                            self.contrib[xs[tid]].append(((m, nn, v, w), t))
        self.names = sorted(xs, key=lambda s: xs[s])

    def n_circuits(self) -> int:
        return len(self.xmask)

    def generate(self, be: Backend, slot: int = 0) -> np.ndarray:
        n = self.n
        ny = np.array([bin(x & z).count("1") & 3 for x, z in zip(self.xmask, self.zmask)])
        be.set_hamiltonian(np.array(self.xmask, np.uint64), np.array(self.zmask, np.uint64), (1j ** ny).astype(complex))
        _, per = be.expectation(slot=slot, per_term=True)
        sums: Dict[tuple, float] = {}
        for k, lst in enumerate(self.contrib):
            for el, t in lst:
                sums[el] = sums.get(el, 0.0) + t * float(per[k])
        for el, t in self.identity.items():
            sums[el] = sums.get(el, 0.0) + t
        rho = np.zeros((n, n, n, n), complex)

        def put(a, b, c, d, val):
            rho[a, b, c, d] = val
            rho[b, a, c, d] = -val
            rho[a, b, d, c] = -val
            rho[b, a, d, c] = val

        for (a, b, c, d), val in sums.items():
            put(a, b, c, d, val)
            put(c, d, a, b, val)
        return rho


def rdm_energy(rho: np.ndarray, e_nuc: float, hpq: np.ndarray, hpqrs: np.ndarray) -> float:
    """Energy from the 2-RDM as decorators/tests/RDMGeneratorTester.cpp:134-158 contracts it."""
    rho_pq = np.einsum("pqrq->pr", rho)
    e2 = 0.5 * np.einsum("pqrs,pqsr->", hpqrs, rho) + 0.5 * np.einsum("pqrs,rsqp->", hpqrs, rho)
    e1 = 0.5 * np.einsum("pq,pq->", hpq, rho_pq) + 0.5 * np.einsum("pq,qp->", hpq, rho_pq)
    return float(e_nuc + e2.real + e1.real)
