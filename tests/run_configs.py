#!/usr/bin/env python
"""Configs 1 and 3 of BASELINE.json on one B200, each checked against the CPU oracle in the same run (verification
script: imports oracle/, not product code).  Prints one JSON line per config.

  config 1: H2 STO-3G, 4 qubits, JW: DiagonalizeTask + ComputeEnergyVQETask with the UCCSD ansatz at the reference's test
            point (task/tests/DiagonalizeTaskTester.cpp:103-107, ComputeEnergyVQETaskTester.cpp:99-106); latency per call.
  config 3: 14-qubit JW molecule-shaped Hamiltonian (7 spatial orbitals, seed 14; the reference ships no H2O integrals,
            SURVEY.md section 7), Lanczos ground state, tol 1e-12 on the Ritz residual, against ARPACK over the C oracle;
            with and without the particle-number sector (diag-number-symmetry).
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)
import xaccvqe_b200 as xb  # noqa: E402
import c_oracle as co  # noqa: E402
import pauli_oracle as po  # noqa: E402
from bench_sweep import lih_shaped_source  # noqa: E402


def config1():
    golden = json.load(open(os.path.join(ROOT, "tests", "golden", "reference_fixtures.json")))
    src = golden["h2_source"]
    op = xb.compile(src)
    r = xb.execute(op, None, task="vqe-diagonalize")            # warm-up (context, tables)
    t0 = time.perf_counter()
    reps = 20
    for _ in range(reps):
        r = xb.execute(op, None, task="vqe-diagonalize")
    dt_diag = (time.perf_counter() - t0) / reps
    theta = golden["h2_uccsd_params"]
    kw = {"task": "compute-energy", "ansatz": "uccsd", "n-electrons": 2, "vqe-params": ",".join(repr(float(t)) for t in theta)}
    e = xb.execute(op, None, **kw)
    t0 = time.perf_counter()
    for _ in range(reps):
        e = xb.execute(op, None, **kw)
    dt_energy = (time.perf_counter() - t0) / reps
    oop, n = po.compile_fermion_source(src, "jw")
    want_e = po.uccsd_energy(oop, n, 2, theta)
    print(json.dumps({"config": "H2 STO-3G 4-qubit JW: DiagonalizeTask + ComputeEnergyVQETask (UCCSD)", "n_qubits": 4,
                      "ground_energy": r.energy, "reference_golden": golden["h2_ground_energy_exact"],
                      "rel_err_ground": abs(r.energy - golden["h2_ground_energy_exact"]) / abs(r.energy),
                      "uccsd_energy": e.energy, "oracle_uccsd_energy": want_e, "abs_err_uccsd": abs(e.energy - want_e),
                      "reference_test_value": golden["h2_uccsd_energy_near"], "nQpuCalls": e.nQpuCalls,
                      "ms_per_diagonalize_call": 1e3 * dt_diag, "ms_per_energy_call": 1e3 * dt_energy,
                      "note": "each call includes handle creation, table upload, ansatz and one D2H read"}))


def config3():
    n, ne = 14, 6
    src = lih_shaped_source(n, seed=14)
    op = xb.compile(src, "jw", n)
    xm, zm, cf = op.pack()
    be = xb.Backend(n)
    be.set_hamiltonian(xm, zm, cf)
    T, G = be.hamiltonian_stats()
    out = {"config": "14-qubit JW molecule-shaped Hamiltonian (7 spatial orbitals, seed 14), Lanczos ground state, 1 B200",
           "n_qubits": n, "n_terms": int(T), "n_xmask_groups": int(G)}
    want, _ = po.ground_state_eigsh(xm, zm, cf, n, tol=1e-13, apply=lambda v: co.apply(n, xm, zm, cf, v))
    for label, sector in (("full_space", None), ("number_sector", ne)):
        if sector is not None:
            be.set_number_sector(sector, "jw")
        be.lanczos_ground(tol=1e-12, max_iter=2000, seed=14)        # warm-up
        be.synchronize()
        l0 = be.kernel_launches()
        t0 = time.perf_counter()
        e0, it, res = be.lanczos_ground(tol=1e-12, max_iter=2000, seed=14)
        be.synchronize()
        dt = time.perf_counter() - t0
        out[label] = {"eigenvalue": e0, "iterations": int(it), "ritz_residual": res, "seconds": dt, "ms_per_iteration": 1e3 * dt / it,
                      "hpsi_updates_per_s": T * (1 << n) * it / dt, "gpu_launches": int(be.kernel_launches() - l0)}
    be.set_number_sector(-1)
    # oracle: full-space ground state by ARPACK over the C oracle; sector ground state = lowest eigenvalue of P H P
    out["oracle_full_space"] = want
    out["rel_err_full_space"] = abs(out["full_space"]["eigenvalue"] - want) / abs(want)
    idx = np.array([j for j in range(1 << n) if bin(j).count("1") == ne])
    from scipy.sparse.linalg import LinearOperator, eigsh

    def mv(v):
        full = np.zeros(1 << n, complex)
        full[idx] = v
        return co.apply(n, xm, zm, cf, full)[idx]
    ws = eigsh(LinearOperator((len(idx), len(idx)), matvec=mv, dtype=complex), k=1, which="SA", tol=1e-13)[0][0]
    out["oracle_number_sector"] = float(ws)
    out["sector_dimension"] = int(len(idx))
    out["rel_err_number_sector"] = abs(out["number_sector"]["eigenvalue"] - ws) / abs(ws)
    out["pass_1e-10"] = bool(out["rel_err_full_space"] < 1e-10 and out["rel_err_number_sector"] < 1e-10)
    # the oracle's CPU time for one H psi (all host threads), for scale
    psi = np.random.default_rng(0).normal(size=1 << n) + 0j
    t0 = time.perf_counter()
    for _ in range(5):
        co.apply(n, xm, zm, cf, psi)
    out["cpu_oracle_ms_per_hpsi"] = 1e3 * (time.perf_counter() - t0) / 5
    out["cpu_threads"] = co.num_threads()
    print(json.dumps(out))
    be.close()


def faithful_baseline():
    """BASELINE.md section 4, "B-faithful": the reference's own way -- string-based row builder (DiagonalizeTask.cpp:189-203) +
    dense Hermitian solve (:205), ONE thread, as the reference is -- timed on this box's host for configs 1-3 beside the device
    numbers above.  Config 3's 2^14 x 2^14 dense matrix (4.3 GB, ~4e12 flops x c to diagonalise) is not formed: the row builder is
    timed on a bounded block of rows and scaled, the dense solve is reported as not run."""
    from threadpoolctl import threadpool_limits
    golden = json.load(open(os.path.join(ROOT, "tests", "golden", "reference_fixtures.json")))
    out = {"baseline": "B-faithful: string-based dense build + dense eigen-solve, 1 thread (oracle/hpsi_oracle.c oracle_faithful_*)",
           "host_cores": os.cpu_count(), "threads_used": 1}
    with threadpool_limits(limits=1):
        # config 1: H2, 4 qubits
        op, n = po.compile_fermion_source(golden["h2_source"], "jw")
        t0 = time.perf_counter()
        for _ in range(200):
            A = co.faithful_dense(op, n)
            w = np.linalg.eigvalsh(A)
        out["config1_h2_4q"] = {"ms_per_diagonalize": 1e3 * (time.perf_counter() - t0) / 200, "E0": float(w[0])}
        # config 2: 12-qubit BK molecule-shaped operator
        op, n = po.compile_fermion_source(lih_shaped_source(12, seed=12), "bk", 12)
        t0 = time.perf_counter()
        A = co.faithful_dense(op, n)
        t_build = time.perf_counter() - t0
        t0 = time.perf_counter()
        w = np.linalg.eigvalsh(A)
        t_eig = time.perf_counter() - t0
        out["config2_12q_bk"] = {"n_terms": len(op), "dense_build_s": t_build, "dense_eigensolve_s": t_eig, "E0": float(w[0])}
        del A
        # config 3: 14-qubit JW molecule-shaped operator: 512 of the 16384 rows of the row builder
        op, n = po.compile_fermion_source(lih_shaped_source(14, seed=14), "jw", 14)
        psi = np.random.default_rng(0).normal(size=1 << n) + 0j
        rows = 512
        t0 = time.perf_counter()
        co.faithful_apply_rows(op, n, psi, 0, rows)
        dt = time.perf_counter() - t0
        out["config3_14q_jw"] = {"n_terms": len(op), "row_builder_s_for_%d_rows" % rows: dt, "row_builder_s_scaled_to_all_rows": dt * (1 << n) / rows,
                                 "dense_eigensolve": "not run: 16384 x 16384 complex (4.3 GB), O(N^3)"}
    print(json.dumps(out))


if __name__ == "__main__":
    config1()
    config3()
    faithful_baseline()
