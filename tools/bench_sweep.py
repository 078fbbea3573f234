#!/usr/bin/env python
"""Config 2 of BASELINE.json (shape): 12-qubit BK Hamiltonian, UCCSD (4 electrons, 44 parameters, 2080 fused rotations),
energy sweep over 64 parameter vectors on one B200.  Times the shared-memory sweep kernel (k_sweep_small, one launch for all
vectors) against the launch-per-rotation path, both through the C ABI.  Synthetic molecule-shaped integrals (the reference
ships no LiH integrals).  Prints one JSON line."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import xaccvqe_b200 as xb  # noqa: E402


def lih_shaped_source(n=12, seed=12):
    rng = np.random.default_rng(seed)
    no = n // 2
    h1 = rng.normal(size=(no, no)) * 0.3
    h1 = 0.5 * (h1 + h1.T) - np.diag(np.linspace(2.0, 0.2, no))
    lines = ["__qpu__ k() {", "0.9"]
    for p in range(no):
        for q in range(no):
            if abs(h1[p, q]) < 0.05:
                continue
            for s in (0, 1):
                lines.append(f"{h1[p, q]:.10f} {2 * p + s} 1 {2 * q + s} 0")
    for _ in range(60):
        p, q, r, s = rng.integers(0, no, 4)
        c = 0.1 * rng.normal()
        for s1 in (0, 1):
            for s2 in (0, 1):
                a, b, cc, d = 2 * p + s1, 2 * q + s2, 2 * r + s2, 2 * s + s1
                if a == b or cc == d:
                    continue
                lines.append(f"{c:.10f} {a} 1 {b} 1 {cc} 0 {d} 0")
                lines.append(f"{c:.10f} {d} 1 {cc} 1 {b} 0 {a} 0")
    lines.append("}")
    return "\n".join(lines)


def main():
    n, ne, nvec = 12, 4, 64
    op = xb.compile(lih_shaped_source(n), "bk", n)
    xm, zm, cf = op.pack()
    ans = xb.Ansatz.uccsd(n, ne, "bk")
    rng = np.random.default_rng(12)
    thetas = rng.uniform(-0.1, 0.1, size=(nvec, ans.n_params))
    be = xb.Backend(n)
    be.set_hamiltonian(xm, zm, cf)
    T, G = be.hamiltonian_stats()
    args = (ans.basis_index, ans.xmask, ans.zmask, ans.imag_coeff, ans.param, ans.n_params, thetas)
    e_sweep = be.sweep_energies(*args)
    be.synchronize()
    t0 = time.perf_counter()
    reps = 5
    for _ in range(reps):
        e_sweep = be.sweep_energies(*args)
    dt_sweep = (time.perf_counter() - t0) / reps
    l0 = be.kernel_launches()
    t0 = time.perf_counter()
    e_loop = np.zeros(nvec)
    for v in range(nvec):
        ans.prepare(be, thetas[v])
        e_loop[v] = be.expectation().real
    dt_loop = time.perf_counter() - t0
    launches_loop = be.kernel_launches() - l0
    print(json.dumps({"workload": "12-qubit BK molecule-shaped Hamiltonian, UCCSD(4e) energy sweep, 64 parameter vectors",
                      "n_terms": int(T), "n_xmask_groups": int(G), "n_rotations": int(len(ans.xmask)), "n_params": ans.n_params,
                      "sweep_kernel_s": dt_sweep, "energies_per_s_sweep": nvec / dt_sweep, "per_rotation_path_s": dt_loop,
                      "energies_per_s_per_rotation_path": nvec / dt_loop, "speedup": dt_loop / dt_sweep,
                      "launches_sweep": 1, "launches_per_rotation_path": int(launches_loop),
                      "max_abs_diff": float(np.abs(e_sweep - e_loop).max())}))


if __name__ == "__main__":
    main()
