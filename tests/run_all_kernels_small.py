#!/usr/bin/env python
"""Small invocation of every tile-engine kernel, checked against the CPU oracle as it goes (verification script, not
product code).  Written for `compute-sanitizer --tool racecheck|memcheck python tests/run_all_kernels_small.py` (SURVEY.md
section 5); compute-sanitizer is closed on this pool's boxes, so in round 1 it only ran bare."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import c_oracle as co  # noqa: E402
import xaccvqe_b200 as xb  # noqa: E402
from test_gpu_parity import random_hamiltonian, random_state  # noqa: E402


def main():
    n, T = 13, 24
    xm, zm, cf = random_hamiltonian(n, T, 7)
    psi = random_state(n, 8)
    eo, pero = co.expect(n, xm, zm, cf, psi)
    ref = co.apply(n, xm, zm, cf, psi)
    scale = np.abs(cf).sum()
    for tile_bits in (12, 11):
        be = xb.Backend(n)
        be.n_slots(2)
        be.set_option("expect-engine", "tile")
        be.set_option("apply-engine", "tile")
        be.set_option("tile-bits", tile_bits)
        be.set_hamiltonian(xm, zm, cf)
        be.upload(psi)
        assert abs(be.expectation() - eo) < 1e-10 * scale                       # k_tile_expect_pipe (views, fused blocks) / k_tile_expect_db<3>
        e, per = be.expectation(per_term=True)                                  # k_tile_expect<2, true> (tile_bits 12)
        assert np.abs(per - pero.real).max() < 1e-11
        be.apply_h(0, 1)                                                        # k_tile_apply_pipe8
        assert np.abs(be.download(1) - ref).max() < 1e-12 * scale
        be.set_number_sector(n // 2, "bk")                                      # k_project_sector
        be.project_sector(0)
        e0, it, _ = be.lanczos_ground(tol=1e-8, max_iter=1000, seed=3)            # k_lanczos_update, k_alpha, ...
        be.close()
    # shared-memory sweep kernel
    n2 = 8
    xs, zs, cs = random_hamiltonian(n2, 12, 9)
    be = xb.Backend(n2)
    be.set_hamiltonian(xs, zs, cs)
    x, z, ic, pa, npar = xb.uccsd_rotations(n2, 4, "jw")
    th = np.random.default_rng(1).uniform(-0.1, 0.1, size=(3, npar))
    es = be.sweep_energies(0b1111, x, z, ic, pa, npar, th)                      # k_sweep_small
    for v in range(3):
        be.reset(0, 0b1111)
        be.apply_pauli_rotations(x, z, ic * th[v][pa])                          # k_pauli_rotation (below one tile: no fused runs)
        assert abs(be.expectation().real - es[v]) < 1e-12 * np.abs(cs).sum()
    be.close()
    print("all kernels small: ok")


if __name__ == "__main__":
    main()
