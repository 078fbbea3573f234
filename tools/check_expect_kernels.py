"""GPU check: every <psi|H|psi> kernel variant of the tile engine against the C oracle on one random state.
usage: python tools/check_expect_kernels.py [n_qubits] [n_terms]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import c_oracle as co  # noqa: E402
import xaccvqe_b200 as xb  # noqa: E402
from test_tile_plan_cpu import random_hamiltonian  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20
T = int(sys.argv[2]) if len(sys.argv) > 2 else 200
xm, zm, cf = random_hamiltonian(n, T, 5)
rng = np.random.default_rng(6)
psi = rng.normal(size=1 << n) + 1j * rng.normal(size=1 << n)
psi /= np.linalg.norm(psi)
eo, _ = co.expect(n, xm, zm, cf, psi)
scale = np.abs(cf).sum()
ok = True
for kern, planner in (("pipe", 3), ("pipe", 2), ("pipe", 1), ("classic", 2), ("classic", 1)):
    be = xb.Backend(n, device=0)
    be.set_option("expect-kernel", kern)
    be.set_option("tile-planner", planner)
    be.set_hamiltonian(xm, zm, cf)
    be.upload(psi)
    e = be.expectation()
    err = abs(e - eo) / scale
    print(f"{kern:8s} planner {planner}: E = {e.real:+.15f}  |dE|/sum|c| = {err:.2e}  {be.last_kernel_ms():.3f} ms", flush=True)
    ok &= err < 1e-12
    del be
print("OK" if ok else "MISMATCH")
sys.exit(0 if ok else 1)
