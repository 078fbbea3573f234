#!/usr/bin/env python
"""UCCSD state preparation above 12 qubits on one B200: the rotation list of ir/algorithms/uccsd/UCCSD.cpp applied to the resident
state, one launch (= one HBM pass, 32 B per amplitude) per rotation against fused tile runs (kernels_rot_tile.cuh).
Prints one JSON line per register size: rotations, HBM passes and CUDA-event ms of both paths, GB/s against the 32 B model.
usage: python tools/bench_ansatz.py [qubits ...]   (default 16 20 24)"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import xaccvqe_b200 as xb  # noqa: E402


def main():
    sizes = [int(a) for a in sys.argv[1:]] or [16, 20, 24]
    for n in sizes:
        ne = 4
        x, z, ic, pa, npar = xb.uccsd_rotations(n, ne, "jw")
        rng = np.random.default_rng(n)
        phi = ic * rng.uniform(-0.1, 0.1, npar)[pa]
        if len(x) > 4000:                      # a contiguous slice keeps the structure of the list (8 strings per double excitation)
            x, z, phi = x[:4000], z[:4000], phi[:4000]
        out = {"ansatz": f"UCCSD, {n} qubits, {ne} electrons, JW", "n_qubits": n, "n_params": int(npar), "rotations": int(len(x))}
        states = {}
        for mode in ("off", "on"):
            be = xb.Backend(n)
            be.set_option("fuse-rotations", mode)
            be.reset(0, (1 << ne) - 1)
            be.apply_pauli_rotations(x[:64], z[:64], phi[:64])       # warm-up
            be.reset(0, (1 << ne) - 1)
            be.synchronize()
            l0 = be.kernel_launches()
            be.timer_start()
            be.apply_pauli_rotations(x, z, phi)
            ms = be.timer_stop()
            passes = be.kernel_launches() - l0
            states[mode] = be.download()
            key = "fused" if mode == "on" else "per_rotation"
            out[key] = {"hbm_passes": int(passes), "ms": ms, "ms_per_rotation": ms / len(x),
                        "GBps_of_32B_per_amplitude_per_pass": 32.0 * (1 << n) * passes / (ms * 1e-3) / 1e9,
                        "rotations_per_pass": len(x) / passes}
            be.close()
        out["max_abs_diff_between_paths"] = float(np.abs(states["on"] - states["off"]).max())
        out["pass_reduction"] = out["per_rotation"]["hbm_passes"] / out["fused"]["hbm_passes"]
        out["speedup"] = out["per_rotation"]["ms"] / out["fused"]["ms"]
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
