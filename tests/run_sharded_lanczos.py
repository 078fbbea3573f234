#!/usr/bin/env python
"""Config 5 of BASELINE.json: sharded Lanczos ground state of a synthetic Pauli sum on N GPUs, verified against a
decoupled-block oracle.  Verification script (imports oracle/), not product code.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 \
        tests/run_sharded_lanczos.py --qubits 32

H = 1 (x) H_B(q0..q_{m-1}) + H_A(q_m..q_{2m-1}) (x) 1 with m = n/2: each block is a transverse-field-Ising backbone plus seeded
random Pauli strings of weight <= 4 (T_A = T_B = 500 with --random-terms 468, seeds 3201 / 3202, tol 1e-12).  Block A lives on the high (sharded) index bits, so its terms read peer shards.
E0(H) = E0(A) + E0(B); the blocks are solved on the host by ARPACK over the C oracle (2^m amplitudes).
Prints one JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def block_hamiltonian(m: int, n_random: int, seed: int):
    """(xmask, zmask, plain real coefficient) on m qubits: -J sum Z_i Z_{i+1} - h sum X_i + random weight<=4 strings."""
    rng = np.random.default_rng(seed)
    terms = {}
    for i in range(m):
        terms[(0, (1 << i) | (1 << ((i + 1) % m)))] = -0.5
        terms[(1 << i, 0)] = -1.0 - 0.1 * rng.random()
    while len(terms) < 2 * m + n_random:
        w = int(rng.integers(1, 5))
        qs = rng.choice(m, w, replace=False)
        x = z = 0
        for q in qs:
            p = int(rng.integers(1, 4))
            if p in (1, 2):
                x |= 1 << int(q)
            if p in (2, 3):
                z |= 1 << int(q)
        terms.setdefault((x, z), 0.15 * float(rng.normal()))
    keys = sorted(terms)
    xm = np.array([k[0] for k in keys], np.uint64)
    zm = np.array([k[1] for k in keys], np.uint64)
    c = np.array([terms[k] for k in keys])
    return xm, zm, c


def packed(xm, zm, c):
    ny = np.array([bin(int(a) & int(b)).count("1") for a, b in zip(xm, zm)])
    return c * (1j ** (ny % 4))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--qubits", type=int, default=32)
    ap.add_argument("--random-terms", type=int, default=468, help="random strings per block on top of the 2m backbone terms (468: T_A = T_B = 500 at 32 qubits)")
    ap.add_argument("--tol", type=float, default=1e-12)
    ap.add_argument("--max-iter", type=int, default=1000)
    ap.add_argument("--out", default="")
    ap.add_argument("--apply-engine", default="auto", choices=["auto", "tile", "stream"])
    ap.add_argument("--shard-exchange", default="auto", choices=["auto", "allgather", "needed", "pairwise"])
    args = ap.parse_args()
    import torch
    import torch.distributed as dist

    import xaccvqe_b200 as xb
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    comm_id = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        ids = [xb.Backend.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        comm_id = ids[0]
    n = args.qubits
    m = n // 2
    assert n == 2 * m
    xb_, zb_, cb = block_hamiltonian(m, args.random_terms, 3202)      # block B: low qubits
    xa_, za_, ca = block_hamiltonian(m, args.random_terms, 3201)      # block A: high qubits (sharded bits)
    xm = np.concatenate([xb_, xa_ << np.uint64(m)])
    zm = np.concatenate([zb_, za_ << np.uint64(m)])
    cf = np.concatenate([packed(xb_, zb_, cb), packed(xa_, za_, ca)])

    be = xb.Backend(n, device=local, rank=rank, world=world, comm_id=comm_id)
    be.set_option("apply-engine", args.apply_engine)
    be.set_option("shard-exchange", args.shard_exchange)
    be.set_hamiltonian(xm, zm, cf)
    T, G = be.hamiltonian_stats()
    be.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    l0 = be.kernel_launches()
    e0, it, res = be.lanczos_ground(tol=args.tol, max_iter=args.max_iter, seed=32)
    be.synchronize()
    if world > 1:
        dist.barrier()
    dt = time.perf_counter() - t0
    launches = be.kernel_launches() - l0
    x_ms, x_bytes = be.last_exchange()                     # the last H psi of the run: NVLink transfers and the wait for them
    x_exposed = be.last_exchange_exposed()
    if world > 1:
        t = torch.tensor([dt], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    if rank == 0:
        import c_oracle as co
        import pauli_oracle as po
        ea, _ = po.ground_state_eigsh(xa_, za_, packed(xa_, za_, ca), m, tol=1e-13, apply=lambda v: co.apply(m, xa_, za_, packed(xa_, za_, ca), v))
        eb, _ = po.ground_state_eigsh(xb_, zb_, packed(xb_, zb_, cb), m, tol=1e-13, apply=lambda v: co.apply(m, xb_, zb_, packed(xb_, zb_, cb), v))
        want = ea + eb
        N = 1 << n
        line = {"config": f"synthetic {n}-qubit Pauli sum H_A(x)1 + 1(x)H_B, sharded Lanczos ground state", "n_qubits": n, "n_gpus": world, "apply_engine": args.apply_engine, "shard_exchange": args.shard_exchange,
                "n_terms": int(T), "n_xmask_groups": int(G), "state_GiB": N * 16 / 2**30, "eigenvalue": e0, "oracle_E0A_plus_E0B": want,
                "rel_err": abs(e0 - want) / abs(want), "iterations": int(it), "ritz_residual": res, "seconds": dt,
                "ms_per_iteration": 1e3 * dt / max(1, it), "hpsi_updates_per_s": T * N * it / dt, "gpu_launches": int(launches),
                "nvlink_GB_received_per_gpu_per_iteration": x_bytes / 1e9, "exchange_ms_per_iteration": x_ms,
                "nvlink_GB_per_s": (x_bytes / 1e9) / (x_ms * 1e-3) if x_ms > 0 else 0.0,
                "nvlink_frac_of_900": ((x_bytes / 1e9) / (x_ms * 1e-3) / 900.0) if x_ms > 0 else 0.0,
                "exposed_exchange_ms_per_iteration": x_exposed, "plan": be.plan_stats(),
                "ritz_residual_over_abs_E0": res / abs(e0),
                "pass_1e-10": bool(abs(e0 - want) <= 1e-10 * abs(want) and res <= 1e-10 * abs(e0))}
        print(json.dumps(line))
        if args.out:
            with open(args.out, "w") as f:
                f.write(json.dumps(line) + "\n")
    be.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
