#!/usr/bin/env python
"""bench.py -- headline benchmark of the hot path (BASELINE.json: term.amplitude updates/s).

Workload at N = 1 (configs[3]): synthetic 28-qubit random Pauli sum, 10^4 terms, fp64 <psi|H|psi>.  A step is one full
<psi|H|psi> over all terms.  At N > 1 the register grows by log2(N) qubits (weak scaling: 2^28 amplitudes per GPU), sharded
by the high index bits.  The same JSON line also carries H psi and one Lanczos step on the same operator (`extra`).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Prints ONE JSON line (rank 0); exits non-zero when a correctness guard fails.  See DESIGN.md "Measurement" for every field.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "H.psi term.amplitude updates/s"
UNIT = "updates/s"
N_QUBITS_1GPU = 28
N_TERMS = 10_000
SEED = 28
GUARD_TOL = 1e-10          # north_star: energies to 1e-10 relative
NVLINK_NOMINAL_GBS, NVLINK_MEASURED_GBS = 900.0, 770.0     # per direction per GPU (B200_PROFILING.md)


def synth_pauli_sum(n: int, n_terms: int, seed: int):
    """Config 4 of SURVEY.md 8d: every qubit's Pauli i.i.d. uniform over {I,X,Y,Z} (all-I rejected,
    duplicates merged), real N(0,1) coefficients.  Returns packed (xmask, zmask, coef = c*i^nY)."""
    rng = np.random.default_rng(seed)
    pa = rng.integers(0, 4, size=(n_terms, n), dtype=np.int8)
    pa[np.all(pa == 0, axis=1), 0] = 3
    xm = np.zeros(n_terms, np.uint64)
    zm = np.zeros(n_terms, np.uint64)
    for q in range(n):
        xm |= ((pa[:, q] == 1) | (pa[:, q] == 2)).astype(np.uint64) << np.uint64(q)
        zm |= ((pa[:, q] == 3) | (pa[:, q] == 2)).astype(np.uint64) << np.uint64(q)
    c = rng.normal(size=n_terms)
    order = np.lexsort((zm, xm))
    xm, zm, c = xm[order], zm[order], c[order]
    keep = np.ones(n_terms, bool)
    keep[1:] = (xm[1:] != xm[:-1]) | (zm[1:] != zm[:-1])
    # merge duplicates into the first occurrence
    idx = np.cumsum(keep) - 1
    cm = np.zeros(int(keep.sum()))
    np.add.at(cm, idx, c)
    xm, zm = xm[keep], zm[keep]
    ny = np.array([bin(int(a) & int(b)).count("1") for a, b in zip(xm, zm)])
    return xm, zm, cm * (1j ** (ny % 4)), cm


def product_state_expectations(xm, zm, a, b):
    ex, ey, ez = np.sin(2 * a) * np.cos(b), np.sin(2 * a) * np.sin(b), np.cos(2 * a)
    n = len(a)
    want = np.ones(len(xm))
    for q in range(n):
        xb = ((xm >> np.uint64(q)) & np.uint64(1)).astype(bool)
        zb = ((zm >> np.uint64(q)) & np.uint64(1)).astype(bool)
        want *= np.where(xb & zb, ey[q], np.where(xb, ex[q], np.where(zb, ez[q], 1.0)))
    return want


class ClockSampler:
    """Samples nvidia-smi during the timed region (recipe of B200_PROFILING.md)."""

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.rows = []
        self.stop = threading.Event()
        self.thread = None

    def _run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(self.gpu)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([v.strip() for v in out.split(",")])
            except Exception:
                pass
            self.stop.wait(0.2)

    def __enter__(self):
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.thread.join(timeout=6)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [nm for i, nm in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def load_oracle():
    """The C/OpenMP oracle with every host core: torchrun exports OMP_NUM_THREADS=1, which must not reach the CPU arm."""
    os.environ["OMP_NUM_THREADS"] = str(host_cores())
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import c_oracle as co
    co.lib()
    return co


_CPU_PSI = {}


def host_random_state(n: int):
    """Normalised complex Gaussian on the host (seeded per 2^24 block), cached: the CPU arm and the GPU cross-check share it."""
    psi = _CPU_PSI.get(n)
    if psi is None:
        N = 1 << n
        rng = np.random.default_rng(SEED)
        psi = np.empty(N, dtype=np.complex128)
        blk = 1 << 24
        for s in range(0, N, blk):
            e = min(N, s + blk)
            psi[s:e] = rng.standard_normal(e - s) + 1j * rng.standard_normal(e - s)
        psi *= 1.0 / np.sqrt(2.0 * N)
        _CPU_PSI.clear()
        _CPU_PSI[n] = psi
    return psi


def cpu_baseline(co, n: int, xm, zm, cf, n_sample_terms: int):
    """The oracle (C/OpenMP mask-based port of the reference algorithm) timed on the host cores on a
    bounded sample: the same register size, a seeded random subset of `n_sample_terms` terms.  Returns (record, seconds, selected indices, energy)."""
    N = 1 << n
    psi = host_random_state(n)
    k = min(n_sample_terms, len(xm))
    # a seeded RANDOM subset: the terms arrive sorted by xmask, so the first k would be the k smallest flip masks (top qubits never flipped)
    sel = np.sort(np.random.default_rng(SEED + 2).choice(len(xm), size=k, replace=False))
    co.expect(n, xm[:1], zm[:1], cf[:1], psi)          # warm up threads / page in
    t0 = time.perf_counter()
    e, _ = co.expect(n, xm[sel], zm[sel], cf[sel], psi)
    dt = time.perf_counter() - t0
    return ({"value": k * N / dt, "unit": UNIT, "cores": co.num_threads(), "kind": "port",
             "sample": f"seeded random {k} of {len(xm)} terms on the full 2^{n} register, C/OpenMP mask-based oracle (oracle/hpsi_oracle.c), {dt:.2f} s"},
            dt, sel, e)


def mem_available_bytes() -> int:
    try:
        with open("/proc/meminfo") as f:
            for line in f:
                if line.startswith("MemAvailable:"):
                    return int(line.split()[1]) * 1024
    except Exception:
        pass
    return 32 << 30


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    co = load_oracle()
    # the GPU arm grows the register with the GPU count (weak scaling); the CPU arm follows while the host RAM holds the state
    wl = int(np.log2(max(1, args.gpus)))
    n = args.qubits + wl
    cap = ""
    while n > args.qubits and (16 << n) > 0.6 * mem_available_bytes():
        n -= 1
        cap = f"; register capped at {n} qubits by host RAM"
    xm, zm, cf, _ = synth_pauli_sum(n, args.terms, SEED)
    # the same ~15-30 s of CPU work per step at every size
    sample_terms = max(2, args.cpu_terms >> (n - args.qubits))
    vals, times = [], []
    base = None
    for i in range(args.warmup + args.steps):
        base, dt, k, _ = cpu_baseline(co, n, xm, zm, cf, sample_terms)
        if i >= args.warmup:
            vals.append(base["value"])
            times.append(dt)
        if i == 0 and dt * (args.warmup + args.steps) > 240:     # keep the whole run within minutes
            sample_terms = max(2, int(sample_terms * 240 / (dt * (args.warmup + args.steps))))
    v = float(np.mean(vals))
    base["value"] = v
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * float(np.mean(times)), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"synthetic {n}-qubit random Pauli sum, {len(xm)} terms, fp64 <psi|H|psi> (bounded sample of terms per step{cap})",
                       "n_qubits": n, "n_terms": int(len(xm))},
            "cpu_baseline": base,
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def sharded_parity_check(xb, dist, torch, rank, world, local_rank):
    """N >= 2 only: H psi, <psi|H|psi> and a short Lanczos run of a SHARDED small register against the CPU oracle, so that every
    multi-GPU driver run exercises the exchange paths (the pytest multi-GPU cases skip on a 1-GPU box)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import c_oracle as co
    wl = int(np.log2(world))
    n = 18 + wl
    ids = [xb.Backend.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, src=0)
    be = xb.Backend(n, device=local_rank, rank=rank, world=world, comm_id=ids[0])
    be.n_slots(3)
    xm, zm, cf, _ = synth_pauli_sum(n, 96, SEED + 7)
    be.set_option("apply-engine", "tile")
    be.set_hamiltonian(xm, zm, cf)
    rng = np.random.default_rng(SEED + 8)
    psi = rng.normal(size=1 << n) + 1j * rng.normal(size=1 << n)
    psi /= np.linalg.norm(psi)
    lo, hi = be.first, be.first + be.n_local
    be.upload(psi[lo:hi])
    scale = float(np.abs(cf).sum())
    e = be.expectation()
    be.apply_h(0, 1, want_alpha=False)
    got = be.download(1)
    out = {}
    if rank == 0:
        eo, _ = co.expect(n, xm, zm, cf, psi)
        out["expect_rel_err"] = abs(e - eo) / max(abs(eo), 1e-300)
        out["expect_err_over_sum_abs_c"] = abs(e - eo) / scale
    want = co.apply(n, xm, zm, cf, psi)[lo:hi]
    err = torch.tensor([float(np.abs(got - want).max()) / scale], dtype=torch.float64, device="cuda")
    dist.all_reduce(err, op=dist.ReduceOp.MAX)
    out["apply_max_err_over_sum_abs_c"] = float(err.item())
    # Lanczos on a block-decoupled operator H_A (x) 1 + 1 (x) H_B with H_A on the sharded (high) qubits: E0 = E0(A) + E0(B)
    nA = wl + 3
    xa, za, ca, _ = synth_pauli_sum(nA, 24, SEED + 9)
    xbm, zbm, cb, _ = synth_pauli_sum(n - nA, 24, SEED + 10)
    xs = np.concatenate([xa << np.uint64(n - nA), xbm])
    zs = np.concatenate([za << np.uint64(n - nA), zbm])
    cs = np.concatenate([ca, cb])
    be.set_hamiltonian(xs, zs, cs)
    e0, it, res = be.lanczos_ground(tol=1e-10, max_iter=3000, seed=3)
    if rank == 0:
        from scipy.sparse.linalg import LinearOperator, eigsh

        def e0_of(nq, x, z, c):
            op = LinearOperator((1 << nq, 1 << nq), matvec=lambda v: co.apply(nq, x, z, c, v.astype(np.complex128)), dtype=np.complex128)
            return float(eigsh(op, k=1, which="SA", tol=1e-13, return_eigenvectors=False)[0])
        ea, eb = e0_of(nA, xa, za, ca), e0_of(n - nA, xbm, zbm, cb)
        out["lanczos_rel_err"] = abs(e0 - (ea + eb)) / abs(ea + eb)
        out["lanczos_iterations"] = int(it)
    be.close()
    return out


def run_gpu(args):
    import torch
    import torch.distributed as dist

    import xaccvqe_b200 as xb

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node N for --gpus N")
    torch.cuda.set_device(local_rank)
    comm_id = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        ids = [xb.Backend.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        comm_id = ids[0]
    wl = int(np.log2(world))
    n = args.qubits + wl
    n_terms = args.terms
    xm, zm, cf, c_plain = synth_pauli_sum(n, n_terms, SEED)
    N = 1 << n
    failures = []

    be = xb.Backend(n, device=local_rank, rank=rank, world=world, comm_id=comm_id)
    be.set_option("expect-engine", args.engine)
    be.set_option("apply-engine", args.engine)
    if args.chunk_bits is not None:
        be.set_option("tile-chunk-bits", args.chunk_bits)
    if args.apply_chunk_bits is not None:
        be.set_option("apply-chunk-bits", args.apply_chunk_bits)
    be.set_option("tile-bits", args.tile_bits)
    be.set_option("tile-planner", args.planner)
    be.set_option("expect-kernel", args.expect_kernel)
    t0 = time.perf_counter()
    be.set_hamiltonian(xm, zm, cf)
    set_ham_ms = (time.perf_counter() - t0) * 1e3
    Tn, G = be.hamiltonian_stats()
    plan = be.plan_stats()

    def barrier():
        if world > 1:
            torch.cuda.synchronize()
            dist.barrier()
        be.synchronize()

    # ---- correctness guards (untimed; a failure makes the run exit 1) ----
    # (i) a seeded product state has an exact, vector-free answer: sum_k c_k prod_q <sigma_q>
    #     Every qubit sits near the Bloch direction (1,1,1)/sqrt(3) (seeded jitter of +-0.15 rad on both angles), where <X>, <Y>, <Z> are all
    #     ~0.58: a weight-w string then contributes ~0.58^w instead of the ~0.45^w of uniformly random angles, i.e. the exact answer is
    #     hundreds of times further from pure cancellation at 28-33 qubits and the RELATIVE error below measures the kernel, not the
    #     conditioning of the sum (with uniform angles the same absolute error read 1.7e-11 of the 1e-10 budget at 29 qubits).
    rng = np.random.default_rng(SEED + 1)
    a = 0.5 * np.arccos(1.0 / np.sqrt(3.0)) + rng.uniform(-0.15, 0.15, n)
    b = np.pi / 4 + rng.uniform(-0.15, 0.15, n)
    be.product_state(a, b)
    e_prod = be.expectation().real
    want = float(np.dot(c_plain, product_state_expectations(xm, zm, a, b)))
    if args.inject_sign_fault:
        # self-test of the guard: the expected value is computed for an operator in which ONE term that flips the top qubit has the
        # opposite sign; the run must exit 1
        k_bad = int(np.nonzero((xm >> np.uint64(n - 1)) & np.uint64(1))[0][0])
        want -= 2.0 * float(c_plain[k_bad] * product_state_expectations(xm[k_bad:k_bad + 1], zm[k_bad:k_bad + 1], a, b)[0])
    check = {"product_state_rel_err": abs(e_prod - want) / abs(want), "product_state_energy": [e_prod, want],
             "product_state_err_over_sum_abs_c": abs(e_prod - want) / float(np.abs(c_plain).sum()),
             "product_state_plan": "the timed plan (all terms)"}
    if not check["product_state_rel_err"] <= GUARD_TOL:
        failures.append(f"product-state <psi|H|psi> off by {check['product_state_rel_err']:.3e} relative")
    # (ii) N = 1: a host-generated normalised complex Gaussian against the CPU oracle on the bounded term sample (SURVEY.md 8d config 4 ii);
    #      the oracle run is also the timed CPU baseline
    cpu_base = None
    if world == 1 and not args.no_cpu:
        co = load_oracle()
        cpu_base, _, sel, e_cpu = cpu_baseline(co, n, xm, zm, cf, args.cpu_terms)
        be.set_hamiltonian(xm[sel], zm[sel], cf[sel])
        be.upload(host_random_state(n))
        e_gpu = be.expectation()
        check["random_state_vs_oracle_rel_err"] = abs(e_gpu - e_cpu) / abs(e_cpu)
        check["random_state_vs_oracle_terms"] = int(len(sel))
        check["random_state_vs_oracle_plan"] = ("a plan built for the seeded random subset of terms (not the timed plan); the timed plan is exercised by "
                                                "product_state_rel_err (all terms, exact answer) and expect_vs_apply_rel_err (all terms, random state)")
        check["random_state_energy"] = [e_gpu.real, e_cpu.real]
        if not check["random_state_vs_oracle_rel_err"] <= GUARD_TOL:
            failures.append(f"random-state <psi|H|psi> differs from the oracle by {check['random_state_vs_oracle_rel_err']:.3e} relative")
        _CPU_PSI.clear()
        be.set_hamiltonian(xm, zm, cf)
    # (iii) N >= 2: sharded H psi / <psi|H|psi> / Lanczos of a small register against the oracle
    if world > 1 and not args.no_shard_check:
        sc = sharded_parity_check(xb, dist, torch, rank, world, local_rank)
        check["sharded"] = sc
        if rank == 0:
            if not sc["expect_err_over_sum_abs_c"] <= 1e-12:
                failures.append(f"sharded <psi|H|psi> off by {sc['expect_err_over_sum_abs_c']:.3e} (sum |c|)")
            if not sc["lanczos_rel_err"] <= GUARD_TOL:
                failures.append(f"sharded Lanczos E0 off by {sc['lanczos_rel_err']:.3e} relative")
        if not sc["apply_max_err_over_sum_abs_c"] <= 1e-12:
            failures.append(f"sharded H psi off by {sc['apply_max_err_over_sum_abs_c']:.3e} (sum |c|)")

    # ---- the timed region: K steps of <psi|H|psi> with the state resident in HBM (2^28 * 16 B = 4.3 GB per GPU >> 126 MB L2) ----
    be.random_state(0, SEED)
    op = args.op

    def step():
        if op == "apply":
            be.apply_h(0, 1, want_alpha=False)      # dst = H psi (one Lanczos matrix-vector product)
            return 0j
        if op == "terms":
            return be.expectation(per_term=True)[0]     # energy + per-term <P_k> (the reference's per-kernel exp-val-z)
        return be.expectation()

    if op == "apply":
        be.n_slots(2)
    for _ in range(args.warmup):
        step()
    barrier()
    l0 = be.kernel_launches()
    kernel_ms, xch = [], []
    with ClockSampler(local_rank) as clk:
        be.timer_start()
        for _ in range(args.steps):
            e = step()
            kernel_ms.append(be.last_kernel_ms())
            if world > 1 and op == "expect":
                xch.append(be.last_exchange() + (be.last_exchange_exposed(),))
        ms = be.timer_stop()
        barrier()
    launches = be.kernel_launches() - l0
    free_b, total_b = torch.cuda.mem_get_info()
    mem_used_gib = (total_b - free_b) / 2**30
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / args.steps
    value = Tn * N / (ms_per_step * 1e-3)
    clocks = clk.summary()

    # ---- end to end through the C ABI with host buffers: state from pinned host memory, energy back to the host ----
    # The Hamiltonian tables stay resident, as the reference builds its per-term kernels once per VQE run (VQEProgram::build) and
    # then calls execute() per parameter vector; set_hamiltonian (host planning + table upload) is reported separately.
    e2e = {"skipped": "--no-e2e (capacity run: the pinned host copy of the shard was not allocated)"}
    if not args.no_e2e:
        host = torch.empty(be.n_local * 2, dtype=torch.float64).pin_memory()
        host_np = host.numpy().view(np.complex128)
        be.download(0, out=host_np)
        be.upload(host_np)
        step()
        barrier()
        reps = max(1, min(args.steps, 2))
        t0 = time.perf_counter()
        for _ in range(reps):
            be.upload(host_np)
            step()
            if op == "apply":
                be.download(1, out=host_np)
        barrier()
        dt = (time.perf_counter() - t0) / reps
        if world > 1:
            tt = torch.tensor([dt], dtype=torch.float64, device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dt = float(tt.item())
        e2e = {"value": Tn * N / dt, "unit": UNIT, "h2d_bytes_per_step": int(be.n_local * 16),
               "d2h_bytes_per_step": int(be.n_local * 16) if op == "apply" else 16, "ms_per_step": dt * 1e3,
               "includes": "per step: state H2D from pinned host memory + the op + result D2H; the Hamiltonian tables are resident "
                           "(set once per VQE run, as the reference's VQEProgram::build)",
               "set_hamiltonian_ms": set_ham_ms,
               "value_with_set_hamiltonian_every_step": Tn * N / (dt + set_ham_ms * 1e-3)}
        del host

    # ---- extra: H psi and one Lanczos step on the same operator at every N (the headline stays <psi|H|psi>) ----
    # Sharded: times are the max over ranks; the slab exchange of H psi (option shard-exchange, default one ncclAllGather per H psi)
    # is reported with its CUDA-event transfer time and the time the compute stream waited for it (exposed).
    extra = {}
    if op == "expect" and not args.no_extra:
        def max_over_ranks(v):
            if world == 1:
                return float(v)
            tv = torch.tensor([float(v)], dtype=torch.float64, device="cuda")
            dist.all_reduce(tv, op=dist.ReduceOp.MAX)
            return float(tv.item())

        be.n_slots(3)
        be.random_state(0, SEED)
        alpha = be.apply_h(0, 1, want_alpha=True)      # plans H psi, warms up; alpha = <psi|H psi> through the H psi engine
        # (iv) the TIMED <psi|H|psi> plan on a random state, all terms: against <psi|(H psi)> from the independently planned H psi engine
        e_expect = be.expectation()
        check["expect_vs_apply_rel_err"] = abs(e_expect.real - alpha.real) / abs(alpha.real)
        check["expect_vs_apply_energy"] = [e_expect.real, alpha.real]
        if not check["expect_vs_apply_rel_err"] <= GUARD_TOL:
            failures.append(f"random-state <psi|H|psi> (timed plan) and <psi|(H psi)> differ by {check['expect_vs_apply_rel_err']:.3e} relative")
        barrier()
        la = be.kernel_launches()
        t_ms, x_ms_a, x_by_a, x_exp_a = [], [], [], []
        for _ in range(2):
            be.apply_h(0, 1, want_alpha=False)
            t_ms.append(be.last_kernel_ms())
            if world > 1:
                xm_, xb_ = be.last_exchange()
                x_ms_a.append(xm_)
                x_by_a.append(xb_)
                x_exp_a.append(be.last_exchange_exposed())
        la = (be.kernel_launches() - la) // 2
        a_ms = max_over_ranks(np.mean(t_ms))
        nb_a = be.plan_stats()["apply_batches"]
        peak, _ = measured_peak_gbs()
        alg = 16.0 * be.n_local * (G + 1)
        model = 16.0 * be.n_local * (3 * nb_a - 1)     # per batch: src read + dst read-modify-write (the first batch only writes)
        extra["apply"] = {"op": "dst = H psi (b200vqe_apply_h), same operator", "ms": a_ms, "value": Tn * N / (a_ms * 1e-3), "unit": UNIT,
                          "launches": int(la), "n_batches": int(nb_a),
                          "roofline": {"bound": "hbm", "achieved": alg / (a_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                                       "frac": alg / (a_ms * 1e-3) / 1e9 / peak,
                                       "note": "per GPU: ALGORITHMIC bytes 16 B x 2^n_local x (G + 1) (SURVEY.md 8d) / CUDA-event time (max over ranks)",
                                       "dram_model": {"bytes": model, "achieved": model / (a_ms * 1e-3) / 1e9, "frac": model / (a_ms * 1e-3) / 1e9 / peak,
                                                      "note": "16 B x 2^n_local x (3 n_batches - 1): src read + dst read + dst write per batch of xmask groups (rank 0's plan)"}}}
        if world > 1:
            xm_ms = float(np.mean(x_ms_a))
            extra["apply"]["nvlink"] = {"exchange": "nccl (shard-exchange = auto: one ncclAllGather of the register per H psi, or send/recv of the needed slabs)",
                                        "bytes_per_gpu": float(np.mean(x_by_a)), "exchange_ms": xm_ms, "exposed_ms": float(np.mean(x_exp_a)),
                                        "achieved": float(np.mean(x_by_a)) / (xm_ms * 1e-3) / 1e9 if xm_ms > 0 else None, "peak": NVLINK_NOMINAL_GBS,
                                        "unit": "GB/s", "note": "rank 0's figures; exposed_ms = CUDA-event time the compute stream waited for the slabs"}
        # one Lanczos iteration = H psi (alpha fused or one k_alpha pass) + k_lanczos_update; measured as the slope of two short runs
        m1, m2 = 3, 9

        def lanczos_steps(m):
            try:
                be.lanczos_ground(tol=1e-300, max_iter=m, seed=5)
            except xb.B200Error as ex:
                if ex.code != -6:          # B200VQE_ENOCONV: the run is meant to stop at max_iter
                    raise
            be.synchronize()

        barrier()
        t0 = time.perf_counter()
        lanczos_steps(m1)
        barrier()
        d1 = time.perf_counter() - t0
        l1 = be.kernel_launches()
        t0 = time.perf_counter()
        lanczos_steps(m2)
        barrier()
        d2 = time.perf_counter() - t0
        l2 = be.kernel_launches() - l1
        it_ms = max_over_ranks((d2 - d1) / (m2 - m1) * 1e3)
        alg_l = alg + 80.0 * be.n_local
        extra["lanczos_step"] = {"op": "one Lanczos iteration: H psi + alpha + three-term update with fused norm", "ms": it_ms,
                                 "value": Tn * N / (it_ms * 1e-3), "unit": UNIT, "launches_per_iteration": int(round(l2 / m2)),
                                 "roofline": {"bound": "hbm", "achieved": alg_l / (it_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                                              "frac": alg_l / (it_ms * 1e-3) / 1e9 / peak,
                                              "note": "per GPU: ALGORITHMIC bytes of H psi + 5 x 16 B x 2^n_local for the fused vector update / host-timed slope of two runs "
                                                      "(barrier on both sides, max over ranks)"}}
        be.n_slots(1)

    if world > 1:
        ok = torch.tensor([0 if failures else 1], dtype=torch.int32, device="cuda")
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) == 0 and not failures:
            failures.append("a correctness guard failed on another rank")
    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        if failures:
            sys.exit(1)
        return

    peak, peak_src = measured_peak_gbs()
    apply_op = op == "apply"
    # algorithmic bytes of one <psi|H|psi> on one GPU: 16 B x local amplitudes x groups
    alg_bytes = 16.0 * be.n_local * (G + 1 if apply_op else G)
    k_ms = ms_per_step if op == "terms" else float(np.mean(kernel_ms))     # terms: energy launch + per-term launch
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9
    tile = args.engine != "stream"
    kernel_name = (("k_tile_apply_pipe8" if tile else "k_apply<8,true>") if apply_op else
                   (("k_tile_expect_db<3>" if args.tile_bits == 11 else
                     {"auto": "k_tile_expect_pipe", "pipe": "k_tile_expect_pipe", 
                      "classic": "k_tile_expect<2,false>"}[args.expect_kernel])
                    if tile else "k_expect_energy<8,true>"))
    # DRAM traffic of the step: the tile engine streams this rank's amplitudes once per batch (pair registers: the peer slab too),
    # modelled from the LIVE plan; the ncu capture of the same launch (profiles/traffic.json) is quoted when it matches the plan
    traffic = traffic_src = None
    if tile and op == "expect":
        passes = plan["expect_batches"]
        traffic = 16.0 * be.n_local * passes
        traffic_src = f"model: 16 B x 2^{int(np.log2(be.n_local))} amplitudes x {passes} batches of the live plan (b200vqe_plan_stats)"
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["k_tile_expect_pipe"]
            if tj.get("n_qubits") == n and tj.get("n_terms") == int(Tn) and tj.get("n_batches") == passes and world == 1:
                traffic = float(tj["dram_bytes_per_launch"])
                traffic_src = "ncu dram__bytes_read.sum + dram__bytes_write.sum of this launch (profiles/traffic.json: " + tj.get("how", "") + ")"
        except Exception:
            pass
    roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
            "traffic_source": traffic_src, "kernel": kernel_name, "peak_source": peak_src,
            "launches_per_step": int(launches // max(1, args.steps)),
            "note": ("achieved = ALGORITHMIC bytes 16 B x 2^n_local x (G + 1) (SURVEY.md 8d: G partner reads + one write) / CUDA-event "
                     "time of the step's H.psi launches; the tile engine reads src once and read-modify-writes dst once per "
                     "BATCH of xmask groups (48 B per amplitude per batch), so frac > 1 is expected") if apply_op else
                    ("achieved = ALGORITHMIC bytes 16 B x 2^n_local x G (SURVEY.md 8d) / CUDA-event time of the step's <psi|H|psi> launch(es). "
                     "The tile engine evaluates ~14 xmask groups per HBM read of the state (GF(2)-subspace tiles in shared memory), so frac > 1 "
                     "by design; the physical bounds of the kernel are in 'dram' (bytes it really moves), 'lsu' (shared-memory pipe) and 'fp64'") if tile else
                    ("achieved = 16 B x 2^n_local x G per step / CUDA-event time of the step's k_expect_energy launches "
                     "(one launch per xmask bucket); the pair form reads each pair once, so DRAM traffic is about half the model")}
    if tile and op == "expect" and traffic:
        n_sm = torch.cuda.get_device_properties(local_rank).multi_processor_count
        roof["dram"] = {"achieved": traffic / (k_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s", "frac": traffic / (k_ms * 1e-3) / 1e9 / peak,
                        "n_batches": plan["expect_batches"]}
        clk_hz = float(clocks.get("sm_mhz") or 0) * 1e6
        if clk_hz > 0:
            # LSU data-pipe wavefronts the design cannot avoid (one 128-byte wavefront per clock per SM), from the LIVE plan
            # (b200vqe_plan_costs): per tile and batch the cp.async fill (64 KB = 512 wavefronts), per view one register load of the
            # 8 warps (128 LDS.128 x 4), per partner-reading record 64 LDS.128 x 4; fused blocks and pure records read nothing
            costs = be.plan_costs()
            tiles = be.n_local / 4096.0
            wave = tiles * (plan["expect_batches"] * 512 + costs["view_loads"] * 512 + costs["partner_records"] * 256)
            roof["lsu"] = {"achieved": wave / (k_ms * 1e-3) / 1e9, "peak": n_sm * clk_hz / 1e9, "unit": "G wavefronts/s",
                           "frac": wave / (k_ms * 1e-3) / (n_sm * clk_hz), "plan_costs": costs,
                           "note": "design floor of shared-memory wavefronts of the live plan / (SMs x sampled SM clock); the kernel is bound by "
                                   "instruction issue and LDS latency at 16 warps per SM, not by this pipe: "
                                   "profiles/r2j_ncu_full_k_tile_expect_pipe_fused_2000terms.txt"}
            # fp64 pipe: per xmask group and thread 16 pair-product + 8 butterfly/accumulate DP instructions (64 lanes per clock per SM)
            dp = G * (be.n_local / 16.0) * 24.0
            roof["fp64"] = {"achieved": dp / (k_ms * 1e-3) / 1e12, "peak": n_sm * 64 * clk_hz / 1e12, "unit": "T DP instr/s",
                            "frac": dp / (k_ms * 1e-3) / (n_sm * 64 * clk_hz),
                            "note": "DMUL/DFMA thread-instructions of the record bodies / (SMs x 64 fp64 lanes x sampled SM clock)"}
    if world > 1 and xch:
        x_ms = float(np.mean([v[0] for v in xch]))
        x_by = float(np.mean([v[1] for v in xch]))
        gbs = x_by / (x_ms * 1e-3) / 1e9 if x_ms > 0 else None
        mode = be.exchange_mode()
        roof["nvlink"] = {"bytes_per_gpu_per_step": x_by, "exchange_ms_per_step": x_ms, "achieved": gbs,
                          "exposed_ms_per_step": float(np.mean([v[2] for v in xch])), "exchange": mode,
                          "peak": NVLINK_NOMINAL_GBS, "unit": "GB/s", "frac": gbs / NVLINK_NOMINAL_GBS if gbs else None,
                          "frac_of_measured_770": gbs / NVLINK_MEASURED_GBS if gbs else None,
                          "exchange_share_of_step": x_ms / k_ms if k_ms > 0 else None,
                          "note": "slab swaps of the x_rank classes (" + ("peer-memory pushes through CUDA IPC staging, copy engine over NVLink, ordered by flag words"
                                                                               if mode == "peer-copy" else "grouped ncclSend/ncclRecv") +
                                  ", both directions at once): bytes this rank RECEIVES per step / CUDA-event time of the swaps on the communication stream, "
                                  "summed over the classes (rank 0's figures).  The swaps run concurrently with the tile kernels of the previous class: "
                                  "exchange_share_of_step is overlapped time; exposed_ms_per_step is the CUDA-event time the compute stream spent "
                                  "waiting for a slab (b200vqe_last_exchange_exposed), the part that was not hidden"}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": f"synthetic {n}-qubit random Pauli sum, {Tn} terms, fp64 " + ("H.psi" if apply_op else ("<psi|H|psi> + per-term <P_k>" if op == "terms" else "<psi|H|psi>")),
                       "n_qubits": n, "n_terms": int(Tn), "n_xmask_groups": int(G), "amplitudes_per_gpu": int(be.n_local),
                       "parallelism": f"state sharded over {world} GPU(s) by high index bits" + (
                           "; per x_rank class the two ranks of a pair swap slabs (" + be.exchange_mode() + ") and evaluate half of the pair register each, "
                           "one ncclAllReduce of the partials" if world > 1 else ""),
                       "l2": "inputs larger than L2 (4.3 GB state per GPU vs 126 MB)"},
            "roofline": roof, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "check": check, "plan": plan,
            "memory": {"device_used_gib_rank0": mem_used_gib, "state_gib_per_gpu": be.n_local * 16 / 2**30,
                       "note": "cudaMemGetInfo total - free on rank 0 right after the timed region (state slot + staging slabs + tables + CUDA/NCCL context)"}}
    line["check"]["energy"] = [e.real, e.imag]
    if extra:
        line["extra"] = extra
    if cpu_base is not None:
        line["cpu_baseline"] = cpu_base
    if failures:
        line["failed"] = failures
    emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if failures:
        print("bench.py: correctness guard failed: " + "; ".join(failures), file=sys.stderr)
        sys.exit(1)


_REAL_STDOUT = None


def _guard_stdout():
    """stdout carries exactly ONE JSON line: everything libraries print (NCCL's version banner and INFO lines, torch warnings)
    goes to stderr; emit() writes the line to the saved descriptor."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    _guard_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--qubits", type=int, default=N_QUBITS_1GPU, help="qubits per GPU shard (default: the named config)")
    ap.add_argument("--terms", type=int, default=N_TERMS)
    ap.add_argument("--cpu-terms", type=int, default=256, help="terms in the bounded CPU-baseline sample")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU oracle legs (cpu_baseline and the random-state cross-check)")
    ap.add_argument("--no-extra", action="store_true", help="skip the H psi / Lanczos-step measurements")
    ap.add_argument("--no-e2e", action="store_true", help="skip the end-to-end leg (host copy of the shard in pinned memory); for capacity runs beyond the named config")
    ap.add_argument("--inject-sign-fault", action="store_true", help="self-test: expect the product-state guard to fail (one top-qubit term mis-signed in the expected value)")
    ap.add_argument("--no-shard-check", action="store_true", help="N > 1: skip the small sharded parity check")
    ap.add_argument("--op", default="expect", choices=["expect", "apply", "terms"], help="expect: <psi|H|psi> (headline); apply: dst = H psi; terms: <psi|H|psi> plus every <P_k>")
    ap.add_argument("--engine", default="auto", choices=["auto", "tile", "stream"], help="<psi|H|psi> kernel family")
    ap.add_argument("--tile-bits", type=int, default=12, choices=[11, 12], help="tile engine: 12 = 4096-amplitude tiles, 11 = 2048 double-buffered")
    ap.add_argument("--expect-kernel", default="auto", choices=["auto", "pipe", "classic"])
    ap.add_argument("--planner", type=int, default=3, choices=[1, 2, 3], help="tile planner version (csrc/tile_plan.hpp)")
    ap.add_argument("--chunk-bits", type=int, default=None, help="tile engine: log2 contiguous amplitudes per HBM chunk (library default: 3 for <psi|H|psi>, 4 for H psi)")
    ap.add_argument("--apply-chunk-bits", type=int, default=None, help="the same for the H psi plans only")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
