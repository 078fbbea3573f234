#!/usr/bin/env python
"""bench.py -- headline benchmark of the hot path (BASELINE.json: term.amplitude updates/s).

Workload at N = 1 (configs[3]): synthetic 28-qubit random Pauli sum, 10^4 terms, fp64
<psi|H|psi>.  A step is one full <psi|H|psi> over all terms.  At N > 1 the register grows by
log2(N) qubits (weak scaling: 2^28 amplitudes per GPU), sharded by the high index bits.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for every field.
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
# NCCL writes its banner ("NCCL version ...") to stdout at NCCL_DEBUG=VERSION/INFO; stdout carries exactly one JSON line
if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION", "INFO"):
    os.environ["NCCL_DEBUG"] = "WARN"

METRIC = "H.psi term.amplitude updates/s"
UNIT = "updates/s"
N_QUBITS_1GPU = 28
N_TERMS = 10_000
SEED = 28


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


_CPU_PSI = {}


def cpu_baseline(n: int, xm, zm, cf, n_sample_terms: int):
    """The oracle (C/OpenMP mask-based port of the reference algorithm) timed on the host cores on a
    bounded sample: the same register size, the first `n_sample_terms` terms."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import c_oracle as co
    N = 1 << n
    psi = _CPU_PSI.get(n)
    if psi is None:                      # the reference arm times several samples on the same state
        rng = np.random.default_rng(SEED)
        psi = np.empty(N, dtype=np.complex128)
        blk = 1 << 24
        for s in range(0, N, blk):
            e = min(N, s + blk)
            psi[s:e] = rng.standard_normal(e - s) + 1j * rng.standard_normal(e - s)
        psi *= 1.0 / np.sqrt(2.0 * N)
        _CPU_PSI.clear()
        _CPU_PSI[n] = psi
    k = min(n_sample_terms, len(xm))
    co.expect(n, xm[:1], zm[:1], cf[:1], psi)          # warm up threads / page in
    t0 = time.perf_counter()
    co.expect(n, xm[:k], zm[:k], cf[:k], psi)
    dt = time.perf_counter() - t0
    return {"value": k * N / dt, "unit": UNIT, "cores": co.num_threads(), "kind": "port",
            "sample": f"{k} of {len(xm)} terms on the full 2^{n} register, C/OpenMP mask-based oracle (oracle/hpsi_oracle.c), {dt:.2f} s"}, dt, k


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = N_QUBITS_1GPU       # host RAM bounds the CPU arm to the 1-GPU register
    xm, zm, cf, _ = synth_pauli_sum(n, N_TERMS, SEED)
    sample_terms = args.cpu_terms
    vals, times = [], []
    base = None
    for i in range(args.warmup + args.steps):
        base, dt, k = cpu_baseline(n, xm, zm, cf, sample_terms) if i == 0 or True else (base, 0, 0)
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
            "config": {"workload": f"synthetic {n}-qubit random Pauli sum, {len(xm)} terms, fp64 <psi|H|psi> (bounded sample of terms per step)",
                       "n_qubits": n, "n_terms": int(len(xm))},
            "cpu_baseline": base,
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


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
    T = len(xm)
    N = 1 << n

    be = xb.Backend(n, device=local_rank, rank=rank, world=world, comm_id=comm_id)
    be.set_option("expect-engine", args.engine)
    be.set_option("apply-engine", args.engine)
    be.set_option("tile-chunk-bits", args.chunk_bits)
    be.set_option("tile-bits", args.tile_bits)
    be.set_option("tile-ctas-per-sm", args.tile_ctas)
    be.set_option("tile-planner", args.planner)
    be.set_option("tile-dispatch", args.dispatch)
    be.set_option("expect-kernel", args.expect_kernel)
    be.set_hamiltonian(xm, zm, cf)
    Tn, G = be.hamiltonian_stats()

    # correctness guard (untimed): a seeded product state has an exact, vector-free answer
    rng = np.random.default_rng(SEED + 1)
    a, b = rng.uniform(0, np.pi, n), rng.uniform(0, 2 * np.pi, n)
    be.product_state(a, b)
    e_prod = be.expectation().real
    want = float(np.dot(c_plain, product_state_expectations(xm, zm, a, b)))
    check_rel = abs(e_prod - want) / max(1.0, float(np.abs(c_plain).sum()))

    be.random_state(0, SEED)        # input resident in HBM; 2^28 * 16 B = 4.3 GB >> 126 MB L2
    apply_op = args.op == "apply"
    if apply_op:
        be.n_slots(2)

    def step():
        if apply_op:
            be.apply_h(0, 1, want_alpha=False)      # dst = H psi (one Lanczos matrix-vector product)
            return 0j
        if args.op == "terms":
            return be.expectation(per_term=True)[0]     # energy + per-term <P_k> (the reference's per-kernel exp-val-z)
        return be.expectation()

    def barrier():
        if world > 1:
            torch.cuda.synchronize()
            dist.barrier()
        be.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    l0 = be.kernel_launches()
    kernel_ms = []
    with ClockSampler(local_rank) as clk:
        be.timer_start()
        for _ in range(args.steps):
            e = step()
            kernel_ms.append(be.last_kernel_ms())
        ms = be.timer_stop()
        barrier()
    launches = be.kernel_launches() - l0
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / args.steps
    value = Tn * N / (ms_per_step * 1e-3)

    # end to end through the C ABI with host buffers: tables + state from pinned host memory, energy back
    e2e = None
    if world == 1:
        host = torch.empty(be.n_local * 2, dtype=torch.float64).pin_memory()
        host_np = host.numpy().view(np.complex128)
        be.download(0, out=host_np)
        be.upload(host_np)
        step()
        be.synchronize()
        t0 = time.perf_counter()
        reps = max(1, min(args.steps, 2))
        for _ in range(reps):
            be.set_hamiltonian(xm, zm, cf)
            be.upload(host_np)
            e2 = step()
            if apply_op:
                be.download(1, out=host_np)
        be.synchronize()
        dt = (time.perf_counter() - t0) / reps
        e2e = {"value": Tn * N / dt, "unit": UNIT, "h2d_bytes_per_step": int(be.n_local * 16 + T * 32),
               "d2h_bytes_per_step": int(be.n_local * 16) if apply_op else 16, "ms_per_step": dt * 1e3}
        del host
    else:
        host = torch.empty(be.n_local * 2, dtype=torch.float64).pin_memory()
        host_np = host.numpy().view(np.complex128)
        be.download(0, out=host_np)
        barrier()
        t0 = time.perf_counter()
        be.set_hamiltonian(xm, zm, cf)
        be.upload(host_np)
        step()
        if apply_op:
            be.download(1, out=host_np)
        barrier()
        dt = time.perf_counter() - t0
        tt = torch.tensor([dt], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
        e2e = {"value": Tn * N / dt, "unit": UNIT, "h2d_bytes_per_step": int(be.n_local * 16 + T * 32),
               "d2h_bytes_per_step": int(be.n_local * 16) if apply_op else 16, "ms_per_step": dt * 1e3}

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peak_gbs()
    # algorithmic bytes of one <psi|H|psi> on one GPU: 16 B x local amplitudes x groups
    alg_bytes = 16.0 * be.n_local * (G + 1 if apply_op else G)
    k_ms = ms_per_step if args.op == "terms" else float(np.mean(kernel_ms))     # terms: energy launch + per-term launch
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9
    traffic = traffic_tile = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            tj = json.load(open(tpath))
            traffic = tj.get("k_expect_energy_bytes_per_step")
            # ncu --set full capture of k_tile_expect on this exact workload (dram__bytes_read.sum + dram__bytes_write.sum, one launch)
            if tj.get("k_tile_expect", {}).get("n_qubits") == n and tj["k_tile_expect"].get("n_terms") == int(Tn):
                traffic_tile = tj["k_tile_expect"]["dram_bytes_per_launch"]
        except Exception:
            traffic = None
    tile = args.engine != "stream"
    if tile and traffic_tile is not None:
        traffic = traffic_tile
    roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
            "kernel": ("k_tile_apply" if tile else "k_apply<8,true>") if apply_op else (("k_tile_expect_db<3>" if args.tile_bits == 11 else "k_tile_expect<2>") if tile else "k_expect_energy<8,true>"),
            "peak_source": peak_src,
            "launches_per_step": int(launches // max(1, args.steps)),
            "note": ("achieved = ALGORITHMIC bytes 16 B x 2^n_local x (G + 1) (SURVEY.md 8d: G partner reads + one write) / CUDA-event "
                     "time of the step's H.psi launch(es); the tile engine reads src once and read-modify-writes dst once per "
                     "BATCH of xmask groups (48 B per amplitude per batch), so frac > 1 is expected") if apply_op else
                    ("achieved = ALGORITHMIC bytes 16 B x 2^n_local x G (SURVEY.md 8d) / CUDA-event time of the step's single "
                     "k_tile_expect launch. The tile engine evaluates >= 9 xmask groups per HBM read of the state (GF(2)-subspace "
                     "tiles in shared memory), so its DRAM traffic is ~16 B x 2^n x n_batches and frac > 1 is expected; the "
                     "kernel is bound by shared-memory bandwidth (see 'shared_memory' and DESIGN.md 3.1)") if tile else
                    ("achieved = 16 B x 2^n_local x G per step / CUDA-event time of the step's k_expect_energy launches "
                     "(one launch per xmask bucket); the pair form reads each pair once, so DRAM traffic is about half the model")}
    clocks = clk.summary()
    if tile and not apply_op and args.op == "expect" and traffic:
        # the two physical bounds of the tile kernel next to the algorithmic one: measured DRAM bytes of this launch and the
        # shared-memory bytes its design cannot avoid (16 B partner read per pair per group, 17/16 for the all-registers class,
        # plus the tile fill of every batch) against 128 B/clk/SM at the SM clock sampled during the run
        n_batches = traffic / (16.0 * be.n_local)
        smem_bytes = 16.0 * (G * (be.n_local / 2) * (17.0 / 16.0) + n_batches * be.n_local)
        smem_peak = torch.cuda.get_device_properties(local_rank).multi_processor_count * 128.0 * float(clocks.get("sm_mhz") or 0) * 1e6 / 1e9
        roof["dram"] = {"achieved": traffic / (k_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s", "frac": traffic / (k_ms * 1e-3) / 1e9 / peak}
        if smem_peak > 0:
            roof["shared_memory"] = {"achieved": smem_bytes / (k_ms * 1e-3) / 1e9, "peak": smem_peak, "unit": "GB/s",
                                     "frac": smem_bytes / (k_ms * 1e-3) / 1e9 / smem_peak,
                                     "note": "design floor of shared-memory bytes / (SMs x 128 B/clk x sampled SM clock)"}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": f"synthetic {n}-qubit random Pauli sum, {Tn} terms, fp64 " + ("H.psi" if apply_op else ("<psi|H|psi> + per-term <P_k>" if args.op == "terms" else "<psi|H|psi>")),
                       "n_qubits": n, "n_terms": int(Tn), "n_xmask_groups": int(G), "amplitudes_per_gpu": int(be.n_local),
                       "parallelism": f"state sharded over {world} GPU(s) by high index bits",
                       "l2": "inputs larger than L2 (4.3 GB state per GPU vs 126 MB)"},
            "roofline": roof, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
            "check": {"product_state_rel_err": check_rel, "energy": [e.real, e.imag]}}
    if world == 1 and not args.no_cpu:
        base, _, _ = cpu_baseline(n, xm, zm, cf, args.cpu_terms)
        line["cpu_baseline"] = base
    emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


_REAL_STDOUT = None


def _guard_stdout():
    """stdout carries exactly ONE JSON line: everything libraries print (NCCL's version banner, torch warnings) goes to
    stderr; emit() writes the line to the saved descriptor."""
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
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--op", default="expect", choices=["expect", "apply", "terms"], help="expect: <psi|H|psi> (headline); apply: dst = H psi; terms: <psi|H|psi> plus every <P_k>")
    ap.add_argument("--engine", default="auto", choices=["auto", "tile", "stream"], help="<psi|H|psi> kernel family")
    ap.add_argument("--tile-ctas", type=int, default=2, choices=[2, 3], help="tile engine: resident CTAs per SM")
    ap.add_argument("--tile-bits", type=int, default=12, choices=[11, 12], help="tile engine: 12 = 4096-amplitude tiles, 11 = 2048 double-buffered")
    ap.add_argument("--expect-kernel", default="auto", choices=["auto", "pipe", "ring", "classic"])
    ap.add_argument("--dispatch", type=int, default=1, choices=[0, 1, 2, 3])
    ap.add_argument("--planner", type=int, default=2, choices=[1, 2], help="tile planner version (csrc/tile_plan.hpp)")
    ap.add_argument("--chunk-bits", type=int, default=3, help="tile engine: log2 contiguous amplitudes per HBM chunk")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
