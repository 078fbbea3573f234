"""Sharded register (state split by the high index bits over several GPUs, one process per GPU, NCCL)
against the CPU oracle.  Needs >= 2 GPUs; skipped otherwise (the round-end single-GPU run skips it,
`gpurun --gpus 2 -- python -m pytest tests/test_gpu_sharded.py -m gpu` runs it)."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def _worker(rank, world, comm_id, n, T, seed, engine, lanczos, out_q):
    for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import xaccvqe_b200 as xb
    from test_gpu_parity import random_hamiltonian, random_state
    xm, zm, cf = random_hamiltonian(n, T, seed)
    psi = random_state(n, seed + 100)
    be = xb.Backend(n, device=rank, rank=rank, world=world, comm_id=comm_id)
    be.set_option("expect-engine", engine)
    be.set_option("apply-engine", engine)
    be.set_option("shard-exchange", ("needed", "pairwise", "allgather")[seed % 3])     # both ways of fetching peer slabs (tile H psi)
    be.n_slots(2)
    be.set_hamiltonian(xm, zm, cf)
    lo, cnt = be.first, be.n_local
    be.upload(psi[lo:lo + cnt])
    e, per = be.expectation(per_term=True)
    # the energy alone takes the capacity-scaling path: x_rank classes over pair registers, slab swaps as peer-memory pushes
    # (twice: the second call reuses the staging buffers and their flag words)
    e_only = [be.expectation(), be.expectation()]
    ref_part = be.download(0, order=xb.capi.ORDER_REFERENCE)       # collective: this rank's range of the reference index order
    alpha = be.apply_h(0, 1)
    hpsi = be.download(1)
    nrm = be.norm2(0)
    e0, vec = 0.0, None
    if lanczos:
        e0, it, res = be.lanczos_ground(tol=1e-10, max_iter=3000, seed=seed, want_vector=True)
        vec = be.download(0)
    out_q.put((rank, e, per, alpha, lo, hpsi, nrm, e0, vec, e_only, ref_part))
    be.close()


@pytest.mark.skipif(_n_gpus() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("n,T,seed,engine,lanczos", [(14, 40, 3, "tile", True), (16, 60, 4, "stream", False),
                                                     (15, 25, 5, "tile", True), (18, 30, 6, "tile", False)])
def test_shards_match_oracle(n, T, seed, engine, lanczos, world):
    import torch.multiprocessing as mp

    import c_oracle as co
    import xaccvqe_b200 as xb
    from test_gpu_parity import random_hamiltonian, random_state
    if _n_gpus() < world:
        pytest.skip("needs %d GPUs" % world)
    comm_id = xb.Backend.comm_unique_id()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, comm_id, n, T, seed, engine, lanczos, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    xm, zm, cf = random_hamiltonian(n, T, seed)
    psi = random_state(n, seed + 100)
    scale = np.abs(cf).sum()
    eo, pero = co.expect(n, xm, zm, cf, psi)
    ref = co.apply(n, xm, zm, cf, psi)
    full = np.zeros(1 << n, complex)
    vec = np.zeros(1 << n, complex)
    # reference order: qubit q <-> bit n-1-q (DiagonalizeTask.cpp:189-193)
    rev = np.array([int(format(i, "0%db" % n)[::-1], 2) for i in range(1 << n)])
    psi_ref = psi[rev]
    for rank, e, per, alpha, lo, hpsi, nrm, e0, v, e_only, ref_part in res:
        assert abs(e - eo) < 1e-10 * scale
        assert max(abs(x - eo) for x in e_only) < 1e-10 * scale and e_only[0] == e_only[1]
        assert np.array_equal(ref_part, psi_ref[lo:lo + len(ref_part)])
        assert np.abs(per - pero.real).max() < 1e-11
        assert abs(alpha - np.vdot(psi, ref)) < 1e-10 * scale
        assert abs(nrm - 1.0) < 1e-12
        full[lo:lo + len(hpsi)] = hpsi
        if lanczos:
            vec[lo:lo + len(v)] = v
    assert np.abs(full - ref).max() < 1e-12 * scale
    if lanczos:
        # both ranks report the same eigenvalue; the gathered Ritz vector is an eigenvector
        assert max(abs(r[7] - res[0][7]) for r in res) < 1e-12 * scale
        hv = co.apply(n, xm, zm, cf, vec)
        assert np.linalg.norm(hv - res[0][7] * vec) < 1e-6 * scale
        import pauli_oracle as po
        want, _ = po.ground_state_eigsh(xm, zm, cf, n, tol=1e-13, apply=lambda v: co.apply(n, xm, zm, cf, v))
        assert abs(res[0][7] - want) < 1e-10 * abs(want)          # truly relative


def _gate_program(n, seed):
    """A seeded gate list that puts controls / targets / rotations on local AND sharded qubits, plus Pauli rotations
    whose flip mask reaches the shard bits (UCCSD.cpp:282-352 uses exactly this gate set)."""
    rng = np.random.default_rng(seed)
    gates = []
    names = ["X", "Y", "Z", "H", "Rx", "Ry", "Rz"]
    for q in range(n):
        gates.append(("H", (q,), 0.0))
    for _ in range(40):
        if rng.random() < 0.4:
            c, t = rng.choice(n, size=2, replace=False)
            gates.append(("CNOT", (int(c), int(t)), 0.0))
        else:
            gates.append((names[int(rng.integers(len(names)))], (int(rng.integers(n)),), float(rng.uniform(-2, 2))))
    # every CNOT placement relative to the top (sharded) qubits
    top = n - 1
    gates += [("CNOT", (top, 0), 0.0), ("CNOT", (1, top), 0.0), ("CNOT", (top, top - 1), 0.0), ("CNOT", (top - 1, top), 0.0),
              ("Rx", (top,), 0.7), ("Ry", (top,), -0.4), ("Rz", (top,), 1.1), ("X", (top,), 0.0), ("Y", (top,), 0.0),
              ("Z", (top,), 0.0), ("H", (top,), 0.0), ("H", (top - 1,), 0.0)]
    rots = []
    for _ in range(12):
        x = int(rng.integers(1, 1 << n)) | (1 << top if rng.random() < 0.7 else 0)
        z = int(rng.integers(0, 1 << n))
        rots.append((x, z, float(rng.uniform(-1, 1))))
    return gates, rots


def _gate_worker(rank, world, comm_id, n, seed, out_q):
    for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import xaccvqe_b200 as xb
    gates, rots = _gate_program(n, seed)
    be = xb.Backend(n, device=rank, rank=rank, world=world, comm_id=comm_id)
    be.reset(0, 5)
    be.apply_gates(gates)
    be.apply_pauli_rotations([r[0] for r in rots], [r[1] for r in rots], [r[2] for r in rots])
    out_q.put((rank, be.first, be.download(0), be.norm2(0)))
    be.close()


@pytest.mark.skipif(_n_gpus() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("world", [2, 4, 8])
def test_gates_on_sharded_qubits_match_oracle(world):
    """Ansatz gates and fused Pauli rotations that touch the shard bits (pair exchange over NCCL send/recv)."""
    import torch.multiprocessing as mp

    import c_oracle as co
    import xaccvqe_b200 as xb
    if _n_gpus() < world:
        pytest.skip("needs %d GPUs" % world)
    n, seed = 12, 21
    comm_id = xb.Backend.comm_unique_id()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gate_worker, args=(r, world, comm_id, n, seed, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    gates, rots = _gate_program(n, seed)
    psi = np.zeros(1 << n, complex)
    psi[5] = 1.0
    for name, qs, th in gates:
        co.apply_gate(n, name, qs, th, psi)
    for x, z, phi in rots:
        co.pauli_rotation(n, x, z, phi, psi)
    got = np.zeros(1 << n, complex)
    for rank, first, amps, nrm in res:
        got[first:first + len(amps)] = amps
        assert abs(nrm - 1.0) < 1e-12
    assert np.abs(got - psi).max() < 1e-12
