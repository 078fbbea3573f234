"""N > 1 host logic on the CPU, world_size 2 over gloo: the shard layout the library uses (slabs, tile shares), the
unique-id broadcast bench.py performs, and the sharded <psi|H|psi> decomposition itself -- each rank replays only ITS
share of the tile-engine plan (numpy replay of k_tile_expect) and the gloo all_reduce of the partial sums must equal the
oracle, exactly as ncclAllReduce combines the per-GPU partials on the device path; and the sharded H psi: each rank replays
its own apply plan (partners from the peer's slab for x_rank != 0) and the all-gathered slabs must equal the oracle's H psi."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _layout(lib, n, rank, world, tile_bits=12):
    nl, first, tb, tc = C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_uint64()
    rc = lib.b200vqe_debug_shard_layout(C.c_int(n), C.c_int(rank), C.c_int(world), C.c_int(tile_bits), C.byref(nl), C.byref(first), C.byref(tb), C.byref(tc))
    assert rc == 0
    return nl.value, first.value, tb.value, tc.value


def _worker(rank, world, port, n, T, seed, q):
    for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch
    import torch.distributed as dist

    import test_tile_plan_cpu as tp
    import xaccvqe_b200 as xb
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lib = xb.capi.lib()
    # the id exchange of bench.py / run_sharded_lanczos.py (rank 0 creates, everyone receives the same 128 bytes)
    ids = [bytes(range(128)) if rank == 0 else None]
    dist.broadcast_object_list(ids, src=0)
    assert ids[0] == bytes(range(128))
    nloc, first, tb, tc = _layout(lib, n, rank, world)
    lay = torch.tensor([nloc, first, tb, tc], dtype=torch.int64)
    lays = [torch.zeros(4, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(lays, lay)
    # every rank holds the same Hamiltonian and (here) the gathered register; it evaluates only its tile share
    xm, zm, cf = tp.random_hamiltonian(n, T, seed)
    rng = np.random.default_rng(seed + 50)
    psi = rng.normal(size=1 << n) + 1j * rng.normal(size=1 << n)
    psi /= np.linalg.norm(psi)
    B, G, Tt = tp.get_plan(n, 3, xm, zm, cf)
    part = tp.replay_records(n, 3, B, tp.get_records(n, 3, xm, zm, cf), Tt, psi, tile_begin=tb, tile_count=tc)
    tot = torch.tensor([part], dtype=torch.float64)
    dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    # the capacity-scaling decomposition b200vqe_expectation uses on a sharded handle: per x_rank class the pair register
    # {own slab, slab of rank ^ x_rank}, half of its tiles per rank; partials all-reduced
    tp.replay_expect_classes.ham = (xm, zm, cf)
    part_c = tp.replay_expect_classes(n, 3, psi, rank, world)
    tot_c = torch.tensor([part_c], dtype=torch.float64)
    dist.all_reduce(tot_c, op=dist.ReduceOp.SUM)
    assert abs(tot_c.item() - tot.item()) < 1e-11 * float(np.abs(cf).sum())
    # the same with the 2048-amplitude tiles / per-key records of the double-buffered kernel
    _, _, tb11, tc11 = _layout(lib, n, rank, world, tile_bits=11)
    B11, _, T11 = tp.get_plan(n, 3, xm, zm, cf, tb=11)
    part11 = tp.replay_records(n, 3, B11, tp.get_records(n, 3, xm, zm, cf, tb=11), T11, psi, tb=11, tile_begin=tb11, tile_count=tc11)
    tot11 = torch.tensor([part11], dtype=torch.float64)
    dist.all_reduce(tot11, op=dist.ReduceOp.SUM)
    assert abs(tot11.item() - tot.item()) < 1e-11 * float(np.abs(cf).sum())
    # norm^2 of the sharded state: local slab sums, all-reduced (b200vqe_state_norm2 on a sharded handle)
    nrm = torch.tensor([float(np.vdot(psi[first:first + nloc], psi[first:first + nloc]).real)], dtype=torch.float64)
    dist.all_reduce(nrm, op=dist.ReduceOp.SUM)
    # H psi: this rank's plan (x_rank classes over the peers' slabs) gives its slab; all_gather assembles the vector, and
    # the Lanczos scalar alpha = <psi|H psi> is the all-reduced sum of the per-slab inner products (ncclAllReduce on the device)
    import c_oracle as co
    Ba, Ra, Ta, classes = tp.get_apply_plan(n, 3, xm, zm, cf, rank, world)
    mine = tp.replay_apply(n, 3, Ba, Ra, Ta, classes, psi, rank, world)
    slabs = [torch.zeros(nloc, dtype=torch.complex128) for _ in range(world)]
    dist.all_gather(slabs, torch.from_numpy(mine))
    hpsi = torch.cat(slabs).numpy()
    want = co.apply(n, xm, zm, cf, psi)
    assert np.abs(hpsi - want).max() < 1e-12 * float(np.abs(cf).sum())
    alpha = torch.tensor([float(np.vdot(psi[first:first + nloc], mine).real)], dtype=torch.float64)
    dist.all_reduce(alpha, op=dist.ReduceOp.SUM)
    assert abs(alpha.item() - float(np.vdot(psi, want).real)) < 1e-11 * float(np.abs(cf).sum())
    assert any(int(c[0]) != 0 for c in classes)               # a class that reads the peer's slab exists
    q.put((rank, [t.tolist() for t in lays], float(tot.item()), part, float(nrm.item())))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n,T,seed", [(13, 30, 3), (14, 60, 4)])
def test_two_rank_expectation_decomposition(n, T, seed):
    import torch.multiprocessing as mp

    import c_oracle as co
    import test_tile_plan_cpu as tp
    world = 2
    port = 29600 + seed
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, T, seed, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=240) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    xm, zm, cf = tp.random_hamiltonian(n, T, seed)
    rng = np.random.default_rng(seed + 50)
    psi = rng.normal(size=1 << n) + 1j * rng.normal(size=1 << n)
    psi /= np.linalg.norm(psi)
    eo, _ = co.expect(n, xm, zm, cf, psi)
    lays = res[0][1]
    assert lays == res[1][1]
    # slabs tile the register, tile shares tile the cosets
    assert [l[1] for l in lays] == [0, 1 << (n - 1)] and all(l[0] == 1 << (n - 1) for l in lays)
    assert lays[0][2] == 0 and lays[1][2] == lays[0][3] and lays[0][3] + lays[1][3] == 1 << (n - 12)
    for rank, _, tot, part, nrm in res:
        assert abs(tot - eo.real) < 1e-11 * np.abs(cf).sum()
        assert abs(nrm - 1.0) < 1e-12
    assert abs(res[0][3] + res[1][3] - res[0][2]) < 1e-13 * np.abs(cf).sum()
    assert abs(res[0][3]) > 0 and abs(res[1][3]) > 0            # both ranks really contribute


def test_shard_layout_covers_every_tile_once():
    import xaccvqe_b200 as xb
    lib = xb.capi.lib()
    for n, tile_bits in [(n, t) for n in (12, 13, 14, 20, 31, 32) for t in (12, 11)]:
        for world in (1, 2, 4, 8):
            if n - int(np.log2(world)) < 1:
                continue
            spans, slabs = [], []
            for r in range(world):
                nloc, first, tb, tc = _layout(lib, n, r, world, tile_bits)
                slabs.append((first, nloc))
                if tc:
                    spans.append((tb, tc))
            assert sum(c for _, c in slabs) == 1 << n and [f for f, _ in slabs] == [r << (n - int(np.log2(world))) for r in range(world)]
            spans.sort()
            assert spans[0][0] == 0 and sum(c for _, c in spans) == 1 << (n - tile_bits)
            for (b0, c0), (b1, _) in zip(spans, spans[1:]):
                assert b0 + c0 == b1
    assert lib.b200vqe_debug_shard_layout(C.c_int(20), C.c_int(2), C.c_int(2), C.c_int(12), None, None, None, None) != 0
    assert lib.b200vqe_debug_shard_layout(C.c_int(20), C.c_int(0), C.c_int(3), C.c_int(12), None, None, None, None) != 0
