"""Model check (CPU) of the slab-push protocol of the sharded <psi|H|psi> (csrc/backend_ham.inc, b200vqe_expectation; csrc/backend.cu,
ensure_peer_staging / flag_wait / flag_push): every rank owns two staging buffers and flag words arrived[b], free_from[p][b]; per
x_rank class i the pair (r, r ^ c_i) push their slabs into each other's buffer i & 1 on their communication streams and consume
them on their compute streams.  The streams of all ranks are replayed as in-order queues under RANDOM interleavings; the model
asserts that no buffer is overwritten before its previous content was read, that every class reads the slab of the right peer and
call, and that no interleaving deadlocks -- over several back-to-back calls WITHOUT any barrier between them (the device code has
one, the all-reduce; the flags alone must suffice)."""
import random

import pytest


def build_streams(world, classes, n_calls):
    """The operations each rank enqueues, in program order, exactly as b200vqe_expectation does: request(0); local class;
    for i: request(i + 1); wait arrived; class kernel; tell every peer the buffer is free."""
    comm = {r: [] for r in range(world)}
    main = {r: [] for r in range(world)}
    for r in range(world):
        seq, last = 0, [0, 0]
        for call in range(n_calls):
            seq_of = {}

            def request(i):
                nonlocal seq
                b, peer = i & 1, r ^ classes[i]
                seq += 1
                seq_of[i] = seq
                comm[r].append(("wait", ("free_from", r, peer, b), last[b]))       # the peer is done with ITS buffer b
                comm[r].append(("copy", peer, b, (r, call, classes[i])))
                comm[r].append(("flag", ("arrived", peer, b), seq))
                last[b] = seq

            if classes:
                request(0)
            main[r].append(("kernel_local", call))
            for i in range(len(classes)):
                if i + 1 < len(classes):
                    request(i + 1)
                b = i & 1
                main[r].append(("wait", ("arrived", r, b), seq_of[i]))
                main[r].append(("kernel_read", b, (r ^ classes[i], call, classes[i])))
                for q in range(world):
                    if q != r:
                        main[r].append(("flag", ("free_from", q, r, b), seq_of[i]))
    return comm, main


def run(world, classes, n_calls, rng):
    comm, main = build_streams(world, classes, n_calls)
    flags = {}
    buf = {(r, b): None for r in range(world) for b in range(2)}         # content of staging buffer b of rank r
    unread = {(r, b): False for r in range(world) for b in range(2)}
    heads = {("c", r): 0 for r in range(world)}
    heads.update({("m", r): 0 for r in range(world)})
    queues = {("c", r): comm[r] for r in range(world)}
    queues.update({("m", r): main[r] for r in range(world)})
    remaining = sum(len(q) for q in queues.values())
    while remaining:
        ready = []
        for key, q in queues.items():
            if heads[key] == len(q):
                continue
            op = q[heads[key]]
            if op[0] == "wait" and flags.get(op[1], 0) < op[2]:
                continue
            ready.append(key)
        assert ready, "deadlock"
        key = rng.choice(ready)
        op = queues[key][heads[key]]
        heads[key] += 1
        remaining -= 1
        if op[0] == "copy":
            _, peer, b, payload = op
            assert not unread[(peer, b)], f"buffer {b} of rank {peer} overwritten before it was read"
            buf[(peer, b)] = payload
            unread[(peer, b)] = True
        elif op[0] == "flag":
            flags[op[1]] = max(flags.get(op[1], 0), op[2])
        elif op[0] == "kernel_read":
            _, b, want = op
            assert buf[(key[1], b)] == want, (key, buf[(key[1], b)], want)
            unread[(key[1], b)] = False


@pytest.mark.parametrize("world,classes", [(2, [1]), (4, [1, 2, 3]), (4, [2]), (8, [1, 2, 3, 4, 5, 6, 7]), (8, [1, 4, 5]), (8, [])])
def test_slab_push_protocol_is_race_and_deadlock_free(world, classes):
    for seed in range(40):
        run(world, classes, n_calls=3, rng=random.Random(seed))


def test_model_catches_a_missing_free_wait(monkeypatch):
    """The model is not vacuous: without the wait for the peer's 'free' flag some interleaving overwrites an unread buffer."""
    import sys
    me = sys.modules[__name__]
    orig = build_streams

    def stripped(w, c, n):
        cm, mn = orig(w, c, n)
        return {r: [op for op in q if not (op[0] == "wait" and op[1][0] == "free_from")] for r, q in cm.items()}, mn
    monkeypatch.setattr(me, "build_streams", stripped)
    hit = False
    for seed in range(200):
        try:
            run(4, [1, 2, 3], 2, random.Random(seed))
        except AssertionError as ex:
            if "overwritten" in str(ex):
                hit = True
                break
    assert hit
