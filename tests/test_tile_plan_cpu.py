"""Host logic of the tile engine (csrc/tile_plan.hpp), checked on the CPU: the plan exported by
b200vqe_debug_tile_plan is replayed in numpy with exactly the index arithmetic of k_tile_expect
(coset representative, chunk combos, virtual masks, sign words, pair selection) and the resulting
<psi|H|psi> must equal the C oracle's."""
import ctypes as C

import numpy as np
import pytest

import c_oracle as co
import xaccvqe_b200 as xb

TILE_BITS = 12
BATCH_DT = np.dtype([("basis", "<u4", 10), ("seg", "<u4", 11), ("g_begin", "<i4"), ("g_end", "<i4"),
                     ("t_begin", "<i4"), ("t_end", "<i4"), ("cls_end", "<i4", 5), ("r_begin", "<i4"), ("r_end", "<i4"),
                     ("lam", "<u4", 4), ("pad", "<u4", 4), ("vcol", "<u2", (6, 12)), ("n_views", "<u4"), ("pad2", "<u4", 3)])
GROUP_DT = np.dtype([("kb", "<u4"), ("n_even", "<u4"), ("n_odd", "<u4"), ("xv", "<u2"), ("hsel", "<i2")])
TERM_DT = np.dtype([("packed", "<u8"), ("u", "<f8")])
REC_DT = np.dtype([("kb", "<u4"), ("count", "<u4"), ("xv", "<u2"), ("key", "u1"), ("pad0", "u1"), ("view", "<u4"),
                   ("packed0", "<u8"), ("u0", "<f8")])
STAGE_RECS, STAGE_TERMS = 128, 256          # capacities of one batch (kStageRecs / kStageTerms in csrc/tile_plan.hpp)
MAX_VIEWS, KEY_VIEW, KEY_FUSED = 6, 0xFE, 0xFD               # kTileMaxViews, kTileRecKeyView, kTileRecKeyFused
SLOT_DT = np.dtype([("zl", "<u4"), ("zt", "<u4"), ("u", "<f8"), ("s0", "<f8"), ("s1", "<f8"), ("kb", "<u4"), ("pad", "<u4"), ("xv", "<u2"),
                    ("key", "u1"), ("pad0", "u1"), ("view", "<u4"), ("s2", "<f8"), ("pad2", "<u8")])      # TileFusedSlot


def get_records(n, c, xm, zm, cf, tb=12):
    lib = xb.capi.lib()
    xm = np.ascontiguousarray(xm, np.uint64)
    zm = np.ascontiguousarray(zm, np.uint64)
    cr, ci = np.ascontiguousarray(cf.real), np.ascontiguousarray(cf.imag)
    nr = C.c_int64()
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    args = (C.c_int(n), C.c_int(c), C.c_int(tb), C.c_int64(len(xm)), vp(xm), vp(zm), vp(cr), vp(ci), C.byref(nr))
    assert lib.b200vqe_debug_tile_records(*args, None) == 0
    R = np.zeros(nr.value, REC_DT)
    assert REC_DT.itemsize == 32
    assert lib.b200vqe_debug_tile_records(*args, vp(R)) == 0
    return R


def get_plan(n, c, xm, zm, cf, tb=12):
    lib = xb.capi.lib()
    xm = np.ascontiguousarray(xm, np.uint64)
    zm = np.ascontiguousarray(zm, np.uint64)
    cr, ci = np.ascontiguousarray(cf.real), np.ascontiguousarray(cf.imag)
    nb, ng, nt = C.c_int64(), C.c_int64(), C.c_int64()
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    args = (C.c_int(n), C.c_int(c), C.c_int(tb), C.c_int64(len(xm)), vp(xm), vp(zm), vp(cr), vp(ci), C.byref(nb), C.byref(ng), C.byref(nt))
    assert lib.b200vqe_debug_tile_plan(*args, None, None, None) == 0
    B, G, T = np.zeros(nb.value, BATCH_DT), np.zeros(ng.value, GROUP_DT), np.zeros(nt.value, TERM_DT)
    assert BATCH_DT.itemsize == 320 and GROUP_DT.itemsize == 16 and TERM_DT.itemsize == 16
    assert lib.b200vqe_debug_tile_plan(*args, vp(B), vp(G), vp(T)) == 0
    return B, G, T


def parity(a):
    a = a.astype(np.uint64)
    for s in (32, 16, 8, 4, 2, 1):
        a = a ^ (a >> np.uint64(s))
    return (a & np.uint64(1)).astype(np.int64)


def shear(b, v, rb=4):
    """Lambda(register of v): register r of a thread holds the amplitude whose low index bits are lo ^ Lambda(r) (TileBatch.lam)."""
    r = v >> np.uint64(8)
    out = np.zeros(len(v), np.uint64)
    for k in range(rb):
        out ^= np.where((r >> np.uint64(k)) & np.uint64(1), np.uint64(b["lam"][k]), np.uint64(0)).astype(np.uint64)
    return out


def replay(n, c, B, G, T, psi, tile_begin=0, tile_count=None):
    """What k_tile_expect computes, vectorised over (tile, thread, register), for tiles [tile_begin, tile_begin + tile_count)."""
    M = TILE_BITS - c
    n_tiles = 1 << (n - TILE_BITS)
    tile_count = n_tiles if tile_count is None else tile_count
    v = np.arange(1 << TILE_BITS, dtype=np.uint64)          # virtual index: r<<8 | thread
    mu = v >> np.uint64(c)
    lo = v & np.uint64((1 << c) - 1)
    total = 0.0
    for b in B:
        combo = np.zeros(1 << TILE_BITS, np.uint64)
        for i in range(M):
            combo ^= np.where((mu >> np.uint64(i)) & np.uint64(1), np.uint64(b["basis"][i]), np.uint64(0)).astype(np.uint64)
        for t in range(tile_begin, tile_begin + tile_count):
            rep = 0
            for i in range(11):
                rep |= (t & int(b["seg"][i])) << i
            j = ((np.uint64(rep) ^ combo) << np.uint64(c)) | (lo ^ shear(b, v))
            tile = psi[j.astype(np.int64)]
            cls = [int(g["hsel"]) for g in G[b["g_begin"]:b["g_end"]]]
            want_cls = sum(([c] * (int(e) - (int(b["cls_end"][i - 1]) if i else 0)) for i, (c, e) in
                            enumerate(zip((0, 1, 2, 3, -1), b["cls_end"]))), [])
            assert cls == want_cls and int(b["cls_end"][4]) == len(cls)
            for g in G[b["g_begin"]:b["g_end"]]:
                xv, hs = int(g["xv"]), int(g["hsel"])
                r = v >> np.uint64(8)
                if hs >= 0:
                    assert (xv >> (8 + hs)) & 1
                    act = ((r >> np.uint64(hs)) & np.uint64(1)) == 0
                else:
                    assert (xv >> 8) == 0
                    act = np.ones(len(v), bool)
                va = v[act]
                a, bb = tile[va.astype(np.int64)], tile[(va ^ np.uint64(xv)).astype(np.int64)]
                w = np.conj(bb) * a
                fe, go = np.zeros(len(va)), np.zeros(len(va))
                k = int(g["kb"])
                for which, cnt in ((0, int(g["n_even"])), (1, int(g["n_odd"]))):
                    for _ in range(cnt):
                        packed, u = int(T[k]["packed"]), float(T[k]["u"])
                        p0 = (bin(rep & packed & 0xFFFFFFFF).count("1") & 1) ^ parity(va & np.uint64(0xFF) & np.uint64((packed >> 32) & 0xFF))
                        zr = (packed >> 40) & 0xF
                        sr = parity((va >> np.uint64(8)) & np.uint64(zr))
                        sign = 1.0 - 2.0 * (p0 ^ sr)
                        if which == 0:
                            fe += sign * u
                        else:
                            go += sign * u
                        k += 1
                total += float(np.sum(fe * w.real + go * w.imag))
    return total


def replay_records(n, c, B, R, T, psi, tb=12, tile_begin=0, tile_count=None):
    """What k_tile_expect / k_tile_expect_pipe compute from the TileRec records: per record, the key fixes WHICH (even / odd nY)
    and XR (the register part of the flip mask); XR != 0 visits the 8 registers with bit HS = top bit of XR clear, XR == 0 visits
    all.  A marker record (key 0xfe) switches to another VIEW of the tile: (thread t, register r) then holds the element at byte
    XOR_{i in t} vcol[view][i] ^ XOR_{j in r} vcol[view][8 + j] of the tile as laid out in shared memory (view 0)."""
    M = tb - c
    n_tiles = 1 << (n - tb)
    tile_count = n_tiles if tile_count is None else tile_count
    v = np.arange(1 << tb, dtype=np.uint64)
    mu = v >> np.uint64(c)
    lo = v & np.uint64((1 << c) - 1)
    total = 0.0
    for b in B:
        recs = R[b["r_begin"]:b["r_end"]]
        assert len(recs) <= STAGE_RECS + MAX_VIEWS and b["t_end"] - b["t_begin"] <= STAGE_TERMS
        assert 1 <= b["n_views"] <= MAX_VIEWS
        assert [int(x) >> 4 for x in b["vcol"][0][:tb]] == [1 << i for i in range(tb)]  # view 0: the shared-memory layout itself
        # split the stream: markers, fused view blocks (4 slots = 8 entries after a marker with count > 0), plain records
        stream, k = [], 0
        while k < len(recs):
            rec = recs[k]
            if rec["key"] == KEY_VIEW and rec["count"] > 0:
                slots = recs[k + 1:k + 9].view(SLOT_DT)
                assert len(slots) == 4 and np.all(slots["key"] == KEY_FUSED) and np.all(slots["view"] == rec["view"])
                stream.append(("view", rec, slots))
                k += 9
            elif rec["key"] == KEY_VIEW:
                stream.append(("view", rec, None))
                k += 1
            else:
                stream.append(("rec", rec, None))
                k += 1
        real = np.array([r for kind, r, _ in stream if kind == "rec"], dtype=REC_DT)
        order = list(zip(real["view"], real["key"]))
        views_seen = sorted(set(int(r["view"]) for kind, r, _ in stream) | {0})
        assert order == sorted(order) and views_seen == list(range(int(b["n_views"]))) and np.all(real["pad0"] == 0)
        vmap = []                                   # per view: virtual index (t | r << 8) -> index into the tile as laid out
        for k in range(int(b["n_views"])):
            m = np.zeros(1 << tb, np.uint64)
            for i in range(tb):
                m ^= np.where((v >> np.uint64(i)) & np.uint64(1), np.uint64(b["vcol"][k][i]), np.uint64(0)).astype(np.uint64)
            assert np.all(m & np.uint64(15) == 0) and len(set(m.tolist())) == 1 << tb
            # the 8 lanes of a quarter-warp (thread bits 0..2) read 8 different 16-byte columns of one row: no bank conflicts
            assert np.all(((m[1] ^ m[0]) >> np.uint64(4)) & np.uint64(7) != 0) if c >= 3 else True
            vmap.append((m >> np.uint64(4)).astype(np.int64))
        combo = np.zeros(1 << tb, np.uint64)
        for i in range(M):
            combo ^= np.where((mu >> np.uint64(i)) & np.uint64(1), np.uint64(b["basis"][i]), np.uint64(0)).astype(np.uint64)
        for t in range(tile_begin, tile_begin + tile_count):
            rep = 0
            for i in range(11):
                rep |= (t & int(b["seg"][i])) << i
            j = ((np.uint64(rep) ^ combo) << np.uint64(c)) | (lo ^ shear(b, v, tb - 8))
            tile0 = psi[j.astype(np.int64)]
            view = 0
            tile = tile0
            for kind, rec, slots in stream:
                key, xv = int(rec["key"]), int(rec["xv"])
                if kind == "view":
                    assert int(rec["view"]) >= view
                    view = int(rec["view"])
                    tile = tile0[vmap[view]]
                    if slots is None:
                        continue
                    # fused block: slot j pairs register r (bit j clear) with r | 1 << j of the SAME thread; the first NE slots
                    # take Re w, the others Im w; the sign is (thread constant) x (+-1 per free register bit)
                    ne = int(rec["count"]) - 1
                    th, rg = v & np.uint64(0xFF), (v >> np.uint64(8)).astype(np.int64)
                    for jj, sl in enumerate(slots):
                        if sl["u"] == 0.0:
                            continue
                        assert int(sl["xv"]) == 1 << (8 + jj) and T[int(b["t_begin"]) + int(sl["kb"])]["u"] == sl["u"]
                        act = ((rg >> jj) & 1) == 0
                        va = v[act]
                        a, bb = tile[va.astype(np.int64)], tile[(va | np.uint64(1 << (8 + jj))).astype(np.int64)]
                        w = np.conj(bb) * a
                        par = (bin(rep & int(sl["zl"])).count("1") & 1) ^ parity(th[act] & np.uint64(sl["zt"]))
                        sg = 1.0 - 2.0 * par
                        free = [q for q in range(4) if q != jj]
                        for kk, q in enumerate(free):
                            sg = sg * np.where((rg[act] >> q) & 1, float(sl["s%d" % kk]), 1.0)
                        total += float(np.sum(sg * float(sl["u"]) * (w.real if jj < ne else w.imag)))
                    continue
                which, xr, pure = (key >> 4) & 1, ((key & 15) + 1) & 15, key >> 5
                assert int(rec["view"]) == view and (pure or view == 0)
                assert xr == xv >> 8 and rec["count"] >= 1 and pure == int((xv & 0xFF) == 0)
                kb = int(b["t_begin"]) + int(rec["kb"])              # record term indices are relative to the batch
                assert kb + rec["count"] <= b["t_end"]
                assert T[kb]["packed"] == rec["packed0"] and T[kb]["u"] == rec["u0"]
                r = v >> np.uint64(8)
                if xr:
                    hs = xr.bit_length() - 1
                    act = ((r >> np.uint64(hs)) & np.uint64(1)) == 0
                else:
                    act = np.ones(len(v), bool)
                va = v[act]
                a, bb = tile[va.astype(np.int64)], tile[(va ^ np.uint64(xv)).astype(np.int64)]
                w = np.conj(bb) * a
                f = np.zeros(len(va))
                for k in range(kb, kb + int(rec["count"])):
                    packed, u = int(T[k]["packed"]), float(T[k]["u"])
                    p0 = (bin(rep & packed & 0xFFFFFFFF).count("1") & 1) ^ parity(va & np.uint64(0xFF) & np.uint64((packed >> 32) & 0xFF))
                    sr = parity((va >> np.uint64(8)) & np.uint64((packed >> 40) & 0xF))
                    f += (1.0 - 2.0 * (p0 ^ sr)) * u
                total += float(np.sum(f * (w.real if which == 0 else w.imag)))
    return total


def get_apply_plan(n, c, xm, zm, cf, rank=0, world=1):
    """b200vqe_debug_apply_plan: the H psi tables of one shard, as ensure_apply_plan uploads them."""
    lib = xb.capi.lib()
    xm = np.ascontiguousarray(xm, np.uint64)
    zm = np.ascontiguousarray(zm, np.uint64)
    cr, ci = np.ascontiguousarray(cf.real), np.ascontiguousarray(cf.imag)
    nb, nr, nt, nc = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    args = (C.c_int(n), C.c_int(rank), C.c_int(world), C.c_int(c), C.c_int64(len(xm)), vp(xm), vp(zm), vp(cr), vp(ci),
            C.byref(nb), C.byref(nr), C.byref(nt), C.byref(nc))
    assert lib.b200vqe_debug_apply_plan(*args, None, None, None, None) == 0
    B, R, T = np.zeros(nb.value, BATCH_DT), np.zeros(nr.value, REC_DT), np.zeros(nt.value, TERM_DT)
    cls = np.zeros(3 * nc.value, np.int32)
    assert lib.b200vqe_debug_apply_plan(*args, vp(B), vp(R), vp(T), vp(cls)) == 0
    return B, R, T, cls.reshape(-1, 3)


def replay_apply(n, c, B, R, T, classes, psi, rank=0, world=1):
    """What k_tile_apply_pipe8 computes for shard `rank`: for every x_rank class the partners come from the slab of rank ^ x_rank;
    per record, output v of a tile receives f(v) * src[v ^ xv] (times i for odd nY), f(v) = sum_e u_e (-1)^sign_e(v) with the
    sign evaluated at the thread's OWN index.  Returns the local slab of H psi."""
    nl = n - int(np.log2(world))
    nloc = 1 << nl
    M = TILE_BITS - c
    v = np.arange(1 << TILE_BITS, dtype=np.uint64)
    mu = v >> np.uint64(c)
    lo = v & np.uint64((1 << c) - 1)
    out = np.zeros(nloc, complex)
    for x_rank, b0, b1 in classes:
        peer = rank ^ int(x_rank)
        slab = psi[peer * nloc:(peer + 1) * nloc]
        for b in B[b0:b1]:
            recs = R[b["r_begin"]:b["r_end"]]
            assert len(recs) <= STAGE_RECS and b["t_end"] - b["t_begin"] <= STAGE_TERMS
            assert list(recs["key"]) == sorted(recs["key"])
            combo = np.zeros(1 << TILE_BITS, np.uint64)
            for i in range(M):
                combo ^= np.where((mu >> np.uint64(i)) & np.uint64(1), np.uint64(b["basis"][i]), np.uint64(0)).astype(np.uint64)
            for t in range(1 << (nl - TILE_BITS)):
                rep = 0
                for i in range(11):
                    rep |= (t & int(b["seg"][i])) << i
                j = (((np.uint64(rep) ^ combo) << np.uint64(c)) | lo).astype(np.int64)
                tile = slab[j]
                acc = np.zeros(1 << TILE_BITS, complex)
                for rec in recs:
                    key, xv = int(rec["key"]), int(rec["xv"])
                    which, xr = key >> 3, xv >> 8              # H psi records: key = which * 8 + (XR & 7), XR bit 3 is read from xv
                    assert key < 16 and (key & 7) == (xr & 7) and not np.any(b["lam"]) and b["n_views"] == 1      # no shear, no views
                    f = np.zeros(1 << TILE_BITS)
                    kb = int(b["t_begin"]) + int(rec["kb"])
                    for k in range(kb, kb + int(rec["count"])):
                        packed, u = int(T[k]["packed"]), float(T[k]["u"])
                        par = (bin(rep & packed & 0xFFFFFFFF).count("1") & 1) ^ parity(v & np.uint64(0xFF) & np.uint64((packed >> 32) & 0xFF))
                        word = (packed >> 48) & 0xFFFF                    # sign word: bit r = parity(r & zr)
                        sr = (np.uint64(word) >> (v >> np.uint64(8))) & np.uint64(1)
                        f += (1.0 - 2.0 * (par ^ sr.astype(np.int64))) * u
                    acc += (f if which == 0 else 1j * f) * tile[(v ^ np.uint64(xv)).astype(np.int64)]
                out[j] += acc
    return out


def get_expect_class_plan(n, c, xm, zm, cf, rank, world):
    """b200vqe_debug_expect_class_plan: the <psi|H|psi> tables of one shard (one plan per x_rank class over the pair register)."""
    lib = xb.capi.lib()
    xm = np.ascontiguousarray(xm, np.uint64)
    zm = np.ascontiguousarray(zm, np.uint64)
    cr, ci = np.ascontiguousarray(cf.real), np.ascontiguousarray(cf.imag)
    nb, nr, nt, nc = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    args = (C.c_int(n), C.c_int(rank), C.c_int(world), C.c_int(c), C.c_int64(len(xm)), vp(xm), vp(zm), vp(cr), vp(ci),
            C.byref(nb), C.byref(nr), C.byref(nt), C.byref(nc))
    assert lib.b200vqe_debug_expect_class_plan(*args, None, None, None, None) == 0
    B, R, T = np.zeros(nb.value, BATCH_DT), np.zeros(nr.value, REC_DT), np.zeros(nt.value, TERM_DT)
    cls = np.zeros(4 * nc.value, np.int32)
    assert lib.b200vqe_debug_expect_class_plan(*args, vp(B), vp(R), vp(T), vp(cls)) == 0
    return B, R, T, cls.reshape(-1, 4)


def replay_expect_classes(n, c, psi, rank, world):
    """What b200vqe_expectation evaluates on shard `rank`: class 0 on the own slab, class x_rank on HALF of the tiles of the pair
    register {slab of the lower rank, slab of the higher rank} of the pair (rank, rank ^ x_rank); the lower rank takes the lower
    half of the tiles.  Returns this rank's partial."""
    nl = n - int(np.log2(world))
    nloc = 1 << nl
    own = psi[rank * nloc:(rank + 1) * nloc]
    xm, zm, cf = replay_expect_classes.ham
    B, R, T, classes = get_expect_class_plan(n, c, xm, zm, cf, rank, world)
    total, seen = 0.0, []
    for x_rank, n_reg, b0, b1 in classes:
        x_rank, n_reg = int(x_rank), int(n_reg)
        seen.append(x_rank)
        if x_rank == 0:
            assert n_reg == nl
            total += replay_records(nl, c, B[b0:b1], R, T, own)
        else:
            assert n_reg == nl + 1
            peer = rank ^ x_rank
            lo_r, hi_r = min(rank, peer), max(rank, peer)          # the same pair register on both ranks of the pair
            reg = np.concatenate([psi[lo_r * nloc:(lo_r + 1) * nloc], psi[hi_r * nloc:(hi_r + 1) * nloc]])
            tiles = 1 << (n_reg - TILE_BITS)
            per = tiles // 2
            role = 0 if rank < peer else 1
            tb, tc = (per * role, per) if per else (0, tiles if role == 0 else 0)
            total += replay_records(n_reg, c, B[b0:b1], R, T, reg, tile_begin=tb, tile_count=tc)
    assert seen == sorted(seen)
    return total


def random_hamiltonian(n, T, seed, weights=None):
    rng = np.random.default_rng(seed)
    pa = rng.integers(0, 4, size=(T, n)) if weights is None else rng.choice(4, size=(T, n), p=weights)
    pa[np.all(pa == 0, axis=1), 0] = 3
    xm = np.zeros(T, np.uint64)
    zm = np.zeros(T, np.uint64)
    for q in range(n):
        xm |= ((pa[:, q] == 1) | (pa[:, q] == 2)).astype(np.uint64) << np.uint64(q)
        zm |= ((pa[:, q] == 3) | (pa[:, q] == 2)).astype(np.uint64) << np.uint64(q)
    keys = {}
    for k in range(T):
        keys.setdefault((int(xm[k]), int(zm[k])), k)
    idx = sorted(keys.values())
    xm, zm = xm[idx], zm[idx]
    ny = np.array([bin(int(x) & int(z)).count("1") for x, z in zip(xm, zm)])
    return xm, zm, rng.normal(size=len(xm)) * (1j ** (ny % 4))


@pytest.mark.parametrize("n,T,c,seed,weights", [
    (12, 20, 3, 1, None), (13, 60, 3, 2, None), (14, 50, 2, 3, None), (14, 40, 5, 4, None), (15, 30, 8, 5, None),
    (14, 120, 3, 6, (0.7, 0.1, 0.1, 0.1)),      # low-weight strings: many groups share a small subspace
    (13, 80, 4, 7, (0.55, 0.0, 0.0, 0.45)),     # diagonal only
])
@pytest.mark.parametrize("planner", [3, 2, 1])
def test_plan_replay_matches_oracle(n, T, c, seed, weights, planner):
    """planner 3 (default of the pipelined kernel): planner 2 + several register subspaces (views) per batch;
    planner 2: collision-greedy subspaces, register subspace spanned by member masks (shear + pure records);
    planner 1: next-unassigned subspaces, reduced-basis coordinates."""
    xb.capi.lib().b200vqe_debug_set_planner(planner)
    try:
        _plan_replay_matches_oracle(n, T, c, seed, weights, planner)
    finally:
        xb.capi.lib().b200vqe_debug_set_planner(3)


def _plan_replay_matches_oracle(n, T, c, seed, weights, planner):
    xm, zm, cf = random_hamiltonian(n, T, seed, weights)
    rng = np.random.default_rng(seed + 50)
    psi = rng.normal(size=1 << n) + 1j * rng.normal(size=1 << n)
    psi /= np.linalg.norm(psi)
    B, G, Tt = get_plan(n, c, xm, zm, cf)
    assert len(G) == len(set(int(x) for x in xm)) and len(Tt) == len(xm)
    assert B["g_end"][-1] == len(G) and B["g_begin"][0] == 0
    eo, _ = co.expect(n, xm, zm, cf, psi)
    if np.all(B["n_views"] == 1):       # the group tables are in view coordinates; the group replay knows view 0 only
        e = replay(n, c, B, G, Tt, psi)
        assert abs(e - eo.real) < 1e-11 * np.abs(cf).sum()
    else:
        assert planner == 3
    R = get_records(n, c, xm, zm, cf)
    er = replay_records(n, c, B, R, Tt, psi)
    assert abs(er - eo.real) < 1e-11 * np.abs(cf).sum()
    if planner == 1:
        assert not np.any(B["lam"])
    if planner < 3:
        assert np.all(B["n_views"] == 1) and not np.any(R["key"] == KEY_VIEW)
    elif weights is None and T >= 40 and c <= 3:
        # unstructured masks: further views turn most of the remaining records into pure-register records
        assert np.any(B["n_views"] > 1) and 2 * np.sum(R["key"] == KEY_FUSED) > 0.5 * len(xm)       # most groups sit in fused view blocks
    if planner >= 2 and weights is None and T >= 30 and c <= 5:
        # unstructured masks: the register subspace is spanned by member masks, so some records need no partner read
        assert np.sum(R["key"] >= 32) >= len(B)
    # 2048-amplitude tiles (3 register bits) of the double-buffered kernel
    if c <= 8:
        B2, G2, T2 = get_plan(n, c, xm, zm, cf, tb=11)
        e2 = replay_records(n, c, B2, get_records(n, c, xm, zm, cf, tb=11), T2, psi, tb=11)
        assert abs(e2 - eo.real) < 1e-11 * np.abs(cf).sum()


@pytest.mark.parametrize("n,T,c,seed,world,weights", [(13, 40, 3, 21, 1, None), (14, 60, 2, 22, 2, None), (14, 50, 3, 23, 4, None),
                                                     (13, 300, 3, 24, 2, (0.6, 0.05, 0.05, 0.3))])
def test_apply_plan_replay_matches_oracle(n, T, c, seed, world, weights):
    """Host logic of the sharded tile H psi on the CPU: every rank's plan (x_rank classes, rank-bit signs folded into the
    coefficients, records) replayed in numpy over the peers' slabs gives that rank's slab of the oracle's H psi."""
    xm, zm, cf = random_hamiltonian(n, T, seed, weights)
    xm = np.concatenate([xm, np.zeros(1, np.uint64)])        # + the identity term (kept in the H psi tables)
    zm = np.concatenate([zm, np.zeros(1, np.uint64)])
    cf = np.concatenate([cf, np.array([0.37 + 0j])])
    rng = np.random.default_rng(seed + 50)
    psi = rng.normal(size=1 << n) + 1j * rng.normal(size=1 << n)
    psi /= np.linalg.norm(psi)
    want = co.apply(n, xm, zm, cf, psi)
    nloc = (1 << n) // world
    for rank in range(world):
        B, R, Tt, classes = get_apply_plan(n, c, xm, zm, cf, rank, world)
        assert len(Tt) == len(xm) and sorted(set(classes[:, 0])) == list(classes[:, 0])
        got = replay_apply(n, c, B, R, Tt, classes, psi, rank, world)
        assert np.abs(got - want[rank * nloc:(rank + 1) * nloc]).max() < 1e-12 * np.abs(cf).sum()


@pytest.mark.parametrize("n,T,c,seed,world", [(14, 60, 3, 31, 2), (15, 80, 3, 32, 4), (16, 120, 3, 33, 8), (15, 50, 2, 34, 4)])
def test_sharded_expectation_class_plans_replay_matches_oracle(n, T, c, seed, world):
    """Host logic of the capacity-scaling sharded <psi|H|psi> on the CPU: every rank's class plans replayed in numpy over {own slab,
    one peer slab} -- never the whole register -- and summed over the ranks (the device all-reduces them) give the oracle's energy."""
    xm, zm, cf = random_hamiltonian(n, T, seed)
    rng = np.random.default_rng(seed + 50)
    psi = rng.normal(size=1 << n) + 1j * rng.normal(size=1 << n)
    psi /= np.linalg.norm(psi)
    eo, _ = co.expect(n, xm, zm, cf, psi)
    replay_expect_classes.ham = (xm, zm, cf)
    parts = [replay_expect_classes(n, c, psi, r, world) for r in range(world)]
    assert abs(sum(parts) - eo.real) < 1e-11 * np.abs(cf).sum()
    assert sum(abs(p) > 0 for p in parts) == world


def test_batch_capacity_splits_large_groups_and_batches():
    """A diagonal group with more terms than one batch stages (kStageTerms) is cut into several groups of the same x, and a
    subspace holding more groups than kStageRecs / 2 is spread over several batches; the records still give <psi|H|psi>."""
    n, c = 13, 3
    rng = np.random.default_rng(11)
    zs = rng.choice(np.arange(1, 1 << n), size=600, replace=False).astype(np.uint64)          # 600 Z-only strings: one x = 0 group
    gens = [int(g) for g in rng.integers(1, 1 << 6, 200, dtype=np.uint64)]                     # flips within the low 6 bits
    xs = np.array(sorted(set(gens)), np.uint64)
    xm = np.concatenate([np.zeros(len(zs), np.uint64), xs])
    zm = np.concatenate([zs, np.zeros(len(xs), np.uint64)])
    cf = rng.normal(size=len(xm)).astype(complex)
    psi = rng.normal(size=1 << n) + 1j * rng.normal(size=1 << n)
    psi /= np.linalg.norm(psi)
    xb.capi.lib().b200vqe_debug_set_planner(2)       # the group replay below knows one register subspace per batch
    try:
        B, G, Tt = get_plan(n, c, xm, zm, cf)
        R = get_records(n, c, xm, zm, cf)
    finally:
        xb.capi.lib().b200vqe_debug_set_planner(3)
    assert len(Tt) == len(xm) and len(G) > len(xs) + 1 and len(B) >= 3
    assert max(int(b["t_end"] - b["t_begin"]) for b in B) <= STAGE_TERMS
    eo, _ = co.expect(n, xm, zm, cf, psi)
    assert abs(replay(n, c, B, G, Tt, psi) - eo.real) < 1e-11 * np.abs(cf).sum()
    assert abs(replay_records(n, c, B, R, Tt, psi) - eo.real) < 1e-11 * np.abs(cf).sum()


def test_plan_batches_many_groups_for_structured_masks():
    """Flip masks spanning a small space (as JW molecular Hamiltonians do) collapse into few batches."""
    n = 16
    rng = np.random.default_rng(0)
    gens = [int(g) for g in rng.integers(1, 1 << n, 7, dtype=np.uint64)]
    xs = set()
    for m in range(1, 1 << 7):
        x = 0
        for i in range(7):
            if (m >> i) & 1:
                x ^= gens[i]
        xs.add(x)
    xm = np.array(sorted(xs), np.uint64)
    zm = np.zeros(len(xm), np.uint64)
    B, G, _ = get_plan(n, 3, xm, zm, np.ones(len(xm), complex))
    assert len(G) == len(xm) and len(B) <= 2


ROT_DT = np.dtype([("zl", "<u4"), ("xv", "<u2"), ("zv", "<u2"), ("hb", "<u4"), ("flip", "<u4"), ("c", "<f8"), ("mre", "<f8"), ("mim", "<f8"),
                   ("pad", "<u8")])          # RotRec (csrc/kernels_rot_tile.cuh)


def get_rotation_runs(n, xm, zm, phi):
    lib = xb.capi.lib()
    xm, zm = np.ascontiguousarray(xm, np.uint64), np.ascontiguousarray(zm, np.uint64)
    phi = np.ascontiguousarray(phi, np.float64)
    nb, nr = C.c_int64(), C.c_int64()
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    args = (C.c_int(n), C.c_int64(len(xm)), vp(xm), vp(zm), vp(phi), C.byref(nb), C.byref(nr))
    assert lib.b200vqe_debug_rotation_runs(*args, None, None) == 0
    B, R = np.zeros(nb.value, BATCH_DT), np.zeros(nr.value, ROT_DT)
    assert ROT_DT.itemsize == 48
    assert lib.b200vqe_debug_rotation_runs(*args, vp(B), vp(R)) == 0
    return B, R


def replay_rotation_runs(n, B, R, psi, c=3):
    """What k_rot_tile does: per run, every tile (coset of the run's subspace) is loaded, the run's rotations are applied to it in
    order -- pairs {v, v ^ xv} of the virtual index with bit hb of v clear; a' = c a + m s(w) b, b' = c b + m s(v) a; diagonal
    strings multiply by c + m s(v) -- and the tile is written back."""
    M = TILE_BITS - c
    v = np.arange(1 << TILE_BITS, dtype=np.uint64)
    mu, lo = v >> np.uint64(c), v & np.uint64((1 << c) - 1)
    psi = psi.copy()
    for b in B:
        combo = np.zeros(1 << TILE_BITS, np.uint64)
        for i in range(M):
            combo ^= np.where((mu >> np.uint64(i)) & np.uint64(1), np.uint64(b["basis"][i]), np.uint64(0)).astype(np.uint64)
        for t in range(1 << (n - TILE_BITS)):
            rep = 0
            for i in range(11):
                rep |= (t & int(b["seg"][i])) << i
            j = (((np.uint64(rep) ^ combo) << np.uint64(c)) | lo).astype(np.int64)
            assert len(set(j.tolist())) == 1 << TILE_BITS
            tile = psi[j]
            for rec in R[b["r_begin"]:b["r_end"]]:
                xv, zv = int(rec["xv"]), np.uint64(rec["zv"])
                cpar = bin(rep & int(rec["zl"])).count("1") & 1
                s = 1.0 - 2.0 * (cpar ^ parity(v & zv))                 # s(v) for every virtual index
                m = complex(rec["mre"], rec["mim"])
                if xv == 0:
                    tile = (rec["c"] + m * s) * tile
                    continue
                hb = int(rec["hb"])
                assert xv >> hb == 1
                va = v[((v >> np.uint64(hb)) & np.uint64(1)) == 0].astype(np.int64)
                wa = va ^ xv
                a, bb = tile[va].copy(), tile[wa].copy()
                assert np.all(s[wa] == s[va] * (1.0 - 2.0 * int(rec["flip"])))
                tile[va] = rec["c"] * a + m * s[wa] * bb
                tile[wa] = rec["c"] * bb + m * s[va] * a
            psi[j] = tile
    return psi


@pytest.mark.parametrize("n,kind,seed", [(13, "random", 1), (15, "low-weight", 2), (14, "same-support", 3), (16, "uccsd", 4)])
def test_rotation_runs_replay_matches_oracle(n, kind, seed):
    """Host logic of the fused ansatz rotations (b200vqe_apply_pauli_rotations -> k_rot_tile) on the CPU: the runs the host cuts
    a rotation list into -- consecutive rotations whose flip masks span <= 9 high-index dimensions -- replayed in numpy with the
    kernel's index arithmetic give the oracle's state after the same rotations applied one by one, in order."""
    rng = np.random.default_rng(seed)
    if kind == "uccsd":
        x, z, ic, pa, npar = xb.uccsd_rotations(n, 4, "jw")
        keep = np.sort(rng.choice(len(x), 120, replace=False))
        xm, zm, phi = x[keep], z[keep], (ic * rng.uniform(-0.4, 0.4, npar)[pa])[keep]
    elif kind == "same-support":
        sup = [int(s) for s in rng.integers(1, 1 << n, 6, dtype=np.uint64)]
        xm = np.array([sup[i // 8 % 6] for i in range(96)], np.uint64)
        zm = rng.integers(0, 1 << n, 96, dtype=np.uint64)
        phi = rng.normal(size=96)
    else:
        xm, zm, _ = random_hamiltonian(n, 90, 40 + seed, None if kind == "random" else (0.8, 0.07, 0.06, 0.07))
        xm = xm.copy()
        xm[::11] = 0
        phi = rng.normal(size=len(xm))
        phi[3] = 0.0
    B, R = get_rotation_runs(n, xm, zm, phi)
    assert len(R) == int(np.count_nonzero(phi)) and B["r_end"][-1] == len(R) and B["r_begin"][0] == 0
    if kind != "random":
        assert len(B) * 4 <= len(R)                      # structured lists: many rotations per HBM pass
    psi = rng.normal(size=1 << n) + 1j * rng.normal(size=1 << n)
    psi /= np.linalg.norm(psi)
    want = psi.copy()
    for x, z, p in zip(xm, zm, phi):
        if p != 0.0:
            co.pauli_rotation(n, int(x), int(z), float(p), want)
    got = replay_rotation_runs(n, B, R, psi)
    assert np.abs(got - want).max() < 1e-12


@pytest.mark.parametrize("n,ok", [(11, False), (12, True), (32, True), (33, False), (35, False)])
def test_rotation_run_planner_refuses_registers_beyond_32_bit_indices(n, ok):
    """k_rot_tile forms 32-bit element indices, so fused runs exist only for 12 <= local qubits <= 32 (a 33-qubit register fits a
    B200 and must fall back to one k_pauli_rotation launch per rotation).  b200vqe_apply_pauli_rotations and this host-only
    planner export share the predicate (rot_tile_supports, csrc/backend_state.inc)."""
    lib = xb.capi.lib()
    top = np.uint64(1) << np.uint64(n - 1)
    xm, zm = np.array([top | np.uint64(1), top], np.uint64), np.array([np.uint64(2), top], np.uint64)
    phi = np.array([0.3, -0.2])
    nb, nr = C.c_int64(), C.c_int64()
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    rc = lib.b200vqe_debug_rotation_runs(C.c_int(n), C.c_int64(2), vp(xm), vp(zm), vp(phi), C.byref(nb), C.byref(nr), None, None)
    assert (rc == 0) == ok, rc
    if ok:
        assert nb.value >= 1 and nr.value == 2
