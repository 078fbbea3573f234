"""CPU-only checks of the product's host side: the C-ABI library loads and exports every symbol
include/b200vqe.h declares, the front end (FermionCompiler text, JW/BK, packing, UCCSD) is bit-exact
against the oracle, and device entry points fail loudly without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

import pauli_oracle as po
import xaccvqe_b200 as xb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    import glob
    headers = sorted(glob.glob(os.path.join(ROOT, "include", "*.h")))
    assert [os.path.basename(h) for h in headers] == ["b200vqe.h", "b200vqe_debug.h"]
    names = sorted({n for h in headers for n in re.findall(r"\b(b200vqe_[a-z0-9_]+)\s*\(", open(h).read())})
    assert len(names) >= 35
    # the drop-in ABI header carries no planner/debug exports
    assert "b200vqe_debug_" not in re.sub(r"/\*.*?\*/", "", open(headers[0]).read(), flags=re.S)
    lib = xb.capi.lib()
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_has_gpu(), reason="only meaningful without a GPU")
def test_no_cpu_fallback():
    with pytest.raises(xb.B200Error) as ei:
        xb.Backend(4)
    assert "no CPU fallback" in str(ei.value) or "CUDA" in str(ei.value)


@pytest.mark.parametrize("tr", ["jw", "bk"])
def test_front_end_packing_bit_exact(golden, tr):
    for key in ("h2_source", "rdm_h2_source", "jw_easy_source"):
        op = xb.OpHandle.from_fermion_source(golden[key], tr)
        xm, zm, cf = op.pack()
        oop, n = po.compile_fermion_source(golden[key], tr)
        oxm, ozm, ocf, oids = po.pack_terms(oop)
        assert op.n_qubits() == n
        assert np.array_equal(xm, oxm) and np.array_equal(zm, ozm)
        assert np.array_equal(cf, ocf)            # same arithmetic order -> identical doubles
        assert op.term_ids() == oids


def test_text_round_trip(golden):
    for key in ("six_qubit_118", "four_qubit_100", "tapered_h2", "deuteron_h3"):
        op = xb.OpHandle.from_string(golden[key])
        oop = po.op_from_string(golden[key])
        xm, zm, cf = op.pack()
        oxm, ozm, ocf, _ = po.pack_terms(oop)
        assert np.array_equal(xm, oxm) and np.array_equal(zm, ozm) and np.array_equal(cf, ocf)
        again = xb.OpHandle.from_string(op.to_string())
        assert np.array_equal(again.pack()[2], cf)
        assert po.pack_terms(po.op_from_string(op.to_string()))[2].tolist() == ocf.tolist()


def test_operator_algebra_matches_oracle(golden):
    a = xb.OpHandle.from_string(golden["deuteron_h3"])
    b = xb.OpHandle.from_string(golden["four_qubit_100"])
    oa, ob = po.op_from_string(golden["deuteron_h3"]), po.op_from_string(golden["four_qubit_100"])
    for got, want in ((a.mul(b), po.op_mul(oa, ob)), (a.add(b), po.op_add(oa, ob)), (b.mul(b), po.op_mul(ob, ob)),
                      (a.scale(0.5 - 2j), po.op_scale(oa, 0.5 - 2j))):
        xm, zm, cf = got.pack()
        oxm, ozm, ocf, _ = po.pack_terms(want)
        assert np.array_equal(xm, oxm) and np.array_equal(zm, ozm)
        assert np.abs(cf - ocf).max() < 1e-12 * max(1.0, np.abs(ocf).max())


@pytest.mark.parametrize("nq,ne,tr", [(4, 2, "jw"), (4, 2, "bk"), (6, 2, "jw"), (8, 4, "bk"), (12, 4, "jw"), (12, 4, "bk")])
def test_uccsd_rotations_match_oracle(nq, ne, tr):
    x, z, ic, pa, npar = xb.uccsd_rotations(nq, ne, tr)
    rots, onp = po.uccsd_rotations(nq, ne, tr)
    assert npar == onp and len(rots) == len(x)
    assert [int(v) for v in x] == [r[0] for r in rots]
    assert [int(v) for v in z] == [r[1] for r in rots]
    assert ic.tolist() == [r[2] for r in rots]
    assert pa.tolist() == [r[3] for r in rots]


def test_uccsd_gate_list_matches_oracle():
    th = np.array([0.1, -0.2])
    g = xb.uccsd_gates(4, 2, th)
    want = po.uccsd_gates(4, 2, th)
    inv = {v: k for k, v in xb.capi.GATE_ID.items()}
    got = [(inv[x.gate], (x.q0, x.q1) if x.gate == 7 else (x.q0,), x.theta) for x in g]
    assert len(got) == len(want)
    for a, b in zip(got, want):
        assert a[0] == b[0] and tuple(a[1]) == tuple(b[1]) and a[2] == b[2]


def test_parse_errors():
    with pytest.raises(xb.B200Error):
        xb.OpHandle.from_string("(1,0) Q0")
    with pytest.raises(xb.B200Error):
        xb.OpHandle.from_fermion_source("__qpu__ k() {\nabc 0 1\n}", "jw")
    with pytest.raises(xb.B200Error):
        xb.OpHandle.from_fermion_source("__qpu__ k() {\n1.0 0 1 0 0\n}", "nope")
    with pytest.raises(xb.B200Error):
        xb.uccsd_rotations(2, 2, "jw")


@pytest.mark.parametrize("n", [2, 4, 6, 8, 12, 14, 16])
def test_number_sector_rows_bit_exact(n):
    """b200vqe_number_sector_rows against the oracle's restatement of the GF(2) matrix of DiagonalizeTask.cpp:76-121:
    the set of basis indices the rows accept equals the set of re-encoded strings the reference enumerates."""
    assert np.array_equal(xb.capi.number_sector_rows(n, "jw"), np.uint64(1) << np.arange(n, dtype=np.uint64))
    rows = [int(r) for r in xb.capi.number_sector_rows(n, "bk")]
    ne = n // 2 if n <= 12 else 2
    want = set()
    for s in po.number_sector_strings(n, ne, "bk"):
        want.add(sum(1 << q for q, ch in enumerate(s) if ch == "1"))     # string position q <-> qubit q <-> index bit q
    got = {j for j in range(1 << n) if sum(bin(r & j).count("1") & 1 for r in rows) == ne} if n <= 12 else None
    if got is not None:
        assert got == want
    else:
        for j in want:
            assert sum(bin(r & j).count("1") & 1 for r in rows) == ne


def test_header_is_plain_c_and_example_links(tmp_path):
    """include/b200vqe.h compiles as strict C99 (no C++ or torch types at the boundary), and the plain-C example links
    against libb200vqe.so; its host part (FermionCompiler text -> JW -> packing) runs anywhere, the device part needs a GPU."""
    import subprocess
    lib_dir = os.path.join(ROOT, "xacc-vqe_b200")
    exe = str(tmp_path / "h2")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "examples", "h2_ground_state.c"), "-L", lib_dir, "-lb200vqe",
                           "-Wl,-rpath," + lib_dir, "-o", exe])
    probe = tmp_path / "dbg.c"          # the debug header is plain C as well
    probe.write_text('#include "b200vqe_debug.h"\nint main(void) { return b200vqe_debug_set_planner(3); }\n')
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"), str(probe),
                           "-L", lib_dir, "-lb200vqe", "-Wl,-rpath," + lib_dir, "-o", str(tmp_path / "dbg")])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert "H2: 4 qubits, 15 Pauli terms" in r.stdout
    if _has_gpu():
        assert r.returncode == 0 and "E0 = -1.13727042206" in r.stdout
    else:
        assert r.returncode == 1 and "no CPU fallback" in r.stdout


def _random_fermion_source(n_sites, n_terms, seed):
    """Seeded FermionCompiler text (compiler/FermionCompiler.cpp:62-103): a constant, then terms of 2 or 4 ladder operators
    '<coeff> <site> <1|0> ...' with repeated sites and both orders; coefficients are real, as the reference's std::stod
    parse requires (:85-87)."""
    rng = np.random.default_rng(seed)
    lines = ["__qpu__ k() {", repr(float(rng.normal()))]
    for _ in range(n_terms):
        k = 2 if rng.random() < 0.5 else 4
        sites = rng.integers(0, n_sites, k)
        dags = rng.integers(0, 2, k)
        lines.append(repr(float(rng.normal())) + " " + " ".join("%d %d" % (s, d) for s, d in zip(sites, dags)))
    lines.append("}")
    return "\n".join(lines)


@pytest.mark.parametrize("n_sites,n_terms,seed", [(3, 12, 1), (5, 30, 2), (6, 40, 3), (7, 25, 4), (10, 60, 5), (13, 40, 6)])
@pytest.mark.parametrize("tr", ["jw", "bk"])
def test_front_end_random_sources_bit_exact(n_sites, n_terms, seed, tr):
    """C++ front end against the oracle on seeded random sources, including register sizes that are not powers of two
    (the Fenwick-tree BK of bk/Fenwick.hpp:83-176): identical masks, identical IEEE doubles, identical term ids."""
    src = _random_fermion_source(n_sites, n_terms, seed)
    oop, n = po.compile_fermion_source(src, tr, n_sites)
    oxm, ozm, ocf, oids = po.pack_terms(oop)
    op = xb.OpHandle.from_fermion_source(src, tr, n_sites)
    xm, zm, cf = op.pack()
    assert op.n_terms() == len(oxm)
    assert np.array_equal(xm, oxm) and np.array_equal(zm, ozm)
    assert np.array_equal(cf.view(np.float64), ocf.view(np.float64))
    assert op.term_ids() == oids
    # text round trip through the PauliOperator string form keeps every bit of every coefficient
    again = xb.OpHandle.from_string(op.to_string())
    xm2, zm2, cf2 = again.pack()
    assert np.array_equal(xm2, xm) and np.array_equal(zm2, zm) and np.array_equal(cf2.view(np.float64), cf.view(np.float64))
