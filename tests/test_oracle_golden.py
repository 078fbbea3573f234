"""Pins the CPU oracle (oracle/) against every golden value the reference's own tests hold for the
Hamiltonian x statevector path.  CPU only."""
import numpy as np
import pytest

import c_oracle as co
import pauli_oracle as po


def _op_from_terms(terms):
    op = {}
    for ops, c in terms:
        op = po.op_add(op, po.pauli({int(q): p for q, p in ops.items()}, c))
    return op


def _close(a, b, tol):
    keys = set(a) | set(b)
    return all(abs(a.get(k, 0) - b.get(k, 0)) < tol for k in keys)


def test_h2_ground_energy_exact(golden):
    # python/tests/eigen_diagonalize_test.py:43 asserts this double exactly
    op, n = po.compile_fermion_source(golden["h2_source"])
    assert n == 4 and len(op) == 15
    e, _ = po.diagonalize_reference(op, n)
    assert abs(e - golden["h2_ground_energy_exact"]) < 5e-16 * 4
    assert abs(e - golden["h2_ground_energy_near"][0]) < golden["h2_ground_energy_near"][1]


def test_jw_h2_terms(golden):
    # transformations/tests/JordanWignerIRTransformationTester.cpp:176-197 (6-digit coefficients)
    op, _ = po.compile_fermion_source(golden["jw_h2_source"])
    assert len(op) == 15
    assert _close(op, _op_from_terms(golden["jw_h2_terms"]), 1e-6)


def test_jw_ladder(golden):
    # JordanWignerIRTransformationTester.cpp:200-220
    f = [(complex(c), [tuple(s) for s in sites]) for c, sites in golden["jw_ladder"]["fermion"]]
    assert _close(po.jordan_wigner(f), _op_from_terms(golden["jw_ladder"]["terms"]), 1e-12)


def test_bk_ladder_and_h2(golden):
    # transformations/tests/BravyiKitaevIRTransformationTester.cpp:108-111,157-174
    f = [(complex(c), [tuple(s) for s in sites]) for c, sites in golden["bk_ladder"]["fermion"]]
    assert _close(po.bravyi_kitaev(f, golden["bk_ladder"]["n_qubits"]), _op_from_terms(golden["bk_ladder"]["terms"]), 1e-12)
    op, n = po.compile_fermion_source(golden["bk_h2_source"], "bk")
    assert _close(op, _op_from_terms(golden["bk_h2_terms"]), 1e-6)
    assert abs(po.diagonalize_reference(op, n)[0] - golden["h2_ground_energy_exact"]) < 1e-14


def test_fenwick_sets():
    # transformations/tests/FTreeTester.cpp:39-99 (n = 8)
    t = po.FenwickTree(8)
    parity = {0: set(), 1: {0}, 2: {1}, 3: {2, 1}, 4: {3}, 5: {4, 3}, 6: {5, 3}, 7: {6, 5, 3}}
    update = {0: {1, 3, 7}, 1: {3, 7}, 2: {3, 7}, 3: {7}, 4: {5, 7}, 5: {7}, 6: {7}, 7: set()}
    children = {0: set(), 1: {0}, 2: set(), 3: {2, 1}, 4: set(), 5: {4}, 6: set(), 7: {6, 5, 3}}
    for i in range(8):
        assert t.parity_set(i) == parity[i]
        assert t.update_set(i) == update[i]
        assert t.children_set(i) == children[i]


def test_uccsd_h2_energy(golden):
    # task/tests/ComputeEnergyVQETaskTester.cpp:99-106 ; python/tests/compute_energy_test.py:41-43
    op, n = po.compile_fermion_source(golden["h2_source"])
    e = po.uccsd_energy(op, n, 2, golden["h2_uccsd_params"])
    assert abs(e - golden["h2_uccsd_energy_near"][0]) < golden["h2_uccsd_energy_near"][1]
    # same circuit as fused rotations
    rots, npar = po.uccsd_rotations(4, 2)
    assert npar == 2
    psi = np.zeros(16, complex)
    psi[0b0011] = 1
    for x, z, ic, p in rots:
        psi = po.pauli_rotation(psi, x, z, ic * golden["h2_uccsd_params"][p])
    assert abs(po.energy_reference(op, psi)[0] - e) < 1e-12


def test_uccsd_h2_generators(golden):
    # ir/tests/CommutingSetGeneratorTester.cpp:38-61: the 12 generator strings of UCCSD(H2)
    gen, _ = po.uccsd_generator(4, 2)
    got = {k[1] for k in gen}
    want = {tuple(sorted((int(q), p) for q, p in g.items())) for g in golden["uccsd_h2_generators"]}
    assert got == want


def test_vqe_h2(golden):
    # task/tests/VQEMinimizeTaskTester.cpp:98-104 ; python/tests/vqe_test.py:43
    op, n = po.compile_fermion_source(golden["h2_source"])
    rots, _ = po.uccsd_rotations(4, 2)

    def f(th):
        psi = np.zeros(16, complex)
        psi[0b0011] = 1
        for x, z, ic, p in rots:
            psi = po.pauli_rotation(psi, x, z, ic * th[p])
        return po.energy_reference(op, psi)[0]

    x, e, it = po.nelder_mead(f, golden["h2_vqe_start"])
    assert abs(e - golden["h2_vqe_energy_near"][0]) < golden["h2_vqe_energy_near"][1]


def _parse_ansatz(src, theta):
    import re
    gates = []
    for line in src.split("\n")[1:]:
        m = re.match(r"\s*(\w+)\((.*)\)", line)
        if not m:
            continue
        name, args = m.group(1), [a.strip() for a in m.group(2).split(",")]
        if name in ("Rx", "Ry", "Rz"):
            th = theta if args[0] == "theta" else float(args[0])
            gates.append((name, (int(args[1]),), th))
        elif name == "CNOT":
            gates.append((name, (int(args[0]), int(args[1])), 0.0))
        else:
            gates.append((name, (int(args[0]),), 0.0))
    return gates


def test_rdm_h2(golden):
    # decorators/tests/RDMGeneratorTester.cpp:123-162: 54 circuits, E = -1.1371 +- 1e-4
    terms, n = po.parse_fermion_source(golden["rdm_h2_source"])
    e_nuc, hpq, hpqrs = po.hpq_hpqrs(terms, n)
    psi = po.run_circuit(4, _parse_ansatz(golden["rdm_ansatz"], golden["rdm_theta"]))
    rho, ncirc = po.rdm2(psi, 4)
    assert ncirc == golden["rdm_n_circuits"]
    e = po.rdm_energy(rho, e_nuc, hpq, hpqrs)
    assert abs(e - golden["rdm_energy_near"][0]) < golden["rdm_energy_near"][1]
    op = po.jordan_wigner(terms)
    assert abs(e - po.energy_reference(op, psi)[0]) < 1e-9     # 2-RDM contraction == <H>


def test_text_fixtures(golden):
    # compiler/optimizers/tests/QubitTaperingTester.cpp:72,149,156 ; examples/deuteron/python/d2.py, d3.py
    d = golden["derived"]
    for key, n, want in (("six_qubit_118", 6, d["six_qubit_E0"]), ("four_qubit_100", 4, d["four_qubit_E0"]),
                         ("deuteron_h2", 2, d["deuteron_h2_E0"]), ("deuteron_h3", 3, d["deuteron_h3_E0"])):
        op = po.op_from_string(golden[key])
        assert po.op_n_qubits(op) == n
        assert abs(po.diagonalize_reference(op, n)[0] - want) < 1e-10 * abs(want)
    assert len(po.op_from_string(golden["six_qubit_118"])) == 118
    assert len(po.op_from_string(golden["four_qubit_100"])) == 100
    assert len(po.op_from_string(golden["tapered_h2"])) == 3


def test_mask_form_matches_string_form(golden):
    """The packed (xmask, zmask, coef*i^nY) form reproduces the string-based row builder."""
    for key in ("six_qubit_118", "four_qubit_100"):
        op = po.op_from_string(golden[key])
        n = po.op_n_qubits(op)
        A = po.dense_matrix_reference(op, n)
        P = po.bit_reverse_permutation(n)
        D = po.dense_matrix_device_order(op, n)
        assert np.abs(D[np.ix_(P, P)] - A).max() < 1e-13 * max(1.0, np.abs(A).max())
        assert np.abs(co.faithful_dense(op, n) - A).max() < 1e-13 * max(1.0, np.abs(A).max())


@pytest.mark.parametrize("n,T,seed", [(6, 20, 1), (10, 40, 2), (13, 30, 3)])
def test_c_oracle_matches_numpy(n, T, seed):
    rng = np.random.default_rng(seed)
    xm = rng.integers(0, 1 << n, T, dtype=np.uint64)
    zm = rng.integers(0, 1 << n, T, dtype=np.uint64)
    ny = np.array([bin(int(x) & int(z)).count("1") for x, z in zip(xm, zm)])
    cf = rng.normal(size=T) * (1j ** (ny % 4))
    psi = rng.normal(size=1 << n) + 1j * rng.normal(size=1 << n)
    psi /= np.linalg.norm(psi)
    assert np.abs(co.apply(n, xm, zm, cf, psi) - po.apply_masks(xm, zm, cf, psi)).max() < 1e-13
    e1, p1 = co.expect(n, xm, zm, cf, psi)
    e2, p2 = po.expect_masks(xm, zm, cf, psi)
    assert abs(e1 - e2) < 1e-13 and np.abs(p1 - p2).max() < 1e-13
    assert np.abs(p1.imag).max() < 1e-13
    q = psi.copy()
    co.pauli_rotation(n, int(xm[0]), int(zm[0]), 0.37, q)
    assert np.abs(q - po.pauli_rotation(psi, int(xm[0]), int(zm[0]), 0.37)).max() < 1e-14
    for name, qs, th in (("H", (1,), 0.0), ("Rx", (2,), 0.3), ("Ry", (0,), -0.8), ("Rz", (3,), 1.1), ("CNOT", (1, 4), 0.0),
                         ("CNOT", (5, 2), 0.0), ("X", (n - 1,), 0.0), ("Y", (0,), 0.0), ("Z", (2,), 0.0)):
        q = psi.copy()
        co.apply_gate(n, name, qs, th, q)
        assert np.abs(q - po.apply_gate(psi, name, qs, th)).max() < 1e-14


@pytest.mark.parametrize("tr", ["jw", "bk"])
def test_number_symmetry_sectors_tile_the_pinned_spectrum(golden, tr):
    """diag-number-symmetry restatement (DiagonalizeTask.cpp:54-183) against the pinned dense branch: for H2 (n = 4, where
    the reference's BK matrix matches its BK transformation) the n_e-electron sectors are invariant, so the literal
    matrix (std::map quirk included) equals the compression P H P, the union of the sector spectra is the full spectrum,
    and the 2-electron sector holds the reference's golden ground energy."""
    op, n = po.compile_fermion_source(golden["h2_source"], tr)
    full = np.linalg.eigvalsh(po.dense_matrix_reference(op, n))
    parts = []
    for ne in range(n + 1):
        lit = po.number_subspace_matrix(op, n, ne, tr, literal=True)
        proj = po.number_subspace_matrix(op, n, ne, tr, literal=False)
        assert np.abs(lit - proj).max() < 1e-14
        assert np.abs(lit - lit.conj().T).max() < 1e-14
        assert lit.shape[0] == len(po.number_sector_strings(n, ne, tr)) == [1, 4, 6, 4, 1][ne]
        parts.append(np.linalg.eigvalsh(lit))
    assert np.abs(np.sort(np.concatenate(parts)) - full).max() < 1e-12
    assert abs(po.diagonalize_number_symmetry(op, n, 2, tr) - golden["h2_ground_energy_exact"]) < 1e-13
    # the BK re-encoding is a bijection on bit strings and keeps the vacuum
    if tr == "bk":
        allstr = [s for ne in range(n + 1) for s in po.number_sector_strings(n, ne, "bk")]
        assert len(set(allstr)) == 1 << n and po.number_sector_strings(n, 0, "bk") == ["0" * n]
