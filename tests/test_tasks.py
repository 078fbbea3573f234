"""The task layer (xacc-vqe_b200/tasks.py): host logic on the CPU, the reference's own test scenarios on the GPU
(python/tests/eigen_diagonalize_test.py, compute_energy_test.py, vqe_test.py, decorators/tests/RDMGeneratorTester.cpp)."""
import os

import numpy as np
import pytest

import pauli_oracle as po
import xaccvqe_b200 as xb
from test_oracle_golden import _parse_ansatz


# ------------------------------------------------------------------ CPU: host logic
def test_nelder_mead_walks_the_same_simplex_as_the_oracle():
    def rosen(v):
        return float((1 - v[0]) ** 2 + 100 * (v[1] - v[0] ** 2) ** 2 + 0.5 * (v[2] - 0.3) ** 2)
    for x0 in ([-1.2, 1.0, 0.0], [0.5, 0.5, 0.5], [0.000641023496104, 4.76879126994, -2.0]):
        x, f, it = xb.nelder_mead(rosen, x0, 1e-9, 400)
        xo, fo, ito = po.nelder_mead(rosen, x0, 1e-9, 400)
        assert it == ito and f == fo and np.array_equal(x, xo)


def test_nelder_mead_follows_the_reference_solver():
    """neldermeadsolver.h details that change iteration counts and nQpuCalls: the spread test (f and x spreads <= 1e-4) ends a
    walk that fDelta = 1e-9 alone would continue; the cap is `iterations` (checkConvergence), not iterations * DIM; the
    reported energy is the LAST evaluated value; every objective call is counted identically by the mirror and the oracle."""
    def quad(v):
        return float((v[0] - 1.0) ** 2 + 3.0 * (v[1] + 0.5) ** 2) - 2.0
    for fn, arg in ((xb.nelder_mead, None), (po.nelder_mead, None)):
        calls = []
        x, f, it = fn(lambda v: (calls.append(np.array(v)), quad(v))[1], [0.0, 0.0], 1e-9, 10000)
        assert f == quad(calls[-1]) and np.abs(x - [1.0, -0.5]).max() < 2e-3
        # stopped by the 1e-4 spread rule: a tighter simplex than 1e-4 would need many more steps
        assert 10 < it < 80, it
        xc, fc, itc = fn(quad, [0.0, 0.0], 1e-12, 5)
        assert itc == 6                      # iter > iterations is tested before the step: steps 0..5 run
    a = []
    b = []
    xb.nelder_mead(lambda v: (a.append(tuple(v)), quad(v))[1], [0.3, -0.2], 1e-9, 300)
    po.nelder_mead(lambda v: (b.append(tuple(v)), quad(v))[1], [0.3, -0.2], 1e-9, 300)
    assert a == b and len(a) > 20


def test_parameter_generation_and_result_persist(tmp_path):
    from xaccvqe_b200.tasks import generate_parameters
    assert generate_parameters(2, "0.0,-0.0571583356234").tolist() == [0.0, -0.0571583356234]
    assert generate_parameters(3, "1.5").tolist() == [1.5, 0.0, 0.0]
    r = generate_parameters(5, None, seed=3)
    assert len(r) == 5 and np.all(np.abs(r) <= np.pi) and np.array_equal(r, generate_parameters(5, None, seed=3))
    f = tmp_path / "vqe.csv"
    res = xb.VQETaskResult(energy=-1.25, angles=np.array([0.1, 0.2]), expVals={"Z0": 0.5, "X0X1": -0.25}, fileName=str(f))
    res.persist()
    res.energy = -1.5
    res.persist()
    lines = f.read_text().strip().split("\n")
    assert lines[0] == "t0,t1,X0X1,Z0,E" and lines[1] == "0.1,0.2,-0.25,0.5,-1.25" and lines[2].endswith(",-1.5") and len(lines) == 3


def test_buffer_and_operator_wrappers(golden):
    op = xb.compile(golden["h2_source"])
    assert op.nTerms() == 15 and op.nQubits() == 4          # python/tests/compile_test.py:43-62
    again = xb.PauliOperator.fromString(op.toString())
    assert np.array_equal(again.pack()[2], op.pack()[2])
    two = op + op
    assert np.allclose(two.pack()[2], 2 * op.pack()[2])
    b = xb.AcceleratorBuffer("q", 4)
    b.addExtraInfo("vqe-energy", -1.0)
    b.appendChild({"kernel": "Z0", "exp-val-z": 1.0})
    assert b.getInformation("vqe-energy") == -1.0 and b.getChildren("kernel", "Z0")[0]["exp-val-z"] == 1.0


def test_rdm_generator_enumerates_the_reference_circuits(golden):
    g = xb.RDMGenerator(4)
    assert g.n_circuits() == golden["rdm_n_circuits"]       # decorators/tests/RDMGeneratorTester.cpp:132
    terms, _ = po.parse_fermion_source(golden["rdm_h2_source"])
    e_nuc, hpq, hpqrs = po.hpq_hpqrs(terms, 4)
    rho = np.random.default_rng(0).normal(size=(4, 4, 4, 4)) + 0j
    assert abs(xb.rdm_energy(rho, e_nuc, hpq, hpqrs) - po.rdm_energy(rho, e_nuc, hpq, hpqrs)) < 1e-12


def test_execute_rejects_bad_arguments(golden):
    op = xb.compile(golden["h2_source"])
    with pytest.raises(xb.B200Error):
        xb.execute(op, None, task="no-such-task")
    with pytest.raises(xb.B200Error):
        xb.execute(op, None, task="compute-energy", ansatz="uccsd")      # n-electrons missing (UCCSD.cpp:16-22)
    with pytest.raises(xb.B200Error, match="disagree"):                   # one option drives compiler, UCCSD and sector rows
        xb.execute(golden["h2_source"], None, **{"transformation": "jw", "fermion-transformation": "bk"})


# ------------------------------------------------------------------ GPU: the reference's scenarios
@pytest.mark.gpu
def test_diagonalize_task_h2_exact(golden):
    """python/tests/eigen_diagonalize_test.py:43 (exact double) and task/tests/DiagonalizeTaskTester.cpp:103-107."""
    buf = xb.AcceleratorBuffer("q", 4)
    r = xb.execute(xb.compile(golden["h2_source"]), buf, task="vqe-diagonalize")
    assert abs(r.energy - golden["h2_ground_energy_exact"]) < 1e-10 * abs(r.energy)
    assert abs(buf.getInformation("vqe-energy") - r.energy) == 0.0
    r2 = xb.execute(golden["h2_source"], None, **{"task": "vqe-diagonalize", "n-qubits": 4})
    assert abs(r2.energy - r.energy) < 1e-12


@pytest.mark.gpu
@pytest.mark.parametrize("tr", ["jw", "bk"])
def test_both_spellings_of_the_transformation_option_agree(golden, tr):
    """'transformation' (python kwarg) and 'fermion-transformation' (the option it sets, python/xacc-vqe-py.cpp:406-408) select
    the same operator AND the same number-symmetry rows when the source is compiled inside execute()."""
    kw = {"task": "vqe-diagonalize", "diag-number-symmetry": True, "n-electrons": 2}
    a = xb.execute(golden["h2_source"], None, **dict(kw, **{"transformation": tr}))
    b = xb.execute(golden["h2_source"], None, **dict(kw, **{"fermion-transformation": tr}))
    assert a.energy == b.energy and abs(a.energy - golden["h2_ground_energy_exact"]) < 1e-10


@pytest.mark.gpu
@pytest.mark.parametrize("tr", ["jw", "bk"])
def test_diagonalize_task_number_symmetry(golden, tr):
    """Options diag-number-symmetry + n-electrons (+ fermion-transformation) of DiagonalizeTask.cpp:48-57: the 1-electron
    sector of H2 lies above the ground state; the 2-electron sector holds it."""
    import pauli_oracle as po
    oop, n = po.compile_fermion_source(golden["h2_source"], tr)
    op = xb.compile(golden["h2_source"], tr)
    for ne in (1, 2):
        r = xb.execute(op, None, **{"task": "vqe-diagonalize", "diag-number-symmetry": True, "n-electrons": ne,
                                    "fermion-transformation": tr})
        assert abs(r.energy - po.diagonalize_number_symmetry(oop, n, ne, tr)) < 1e-10
    assert abs(r.energy - golden["h2_ground_energy_exact"]) < 1e-10


@pytest.mark.gpu
def test_compute_energy_task_uccsd(golden):
    """task/tests/ComputeEnergyVQETaskTester.cpp:99-106, python/tests/compute_energy_test.py:41-43."""
    op = xb.compile(golden["h2_source"])
    buf = xb.AcceleratorBuffer("q", 4)
    th = golden["h2_uccsd_params"]
    r = xb.execute(op, buf, **{"task": "compute-energy", "ansatz": "uccsd", "n-electrons": 2, "vqe-params": f"{th[0]},{th[1]}"})
    assert abs(r.energy - golden["h2_uccsd_energy_near"][0]) < golden["h2_uccsd_energy_near"][1]
    oop, _ = po.compile_fermion_source(golden["h2_source"])
    assert abs(r.energy - po.uccsd_energy(oop, 4, 2, th)) < 1e-10
    assert r.nQpuCalls == 14 and r.vqeIterations == 1                  # 15 terms, one of them the identity (:134)
    assert len(buf.getChildren()) == 15 and buf.getChildren("kernel", "I")[0]["exp-val-z"] == 1.0
    e_from_children = sum(c["coefficient"] * c["exp-val-z"] for c in buf.getChildren())
    assert abs(e_from_children - r.energy) < 1e-12
    assert buf.getInformation("vqe-nQPU-calls") == 14


@pytest.mark.gpu
def test_vqe_minimize_task(golden):
    """task/tests/VQEMinimizeTaskTester.cpp:98-104, python/tests/vqe_test.py:43; the oracle walks the same simplex."""
    op = xb.compile(golden["h2_source"])
    s = golden["h2_vqe_start"]
    r = xb.execute(op, None, **{"task": "vqe", "ansatz": "uccsd", "n-electrons": 2, "vqe-params": f"{s[0]},{s[1]}"})
    assert abs(r.energy - golden["h2_vqe_energy_near"][0]) < golden["h2_vqe_energy_near"][1]
    oop, _ = po.compile_fermion_source(golden["h2_source"])
    xo, eo, ito = po.nelder_mead(lambda t: po.uccsd_energy(oop, 4, 2, t), s)
    assert abs(r.energy - eo) < 1e-10 and np.abs(r.angles - xo).max() < 1e-6
    assert r.nQpuCalls == 14 * r.vqeIterations and r.vqeIterations >= ito


@pytest.mark.gpu
def test_rdm_generator_on_device(golden):
    """decorators/tests/RDMGeneratorTester.cpp:123-162: 54 circuits, energy from the 2-RDM contraction."""
    gates = _parse_ansatz(golden["rdm_ansatz"], golden["rdm_theta"])
    be = xb.Backend(4)
    be.reset(0, 0)
    be.apply_gates(gates)
    g = xb.RDMGenerator(4)
    rho = g.generate(be)
    psi = po.run_circuit(4, gates)
    rho_o, ncirc = po.rdm2(psi, 4)
    assert g.n_circuits() == ncirc == golden["rdm_n_circuits"]
    assert np.abs(rho - rho_o).max() < 1e-10
    terms, _ = po.parse_fermion_source(golden["rdm_h2_source"])
    e_nuc, hpq, hpqrs = po.hpq_hpqrs(terms, 4)
    e = xb.rdm_energy(rho, e_nuc, hpq, hpqrs)
    assert abs(e - golden["rdm_energy_near"][0]) < golden["rdm_energy_near"][1]
    op = xb.compile(golden["rdm_h2_source"])
    prog = xb.VQEProgram(op, 4, backend=be)              # 2-RDM contraction == <H> on the same state
    assert abs(e - be.expectation().real) < 1e-9


@pytest.mark.gpu
def test_twelve_qubit_bk_uccsd_sweep_matches_oracle():
    """Config 2 shape: 12-qubit BK Hamiltonian, UCCSD (4 electrons, 44 parameters), a few parameter vectors."""
    rng = np.random.default_rng(12)
    n, ne = 12, 4
    src = ["__qpu__ k() {", "0.3"]
    for p in range(n):
        src.append(f"{rng.normal():.6f} {p} 1 {p} 0")
    for _ in range(40):
        p, q, r, s = rng.choice(n, 4, replace=False)
        c = rng.normal() * 0.1
        src.append(f"{c:.6f} {p} 1 {q} 1 {r} 0 {s} 0")
        src.append(f"{c:.6f} {s} 1 {r} 1 {q} 0 {p} 0")
    src.append("}")
    src = "\n".join(src)
    op = xb.compile(src, "bk")
    oop, on = po.compile_fermion_source(src, "bk")
    assert on == n and op.nQubits() == n
    ans = xb.Ansatz.uccsd(n, ne, "bk")
    assert ans.n_params == 44                            # nS + nS(nS+1)/2, nS = 2*4 (UCCSD.cpp:97-101)
    prog = xb.VQEProgram(op, n, ans)
    task = xb.ComputeEnergyVQETask(prog, per_term=False)
    thetas = rng.uniform(-0.1, 0.1, size=(3, 44))
    for th in thetas:
        e = task.execute(th).energy
        eo = po.uccsd_energy(oop, n, ne, th, "bk")
        assert abs(e - eo) < 1e-10 * max(1.0, abs(eo))
    sw = prog.backend.sweep_energies(ans.basis_index, ans.xmask, ans.zmask, ans.imag_coeff, ans.param, ans.n_params, thetas)
    assert np.abs(sw - [po.uccsd_energy(oop, n, ne, th, "bk") for th in thetas]).max() < 1e-10
