"""Extracts the reference's own inline fixtures (fermion sources, Pauli strings, expected values)
from /root/reference test files into tests/golden/reference_fixtures.json.  Run in the build
container only (the GPU box has no /root/reference); the JSON is committed."""
import json
import os
import re

REF = "/root/reference"


def read(p):
    with open(os.path.join(REF, p)) as f:
        return f.read()


def raw_strings(txt, tag):
    return re.findall(r'R"%s\((.*?)\)%s"' % (tag, tag), txt, re.S)


def pauli_terms(block):
    """PauliOperator({{0,"Z"},{1,"Z"}}, .17) lines -> [[{q:P}, coeff], ...]"""
    out = []
    for m in re.finditer(r'PauliOperator(?: expected)?\(\s*\{(.*?)\}\s*,\s*([-+0-9.eE]+)\s*\)', block):
        ops = {int(q): p for q, p in re.findall(r'\{\s*(\d+)\s*,\s*"([XYZ])"\s*\}', m.group(1))}
        out.append([{str(k): v for k, v in ops.items()}, float(m.group(2))])
    return out


fx = {}
diag = read("task/tests/DiagonalizeTaskTester.cpp")
fx["h2_source"] = raw_strings(diag, "src")[0]
fx["h2_ground_energy_exact"] = -1.137270422069755          # python/tests/eigen_diagonalize_test.py:43
fx["h2_ground_energy_near"] = [-1.13727042207, 1e-4]       # task/tests/DiagonalizeTaskTester.cpp:106
fx["h2_uccsd_params"] = [0.0, -0.0571583356234]            # task/tests/ComputeEnergyVQETaskTester.cpp:99-106
fx["h2_uccsd_energy_near"] = [-1.13727042207, 1e-4]
fx["h2_vqe_start"] = [0.000641023496104, 4.76879126994]    # task/tests/VQEMinimizeTaskTester.cpp:98-104
fx["h2_vqe_energy_near"] = [-1.13727, 1e-4]

jw = read("transformations/tests/JordanWignerIRTransformationTester.cpp")
i0 = jw.index("PauliOperator expected({{2,\"Z\"}, {3,\"Z\"}}")
fx["jw_easy_source"] = raw_strings(jw, "code")[0]
fx["jw_h2_source"] = raw_strings(jw, "code")[1]
fx["jw_h2_terms"] = pauli_terms(jw[i0:jw.index("EXPECT_TRUE(expected == result)", i0)])
fx["jw_ladder"] = {"fermion": [[3.17, [[2, 1], [0, 0]]], [3.17, [[0, 1], [2, 0]]]],
                   "terms": [[{"0": "X", "1": "Z", "2": "X"}, 1.585], [{"0": "Y", "1": "Z", "2": "Y"}, 1.585]]}

bk = read("transformations/tests/BravyiKitaevIRTransformationTester.cpp")
fx["bk_ladder"] = {"fermion": [[3.17, [[2, 1], [0, 0]]], [3.17, [[0, 1], [2, 0]]]], "n_qubits": 3,
                   "terms": [[{"0": "Y", "1": "Y"}, -1.585], [{"0": "X", "1": "X", "2": "Z"}, -1.585]]}
i1 = bk.index("PauliOperator expected({{0,\"Z\"}, {1,\"Z\"}, {2,\"Z\"}}")
fx["bk_h2_source"] = raw_strings(bk, "code")[0]
fx["bk_h2_terms"] = pauli_terms(bk[i1:bk.index("EXPECT_TRUE(expected == result)", i1)]) 

rdm = read("decorators/tests/RDMGeneratorTester.cpp")
fx["rdm_h2_source"] = raw_strings(rdm, "src")[0]
fx["rdm_ansatz"] = raw_strings(rdm, "rucc")[0]
fx["rdm_theta"] = 0.22984
fx["rdm_n_circuits"] = 54
fx["rdm_energy_near"] = [-1.1371, 1e-4]

tap = read("compiler/optimizers/tests/QubitTaperingTester.cpp")
strs = re.findall(r'fromString\(\s*"(.*?)"\s*\)', tap, re.S)
fx["tapered_h2"] = strs[0]
long_strs = sorted(set(s for s in strs if len(s) > 500), key=len)
fx["four_qubit_100"] = long_strs[0]
fx["six_qubit_118"] = long_strs[-1]

cs = read("ir/tests/CommutingSetGeneratorTester.cpp")
gens = []
for m in re.finditer(r'PauliOperator inst\d+\(std::map<int, std::string>\s*\{(.*?)\}\s*\)\s*;', cs, re.S):
    gens.append({q: p for q, p in re.findall(r'\{\s*(\d+)\s*,\s*"([XYZ])"\s*\}', m.group(1))})
fx["uccsd_h2_generators"] = gens

fx["deuteron_h2"] = "(5.906709445,0) I + (-2.1433,0) X0 X1 + (-2.1433,0) Y0 Y1 + (0.21829,0) Z0 + (-6.125,0) Z1"
fx["deuteron_h3"] = (fx["deuteron_h2"] + " + (9.625,0) I + (-3.913119,0) X1 X2 + (-3.913119,0) Y1 Y2 + (-9.625,0) Z2")
# derived known answers listed in SURVEY.md section 4 (dense solve of the fixtures above; cross-checks, not reference asserts)
fx["derived"] = {"rdm_h2_E0": -1.1371757101731699, "six_qubit_E0": -160.31197869072162,
                 "four_qubit_E0": -160.3119790682468, "deuteron_h2_E0": -1.7491554691752755,
                 "deuteron_h3_E0": -2.0456659405706388}

out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_fixtures.json")
with open(out, "w") as f:
    json.dump(fx, f, indent=1, sort_keys=True)
print("wrote", out, {k: (len(v) if hasattr(v, "__len__") else v) for k, v in fx.items()})
