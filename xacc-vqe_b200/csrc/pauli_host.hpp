// Host-side Pauli operator algebra in mask form (product code; independent of oracle/).
//
// Stands in for xacc::quantum::PauliOperator as the reference uses it
// (transformations/jw/JordanWignerIRTransformation.cpp:35-54,
//  transformations/bk/BravyiKitaevIRTransformation.cpp:38-74, task/VQEProgram.hpp:204-220).
// A term is (variable name, xmask, zmask) -> complex coefficient, the string being the plain
// tensor product of I/X/Y/Z (Y is Y, not XZ); packing to coef*i^nY happens in pack().
#pragma once
#include <complex>
#include <cstdint>
#include <map>
#include <string>
#include <tuple>
#include <vector>

namespace b200vqe {

using cplx = std::complex<double>;

struct TermKey {
    std::string var;
    uint64_t x = 0, z = 0;
    bool operator<(const TermKey &o) const { return std::tie(var, x, z) < std::tie(o.var, o.x, o.z); }
};

constexpr double kDropTol = 1e-12;   // XACC drops |c| < 1e-12 when accumulating

struct PauliOp {
    std::map<TermKey, cplx> terms;

    static PauliOp identity(cplx c, const std::string &var = "") {
        PauliOp p;
        p.terms[TermKey{var, 0, 0}] = c;
        return p;
    }
    static PauliOp single(uint64_t x, uint64_t z, cplx c) {
        PauliOp p;
        p.terms[TermKey{"", x, z}] = c;
        return p;
    }
    void add_term(const TermKey &k, cplx c) {
        auto it = terms.find(k);
        if (it == terms.end()) {
            if (std::abs(c) >= kDropTol) terms.emplace(k, c);
        } else {
            it->second += c;
            if (std::abs(it->second) < kDropTol) terms.erase(it);
        }
    }
    PauliOp &operator+=(const PauliOp &o) {
        for (auto &kv : o.terms) add_term(kv.first, kv.second);
        return *this;
    }
    PauliOp operator*(const PauliOp &o) const;
    PauliOp scaled(cplx s) const {
        PauliOp r;
        for (auto &kv : terms) r.terms[kv.first] = kv.second * s;
        return r;
    }
    int n_qubits() const {
        uint64_t m = 0;
        for (auto &kv : terms) m |= kv.first.x | kv.first.z;
        return m ? 64 - __builtin_clzll(m) : 0;
    }
};

// var + "X0Z1..." ascending qubit, "I" when empty (python/xaccvqe.py:149-152)
std::string term_id(const TermKey &k);
std::string to_string(const PauliOp &op);
bool from_string(const std::string &text, PauliOp &out, std::string &err);

struct FermionTerm {
    cplx coeff;
    std::vector<std::pair<int, int>> sites;   // (site, isCreation)
    std::string var;
};

// compiler/FermionCompiler.cpp:62-103
bool parse_fermion_source(const std::string &src, std::vector<FermionTerm> &terms, int &n_qubits,
                          std::string &err);
PauliOp jordan_wigner(const std::vector<FermionTerm> &terms);
PauliOp bravyi_kitaev(const std::vector<FermionTerm> &terms, int n_qubits);

struct PackedTerm { uint64_t x, z; cplx coef; TermKey key; };
// sorted by (xmask, zmask); coef = c * i^nY  
std::vector<PackedTerm> pack(const PauliOp &op);

struct Rotation { uint64_t x, z; double imag_coeff; int param; };
// ir/algorithms/uccsd/UCCSD.cpp:96-365 ; order documented in DESIGN.md (ascending term id, then
// the reference's commuting-set grouping, CommutingSetGenerator.hpp:51-93)
bool uccsd_rotations(int n_qubits, int n_electrons, const std::string &transformation,
                     std::vector<Rotation> &rot, int &n_params, std::string &err);

// task/tasks/DiagonalizeTask.cpp:54-144: rows[i] such that occupation bit i of basis index j is parity(rows[i] & j)
bool number_sector_rows(int n, const std::string &transformation, std::vector<uint64_t> &rows, std::string &err);

}  // namespace b200vqe
