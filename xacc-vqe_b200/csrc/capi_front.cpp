// C ABI of the host front end (PauliOperator stand-in, FermionCompiler text, JW/BK, UCCSD).
#include "../../include/b200vqe.h"
#include "pauli_host.hpp"

#include <cmath>
#include <cstring>

using namespace b200vqe;

struct b200vqe_opr {
    PauliOp op;
    int n_qubits_hint = 0;
    std::vector<PackedTerm> packed;   // cache, rebuilt on demand
    bool packed_valid = false;
    const std::vector<PackedTerm> &get_packed() {
        if (!packed_valid) { packed = pack(op); packed_valid = true; }
        return packed;
    }
};

namespace { thread_local std::string f_err; }
extern "C" const char *b200vqe_front_error(void) { return f_err.c_str(); }

static int copy_out(const std::string &s, char *buf, int64_t cap, int64_t *needed) {
    if (needed) *needed = (int64_t)s.size() + 1;
    if (!buf || cap < (int64_t)s.size() + 1) return buf ? B200VQE_EINVAL : 0;
    memcpy(buf, s.c_str(), s.size() + 1);
    return 0;
}

extern "C" {

int b200vqe_op_from_fermion_source(const char *src, const char *transformation, int n_qubits, b200vqe_op *out) {
    if (!src || !out) { f_err = "null argument"; return B200VQE_EINVAL; }
    std::vector<FermionTerm> terms;
    int nq = 0;
    if (!parse_fermion_source(src, terms, nq, f_err)) return B200VQE_EPARSE;
    if (n_qubits > 0) {
        if (n_qubits < nq) { f_err = "n-qubits smaller than the largest site index + 1"; return B200VQE_EINVAL; }
        nq = n_qubits;
    }
    const std::string t = transformation ? transformation : "jw";
    auto *o = new b200vqe_opr();
    if (t == "jw" || t.empty()) o->op = jordan_wigner(terms);
    else if (t == "bk") o->op = bravyi_kitaev(terms, nq);
    else { delete o; f_err = "unknown fermion-transformation '" + t + "' (jw | bk)"; return B200VQE_EINVAL; }
    o->n_qubits_hint = nq;
    *out = o;
    return 0;
}

int b200vqe_op_from_string(const char *text, b200vqe_op *out) {
    if (!text || !out) { f_err = "null argument"; return B200VQE_EINVAL; }
    auto *o = new b200vqe_opr();
    if (!from_string(text, o->op, f_err)) { delete o; return B200VQE_EPARSE; }
    *out = o;
    return 0;
}

int b200vqe_op_from_terms(int64_t n_terms, const uint64_t *xmask, const uint64_t *zmask, const double *c_re,
                          const double *c_im, b200vqe_op *out) {
    if (!out || n_terms < 0 || (n_terms && (!xmask || !zmask || !c_re))) { f_err = "null argument"; return B200VQE_EINVAL; }
    auto *o = new b200vqe_opr();
    for (int64_t k = 0; k < n_terms; ++k)
        o->op.add_term(TermKey{"", xmask[k], zmask[k]}, cplx(c_re[k], c_im ? c_im[k] : 0.0));
    *out = o;
    return 0;
}

int b200vqe_op_destroy(b200vqe_op op) { delete op; return 0; }

int b200vqe_op_n_terms(b200vqe_op op, int64_t *n) {
    if (!op || !n) return B200VQE_EINVAL;
    *n = (int64_t)op->op.terms.size();
    return 0;
}
int b200vqe_op_n_qubits(b200vqe_op op, int *n) {
    if (!op || !n) return B200VQE_EINVAL;
    *n = std::max(op->n_qubits_hint, op->op.n_qubits());
    return 0;
}

int b200vqe_op_pack(b200vqe_op op, uint64_t *xmask, uint64_t *zmask, double *coef_re, double *coef_im) {
    if (!op || !xmask || !zmask || !coef_re || !coef_im) return B200VQE_EINVAL;
    const auto &p = op->get_packed();
    for (size_t k = 0; k < p.size(); ++k) {
        xmask[k] = p[k].x; zmask[k] = p[k].z; coef_re[k] = p[k].coef.real(); coef_im[k] = p[k].coef.imag();
    }
    return 0;
}

int b200vqe_op_to_string(b200vqe_op op, char *buf, int64_t cap, int64_t *needed) {
    if (!op) return B200VQE_EINVAL;
    return copy_out(to_string(op->op), buf, cap, needed);
}

int b200vqe_op_term_id(b200vqe_op op, int64_t k, char *buf, int64_t cap, int64_t *needed) {
    if (!op) return B200VQE_EINVAL;
    const auto &p = op->get_packed();
    if (k < 0 || k >= (int64_t)p.size()) return B200VQE_EINVAL;
    return copy_out(term_id(p[k].key), buf, cap, needed);
}

static b200vqe_op wrap(PauliOp &&p, int hint) {
    auto *o = new b200vqe_opr();
    o->op = std::move(p);
    o->n_qubits_hint = hint;
    return o;
}
int b200vqe_op_add(b200vqe_op a, b200vqe_op b, b200vqe_op *out) {
    if (!a || !b || !out) return B200VQE_EINVAL;
    PauliOp r = a->op;
    r += b->op;
    *out = wrap(std::move(r), std::max(a->n_qubits_hint, b->n_qubits_hint));
    return 0;
}
int b200vqe_op_mul(b200vqe_op a, b200vqe_op b, b200vqe_op *out) {
    if (!a || !b || !out) return B200VQE_EINVAL;
    *out = wrap(a->op * b->op, std::max(a->n_qubits_hint, b->n_qubits_hint));
    return 0;
}
int b200vqe_op_scale(b200vqe_op a, double re, double im, b200vqe_op *out) {
    if (!a || !out) return B200VQE_EINVAL;
    *out = wrap(a->op.scaled(cplx(re, im)), a->n_qubits_hint);
    return 0;
}

int b200vqe_uccsd_rotations(int n_qubits, int n_electrons, const char *transformation, int64_t *count,
                            int *n_params, uint64_t *xmask, uint64_t *zmask, double *imag_coeff, int32_t *param) {
    if (!count) return B200VQE_EINVAL;
    std::vector<Rotation> rot;
    int np = 0;
    if (!uccsd_rotations(n_qubits, n_electrons, transformation ? transformation : "jw", rot, np, f_err)) return B200VQE_EINVAL;
    *count = (int64_t)rot.size();
    if (n_params) *n_params = np;
    if (xmask && zmask && imag_coeff && param)
        for (size_t r = 0; r < rot.size(); ++r) {
            xmask[r] = rot[r].x; zmask[r] = rot[r].z; imag_coeff[r] = rot[r].imag_coeff; param[r] = rot[r].param;
        }
    return 0;
}

// literal gate list of UCCSD.cpp:259-365 for the same ordered rotations
int b200vqe_uccsd_gates(int n_qubits, int n_electrons, const char *transformation, const double *theta, int n_theta,
                        int64_t *count, b200vqe_gate *gates) {
    if (!count || !theta) return B200VQE_EINVAL;
    std::vector<Rotation> rot;
    int np = 0;
    if (!uccsd_rotations(n_qubits, n_electrons, transformation ? transformation : "jw", rot, np, f_err)) return B200VQE_EINVAL;
    if (n_theta != np) { f_err = "UCCSD expects " + std::to_string(np) + " parameters, got " + std::to_string(n_theta); return B200VQE_EINVAL; }
    const double pi = 3.14159265358979323;
    std::vector<b200vqe_gate> g;
    for (int i = 0; i < n_electrons; ++i) g.push_back(b200vqe_gate{B200VQE_X, i, 0, 0, 0.0});
    for (auto &r : rot) {
        std::vector<std::pair<int, char>> t;
        for (int q = 0; q < n_qubits; ++q) {
            const bool x = (r.x >> q) & 1, z = (r.z >> q) & 1;
            if (x || z) t.emplace_back(q, x ? (z ? 'Y' : 'X') : 'Z');
        }
        for (int i = (int)t.size() - 1; i >= 0; --i) {          // inserted at position 0 in ascending i
            if (t[i].second == 'X') g.push_back(b200vqe_gate{B200VQE_H, t[i].first, 0, 0, 0.0});
            else if (t[i].second == 'Y') g.push_back(b200vqe_gate{B200VQE_RX, t[i].first, 0, 0, pi / 2.0});
        }
        for (size_t i = 0; i + 1 < t.size(); ++i) g.push_back(b200vqe_gate{B200VQE_CNOT, t[i].first, t[i + 1].first, 0, 0.0});
        g.push_back(b200vqe_gate{B200VQE_RZ, t.back().first, 0, 0, 2.0 * r.imag_coeff * theta[r.param]});
        for (int i = (int)t.size() - 2; i >= 0; --i) g.push_back(b200vqe_gate{B200VQE_CNOT, t[i].first, t[i + 1].first, 0, 0.0});
        for (int i = (int)t.size() - 1; i >= 0; --i) {
            if (t[i].second == 'X') g.push_back(b200vqe_gate{B200VQE_H, t[i].first, 0, 0, 0.0});
            else if (t[i].second == 'Y') g.push_back(b200vqe_gate{B200VQE_RX, t[i].first, 0, 0, -pi / 2.0});
        }
    }
    *count = (int64_t)g.size();
    if (gates) memcpy(gates, g.data(), g.size() * sizeof(b200vqe_gate));
    return 0;
}

// host-only view of the diag-number-symmetry occupation read-out (task/tasks/DiagonalizeTask.cpp:54-144)
int b200vqe_number_sector_rows(int n_qubits, const char *transformation, uint64_t *rows) {
    if (!rows) { f_err = "rows is NULL"; return B200VQE_EINVAL; }
    std::vector<uint64_t> r;
    if (!number_sector_rows(n_qubits, transformation ? transformation : "jw", r, f_err)) return B200VQE_EINVAL;
    for (int i = 0; i < n_qubits; ++i) rows[i] = r[i];
    return 0;
}

}  // extern "C"
