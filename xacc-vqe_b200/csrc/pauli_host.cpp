// Host-side Pauli algebra, FermionCompiler text parser, JW / BK and the UCCSD generator.
// See pauli_host.hpp for what each function stands in for.
#include "pauli_host.hpp"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <set>
#include <sstream>

namespace b200vqe {

static inline cplx ipow(int k) {
    switch (((k % 4) + 4) % 4) {
    case 0: return {1, 0};
    case 1: return {0, 1};
    case 2: return {-1, 0};
    default: return {0, -1};
    }
}

// (xa,za)*(xb,zb) per qubit: XY=iZ, YZ=iX, ZX=iY and the reversed products with -i.
static inline int product_phase(uint64_t xa, uint64_t za, uint64_t xb, uint64_t zb) {
    const uint64_t plus = (xa & ~za & xb & zb) | (xa & za & ~xb & zb) | (~xa & za & xb & ~zb);
    const uint64_t minus = (xb & ~zb & xa & za) | (xb & zb & ~xa & za) | (~xb & zb & xa & ~za);
    return __builtin_popcountll(plus) - __builtin_popcountll(minus);
}

PauliOp PauliOp::operator*(const PauliOp &o) const {
    PauliOp r;
    std::map<TermKey, cplx> acc;
    for (auto &a : terms)
        for (auto &b : o.terms) {
            TermKey k{a.first.var + b.first.var, a.first.x ^ b.first.x, a.first.z ^ b.first.z};
            cplx c = a.second * b.second * ipow(product_phase(a.first.x, a.first.z, b.first.x, b.first.z));
            acc[k] += c;
        }
    for (auto &kv : acc)
        if (std::abs(kv.second) >= kDropTol) r.terms.emplace(kv.first, kv.second);
    return r;
}

static char pauli_char(const TermKey &k, int q) {
    const bool x = (k.x >> q) & 1, z = (k.z >> q) & 1;
    return x ? (z ? 'Y' : 'X') : (z ? 'Z' : 'I');
}

std::string term_id(const TermKey &k) {
    std::string s = k.var;
    const uint64_t m = k.x | k.z;
    for (int q = 0; q < 64; ++q)
        if ((m >> q) & 1) {
            s += pauli_char(k, q);
            s += std::to_string(q);
        }
    return s.empty() ? "I" : s;
}

std::string to_string(const PauliOp &op) {
    std::vector<std::pair<std::string, const std::pair<const TermKey, cplx> *>> v;
    for (auto &kv : op.terms) v.emplace_back(term_id(kv.first), &kv);
    std::sort(v.begin(), v.end(), [](auto &a, auto &b) { return a.first < b.first; });
    std::string out;
    char buf[96];
    for (size_t i = 0; i < v.size(); ++i) {
        const TermKey &k = v[i].second->first;
        const cplx c = v[i].second->second;
        snprintf(buf, sizeof buf, "(%.17g,%.17g)", c.real(), c.imag());
        if (i) out += " + ";
        out += buf;
        if (!k.var.empty()) out += " " + k.var;
        const uint64_t m = k.x | k.z;
        if (!m) out += " I";
        for (int q = 0; q < 64; ++q)
            if ((m >> q) & 1) {
                out += ' ';
                out += pauli_char(k, q);
                out += std::to_string(q);
            }
    }
    return out;
}

// "(re,im) X0 Z1 + (re,im) I" -- PauliOperator::fromString as used at
// compiler/optimizers/tests/QubitTaperingTester.cpp:72,149,156
bool from_string(const std::string &text, PauliOp &out, std::string &err) {
    out.terms.clear();
    size_t pos = 0;
    const size_t n = text.size();
    auto skip_ws = [&]() { while (pos < n && isspace((unsigned char)text[pos])) ++pos; };
    while (true) {
        skip_ws();
        if (pos >= n) break;
        if (text[pos] == '+') { ++pos; continue; }
        if (text[pos] != '(') { err = "expected '(' at offset " + std::to_string(pos); return false; }
        const size_t close = text.find(')', pos);
        const size_t comma = text.find(',', pos);
        if (close == std::string::npos || comma == std::string::npos || comma > close) {
            err = "malformed coefficient at offset " + std::to_string(pos);
            return false;
        }
        char *e1 = nullptr, *e2 = nullptr;
        const std::string sre = text.substr(pos + 1, comma - pos - 1), sim = text.substr(comma + 1, close - comma - 1);
        const double re = strtod(sre.c_str(), &e1), im = strtod(sim.c_str(), &e2);
        if (e1 == sre.c_str() || e2 == sim.c_str()) { err = "bad number in coefficient '" + sre + "," + sim + "'"; return false; }
        pos = close + 1;
        TermKey k;
        while (true) {
            skip_ws();
            if (pos >= n || text[pos] == '+') break;
            const char p = text[pos];
            if (p == 'I' && (pos + 1 >= n || !isdigit((unsigned char)text[pos + 1]))) { ++pos; continue; }
            if (p != 'X' && p != 'Y' && p != 'Z' && p != 'I') { err = std::string("unexpected character '") + p + "' at offset " + std::to_string(pos); return false; }
            ++pos;
            size_t d = pos;
            while (d < n && isdigit((unsigned char)text[d])) ++d;
            if (d == pos) { err = "missing qubit index at offset " + std::to_string(pos); return false; }
            const int q = atoi(text.substr(pos, d - pos).c_str());
            if (q < 0 || q > 63) { err = "qubit index out of range (0..63): " + std::to_string(q); return false; }
            pos = d;
            // repeated site multiplies (keeps the algebra closed)
            PauliOp cur = PauliOp::single(k.x, k.z, 1.0);
            uint64_t x = 0, z = 0;
            if (p == 'X' || p == 'Y') x = 1ull << q;
            if (p == 'Z' || p == 'Y') z = 1ull << q;
            if (((k.x | k.z) >> q) & 1) { err = "site " + std::to_string(q) + " repeated in one term"; return false; }
            k.x |= x;
            k.z |= z;
        }
        out.add_term(k, cplx(re, im));
    }
    return true;
}

bool parse_fermion_source(const std::string &src, std::vector<FermionTerm> &terms, int &n_qubits, std::string &err) {
    std::vector<std::string> lines;
    {
        std::string cur;
        for (char c : src) {
            if (c == '\n') { lines.push_back(cur); cur.clear(); } else cur += c;
        }
        lines.push_back(cur);
    }
    if (lines.size() < 2) { err = "fermion source needs a '__qpu__ name() {' line and a closing line"; return false; }
    terms.clear();
    int nq = 0;
    for (size_t li = 1; li + 1 < lines.size(); ++li) {
        std::istringstream ss(lines[li]);
        std::vector<std::string> tok;
        std::string t;
        while (ss >> t) tok.push_back(t);
        if (tok.empty()) continue;
        if (lines[li].find_first_of("0123456789") == std::string::npos) continue;
        char *e = nullptr;
        const double c = strtod(tok[0].c_str(), &e);
        if (e == tok[0].c_str()) { err = "bad coefficient '" + tok[0] + "' on line " + std::to_string(li + 1); return false; }
        FermionTerm ft;
        ft.coeff = c;
        for (size_t i = 1; i + 1 < tok.size(); i += 2) {   // FermionCompiler.cpp:84
            char *e1 = nullptr, *e2 = nullptr;
            const long site = strtol(tok[i].c_str(), &e1, 10), flag = strtol(tok[i + 1].c_str(), &e2, 10);
            if (e1 == tok[i].c_str() || e2 == tok[i + 1].c_str() || site < 0 || site > 63) {
                err = "bad site/flag '" + tok[i] + " " + tok[i + 1] + "' on line " + std::to_string(li + 1);
                return false;
            }
            nq = std::max(nq, (int)site);
            ft.sites.emplace_back((int)site, (int)flag);
        }
        terms.push_back(ft);
    }
    n_qubits = nq + 1;
    return true;
}

PauliOp jordan_wigner(const std::vector<FermionTerm> &terms) {
    PauliOp result;
    for (auto &f : terms) {
        PauliOp cur = PauliOp::identity(f.coeff, f.var);
        for (auto &s : f.sites) {
            const int idx = s.first;
            const cplx yc = s.second ? cplx(0, -.5) : cplx(0, .5);
            const uint64_t zs = idx ? ((idx >= 64 ? 0 : (1ull << idx)) - 1) : 0;
            PauliOp ladder = PauliOp::single(1ull << idx, 0, cplx(.5, 0));
            ladder += PauliOp::single(1ull << idx, 1ull << idx, yc);
            cur = cur * (PauliOp::single(0, zs, 1.0) * ladder);
        }
        result += cur;
    }
    return result;
}

namespace {
// transformations/bk/Fenwick.hpp:83-176
struct Fenwick {
    std::vector<int> parent;
    std::vector<std::set<int>> children;
    explicit Fenwick(int n) : parent(n, -1), children(n) {
        if (n > 0) build(0, n - 1, n - 1);
    }
    void build(int i1, int i2, int par) {
        if (i1 >= i2) return;
        const int pvt = (i1 + i2) >> 1;
        children[par].insert(pvt);
        parent[pvt] = par;
        build(i1, pvt, pvt);
        build(pvt + 1, i2, par);
    }
    uint64_t update_set(int i) const {
        uint64_t m = 0;
        while (parent[i] >= 0) { m |= 1ull << parent[i]; i = parent[i]; }
        return m;
    }
    uint64_t children_set(int i) const {
        uint64_t m = 0;
        for (int c : children[i]) m |= 1ull << c;
        return m;
    }
    uint64_t remainder_set(int i) const {
        uint64_t m = 0;
        int a = i;
        while (parent[a] >= 0) {
            a = parent[a];
            for (int c : children[a]) if (c < i) m |= 1ull << c;
        }
        return m;
    }
    uint64_t parity_set(int i) const { return remainder_set(i) | children_set(i); }
};
}  // namespace

PauliOp bravyi_kitaev(const std::vector<FermionTerm> &terms, int n_qubits) {
    Fenwick tree(n_qubits);
    PauliOp result;
    for (auto &f : terms) {
        PauliOp prod = PauliOp::identity(f.coeff, f.var);
        for (auto &s : f.sites) {
            const int idx = s.first;
            const cplx dc = s.second ? cplx(0, -.5) : cplx(0, .5);
            const uint64_t U = tree.update_set(idx), R = tree.remainder_set(idx), P = tree.parity_set(idx);
            const uint64_t bit = 1ull << idx;
            // c = 1/2 X_j X_U Z_P ; d = -/+ i/2 Y_j X_U Z_R   (BravyiKitaevIRTransformation.cpp:46-71;
            // later map assignments win: update set overrides parity/remainder entries)
            PauliOp cd = PauliOp::single(bit | U, P & ~U, cplx(.5, 0));
            cd += PauliOp::single(bit | U, bit | (R & ~U), dc);
            prod = prod * cd;
        }
        result += prod;
    }
    return result;
}

std::vector<PackedTerm> pack(const PauliOp &op) {
    std::vector<PackedTerm> v;
    for (auto &kv : op.terms) {
        const int ny = __builtin_popcountll(kv.first.x & kv.first.z);
        v.push_back(PackedTerm{kv.first.x, kv.first.z, kv.second * ipow(ny), kv.first});
    }
    std::stable_sort(v.begin(), v.end(), [](const PackedTerm &a, const PackedTerm &b) {
        return std::tie(a.x, a.z, a.key.var) < std::tie(b.x, b.z, b.key.var);
    });
    return v;
}

bool uccsd_rotations(int n_qubits, int n_electrons, const std::string &transformation,
                     std::vector<Rotation> &rot, int &n_params, std::string &err) {
    if (n_qubits < 2 || n_qubits > 64 || n_electrons < 1) { err = "UCCSD needs n-qubits >= 2 and n-electrons >= 1"; return false; }
    const int nocc = (int)std::ceil(n_electrons / 2.0);
    const int nvirt = n_qubits / 2 - nocc;
    if (nvirt < 1) { err = "UCCSD: no virtual orbitals for (n-qubits, n-electrons) = (" + std::to_string(n_qubits) + ", " + std::to_string(n_electrons) + ")"; return false; }
    const int nsingle = nocc * nvirt, ndouble = nsingle * (nsingle + 1) / 2;
    n_params = nsingle + ndouble;
    auto pname = [](int i) { return "theta" + std::to_string(i); };
    auto fs = [](int s, int i) { return s == 0 ? 2 * i : 2 * i + 1; };
    std::vector<FermionTerm> ft;
    auto add = [&](double c, std::vector<std::pair<int, int>> sites, int param) {
        ft.push_back(FermionTerm{cplx(c, 0), std::move(sites), pname(param)});
    };
    int count = 0;
    for (int i = 0; i < nvirt; ++i)
        for (int j = 0; j < nocc; ++j) {
            const int vs = nocc + i, os = j;
            for (int s = 0; s < 2; ++s) {
                const int vt = fs(s, vs), vo = fs(1 - s, vs), ot = fs(s, os), oo = fs(1 - s, os);
                add(1.0, {{vt, 1}, {ot, 0}}, count);
                add(-1.0, {{ot, 1}, {vt, 0}}, count);
                add(-1.0, {{vt, 1}, {ot, 0}, {vo, 1}, {oo, 0}}, nsingle + count);
                add(1.0, {{oo, 1}, {vo, 0}, {ot, 1}, {vt, 0}}, nsingle + count);
            }
            ++count;
        }
    std::vector<std::pair<int, int>> tup;
    for (int i = 0; i < nvirt; ++i) for (int j = 0; j < nocc; ++j) tup.emplace_back(i, j);
    count = 0;
    for (size_t a = 0; a < tup.size(); ++a)
        for (size_t b = a + 1; b < tup.size(); ++b) {
            const int vs1 = nocc + tup[a].first, os1 = tup[a].second, vs2 = nocc + tup[b].first, os2 = tup[b].second;
            for (int sa = 0; sa < 2; ++sa)
                for (int sb = 0; sb < 2; ++sb) {
                    const int v1a = fs(sa, vs1), o1a = fs(sa, os1), v2b = fs(sb, vs2), o2b = fs(sb, os2);
                    add(-1.0, {{v1a, 1}, {o1a, 0}, {v2b, 1}, {o2b, 0}}, 2 * nsingle + count);
                    // literal: UCCSD.cpp:210 (op6 is not the h.c. of op5)
                    add(1.0, {{o2b, 1}, {o1a, 0}, {v2b, 1}, {o2b, 0}}, 2 * nsingle + count);
                }
            ++count;
        }
    PauliOp gen;
    if (transformation == "jw" || transformation.empty()) gen = jordan_wigner(ft);
    else if (transformation == "bk") gen = bravyi_kitaev(ft, n_qubits);
    else { err = "unknown fermion-transformation '" + transformation + "'"; return false; }

    // fixed order: ascending term id, then CommutingSetGenerator's grouping (incl. its quirk of
    // testing against allTerms[innerCounter], CommutingSetGenerator.hpp:72-77)
    std::vector<std::pair<std::string, const std::pair<const TermKey, cplx> *>> all;
    for (auto &kv : gen.terms) all.emplace_back(term_id(kv.first), &kv);
    std::sort(all.begin(), all.end(), [](auto &a, auto &b) { return a.first < b.first; });
    auto anticommute = [](const TermKey &a, const TermKey &b) {
        return (__builtin_popcountll(a.x & b.z) + __builtin_popcountll(a.z & b.x)) & 1;
    };
    std::vector<std::vector<int>> sets;
    for (int i = 0; i < (int)all.size(); ++i) {
        if (i == 0) { sets.push_back({0}); continue; }
        bool placed = false;
        for (auto &s : sets) {
            int sum = 0;
            for (int inner = 0; inner < (int)s.size(); ++inner)
                sum += anticommute(all[i].second->first, all[inner].second->first);
            if (sum == 0) { s.push_back(i); placed = true; break; }
        }
        if (!placed) sets.push_back({i});
    }
    rot.clear();
    for (auto &s : sets)
        for (int i : s) {
            const TermKey &k = all[i].second->first;
            if (!(k.x | k.z)) continue;
            rot.push_back(Rotation{k.x, k.z, all[i].second->second.imag(), atoi(k.var.c_str() + 5)});
        }
    return true;
}

// Occupation read-out of a basis index for the diag-number-symmetry branch (task/tasks/DiagonalizeTask.cpp:54-144).
// The reference enumerates occupation strings with n_electrons ones (:58-72) and, for "bk", re-encodes each with the
// GF(2) matrix B built at :76-121 (Seeley's recursion :80-104, bottom-right n x n block :98-103, then rows and
// columns reversed :108-121): y = B x mod 2 (:125-141), string position i <-> qubit i.  The device needs the converse:
// is basis index j (bit q = y_q) in the sector?  x = B^-1 y, i.e. x_i = parity(rows[i] & j), rows = B^-1 over GF(2).
bool number_sector_rows(int n, const std::string &transformation, std::vector<uint64_t> &rows, std::string &err) {
    if (n < 1 || n > 63) { err = "n_qubits must be in [1, 63]"; return false; }
    rows.assign(n, 0);
    if (transformation != "bk") {            // "jw" and everything else: the index is the occupation vector (:74 only tests "bk")
        for (int i = 0; i < n; ++i) rows[i] = 1ull << i;
        return true;
    }
    // Seeley matrix by doubling: old block on the diagonal quadrants, ones in the first row of the upper-right quadrant
    int dim = 1;
    std::vector<std::vector<uint8_t>> B(1, std::vector<uint8_t>(1, 1));
    do {
        std::vector<std::vector<uint8_t>> nb(2 * dim, std::vector<uint8_t>(2 * dim, 0));
        for (int r = 0; r < dim; ++r)
            for (int c = 0; c < dim; ++c) nb[r][c] = nb[dim + r][dim + c] = B[r][c];
        for (int c = 0; c < dim; ++c) nb[0][dim + c] = 1;
        B.swap(nb);
        dim *= 2;
    } while (dim < n);
    const int off = dim - n;                 // bottom-right n x n block (no-op when dim == n)
    // reversing rows and columns (the reference's swap loops terminate only for even n; odd n gets the intended full reversal)
    std::vector<uint64_t> Bt(n, 0);          // Bt[i] bit k = B'(i, k), y_i = parity(Bt[i] & x)
    for (int i = 0; i < n; ++i)
        for (int k = 0; k < n; ++k)
            if (B[off + (n - 1 - i)][off + (n - 1 - k)]) Bt[i] |= 1ull << k;
    // Gauss-Jordan inverse over GF(2)
    std::vector<uint64_t> inv(n);
    for (int i = 0; i < n; ++i) inv[i] = 1ull << i;
    for (int c = 0; c < n; ++c) {
        int p = -1;
        for (int r = c; r < n; ++r) if ((Bt[r] >> c) & 1) { p = r; break; }
        if (p < 0) { err = "BK matrix is singular"; return false; }
        std::swap(Bt[c], Bt[p]);
        std::swap(inv[c], inv[p]);
        for (int r = 0; r < n; ++r)
            if (r != c && ((Bt[r] >> c) & 1)) { Bt[r] ^= Bt[c]; inv[r] ^= inv[c]; }
    }
    rows = inv;                              // x_i = parity(inv[i] & y)
    return true;
}


}  // namespace b200vqe
