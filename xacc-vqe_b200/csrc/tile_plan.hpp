// Host-side planner of the tile engine (kernels_tile.cuh).
//
// Idea.  <psi|P|psi> for a Pauli string with flip mask x pairs amplitude j with j^x.  A CTA that
// holds in shared memory a set of amplitudes closed under j -> j^x for SEVERAL masks x_1..x_d can
// evaluate all of those xmask groups from one HBM read of the set.  The closed sets are the cosets
// of the GF(2) subspace S = span(x_1 >> C, ...) of the high index bits; each coset is 2^M chunks of
// 2^C contiguous amplitudes (M = dim S, tile = 2^(M+C) amplitudes).  Inside a tile the amplitudes
// form a "virtual register" of M+C qubits: chunk coordinate mu (bit i <-> basis vector i) and the
// low C index bits; group x acts there as the virtual mask xv = (mu(x) << C) | (x & (2^C-1)), and
// the Z parity splits into a per-coset part, parity(rep & z_hi), and a virtual part parity(v & zv)
// with zv = (Zv << C) | (z & (2^C-1)),  Zv bit i = parity(basis_i & z_hi).
//
// The planner partitions the xmask groups of a Hamiltonian into batches, one subspace each, and
// emits the device tables.  HBM traffic per group drops from 8 B (pair form of the streaming kernel)
// to 16/d B per amplitude, d = groups per batch (>= M for unstructured Hamiltonians, far more for
// molecular ones whose flip masks span a small space).
//
// Device tables: TileBatch (subspace basis, coset-representative masks, ranges), TileTerm (packed z parts + coefficient)
// and, for the kernels, TileRec records (build_tile_records): one per (group, even/odd nY), sorted by the key that
// selects the kernel's specialised loop.  TileGroup records remain for the natural-layout sweep kernel and the CPU
// replay tests.  Plans for the kernels are built with per-batch capacities (kStageGroupsCap, kStageTerms) so that a
// batch's records and terms always fit the kernels' shared-memory staging.
#pragma once
#include <algorithm>
#include <cstdint>
#include <unordered_map>
#include <vector>

namespace b200vqe {

constexpr int kTileBits = 12;        // largest tile = 4096 amplitudes = 64 KB of double2 (H psi, sweep, single-buffer <psi|H|psi>)
constexpr int kTileRegBits = 4;      // 16 amplitudes per thread (virtual bits 8..11)
constexpr int kTileBitsSmall = 11;   // 2048-amplitude tiles, 8 per thread: the double-buffered <psi|H|psi> kernel
constexpr int kTileThreadBits = 8;   // 256 threads per CTA: virtual bits 0..7 of a tile index are the thread
constexpr int kTileMaxDim = 10;      // M = kTileBits - C, C >= 2

constexpr int kTileClasses = 5;      // groups of a batch are ordered by hsel = 0, 1, 2, 3, -1

struct TileBatch {                   // 128 bytes
    uint32_t basis[kTileMaxDim];     // basis[i] <-> virtual bit C + i; acts on the high index (j >> C), n - C <= 32
    uint32_t seg[kTileMaxDim + 1];   // coset representative of tile t:  rep = OR_i (t & seg[i]) << i  (zeros at the pivots)
    int32_t g_begin, g_end;          // groups of this batch
    int32_t t_begin, t_end;          // terms of this batch
    int32_t cls_end[kTileClasses];   // end of each hsel class, relative to g_begin
    int32_t r_begin, r_end;          // TileRec range of this batch (expectation engine; see build_tile_records)
};
static_assert(sizeof(TileBatch) == 128, "TileBatch layout");

struct TileGroup {                   // 16 bytes
    uint32_t kb;                     // first term
    uint32_t n_even;                 // terms with even nY (use Re w), then n_odd with odd nY (use Im w)
    uint32_t n_odd;
    uint16_t xv;                     // virtual flip mask (kTileBits bits)
    int16_t hsel;                    // register bit (0..3) selecting one element per pair, -1: visit all
};
static_assert(sizeof(TileGroup) == 16, "TileGroup layout");

struct TileTerm {                    // 16 bytes
    uint64_t packed;                 // bits 0..31 z_hi | 32..39 zv & 0xff (thread bits) | 40..43 zv >> 8 (register bits)
                                     // | 48..63 sign word: bit r = parity(r & (zv >> 8)), r = 0..15
    double u;                        // weight * (coef | i*coef): real for a Hermitian H
};
static_assert(sizeof(TileTerm) == 16, "TileTerm layout");

// One record of the expectation kernel: the even-nY or the odd-nY terms of one xmask group, first term inline.
// Records of a batch are sorted by key, so the device runs one straight-line, fully specialised loop per key:
//   key = which * 16 + (xr ? xr - 1 : 15),  xr = xv >> 8 (register bits of the virtual flip mask), which = 0 even nY / 1 odd nY
struct TileRec {                     // 32 bytes
    uint32_t kb;                     // first term (relative to the batch's t_begin once staged)
    uint32_t count;                  // terms of this record (>= 1)
    uint16_t xv;
    uint8_t key;
    uint8_t pad0;
    uint32_t pad1;
    uint64_t packed0;                // terms[kb].packed
    double u0;                       // terms[kb].u
};
static_assert(sizeof(TileRec) == 32, "TileRec layout");

constexpr int kTileRecKeys = 32;
constexpr int kStageRecs = 128;      // records / terms of one batch staged in shared memory by k_tile_expect;
constexpr int kStageTerms = 256;     // expectation plans are built with these as per-batch capacities
constexpr int kStageGroupsCap = kStageRecs / 2;

struct TilePlanTermIn { uint64_t x, z; double u; int odd; int64_t idx; };

struct TilePlan {
    int n = 0, C = 0, M = 0, TB = kTileBits;
    std::vector<TileBatch> batches;
    std::vector<TileGroup> groups;
    std::vector<TileTerm> terms;
    std::vector<int64_t> term_src;   // plan term -> index into the input array
    std::vector<TileRec> recs;       // expectation engine only (build_tile_records)
};

// Which tiles (cosets) of every batch a rank of a sharded register evaluates in <psi|H|psi>: a contiguous 1/world share;
// registers with fewer tiles than ranks are evaluated by rank 0 alone.  The union over ranks is every tile exactly once.
inline void shard_tile_range(int n, int rank, int world, uint64_t &begin, uint64_t &count, int tile_bits = kTileBits) {
    const uint64_t all = 1ull << (n - tile_bits);
    const uint64_t per = all / (uint64_t)world;
    if (per) { begin = per * (uint64_t)rank; count = per; }
    else { begin = 0; count = rank == 0 ? all : 0; }
}

namespace tileplan {

inline int parity64h(uint64_t v) { return __builtin_popcountll(v) & 1; }

// reduced basis with unique pivots, pivot = highest set bit of each vector
struct Rref {
    std::vector<uint64_t> b;
    std::vector<int> piv;
    uint64_t reduce(uint64_t v) const {
        for (size_t i = 0; i < b.size(); ++i) if ((v >> piv[i]) & 1) v ^= b[i];
        return v;
    }
    bool insert(uint64_t v) {
        v = reduce(v);
        if (!v) return false;
        const int p = 63 - __builtin_clzll(v);
        for (size_t i = 0; i < b.size(); ++i) if ((b[i] >> p) & 1) b[i] ^= v;
        b.push_back(v);
        piv.push_back(p);
        return true;
    }
    uint64_t coords(uint64_t v) const {   // valid when reduce(v) == 0
        uint64_t mu = 0;
        for (size_t i = 0; i < b.size(); ++i) mu |= ((v >> piv[i]) & 1ull) << i;
        return mu;
    }
};

inline uint16_t sign_word16(unsigned zr) {   // bit r = parity(r & zr), r = 0..15
    uint16_t w = 0;
    for (unsigned r = 0; r < 16; ++r) w |= (uint16_t)((__builtin_popcount(r & zr) & 1) << r);
    return w;
}

}  // namespace tileplan

// terms: non-identity terms with u already set (coef or i*coef, real part).  Groups of equal x.
// pair_weights: expectation plans double the coefficients of groups evaluated once per pair; apply plans do not.
// natural (n == kTileBits only): one batch whose virtual index equals the amplitude index (basis = unit vectors in order),
// for kernels that keep the whole register in shared memory in its natural order (k_sweep_small).
// max_groups / max_terms (0 = unlimited): capacity of one batch; an xmask group with more terms than max_terms is cut
// into several groups of the same x.
inline TilePlan build_tile_plan(int n, int C, const std::vector<TilePlanTermIn> &in, bool pair_weights = true, bool natural = false,
                                size_t max_groups = 0, size_t max_terms = 0, int tile_bits = kTileBits) {
    using namespace tileplan;
    const int TB = tile_bits, RB = tile_bits - kTileThreadBits;      // tile bits, register bits (4 or 3)
    TilePlan P;
    P.n = n; P.C = C; P.M = TB - C; P.TB = TB;
    const int M = P.M;
    const int nh = n - C;                         // bits of the high index
    const uint64_t lomask = (1ull << C) - 1;

    // ---- groups by x (sorted) ----
    std::vector<int64_t> order(in.size());
    for (size_t i = 0; i < in.size(); ++i) order[i] = (int64_t)i;
    std::stable_sort(order.begin(), order.end(), [&](int64_t a, int64_t b) {
        if (in[a].x != in[b].x) return in[a].x < in[b].x;
        if (in[a].odd != in[b].odd) return in[a].odd < in[b].odd;
        return in[a].z < in[b].z;
    });
    struct G { uint64_t x; size_t tb, te; };
    std::vector<G> gs;
    for (size_t i = 0; i < order.size();) {
        size_t j = i;
        while (j < order.size() && in[order[j]].x == in[order[i]].x && (!max_terms || j - i < max_terms)) ++j;
        gs.push_back(G{in[order[i]].x, i, j});
        i = j;
    }
    const size_t NG = gs.size();
    std::unordered_map<uint64_t, std::vector<size_t>> by_hi;
    by_hi.reserve(NG * 2);
    for (size_t g = 0; g < NG; ++g) by_hi[gs[g].x >> C].push_back(g);
    std::vector<char> done(NG, 0);
    size_t n_done = 0, cursor = 0;

    while (n_done < NG) {
        while (done[cursor]) ++cursor;
        // ---- choose the subspace: greedy over the next unassigned groups ----
        Rref rr;
        size_t scanned = 0;
        if (natural) {
            // unit vectors inserted from the top bit down stay unit vectors: basis index i <-> high bit nh-1-i
            for (int p = nh - 1; p >= 0; --p) rr.insert(1ull << p);
        }
        for (size_t g = cursor; g < NG && (int)rr.b.size() < M && scanned < 8192; ++g) {
            if (done[g]) continue;
            ++scanned;
            rr.insert(gs[g].x >> C);
        }
        for (int p = 0; p < nh && (int)rr.b.size() < M; ++p) rr.insert(1ull << p);   // pad with unit vectors
        // ---- members: every unassigned group whose high mask lies in the span ----
        std::vector<size_t> members;
        const size_t remaining = NG - n_done;
        if (remaining <= (1ull << rr.b.size())) {
            for (size_t g = cursor; g < NG; ++g)
                if (!done[g] && rr.reduce(gs[g].x >> C) == 0) members.push_back(g);
        } else {
            for (uint64_t m = 0; m < (1ull << rr.b.size()); ++m) {
                uint64_t v = 0;
                for (size_t i = 0; i < rr.b.size(); ++i) if ((m >> i) & 1) v ^= rr.b[i];
                auto it = by_hi.find(v);
                if (it == by_hi.end()) continue;
                for (size_t g : it->second) if (!done[g]) members.push_back(g);
            }
            std::sort(members.begin(), members.end());
        }
        if (max_groups || max_terms) {             // batch capacity: the rest stays unassigned for later batches
            size_t keep = 0, nt = 0;
            for (; keep < members.size(); ++keep) {
                const size_t t = gs[members[keep]].te - gs[members[keep]].tb;
                if ((max_groups && keep >= max_groups) || (max_terms && nt + t > max_terms)) break;
                nt += t;
            }
            members.resize(keep);
        }
        // ---- order the basis: the RB coordinates that cover most work go on the register bits ----
        const int dim = (int)rr.b.size();
        std::vector<uint64_t> mu(members.size());
        std::vector<double> cost(members.size());
        for (size_t k = 0; k < members.size(); ++k) {
            mu[k] = rr.coords(gs[members[k]].x >> C);
            cost[k] = 4.0 + (double)(gs[members[k]].te - gs[members[k]].tb);
        }
        std::vector<int> perm;                     // perm[new position] = old index
        std::vector<char> used(dim, 0), covered(members.size(), 0);
        std::vector<int> top;
        for (int pick = 0; pick < RB && pick < dim; ++pick) {
            int best = -1;
            double bestw = -1.0;
            for (int i = 0; i < dim; ++i) {
                if (used[i]) continue;
                double w = 0.0;
                for (size_t k = 0; k < members.size(); ++k) if (!covered[k] && ((mu[k] >> i) & 1)) w += cost[k];
                if (w > bestw) { bestw = w; best = i; }
            }
            used[best] = 1;
            top.push_back(best);
            for (size_t k = 0; k < members.size(); ++k) if ((mu[k] >> best) & 1) covered[k] = 1;
        }
        if (natural) {
            perm.clear();
            for (int i = 0; i < dim; ++i) perm.push_back(dim - 1 - i);     // virtual bit C+i <-> high bit i
        } else {
            for (int i = 0; i < dim; ++i) if (!used[i]) perm.push_back(i);
            for (int t : top) perm.push_back(t);       // register bits are the top RB coordinates
        }

        TileBatch B{};
        for (int i = 0; i < dim; ++i) B.basis[i] = (uint32_t)rr.b[perm[i]];
        std::vector<int> sp(rr.piv.begin(), rr.piv.end());
        std::sort(sp.begin(), sp.end());
        {   // tile index bits [lo_i, hi_i) move up by i = number of pivots below them
            int lo = 0;
            for (int i = 0; i <= dim; ++i) {
                const int hi = i < dim ? sp[i] - i : 32;
                uint64_t m = 0;
                if (hi > lo) m = ((hi >= 32 ? (1ull << 32) : (1ull << hi)) - 1) & ~((1ull << lo) - 1);
                B.seg[i] = (uint32_t)m;
                lo = hi;
            }
        }
        B.g_begin = (int32_t)P.groups.size();
        B.t_begin = (int32_t)P.terms.size();
        auto permute_mu = [&](uint64_t m) {
            uint64_t r = 0;
            for (int i = 0; i < dim; ++i) r |= ((m >> perm[i]) & 1ull) << i;
            return r;
        };
        // order the members by pair-selector class (hsel 0, 1, 2, 3, then -1): one code path per class on the device
        auto class_of = [&](size_t k) {
            const unsigned xr = (unsigned)(((permute_mu(mu[k]) << C) >> (TB - RB)) & ((1u << RB) - 1));
            return xr ? 31 - __builtin_clz(xr) : kTileClasses - 1;
        };
        std::vector<size_t> ord(members.size());
        for (size_t k = 0; k < ord.size(); ++k) ord[k] = k;
        std::stable_sort(ord.begin(), ord.end(), [&](size_t a, size_t b) { return class_of(a) < class_of(b); });
        for (int c = 0; c < kTileClasses; ++c) B.cls_end[c] = 0;
        for (size_t kk = 0; kk < members.size(); ++kk) {
            const size_t k = ord[kk];
            for (int c = class_of(k); c < kTileClasses; ++c) ++B.cls_end[c];
            const G &g = gs[members[k]];
            const uint64_t m = permute_mu(mu[k]);
            TileGroup tg{};
            tg.xv = (uint16_t)((m << C) | (g.x & lomask));
            const unsigned xr = tg.xv >> (TB - RB);
            tg.hsel = xr ? (int16_t)(31 - __builtin_clz(xr)) : (int16_t)-1;
            const double weight = tg.hsel >= 0 ? 2.0 : 1.0;
            tg.kb = (uint32_t)P.terms.size();
            for (size_t t = g.tb; t < g.te; ++t) {
                const TilePlanTermIn &ti = in[order[t]];
                const uint64_t zhi = ti.z >> C;
                uint64_t Zv = 0;
                for (int i = 0; i < dim; ++i) Zv |= (uint64_t)parity64h((uint64_t)B.basis[i] & zhi) << i;
                const unsigned zv = (unsigned)((Zv << C) | (ti.z & lomask));
                TileTerm tt;
                tt.packed = (zhi & 0xffffffffull) | ((uint64_t)(zv & 0xffu) << 32) |
                            ((uint64_t)((zv >> (TB - RB)) & 0xfu) << 40) |
                            ((uint64_t)sign_word16((zv >> (TB - RB)) & 0xfu) << 48);
                tt.u = (pair_weights ? weight : 1.0) * ti.u;
                P.terms.push_back(tt);
                P.term_src.push_back(ti.idx);
                ti.odd ? ++tg.n_odd : ++tg.n_even;
            }
            P.groups.push_back(tg);
            done[members[k]] = 1;
            ++n_done;
        }
        B.g_end = (int32_t)P.groups.size();
        B.t_end = (int32_t)P.terms.size();
        P.batches.push_back(B);
    }
    return P;
}

// Records of the expectation kernel for a plan built with pair_weights = true and the kStage* capacities.
inline void build_tile_records(TilePlan &P) {
    P.recs.clear();
    for (TileBatch &B : P.batches) {
        std::vector<TileRec> rs;
        for (int g = B.g_begin; g < B.g_end; ++g) {
            const TileGroup &tg = P.groups[g];
            const unsigned xr = tg.xv >> kTileThreadBits;
            for (int which = 0; which < 2; ++which) {
                const uint32_t cnt = which ? tg.n_odd : tg.n_even;
                if (!cnt) continue;
                TileRec r{};
                r.kb = tg.kb + (which ? tg.n_even : 0);
                r.count = cnt;
                r.xv = tg.xv;
                r.key = (uint8_t)(which * 16 + (xr ? xr - 1 : 15));
                r.packed0 = P.terms[r.kb].packed;
                r.u0 = P.terms[r.kb].u;
                rs.push_back(r);
            }
        }
        std::stable_sort(rs.begin(), rs.end(), [](const TileRec &a, const TileRec &b) { return a.key < b.key; });
        B.r_begin = (int32_t)P.recs.size();
        P.recs.insert(P.recs.end(), rs.begin(), rs.end());
        B.r_end = (int32_t)P.recs.size();
    }
}

}  // namespace b200vqe
