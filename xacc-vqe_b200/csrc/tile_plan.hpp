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
#include <cmath>
#include <cstdint>
#include <cstring>
#include <unordered_map>
#include <vector>

namespace b200vqe {

constexpr int kTileBits = 12;        // largest tile = 4096 amplitudes = 64 KB of double2 (H psi, sweep, single-buffer <psi|H|psi>)
constexpr int kTileRegBits = 4;      // 16 amplitudes per thread (virtual bits 8..11)
constexpr int kTileBitsSmall = 11;   // 2048-amplitude tiles, 8 per thread: the double-buffered <psi|H|psi> kernel
constexpr int kTileThreadBits = 8;   // 256 threads per CTA: virtual bits 0..7 of a tile index are the thread
constexpr int kTileMaxDim = 10;      // M = kTileBits - C, C >= 2

constexpr int kTileClasses = 5;      // groups of a batch are ordered by hsel = 0, 1, 2, 3, -1
constexpr int kTileMaxViews = 6;     // register subspaces ("views") of one batch, view 0 included (expectation plans, TilePlanOpts::views)

struct TileBatch {                   // 320 bytes
    uint32_t basis[kTileMaxDim];     // basis[i] <-> virtual bit C + i; acts on the high index (j >> C), n - C <= 32
    uint32_t seg[kTileMaxDim + 1];   // coset representative of tile t:  rep = OR_i (t & seg[i]) << i  (zeros at the pivots)
    int32_t g_begin, g_end;          // groups of this batch
    int32_t t_begin, t_end;          // terms of this batch
    int32_t cls_end[kTileClasses];   // end of each hsel class, relative to g_begin
    int32_t r_begin, r_end;          // TileRec range of this batch (expectation engine; see build_tile_records)
    uint32_t lam[4];                 // shear: register r of a thread holds the amplitude with low index bits lo ^ Lambda(r),
                                     // Lambda(r) = XOR of lam[k] over the set bits k of r (0 for plans built without shear)
    uint32_t pad[4];                 // pad[0] = pivot mask of the subspace (bits of the high index the tile number skips)
    // Views.  The tile sits in shared memory in the layout of view 0: element (thread t0, register r0) at byte (r0 << 12) | (t0 << 4).
    // In view k a thread keeps 16 OTHER amplitudes of the tile in its registers: (thread t, register r) of view k is the element at
    // byte  XOR_{i in t} vcol[k][i]  ^  XOR_{j in r} vcol[k][8 + j]  (a GF(2)-linear map; view 0: the identity, vcol[0][i] = 16 << i),
    // chosen so that the flip masks of four more xmask groups lie in the register subspace (pure-register records, no partner
    // reads).  The thread bits that carry the low index bits of a chunk are the same in every view, so the 8 lanes of a
    // quarter-warp still read the 8 different 16-byte columns of one 128-byte line (no bank conflicts).
    uint16_t vcol[kTileMaxViews][12];
    uint32_t n_views;                // >= 1
    uint32_t pad2[3];
};
constexpr int kTileBatchWords = 80;
static_assert(sizeof(TileBatch) == 4 * kTileBatchWords, "TileBatch layout");

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
    uint32_t kb;                     // first term, relative to the batch's t_begin
    uint32_t count;                  // terms of this record (>= 1)
    uint16_t xv;
    uint8_t key;
    uint8_t pad0;
    uint32_t view;                   // the view this record is evaluated in (0 unless the plan has views); a record with
                                     // key kTileRecKeyView is a marker: "load the registers of view `view`", no terms
    uint64_t packed0;                // terms[kb].packed
    double u0;                       // terms[kb].u
};
static_assert(sizeof(TileRec) == 32, "TileRec layout");

// One slot of a FUSED VIEW BLOCK (64 bytes = two entries of the record stream).  The four register bits of a view are the flip
// masks of four member groups ("generators": XR = 1, 2, 4, 8); when those groups have one term each, the pipelined kernel
// evaluates the four of them in one straight-line block right after loading the view's registers -- no per-record dispatch,
// four independent dependency chains -- from these slots.  Slot j: XR = 1 << j, pairs {r, r | 1 << j}; the first NE slots of a
// block hold even-nY terms (Re w), the others odd-nY terms (Im w) or nothing (u = 0).
struct TileFusedSlot {
    uint32_t zl;                     // z on the high index bits
    uint32_t zt;                     // thread bits of z in the view's coordinates
    double u;                        // 2 * coefficient; 0: empty slot
    double s0, s1;                   // +-1.0: (-1)^{z bit} of the three register bits other than j, in ascending order (s2 below)
    uint32_t kb;                     // second half, laid out like a TileRec: the plan term (relative to the batch's t_begin),
    uint32_t pad;
    uint16_t xv; uint8_t key; uint8_t pad0;      // key = kTileRecKeyFused
    uint32_t view;
    double s2;
    uint64_t pad2;
};
static_assert(sizeof(TileFusedSlot) == 2 * sizeof(TileRec), "TileFusedSlot layout");

constexpr int kTileRecKeys = 32;      // (which, XR) combinations; + 32: pure-register records (expectation engine)
constexpr int kTileRecKeyFused = 0xfd; // the entries of a fused view block (never dispatched: the block is skipped as a whole)
constexpr int kTileRecKeyView = 0xfe; // marker record in front of the records of a view: TileRec::view = the view; count = 0: no fused block,
                                      // 1 + NE: a fused block of 4 TileFusedSlot follows, its first NE slots hold even-nY terms
constexpr int kTileRecKeyEnd = 0xff;  // sentinel the kernels append after the last record of a batch
constexpr int kStageRecs = 128;      // records / terms of one batch staged in shared memory by k_tile_expect (+ kTileMaxViews markers);
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
    std::vector<uint8_t> group_view; // per group: the view its xv / its terms' z parts are expressed in (all 0 without views)
    int max_views = 1;               // largest TileBatch::n_views of the plan
    bool fused = false;              // the record stream has view markers in front of EVERY view and fused view blocks
    // what one pass over the plan costs per tile (build_tile_records): register loads (views), fused blocks, plain records and
    // those of them that read partner rows from shared memory
    int64_t n_view_loads = 0, n_fused_blocks = 0, n_plain_recs = 0, n_partner_recs = 0;
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

// How a plan is built.
//   pair_weights: expectation plans double the coefficients of groups evaluated once per pair; apply plans do not.
//   natural (n == tile_bits only): one batch whose virtual index equals the amplitude index (basis = unit vectors in
//     order), for kernels that keep the whole register in shared memory in its natural order (k_sweep_small).
//   max_groups / max_terms (0 = unlimited): capacity of one batch; an xmask group with more terms than max_terms is cut
//     into several groups of the same x.
//   collide: choose every new basis vector of a batch's subspace as the coset (modulo the vectors chosen so far) that
//     holds the most unassigned groups, instead of the next unassigned group: random masks then give ~12 instead of
//     ~9 groups per 9-dimensional batch, structured ones collapse as before.
//   shear: choose the virtual coordinate system so that the register subspace (the RB virtual bits a thread keeps in its
//     own registers) is spanned by the full virtual masks of RB member groups -- their partner amplitude then sits in
//     the SAME thread (no shared-memory read: "pure" records) -- which needs the low-index shear TileBatch::lam, and
//     tilt the remaining coordinates so that the other groups have a register component (8 instead of 16 partner reads).
struct TilePlanOpts {
    bool pair_weights = true;
    bool natural = false;
    size_t max_groups = 0, max_terms = 0;
    int tile_bits = kTileBits;
    bool collide = false;
    bool shear = false;
    int views = 1;                   // (shear only) up to this many register subspaces per batch: after view 0, further views are
                                     // chosen greedily, each spanned by the full virtual masks of RB members that are not pure yet;
                                     // a view is opened only when it saves more partner reads than loading its registers costs
};

namespace tileplan {

// rows of the inverse of the M x M matrix over GF(2) whose COLUMN i is col[i]:  coordinate i of c = parity(inv[i] & c)
inline std::vector<uint64_t> invert_columns(const std::vector<uint64_t> &col, int M) {
    std::vector<uint64_t> row(M, 0), inv(M, 0);
    for (int r = 0; r < M; ++r) {
        for (int i = 0; i < M; ++i) row[r] |= ((col[i] >> r) & 1ull) << i;
        inv[r] = 1ull << r;
    }
    for (int c = 0; c < M; ++c) {
        int p = -1;
        for (int r = c; r < M; ++r) if ((row[r] >> c) & 1) { p = r; break; }
        if (p < 0) return {};                      // singular: the caller passes a basis
        std::swap(row[c], row[p]);
        std::swap(inv[c], inv[p]);
        for (int r = 0; r < M; ++r)
            if (r != c && ((row[r] >> c) & 1)) { row[r] ^= row[c]; inv[r] ^= inv[c]; }
    }
    return inv;
}

// counts equal keys of one pass over at most 65536 positions (open addressing, stamped instead of cleared)
struct KeyCounter {
    struct Slot { uint64_t key; uint32_t stamp; uint16_t cnt; uint16_t first; };
    std::vector<Slot> t;
    uint32_t now = 0;
    size_t mask = 0;
    void reserve(size_t n) {
        size_t cap = 16;
        while (cap < 2 * n) cap <<= 1;
        if (cap > t.size()) { t.assign(cap, Slot{0, 0, 0, 0}); now = 0; }
        mask = t.size() - 1;
    }
    void begin() { ++now; }
    // counts `key`; returns its slot (first = position of its first occurrence)
    const Slot &add(uint64_t key, uint32_t pos) {
        size_t h = (size_t)((key * 0x9E3779B97F4A7C15ull) >> 24) & mask;
        for (;;) {
            Slot &s = t[h];
            if (s.stamp != now) { s.stamp = now; s.key = key; s.cnt = 1; s.first = (uint16_t)pos; return s; }
            if (s.key == key) { ++s.cnt; return s; }
            h = (h + 1) & mask;
        }
    }
};

}  // namespace tileplan

// Tile number -> coset representative.  The subspace of a batch has one pivot bit per basis vector; the representatives of
// its cosets are the high indices with zeros at the pivots, and tile t is the representative whose remaining bits spell t:
// tile index bits [lo_i, hi_i) move up by i = number of pivots below them (TileBatch::seg), pad[0] = the pivot mask.
inline void set_tile_segments(TileBatch &B, const std::vector<int> &pivots) {
    std::vector<int> sp(pivots.begin(), pivots.end());
    std::sort(sp.begin(), sp.end());
    const int dim = (int)sp.size();
    int lob = 0;
    for (int i = 0; i <= kTileMaxDim; ++i) B.seg[i] = 0;
    for (int i = 0; i <= dim; ++i) {
        const int hib = i < dim ? sp[i] - i : 32;
        uint64_t m = 0;
        if (hib > lob) m = ((hib >= 32 ? (1ull << 32) : (1ull << hib)) - 1) & ~((1ull << lob) - 1);
        B.seg[i] = (uint32_t)m;
        lob = hib;
    }
    B.pad[0] = 0;
    for (int p : pivots) B.pad[0] |= 1u << p;
}

// terms: non-identity terms with u already set (coef or i*coef, real part).  Groups of equal x.
inline TilePlan build_tile_plan(int n, int C, const std::vector<TilePlanTermIn> &in, const TilePlanOpts &o) {
    using namespace tileplan;
    const int TB = o.tile_bits, RB = o.tile_bits - kTileThreadBits;      // tile bits, register bits (4 or 3)
    const bool natural = o.natural;
    const size_t max_groups = o.max_groups, max_terms = o.max_terms;
    TilePlan P;
    P.n = n; P.C = C; P.M = TB - C; P.TB = TB;
    const int M = P.M;
    const int TD = M - RB;                        // thread coordinates (virtual bits C .. C+TD-1); may be 0
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
    struct G { uint64_t x; size_t tb, te; int recs; };
    std::vector<G> gs;
    for (size_t i = 0; i < order.size();) {
        size_t j = i;
        while (j < order.size() && in[order[j]].x == in[order[i]].x && (!max_terms || j - i < max_terms)) ++j;
        const int recs = (in[order[i]].odd != in[order[j - 1]].odd) ? 2 : 1;     // even-nY and odd-nY terms are separate records
        gs.push_back(G{in[order[i]].x, i, j, recs});
        i = j;
    }
    const size_t NG = gs.size();
    std::unordered_map<uint64_t, std::vector<uint32_t>> by_hi;
    by_hi.reserve(NG * 2);
    for (size_t g = 0; g < NG; ++g) by_hi[gs[g].x >> C].push_back((uint32_t)g);
    std::vector<char> done(NG, 0);
    size_t n_done = 0, cursor = 0;
    // the subspace of a batch is chosen among the next kWindow unassigned groups (bounds the cost per batch); its members
    // are then collected from ALL unassigned groups
    constexpr size_t kWindow = 4096;
    std::vector<uint32_t> win;
    std::vector<uint64_t> wkey;                   // high mask of a window group reduced modulo the subspace under construction
    KeyCounter kc;
    kc.reserve(kWindow);

    while (n_done < NG) {
        while (done[cursor]) ++cursor;
        win.clear();
        wkey.clear();
        for (size_t g = cursor; g < NG && win.size() < kWindow; ++g)
            if (!done[g]) { win.push_back((uint32_t)g); wkey.push_back(gs[g].x >> C); }
        // ---- choose the subspace ----
        Rref rr;
        auto add = [&](uint64_t v) {              // v: reduced, non-zero
            const int p = 63 - __builtin_clzll(v);
            rr.insert(v);
            uint64_t *k = wkey.data();
            const size_t nw = wkey.size();
            for (size_t i = 0; i < nw; ++i) k[i] ^= v & (0ull - ((k[i] >> p) & 1ull));
        };
        if (natural) {
            // unit vectors inserted from the top bit down stay unit vectors: basis index i <-> high bit nh-1-i
            for (int p = nh - 1; p >= 0; --p) add(1ull << p);
        }
        size_t scan = 0;
        while ((int)rr.b.size() < M) {
            uint64_t pick = 0;
            const int d = (int)rr.b.size();
            // expected number of window pairs that agree modulo the current span; below ~1/2 a counting pass finds nothing
            const double pairs = 0.5 * (double)win.size() * (double)win.size() * std::ldexp(1.0, d - nh);
            uint64_t pick2 = 0;                   // runner-up of a counting pass, when it holds a collision too
            if (o.collide && d >= 1 && pairs >= 0.5) {
                // the coset holding the most window groups (ties: the one met first)
                kc.begin();
                uint32_t best = 0, best_first = 0, second = 0;
                for (size_t i = 0; i < wkey.size(); ++i) {
                    const uint64_t k = wkey[i];
                    if (!k) continue;
                    const KeyCounter::Slot &sl = kc.add(k, (uint32_t)i);
                    if (k == pick) best = sl.cnt;
                    else if (sl.cnt > best || (sl.cnt == best && sl.first < best_first)) {
                        second = best; pick2 = pick;
                        best = sl.cnt; best_first = sl.first; pick = k;
                    } else if (k == pick2) second = sl.cnt;
                    else if (sl.cnt > second) { second = sl.cnt; pick2 = k; }
                }
                if (second < 2) pick2 = 0;
            } else {
                while (scan < wkey.size() && !wkey[scan]) ++scan;
                if (scan < wkey.size()) pick = wkey[scan];
            }
            if (!pick) break;                     // every window group already lies in the span
            add(pick);
            if (pick2 && (int)rr.b.size() < M) {
                const uint64_t v = rr.reduce(pick2);
                if (v) add(v);
            }
        }
        for (int p = 0; p < nh && (int)rr.b.size() < M; ++p) {   // pad with unit vectors
            const uint64_t v = rr.reduce(1ull << p);
            if (v) add(v);
        }
        // ---- members: every unassigned group whose high mask lies in the span, in ascending order ----
        std::vector<size_t> members;
        if (NG - n_done <= (1ull << rr.b.size())) {
            for (size_t g = cursor; g < NG; ++g)
                if (!done[g] && rr.reduce(gs[g].x >> C) == 0) members.push_back(g);
        } else {
            for (uint64_t m = 0; m < (1ull << rr.b.size()); ++m) {
                uint64_t v = 0;
                for (size_t i = 0; i < rr.b.size(); ++i) if ((m >> i) & 1) v ^= rr.b[i];
                auto it = by_hi.find(v);
                if (it == by_hi.end()) continue;
                for (uint32_t g : it->second) if (!done[g]) members.push_back(g);
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
        for (size_t g : members) { done[g] = 1; ++n_done; }
        const int dim = (int)rr.b.size();          // == M (n >= tile_bits)
        const size_t NM = members.size();
        std::vector<uint64_t> cc(NM);              // coordinates in the reduced basis rr.b
        std::vector<uint32_t> lo(NM);
        for (size_t k = 0; k < NM; ++k) { cc[k] = rr.coords(gs[members[k]].x >> C); lo[k] = (uint32_t)(gs[members[k]].x & lomask); }

        // 0: one even-nY term, 1: one odd-nY term, 2: anything else
        auto fuse_type = [&](int k) {
            const G &g = gs[members[k]];
            return g.te - g.tb == 1 ? (in[order[g.tb]].odd ? 1 : 0) : 2;
        };
        // ---- virtual coordinates: newb[i] (in rr coordinates) <-> virtual bit C + i; the top RB are the register bits ----
        std::vector<uint64_t> newb;
        uint32_t lam[4] = {0, 0, 0, 0};
        if (natural) {
            for (int i = 0; i < dim; ++i) newb.push_back(1ull << (dim - 1 - i));      // virtual bit C+i <-> high bit i
        } else if (!o.shear) {
            // the RB reduced-basis coordinates that cover most work go on the register bits
            std::vector<double> cost(NM);
            for (size_t k = 0; k < NM; ++k) cost[k] = 4.0 + (double)(gs[members[k]].te - gs[members[k]].tb);
            std::vector<char> used(dim, 0), covered(NM, 0);
            std::vector<int> top;
            for (int pick = 0; pick < RB && pick < dim; ++pick) {
                int best = -1;
                double bestw = -1.0;
                for (int i = 0; i < dim; ++i) {
                    if (used[i]) continue;
                    double w = 0.0;
                    for (size_t k = 0; k < NM; ++k) if (!covered[k] && ((cc[k] >> i) & 1)) w += cost[k];
                    if (w > bestw) { bestw = w; best = i; }
                }
                used[best] = 1;
                top.push_back(best);
                for (size_t k = 0; k < NM; ++k) if ((cc[k] >> best) & 1) covered[k] = 1;
            }
            for (int i = 0; i < dim; ++i) if (!used[i]) newb.push_back(1ull << i);
            for (int t : top) newb.push_back(1ull << t);
        } else {
            // register subspace R = span of the full virtual masks V = (cc << C | lo) of RB members with independent cc,
            // chosen greedily for the number of member records inside the span (those become pure-register records)
            std::vector<uint64_t> V(NM);
            for (size_t k = 0; k < NM; ++k) V[k] = (cc[k] << C) | lo[k];
            Rref rc, rv;
            std::vector<uint64_t> rho;
            std::vector<int> picked;
            for (int pick = 0; pick < RB; ++pick) {
                int best = -1, best_score = -1;
                for (size_t k = 0; k < NM; ++k) {
                    if (!cc[k] || !rc.reduce(cc[k])) continue;
                    Rref t = rv;
                    t.insert(V[k]);
                    int score = 0;
                    for (size_t m = 0; m < NM; ++m) if (cc[m] && !t.reduce(V[m])) score += gs[members[m]].recs;
                    if (score > best_score) { best_score = score; best = (int)k; }
                }
                if (best < 0) break;
                rv.insert(V[best]);
                rc.insert(cc[best]);
                picked.push_back(best);
            }
            // register bit j <-> j-th generator: single-term even-nY groups first, then single-term odd-nY, then the rest
            // (the fused view body of the pipelined kernel, build_tile_records)
            std::stable_sort(picked.begin(), picked.end(), [&](int a, int b) { return fuse_type(a) < fuse_type(b); });
            for (int k : picked) { lam[rho.size()] = lo[k]; rho.push_back(cc[k]); }
            for (int i = 0; i < dim && (int)rho.size() < RB; ++i)
                if (rc.reduce(1ull << i)) { rc.insert(1ull << i); rho.push_back(1ull << i); }
            std::vector<uint64_t> theta;
            for (int i = 0; i < dim && (int)theta.size() < TD; ++i)
                if (rc.reduce(1ull << i)) { rc.insert(1ull << i); theta.push_back(1ull << i); }
            newb = theta;
            newb.insert(newb.end(), rho.begin(), rho.end());
            // tilt: theta_j += sum_k s_jk rho_k changes the register part of a member with thread part t to a ^ XOR_{j in t} s_j
            const std::vector<uint64_t> inv = invert_columns(newb, dim);
            std::vector<uint32_t> tt(NM), aa(NM);
            for (size_t k = 0; k < NM; ++k) {
                uint64_t m = 0;
                for (int i = 0; i < dim; ++i) m |= (uint64_t)parity64h(inv[i] & cc[k]) << i;
                tt[k] = (uint32_t)(m & ((1ull << TD) - 1));
                aa[k] = (uint32_t)(m >> TD);
            }
            std::vector<uint32_t> s(TD, 0);
            auto total_cost = [&]() {
                int c = 0;
                for (size_t k = 0; k < NM; ++k) {
                    if (!tt[k]) continue;         // pure or register-only: not affected by the tilt
                    uint32_t a = aa[k];
                    for (int j = 0; j < TD; ++j) if ((tt[k] >> j) & 1) a ^= s[j];
                    c += (a ? 1 : 2) * gs[members[k]].recs;
                }
                return c;
            };
            for (int round = 0; round < 2; ++round)
                for (int j = 0; j < TD; ++j) {
                    int best_c = -1;
                    uint32_t best_s = 0;
                    for (uint32_t cand = 0; cand < (1u << RB); ++cand) {
                        s[j] = cand;
                        const int c = total_cost();
                        if (best_c < 0 || c < best_c) { best_c = c; best_s = cand; }
                    }
                    s[j] = best_s;
                }
            for (int j = 0; j < TD; ++j)
                for (int k = 0; k < RB; ++k) if ((s[j] >> k) & 1) newb[j] ^= rho[k];
        }

        // ---- views: (thread t, register r) <-> U-space vector, U = (coordinates in rr.b) << C | low index bits ----
        struct View { std::vector<uint64_t> col, inv; };
        auto make_view = [&](const std::vector<uint64_t> &thr_hi, const std::vector<uint64_t> &reg_u) {
            View v;
            for (int i = 0; i < C; ++i) v.col.push_back(1ull << i);                 // thread bits 0..C-1: the low index bits
            for (int j = 0; j < TD; ++j) v.col.push_back(thr_hi[j] << C);           // thread bits C..7: high-index directions
            for (int k = 0; k < RB; ++k) v.col.push_back(reg_u[k]);                 // register bits
            v.inv = invert_columns(v.col, TB);
            return v;
        };
        auto view_coords = [&](const View &v, uint64_t u) {
            uint64_t m = 0;
            for (int i = 0; i < TB; ++i) m |= (uint64_t)parity64h(v.inv[i] & u) << i;
            return m;
        };
        std::vector<View> views;
        {
            std::vector<uint64_t> thr0(newb.begin(), newb.begin() + TD), reg0;
            for (int k = 0; k < RB; ++k) reg0.push_back((newb[TD + k] << C) | lam[k]);
            views.push_back(make_view(thr0, reg0));
        }
        std::vector<uint64_t> Vu(NM);              // full virtual flip mask of every member
        std::vector<int> mview(NM, 0);
        std::vector<uint32_t> mxv(NM);             // ... in the coordinates of its view: thread part | register part << 8
        for (size_t k = 0; k < NM; ++k) { Vu[k] = (cc[k] << C) | lo[k]; mxv[k] = (uint32_t)view_coords(views[0], Vu[k]); }
        if (o.shear && o.views > 1) {
            constexpr int kViewMinGain = 24;       // partner reads (LDS.128 per thread) a view must save: loading its registers costs 16
            const uint32_t tmask = (1u << (TB - RB)) - 1;
            std::vector<char> open(NM);
            std::vector<int> wgt(NM);
            for (size_t k = 0; k < NM; ++k) {
                open[k] = (mxv[k] & tmask) != 0;
                wgt[k] = gs[members[k]].recs * ((mxv[k] >> (TB - RB)) ? (1 << (RB - 1)) : (1 << RB));
            }
            while ((int)views.size() < std::min(o.views, kTileMaxViews)) {
                Rref rc, rv;
                std::vector<uint64_t> reg_u;
                std::vector<int> picked;
                for (int pick = 0; pick < RB; ++pick) {
                    int best = -1, best_score = -1;
                    for (size_t k = 0; k < NM; ++k) {
                        if (!open[k] || !rc.reduce(cc[k])) continue;
                        Rref t = rv;
                        t.insert(Vu[k]);
                        int score = 0;
                        for (size_t m = 0; m < NM; ++m) if (open[m] && !t.reduce(Vu[m])) score += wgt[m];
                        if (score > best_score) { best_score = score; best = (int)k; }
                    }
                    if (best < 0) break;
                    rv.insert(Vu[best]);
                    rc.insert(cc[best]);
                    picked.push_back(best);
                }
                std::stable_sort(picked.begin(), picked.end(), [&](int a, int b) { return fuse_type(a) < fuse_type(b); });
                for (int k : picked) reg_u.push_back(Vu[k]);
                int gain = 0;
                for (size_t m = 0; m < NM; ++m) if (open[m] && !rv.reduce(Vu[m])) gain += wgt[m];
                if (reg_u.empty() || gain < kViewMinGain) break;
                for (int i = 0; i < dim && (int)reg_u.size() < RB; ++i)
                    if (rc.reduce(1ull << i)) { rc.insert(1ull << i); reg_u.push_back((1ull << i) << C); }
                std::vector<uint64_t> thr;
                for (int i = 0; i < dim && (int)thr.size() < TD; ++i)
                    if (rc.reduce(1ull << i)) { rc.insert(1ull << i); thr.push_back(1ull << i); }
                views.push_back(make_view(thr, reg_u));
                for (size_t m = 0; m < NM; ++m)
                    if (open[m] && !rv.reduce(Vu[m])) {
                        mview[m] = (int)views.size() - 1;
                        mxv[m] = (uint32_t)view_coords(views.back(), Vu[m]);      // thread part 0 by construction
                        open[m] = 0;
                    }
            }
        }

        TileBatch B{};
        for (int i = 0; i < dim; ++i) {
            uint64_t v = 0;
            for (int q = 0; q < dim; ++q) if ((newb[i] >> q) & 1) v ^= rr.b[q];
            B.basis[i] = (uint32_t)v;
        }
        for (int k = 0; k < 4; ++k) B.lam[k] = lam[k];
        B.n_views = (uint32_t)views.size();
        P.max_views = std::max(P.max_views, (int)views.size());
        for (size_t k = 0; k < views.size(); ++k)
            for (int i = 0; i < TB; ++i) B.vcol[k][i] = (uint16_t)(view_coords(views[0], views[k].col[i]) << 4);
        set_tile_segments(B, rr.piv);
        B.g_begin = (int32_t)P.groups.size();
        B.t_begin = (int32_t)P.terms.size();
        // order the members by pair-selector class (hsel 0, 1, 2, 3, then -1): one code path per class on the device
        auto class_of = [&](size_t k) {
            const unsigned xr = (unsigned)mxv[k] >> (TB - RB);
            return xr ? 31 - __builtin_clz(xr) : kTileClasses - 1;
        };
        std::vector<size_t> ord(NM);
        for (size_t k = 0; k < NM; ++k) ord[k] = k;
        std::stable_sort(ord.begin(), ord.end(), [&](size_t a, size_t b) { return class_of(a) < class_of(b); });
        for (int c = 0; c < kTileClasses; ++c) B.cls_end[c] = 0;
        std::vector<uint64_t> Zrr_basis(rr.b.begin(), rr.b.end());
        for (size_t kk = 0; kk < NM; ++kk) {
            const size_t k = ord[kk];
            for (int c = class_of(k); c < kTileClasses; ++c) ++B.cls_end[c];
            const G &g = gs[members[k]];
            const View &vw = views[mview[k]];
            TileGroup tg{};
            tg.xv = (uint16_t)mxv[k];
            const unsigned xr = tg.xv >> (TB - RB);
            tg.hsel = xr ? (int16_t)(31 - __builtin_clz(xr)) : (int16_t)-1;
            const double weight = tg.hsel >= 0 ? 2.0 : 1.0;
            tg.kb = (uint32_t)P.terms.size();
            for (size_t t = g.tb; t < g.te; ++t) {
                const TilePlanTermIn &ti = in[order[t]];
                const uint64_t zhi = ti.z >> C;
                // z in U space: coordinate i of the high part = parity(rr.b[i] & z_hi); then one bit per column of the view
                uint64_t zu = ti.z & lomask;
                for (int i = 0; i < dim; ++i) zu |= (uint64_t)parity64h(Zrr_basis[i] & zhi) << (C + i);
                unsigned zv = 0;
                for (int i = 0; i < TB; ++i) zv |= (unsigned)parity64h(vw.col[i] & zu) << i;
                TileTerm tt;
                tt.packed = (zhi & 0xffffffffull) | ((uint64_t)(zv & 0xffu) << 32) |
                            ((uint64_t)((zv >> (TB - RB)) & 0xfu) << 40) |
                            ((uint64_t)sign_word16((zv >> (TB - RB)) & 0xfu) << 48);
                tt.u = (o.pair_weights ? weight : 1.0) * ti.u;
                P.terms.push_back(tt);
                P.term_src.push_back(ti.idx);
                ti.odd ? ++tg.n_odd : ++tg.n_even;
            }
            P.groups.push_back(tg);
            P.group_view.push_back((uint8_t)mview[k]);
        }
        B.g_end = (int32_t)P.groups.size();
        B.t_end = (int32_t)P.terms.size();
        P.batches.push_back(B);
    }
    return P;
}

// the signature of the first version: reduced-basis coordinates, next-unassigned subspace choice
inline TilePlan build_tile_plan(int n, int C, const std::vector<TilePlanTermIn> &in, bool pair_weights = true, bool natural = false,
                                size_t max_groups = 0, size_t max_terms = 0, int tile_bits = kTileBits) {
    TilePlanOpts o;
    o.pair_weights = pair_weights; o.natural = natural; o.max_groups = max_groups; o.max_terms = max_terms; o.tile_bits = tile_bits;
    return build_tile_plan(n, C, in, o);
}

// The plans the backend builds (planner: 3 = 2 + several register subspaces (views) per batch [default for the pipelined <psi|H|psi>
// kernel], 2 = collision-greedy subspaces + sheared register subspace,
// 1 = the first version: next-unassigned subspaces, reduced-basis coordinates).
inline TilePlanOpts expect_plan_opts(int tile_bits, int planner = 2) {
    TilePlanOpts o;
    o.pair_weights = true; o.max_groups = kStageGroupsCap; o.max_terms = kStageTerms; o.tile_bits = tile_bits;
    o.collide = planner >= 2; o.shear = planner >= 2;
    o.views = (planner >= 3 && tile_bits == kTileBits) ? kTileMaxViews : 1;      // only the pipelined 4096-amplitude kernel walks views
    return o;
}
inline TilePlanOpts apply_plan_opts(int planner = 2) {
    TilePlanOpts o;
    o.pair_weights = false; o.max_groups = kStageGroupsCap; o.max_terms = kStageTerms; o.tile_bits = kTileBits;
    o.collide = planner >= 2; o.shear = false;       // H psi reads all 16 partner rows whatever the coordinates
    return o;
}

// Records of the tile kernels for a plan built with the kStage* capacities.  allow_pure (expectation plans built with
// shear): a record whose sheared flip mask has no thread part gets key + 32 -- the kernel then takes the partner from
// the thread's own registers.
// apply_keys (H psi plans, k_tile_apply_pipe8): key = which * 8 + (XR & 7); bit 3 of XR stays a run-time quantity (xv bit 11).
inline void build_tile_records(TilePlan &P, bool allow_pure = false, bool fuse = false, bool apply_keys = false) {
    P.recs.clear();
    P.fused = fuse;
    P.n_view_loads = P.n_fused_blocks = P.n_plain_recs = P.n_partner_recs = 0;
    constexpr unsigned tmask = (1u << kTileThreadBits) - 1;
    for (TileBatch &B : P.batches) {
        // fused view blocks: per view, the generator groups (xv = register bit j alone, one term)
        struct Block { int slot_group[4]; int ne; bool on; };
        std::vector<Block> blocks(fuse ? B.n_views : 0u);
        std::vector<char> in_block(B.g_end - B.g_begin, 0);
        for (size_t v = 0; v < blocks.size(); ++v) {
            Block &bl = blocks[v];
            int type[4];
            for (int j = 0; j < 4; ++j) { bl.slot_group[j] = -1; type[j] = 2; }
            for (int g = B.g_begin; g < B.g_end; ++g) {
                const TileGroup &tg = P.groups[g];
                if (P.group_view[g] != v || (tg.xv & tmask) || tg.n_even + tg.n_odd != 1) continue;
                const unsigned xr = tg.xv >> kTileThreadBits;
                if (xr == 0 || (xr & (xr - 1))) continue;
                const int j = 31 - __builtin_clz(xr);
                bl.slot_group[j] = g;
                type[j] = tg.n_even ? 0 : 1;
            }
            // the planner orders the generators even, odd, rest; a slot that breaks the order stays a plain record
            int n_used = 0;
            bl.ne = 0;
            for (int j = 0, prev = 0; j < 4; ++j) {
                if (type[j] < prev || type[j] == 2) { bl.slot_group[j] = -1; prev = 2; continue; }
                prev = type[j];
                ++n_used;
                if (type[j] == 0) bl.ne = j + 1;
            }
            bl.on = n_used >= 2;
            if (bl.on) for (int j = 0; j < 4; ++j) if (bl.slot_group[j] >= 0) in_block[bl.slot_group[j] - B.g_begin] = 1;
        }
        std::vector<TileRec> rs;
        for (int g = B.g_begin; g < B.g_end; ++g) {
            if (in_block[g - B.g_begin]) continue;
            const TileGroup &tg = P.groups[g];
            const unsigned xr = tg.xv >> kTileThreadBits;
            const uint32_t view = (size_t)g < P.group_view.size() ? P.group_view[g] : 0u;
            const bool pure = (allow_pure || view) && (tg.xv & tmask) == 0;
            for (int which = 0; which < 2; ++which) {
                const uint32_t cnt = which ? tg.n_odd : tg.n_even;
                if (!cnt) continue;
                TileRec r{};
                const uint32_t kb = tg.kb + (which ? tg.n_even : 0);
                r.kb = kb - (uint32_t)B.t_begin;
                r.count = cnt;
                r.xv = tg.xv;
                r.key = apply_keys ? (uint8_t)(which * 8 + (xr & 7)) : (uint8_t)((pure ? 32 : 0) + which * 16 + (xr ? xr - 1 : 15));
                r.view = view;
                r.packed0 = P.terms[kb].packed;
                r.u0 = P.terms[kb].u;
                rs.push_back(r);
            }
        }
        std::stable_sort(rs.begin(), rs.end(), [](const TileRec &a, const TileRec &b) { return a.view != b.view ? a.view < b.view : a.key < b.key; });
        B.r_begin = (int32_t)P.recs.size();
        ++P.n_view_loads;                       // view 0 of the batch
        P.n_plain_recs += (int64_t)rs.size();
        for (const TileRec &r : rs) if (r.key < kTileRecKeys || apply_keys) ++P.n_partner_recs;
        auto open_view = [&](uint32_t v) {      // marker (+ fused block): the kernel loads its registers in the layout of view v
            TileRec m{};
            m.key = (uint8_t)kTileRecKeyView;
            m.view = v;
            const bool on = fuse && blocks[v].on;
            m.count = on ? 1u + (uint32_t)blocks[v].ne : 0u;
            P.recs.push_back(m);
            if (v) ++P.n_view_loads;
            if (on) ++P.n_fused_blocks;
            if (!on) return;
            for (int j = 0; j < 4; ++j) {
                TileFusedSlot fs{};
                fs.key = (uint8_t)kTileRecKeyFused;
                fs.view = v;
                fs.s0 = fs.s1 = fs.s2 = 1.0;
                const int g = blocks[v].slot_group[j];
                if (g >= 0) {
                    const TileTerm &tt = P.terms[P.groups[g].kb];
                    fs.zl = (uint32_t)tt.packed;
                    fs.zt = (uint32_t)(tt.packed >> 32) & 0xffu;
                    fs.u = tt.u;
                    fs.kb = P.groups[g].kb - (uint32_t)B.t_begin;
                    fs.xv = P.groups[g].xv;
                    const unsigned zr = (unsigned)(tt.packed >> 40) & 0xfu;
                    double *sg[3] = {&fs.s0, &fs.s1, &fs.s2};
                    for (int k = 0; k < 3; ++k) if ((zr >> (k < j ? k : k + 1)) & 1u) *sg[k] = -1.0;
                }
                TileRec half[2];
                memcpy(half, &fs, sizeof(fs));
                P.recs.push_back(half[0]);
                P.recs.push_back(half[1]);
            }
        };
        uint32_t cur_view = 0;
        if (fuse) open_view(0);
        for (const TileRec &r : rs) {
            if (fuse) while (cur_view < r.view) open_view(++cur_view);
            else if (r.view != cur_view) { open_view(r.view); cur_view = r.view; }
            P.recs.push_back(r);
        }
        if (fuse) while (cur_view + 1 < B.n_views) open_view(++cur_view);     // views whose groups all sit in their fused block
        B.r_end = (int32_t)P.recs.size();
    }
}

}  // namespace b200vqe
