// Host-side integer model of element-cyclic distributions on an r x c process grid.
// Restates (as closed forms, 64-bit) the reference's definitions:
//   Shift_/Length_/MaxLength_/Mod_   include/El/core/indexing/impl.hpp:33-66,229-245
//   grid rank maps                   src/core/Grid.cpp:139-150
//   per-distribution stride / rank   src/core/DistMatrix/ElementMatrix/*.cpp (bottom of each file)
//   view alignment                   include/El/core/DistMatrix/... View: align = (align + i) mod stride
// No CUDA in this header: it is also used by the CPU-only C-ABI entry points.
#pragma once
#include <stdint.h>
#include <algorithm>
#include <vector>

namespace hyd {

typedef long long i64;

enum Dist { MC = 0, MD = 1, MR = 2, VC = 3, VR = 4, STAR = 5, CIRC = 6 };  // types.hpp:208-216

inline i64 mod_(i64 a, i64 b) { i64 r = a % b; return r >= 0 ? r : r + b; }
inline i64 shift_(i64 rank, i64 align, i64 stride) { return mod_(rank - align, stride); }
inline i64 length_(i64 n, i64 shift, i64 stride) { return n > shift ? (n - shift - 1) / stride + 1 : 0; }
inline i64 max_length_(i64 n, i64 stride) { return n > 0 ? (n - 1) / stride + 1 : 0; }
inline i64 gcd_(i64 a, i64 b) { while (b) { i64 t = a % b; a = b; b = t; } return a; }

struct GridShape {
    int r = 1, c = 1;      // Height(), Width()
    int col_major = 1;     // GridOrder COLUMN_MAJOR (default)
    int size() const { return r * c; }
    // world (owning-comm) rank <-> (mc, mr); Grid.cpp:121-148
    void coords(int rank, int& mc, int& mr) const {
        if (col_major) { mc = rank % r; mr = rank / r; }
        else           { mc = rank / c; mr = rank % c; }
    }
    int rank_of(int mc, int mr) const { return col_major ? mc + r * mr : mr + c * mc; }
    int vc(int mc, int mr) const { return mc + r * mr; }
    int vr(int mc, int mr) const { return mr + c * mc; }
    int stride(int d) const {
        switch (d) { case MC: return r; case MR: return c; case VC: case VR: return r * c; default: return 1; }
    }
    int dist_rank(int d, int mc, int mr) const {
        switch (d) { case MC: return mc; case MR: return mr; case VC: return vc(mc, mr); case VR: return vr(mc, mr); default: return 0; }
    }
};

inline bool dist_supported(int d) { return d == MC || d == MR || d == VC || d == VR || d == STAR; }
// legal (colDist,rowDist) pairs among the GPU-valid element distributions
inline bool pair_supported(int u, int v) {
    if (!dist_supported(u) || !dist_supported(v)) return false;
    if (u == STAR || v == STAR) return true;
    return (u == MC && v == MR) || (u == MR && v == MC);
}

// Partial(U) (types.hpp:299-307) and the alignment an unconstrained dimension adopts in B = A when it goes src -> dst: the
// alignment argument of the reference's copy:: primitive for that pair (Copy/Translate.hpp:38-42, RowAllGather.hpp:23,
// Partial{Col,Row}AllGather.hpp, {Col,Row}AllToAll{Promote,Demote}.hpp, *Filter.hpp); 0 where no primitive passes one.
inline int partial_dist(int d) { return d == VC ? MC : (d == VR ? MR : d); }
inline int adopted_align(const GridShape& g, int src, int srcAlign, int dst) {
    if (dst == STAR || dst == CIRC || src == STAR || src == CIRC) return 0;
    if (src == dst || partial_dist(dst) == src) return srcAlign;
    if (partial_dist(src) == dst) return srcAlign % g.stride(dst);
    return 0;
}

// GEMM_DEFAULT: which SUMMA variant El::Gemm picks from the shape (Gemm/NN.hpp:903-925, identical in NT/TN/TT.hpp):
// 1 = A, 2 = B, 3 = C, 4 = Dot (the GemmAlgorithm enumerators).
inline int gemm_default_alg(i64 m, i64 n, i64 k) {
    if (10.0 * (double)m <= (double)k && 10.0 * (double)n <= (double)k) return 4;
    if (m <= n && 2.0 * (double)m <= (double)k) return 2;
    if (n <= m && 2.0 * (double)n <= (double)k) return 1;
    return 3;
}

// Metadata of one distributed matrix (the POD the C ABI passes, include/hydrogen_b200.h hyd_dm).
struct DM {
    void* buf = nullptr;
    i64 ldim = 0, height = 0, width = 0;
    int colDist = MC, rowDist = MR, colAlign = 0, rowAlign = 0, root = 0;
};

struct LocalShape { i64 colShift, rowShift, lh, lw; int colStride, rowStride; };

inline LocalShape local_shape(const GridShape& g, const DM& a, int mc, int mr) {
    LocalShape s;
    s.colStride = g.stride(a.colDist);
    s.rowStride = g.stride(a.rowDist);
    s.colShift = shift_(g.dist_rank(a.colDist, mc, mr), a.colAlign, s.colStride);
    s.rowShift = shift_(g.dist_rank(a.rowDist, mc, mr), a.rowAlign, s.rowStride);
    s.lh = length_(a.height, s.colShift, s.colStride);
    s.lw = length_(a.width, s.rowShift, s.rowStride);
    return s;
}

// Row cuts of one column block [j0, j0+w) of a local [MC,MR]-like block for the trapezoidal update (summa.cu local_trrk):
// rows [d0,d1) lie in the `tri` triangle (0 LOWER: gi >= gj, 1 UPPER: gi <= gj) for EVERY column of the block, rows [t0,t1) are
// the strip the diagonal may cross (computed into a temporary and added under a mask), every other row is outside for every
// column.  Cuts between the two ranges are rounded to q rows (16 bytes) towards the strip.
struct TrrkCuts { i64 d0, d1, t0, t1; };
inline TrrkCuts trrk_cuts(int tri, const LocalShape& L, i64 j0, i64 w, i64 q) {
    const i64 r = L.colStride, c = L.rowStride, lh = L.lh;
    auto clip = [&](i64 x) { return x < 0 ? 0 : (x > lh ? lh : x); };
    auto ceil_div = [](i64 a, i64 b) { return a >= 0 ? (a + b - 1) / b : -((-a) / b); };
    auto floor_div = [](i64 a, i64 b) { return a >= 0 ? a / b : -((-a + b - 1) / b); };
    const i64 gj_first = L.rowShift + j0 * c, gj_last = L.rowShift + (j0 + w - 1) * c;
    TrrkCuts k;
    if (tri == 0) {
        k.t0 = clip(ceil_div(gj_first - L.colShift, r)) / q * q;
        const i64 up = (clip(ceil_div(gj_last - L.colShift, r)) + q - 1) / q * q;
        k.t1 = up < lh ? up : lh;
        k.d0 = k.t1; k.d1 = lh;
    } else {
        k.d0 = 0;
        k.d1 = clip(floor_div(gj_first - L.colShift, r) + 1) / q * q;
        k.t0 = k.d1;
        k.t1 = clip(floor_div(gj_last - L.colShift, r) + 1);
    }
    return k;
}

// Column blocks of the local C for the host-resident stationary-C Gemm (summa.cu): C is computed block by block so that a
// finished block downloads under the next block's GEMMs.  Cuts are in 128-column tiles of the LOCAL block (tm x tn tiles).
// Model: a block of w tile columns costs ceil(tm*w/sms) waves of the persistent GEMM and lambda*w waves of download.
// Plans are restricted to those whose every download hides under the NEXT block's GEMMs (lambda*w_j <= waves(w_{j+1})), so
// only the last block's download is exposed; among them minimise   sum_j waves(w_j) + lambda*w_last + mu*(blocks-1).
// Widths therefore shrink geometrically towards the end.  Example: tm = tn = 128, 148 SMs, lambda = 0.17 -> 100|23|4|1.
// Dynamic programme from the right end over (columns left, width of the block to the right); returns the cuts
// 0 = t_0 < ... < t_nblk = tn (at most max_blocks <= 8 blocks; one block if nothing better exists).
inline int host_col_cuts(i64 tm, i64 tn, i64 sms, double lambda, double mu, int max_blocks, i64* cuts /*[max_blocks+1]*/) {
    cuts[0] = 0;
    cuts[1] = tn > 0 ? tn : 0;
    if (tn <= 1 || tm <= 0 || sms <= 0 || max_blocks <= 1 || !(lambda > 0)) return 1;
    if (max_blocks > 8) max_blocks = 8;
    // the plan of the previous call (per thread): a Gemm in a loop asks for the same one every time (the table costs ~10 ms)
    struct Memo { i64 tm, tn, sms; double lambda, mu; int max_blocks, nb; i64 cuts[9]; };
    static thread_local Memo memo = {0, 0, 0, 0, 0, 0, 0, {0}};
    if (memo.nb > 0 && memo.tm == tm && memo.tn == tn && memo.sms == sms && memo.lambda == lambda && memo.mu == mu &&
        memo.max_blocks == max_blocks) {
        for (int i = 0; i <= memo.nb; ++i) cuts[i] = memo.cuts[i];
        return memo.nb;
    }
    auto remember = [&](int nb) {
        memo.tm = tm; memo.tn = tn; memo.sms = sms; memo.lambda = lambda; memo.mu = mu; memo.max_blocks = max_blocks; memo.nb = nb;
        for (int i = 0; i <= nb; ++i) memo.cuts[i] = cuts[i];
        return nb;
    };
    const i64 unit = (tn + 191) / 192;                    // coarsen very wide blocks: the table stays <= 192 x 192 x 8
    const i64 n = (tn + unit - 1) / unit;
    auto waves = [&](i64 w) { return (double)((tm * w * unit + sms - 1) / sms); };
    const double lam = lambda * (double)unit, INF = 1e300;
    // best[b][t][w]: least cost of covering [0, t) with at most b more blocks when the block to the right of t has width w
    static thread_local std::vector<double> best;
    static thread_local std::vector<short> pick;
    const size_t plane = (size_t)(n + 1) * (size_t)(n + 1);
    best.assign(plane * (size_t)max_blocks, INF);
    pick.assign(plane * (size_t)max_blocks, 0);
    auto at = [&](int b, i64 t, i64 w) { return (size_t)b * plane + (size_t)t * (size_t)(n + 1) + (size_t)w; };
    for (i64 w = 1; w <= n; ++w) best[at(0, 0, w)] = 0;   // nothing left: zero further blocks
    for (int b = 1; b < max_blocks; ++b)
        for (i64 t = 0; t <= n; ++t)
            for (i64 w = 1; w + t <= n; ++w) {
                if (t == 0) { best[at(b, 0, w)] = 0; continue; }
                const i64 cap = std::min<i64>(t, (i64)(waves(w) / lam));   // widths whose download hides under the block to the right
                double bc = INF;
                short bp = 0;
                for (i64 v = 1; v <= cap; ++v) {
                    const double sub = best[at(b - 1, t - v, v)];
                    if (sub >= INF) continue;
                    const double c = waves(v) + mu + sub;
                    if (c < bc) { bc = c; bp = (short)v; }
                }
                best[at(b, t, w)] = bc;
                pick[at(b, t, w)] = bp;
            }
    double bc = waves(n) + lam * (double)n;               // one block: everything exposed
    i64 wl = n;
    for (i64 w = 1; w < n; ++w) {
        const double sub = best[at(max_blocks - 1, n - w, w)];
        if (sub >= INF) continue;
        const double c = waves(w) + lam * (double)w + sub;
        if (c < bc - 1e-9) { bc = c; wl = w; }
    }
    if (wl == n) return remember(1);
    i64 rev[9];
    int nb = 0;
    rev[nb++] = wl;
    i64 t = n - wl, w = wl;
    for (int b = max_blocks - 1; t > 0 && b >= 1; --b) {
        const i64 v = pick[at(b, t, w)];
        if (v <= 0) return 1;                             // cannot happen for a finite cost
        rev[nb++] = v;
        t -= v;
        w = v;
    }
    if (t != 0) return 1;
    i64 run = 0;
    for (int i = 0; i < nb; ++i) { run += rev[nb - 1 - i] * unit; cuts[i + 1] = std::min(run, tn); }
    cuts[nb] = tn;
    return remember(nb);
}

// One arithmetic progression of global indices i = first + t*step, t < count,
// together with where it sits in the source's and the destination's local index space.
struct Prog { i64 count = 0, src0 = 0, srcStep = 1, dst0 = 0, dstStep = 1; };

// Intersect { i ≡ s1 (mod n1) } and { i ≡ s2 (mod n2) } on [0, n).  (CRT; n1, n2 small.)
inline Prog intersect(i64 n, i64 s1, i64 n1, i64 s2, i64 n2) {
    Prog p;
    const i64 g = gcd_(n1, n2);
    if (mod_(s1 - s2, g) != 0) return p;
    const i64 L = n1 / g * n2;
    i64 x = s1;
    while (mod_(x - s2, n2) != 0) x += n1;  // at most n2/g steps
    p.count = length_(n, x, L);
    p.src0 = (x - s1) / n1; p.srcStep = L / n1;
    p.dst0 = (x - s2) / n2; p.dstStep = L / n2;
    return p;
}

// Which coordinates of a rank does distribution pair (U,V) leave undetermined?
// (free coordinate == the data is replicated, or -- on the reduce side -- partially summed, along it)
inline void free_coords(int u, int v, bool& mcFree, bool& mrFree) {
    auto uses_mc = [](int d) { return d == MC || d == VC || d == VR; };
    auto uses_mr = [](int d) { return d == MR || d == VC || d == VR; };
    mcFree = !(uses_mc(u) || uses_mc(v));
    mrFree = !(uses_mr(u) || uses_mr(v));
}

}  // namespace hyd
