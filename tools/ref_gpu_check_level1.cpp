// ref_gpu_check_level1 -- second checker against the recompiled reference GPU layer (oracle/_ref/libref_gpu.so), for what
// tools/ref_gpu_check.cpp does not cover: the fp32 pack/unpack families, and the BLAS-1 layer around Gemm:
//   hydrogen::Fill_GPU_impl, Hadamard_GPU_impl, gpu_blas::Scale 2-D, gpu_blas::Copy(NORMAL) (cublas?geam),
//   gpu_blas::Axpy 2-D and gpu_blas::Axpy(TransposeMode, ...) (cublas?geam), in fp64 and fp32.
// Copies, Fill, Hadamard (one multiply) and Scale (one multiply) must be bit-identical; the axpy forms within 1 ulp of the
// larger addend (the reference goes through cublas?geam = alpha*A + 1*B, fused or not is cuBLAS's choice).
// STATUS: written after round 1's GPU budget was spent; has not run on a GPU yet (tools/gpu_first_call.sh runs it).
// Prints one JSON object; exit code 0 iff every check passed, 2 if the reference side could not run.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <string>
#include <vector>

#include "../include/hydrogen_b200.h"

#define CK(x)                                                                                             \
    do {                                                                                                  \
        cudaError_t e_ = (x);                                                                             \
        if (e_ != cudaSuccess) {                                                                          \
            std::fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
            std::exit(2);                                                                                 \
        }                                                                                                 \
    } while (0)

typedef int64_t i64;
static void* g_ref = nullptr;
static int g_fail = 0;
static std::string g_json;
template <typename F> static F sym(const std::string& name) {
    void* p = dlsym(g_ref, name.c_str());
    if (!p) { std::fprintf(stderr, "missing symbol %s\n", name.c_str()); std::exit(2); }
    return (F)p;
}
static void report(const std::string& what, bool ok, const std::string& detail = "") {
    if (!ok) ++g_fail;
    if (!g_json.empty()) g_json += ",\n ";
    g_json += "{\"check\": \"" + what + "\", \"ok\": " + (ok ? "true" : "false") + (detail.empty() ? "" : ", \"detail\": \"" + detail + "\"") + "}";
    std::fprintf(stderr, "%-64s %s %s\n", what.c_str(), ok ? "ok" : "FAILED", detail.c_str());
}
static const char* (*ref_err)() = nullptr;
static void R(int rc, const char* w) { if (rc) { std::fprintf(stderr, "reference %s: %s\n", w, ref_err()); std::exit(2); } }
static void H(int rc, const char* w) { if (rc) { std::fprintf(stderr, "product %s: %s\n", w, hyd_last_error()); std::exit(2); } }

template <typename T> struct Dev {
    T* p = nullptr;
    size_t n;
    explicit Dev(size_t n_) : n(n_) { CK(cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T))); }
    ~Dev() { cudaFree(p); }
    Dev(const Dev&) = delete;
    void set(const std::vector<T>& h) { CK(cudaMemcpy(p, h.data(), n * sizeof(T), cudaMemcpyHostToDevice)); }
    void index() { std::vector<T> h(n); for (size_t i = 0; i < n; ++i) h[i] = (T)(i % (1u << 24)); set(h); }
    void frac(int a, int b, int c, double d) { std::vector<T> h(n); for (size_t i = 0; i < n; ++i) h[i] = (T)((double)((i64)((i * a) % b) - c) / d); set(h); }
    void fill(T v) { std::vector<T> h(n, v); set(h); }
    std::vector<T> host() const { std::vector<T> h(n); CK(cudaMemcpy(h.data(), p, n * sizeof(T), cudaMemcpyDeviceToHost)); return h; }
};
template <typename T> static bool same(const Dev<T>& a, const Dev<T>& b) {
    CK(cudaDeviceSynchronize());
    auto x = a.host(), y = b.host();
    return x.size() == y.size() && std::memcmp(x.data(), y.data(), x.size() * sizeof(T)) == 0;
}
template <typename T> static bool within_ulp(const Dev<T>& a, const Dev<T>& b, double scale) {
    CK(cudaDeviceSynchronize());
    auto x = a.host(), y = b.host();
    const double eps = std::numeric_limits<T>::epsilon();
    for (size_t i = 0; i < x.size(); ++i)
        if (std::fabs((double)x[i] - (double)y[i]) > eps * std::max(scale, std::fabs((double)x[i]))) return false;
    return true;
}
static i64 ceil_div(i64 a, i64 b) { return (a + b - 1) / b; }
static i64 shift_(i64 r, i64 a, i64 s) { return ((r - a) % s + s) % s; }
static i64 length_(i64 n, i64 sh, i64 st) { return n > sh ? (n - sh - 1) / st + 1 : 0; }

// product entry points by element type
template <typename T> struct P;
#define HYD_P(T, sfx)                                                                                     \
    template <> struct P<T> {                                                                             \
        static constexpr const char* s = #sfx;                                                            \
        static constexpr auto fill = hyd_fill_##sfx; static constexpr auto scale2d = hyd_scale2d_##sfx;   \
        static constexpr auto hadamard = hyd_hadamard_##sfx; static constexpr auto copy2d = hyd_copy2d_##sfx; \
        static constexpr auto axpy2d = hyd_axpy2d_##sfx; static constexpr auto axpy_trans = hyd_axpy_trans_##sfx; \
        static constexpr auto interleave = hyd_interleave_##sfx;                                          \
        static constexpr auto row_pack = hyd_row_strided_pack_##sfx; static constexpr auto row_unpack = hyd_row_strided_unpack_##sfx; \
        static constexpr auto col_pack = hyd_col_strided_pack_##sfx; static constexpr auto col_unpack = hyd_col_strided_unpack_##sfx; \
        static constexpr auto prow_pack = hyd_partial_row_strided_pack_##sfx; static constexpr auto prow_unpack = hyd_partial_row_strided_unpack_##sfx; \
        static constexpr auto pcol_pack = hyd_partial_col_strided_pack_##sfx; static constexpr auto pcol_unpack = hyd_partial_col_strided_unpack_##sfx; \
        static constexpr auto pack2d = hyd_strided_pack_##sfx; static constexpr auto unpack2d = hyd_strided_unpack_##sfx; \
    };
HYD_P(double, f64)
HYD_P(float, f32)

struct Shape { i64 h, w; };
static const Shape kShapes[] = {{1, 1}, {7, 5}, {50, 33}, {129, 67}, {260, 9}, {3, 300}, {64, 64}, {1000, 131}};

template <typename T> static void level1() {
    const std::string sfx = P<T>::s, tag = std::string(" ") + P<T>::s;
    typedef int (*fill_fn)(i64, i64, T, T*, i64, void*);
    typedef int (*scale_fn)(i64, i64, T, T*, i64, void*);
    typedef int (*had_fn)(i64, i64, const T*, i64, i64, const T*, i64, i64, T*, i64, i64, void*);
    typedef int (*copy_fn)(int, i64, i64, const T*, i64, T*, i64, void*);
    typedef int (*axpy_fn)(i64, i64, T, const T*, i64, T*, i64, void*);
    typedef int (*axpyt_fn)(int, i64, i64, T, const T*, i64, T*, i64, void*);
    fill_fn rfill = sym<fill_fn>("ref_gpu_fill_" + sfx);
    scale_fn rscale = sym<scale_fn>("ref_gpu_scale2d_" + sfx);
    had_fn rhad = sym<had_fn>("ref_gpu_hadamard_" + sfx);
    copy_fn rcopy = sym<copy_fn>("ref_gpu_copy_" + sfx);
    axpy_fn raxpy = sym<axpy_fn>("ref_gpu_axpy2d_" + sfx);
    axpyt_fn raxpyt = sym<axpyt_fn>("ref_gpu_axpy_trans_" + sfx);
    bool ok_fill = true, ok_scale = true, ok_had = true, ok_copy = true, ok_axpy = true, ok_axpyt = true;
    for (auto s : kShapes) {
        const i64 m = s.h, n = s.w, lda = m + 3, ldb = m + 1, ldc = m;
        {   // Fill
            Dev<T> A1(lda * n), A2(lda * n);
            A1.fill((T)-9); A2.fill((T)-9);
            R(rfill(m, n, (T)2.5, A1.p, lda, nullptr), "Fill"); H(P<T>::fill(m, n, (T)2.5, A2.p, lda, nullptr), "hyd_fill");
            ok_fill = ok_fill && same(A1, A2);
        }
        {   // Scale (one multiply)
            Dev<T> A1(lda * n), A2(lda * n);
            A1.frac(37, 101, 50, 64.0); A2.frac(37, 101, 50, 64.0);
            R(rscale(m, n, (T)0.7, A1.p, lda, nullptr), "Scale"); H(P<T>::scale2d(m, n, (T)0.7, A2.p, lda, nullptr), "hyd_scale2d");
            ok_scale = ok_scale && same(A1, A2);
        }
        {   // Hadamard, contiguous and with a row stride of 2 on X
            Dev<T> X(2 * lda * n), Y(ldb * n), Z1(ldc * n), Z2(ldc * n);
            X.frac(37, 101, 50, 64.0); Y.frac(53, 97, 48, 32.0); Z1.fill((T)-9); Z2.fill((T)-9);
            R(rhad(m, n, X.p, 1, lda, Y.p, 1, ldb, Z1.p, 1, ldc, nullptr), "Hadamard");
            H(P<T>::hadamard(m, n, X.p, 1, lda, Y.p, 1, ldb, Z2.p, 1, ldc, nullptr), "hyd_hadamard");
            ok_had = ok_had && same(Z1, Z2);
            R(rhad(m, n, X.p, 2, 2 * lda, Y.p, 1, ldb, Z1.p, 1, ldc, nullptr), "Hadamard");
            H(P<T>::hadamard(m, n, X.p, 2, 2 * lda, Y.p, 1, ldb, Z2.p, 1, ldc, nullptr), "hyd_hadamard");
            ok_had = ok_had && same(Z1, Z2);
        }
        {   // Copy(NORMAL) = cublas?geam(1, A, 0, B)
            Dev<T> A(lda * n), B1(ldb * n), B2(ldb * n);
            A.index(); B1.fill((T)-9); B2.fill((T)-9);
            R(rcopy(0, m, n, A.p, lda, B1.p, ldb, nullptr), "Copy(NORMAL)"); H(P<T>::copy2d(m, n, A.p, lda, B2.p, ldb, nullptr), "hyd_copy2d");
            ok_copy = ok_copy && same(B1, B2);
        }
        {   // Axpy 2-D and transposing Axpy (geam in the reference)
            Dev<T> A(lda * n), At((n + 2) * m), B1(ldb * n), B2(ldb * n);
            A.frac(37, 101, 50, 64.0); At.frac(29, 103, 51, 64.0);
            B1.frac(53, 97, 48, 32.0); B2.frac(53, 97, 48, 32.0);
            R(raxpy(m, n, (T)0.7, A.p, lda, B1.p, ldb, nullptr), "Axpy"); H(P<T>::axpy2d(m, n, (T)0.7, A.p, lda, B2.p, ldb, nullptr), "hyd_axpy2d");
            ok_axpy = ok_axpy && within_ulp(B1, B2, 2.0);
            B1.frac(53, 97, 48, 32.0); B2.frac(53, 97, 48, 32.0);
            R(raxpyt(1, m, n, (T)0.7, At.p, n + 2, B1.p, ldb, nullptr), "Axpy(TRANSPOSE)");
            H(P<T>::axpy_trans(1, m, n, (T)0.7, At.p, n + 2, B2.p, ldb, nullptr), "hyd_axpy_trans");
            ok_axpyt = ok_axpyt && within_ulp(B1, B2, 2.0);
        }
    }
    report("Fill_GPU_impl" + tag, ok_fill);
    report("gpu_blas::Scale 2-D" + tag, ok_scale);
    report("Hadamard_GPU_impl (contiguous, strided X)" + tag, ok_had);
    report("gpu_blas::Copy(NORMAL) (geam)" + tag, ok_copy);
    report("gpu_blas::Axpy 2-D within 1 ulp (geam)" + tag, ok_axpy);
    report("gpu_blas::Axpy(TRANSPOSE) within 1 ulp (geam)" + tag, ok_axpyt);
}

template <typename T> static void packs() {
    const std::string sfx = P<T>::s, tag = std::string(" ") + P<T>::s;
    typedef int (*pack_fn)(i64, i64, i64, i64, const T*, i64, T*, i64, void*);
    typedef int (*ppack_fn)(i64, i64, i64, i64, i64, i64, i64, i64, const T*, i64, T*, i64, void*);
    typedef int (*pack2d_fn)(i64, i64, i64, i64, i64, i64, const T*, i64, T*, i64, void*);
    typedef int (*inter_fn)(i64, i64, const T*, i64, i64, T*, i64, i64, void*);
    const T SENT = (T)-777;
    struct F1 { const char* name; pack_fn rp, ru, hp, hu; bool row; };
    F1 f1[] = {{"RowStridedPack/Unpack", sym<pack_fn>("ref_gpu_row_strided_pack_" + sfx), sym<pack_fn>("ref_gpu_row_strided_unpack_" + sfx),
                (pack_fn)P<T>::row_pack, (pack_fn)P<T>::row_unpack, true},
               {"ColStridedPack/Unpack", sym<pack_fn>("ref_gpu_col_strided_pack_" + sfx), sym<pack_fn>("ref_gpu_col_strided_unpack_" + sfx),
                (pack_fn)P<T>::col_pack, (pack_fn)P<T>::col_unpack, false}};
    for (auto& f : f1) {
        bool ok = true;
        for (auto s : kShapes)
            for (i64 stride : {1, 2, 3, 4, 8}) {
                const i64 align = (s.h + s.w + stride) % stride, ld = s.h + (s.w % 4);
                const i64 maxl = f.row ? ceil_div(s.w, stride) : ceil_div(s.h, stride);
                const i64 portion = std::max<i64>(1, (f.row ? s.h : s.w) * maxl) + (s.h % 3);
                Dev<T> A(ld * s.w), P1(stride * portion), P2(stride * portion), B1(ld * s.w), B2(ld * s.w);
                A.index(); P1.fill(SENT); P2.fill(SENT); B1.fill(SENT); B2.fill(SENT);
                R(f.rp(s.h, s.w, align, stride, A.p, ld, P1.p, portion, nullptr), f.name);
                H(f.hp(s.h, s.w, align, stride, A.p, ld, P2.p, portion, nullptr), f.name);
                R(f.ru(s.h, s.w, align, stride, P1.p, portion, B1.p, ld, nullptr), f.name);
                H(f.hu(s.h, s.w, align, stride, P1.p, portion, B2.p, ld, nullptr), f.name);
                ok = ok && same(P1, P2) && same(B1, B2);
            }
        report(std::string(f.name) + tag, ok);
    }
    struct F2 { const char* name; ppack_fn rp, ru, hp, hu; bool row; };
    F2 f2[] = {{"PartialRowStridedPack/Unpack", sym<ppack_fn>("ref_gpu_partial_row_strided_pack_" + sfx),
                sym<ppack_fn>("ref_gpu_partial_row_strided_unpack_" + sfx), (ppack_fn)P<T>::prow_pack, (ppack_fn)P<T>::prow_unpack, true},
               {"PartialColStridedPack/Unpack", sym<ppack_fn>("ref_gpu_partial_col_strided_pack_" + sfx),
                sym<ppack_fn>("ref_gpu_partial_col_strided_unpack_" + sfx), (ppack_fn)P<T>::pcol_pack, (ppack_fn)P<T>::pcol_unpack, false}};
    const i64 up[][2] = {{1, 1}, {2, 2}, {4, 2}, {2, 4}, {3, 2}};
    for (auto& f : f2) {
        bool ok = true;
        for (auto s : kShapes)
            for (auto& q : up) {
                const i64 uni = q[0], part = q[1], stride = uni * part;
                const i64 align = (s.h * 3 + s.w) % stride, rankPart = (s.h + s.w) % part;
                const i64 shiftM = shift_(rankPart, align % part, part);
                const i64 ndim = f.row ? s.w : s.h, loc = length_(ndim, shiftM, part);
                const i64 lh = f.row ? s.h : loc, lw = f.row ? loc : s.w;
                const i64 ld = std::max<i64>(1, lh) + (s.w % 4);
                const i64 portion = std::max<i64>(1, (f.row ? s.h : s.w) * ceil_div(ndim, stride)) + (s.h % 3);
                const size_t nm = (size_t)ld * std::max<i64>(lw, 1);
                Dev<T> A(nm), P1(uni * portion), P2(uni * portion), B1(nm), B2(nm);
                A.index(); P1.fill(SENT); P2.fill(SENT); B1.fill(SENT); B2.fill(SENT);
                R(f.rp(s.h, s.w, align, stride, uni, part, rankPart, shiftM, A.p, ld, P1.p, portion, nullptr), f.name);
                H(f.hp(s.h, s.w, align, stride, uni, part, rankPart, shiftM, A.p, ld, P2.p, portion, nullptr), f.name);
                R(f.ru(s.h, s.w, align, stride, uni, part, rankPart, shiftM, P1.p, portion, B1.p, ld, nullptr), f.name);
                H(f.hu(s.h, s.w, align, stride, uni, part, rankPart, shiftM, P1.p, portion, B2.p, ld, nullptr), f.name);
                ok = ok && same(P1, P2) && same(B1, B2);
            }
        report(std::string(f.name) + tag, ok);
    }
    {
        pack2d_fn rp = sym<pack2d_fn>("ref_gpu_strided_pack_" + sfx), ru = sym<pack2d_fn>("ref_gpu_strided_unpack_" + sfx);
        bool ok = true;
        const i64 cr[][2] = {{1, 1}, {2, 2}, {2, 4}, {8, 4}, {3, 2}};
        for (auto s : kShapes)
            for (auto& q : cr) {
                const i64 cs = q[0], rs = q[1], ca = s.h % cs, ra = s.w % rs, ld = s.h + (s.w % 4);
                const i64 portion = std::max<i64>(1, ceil_div(s.h, cs) * ceil_div(s.w, rs)) + 1;
                Dev<T> A(ld * s.w), P1(cs * rs * portion), P2(cs * rs * portion), B1(ld * s.w), B2(ld * s.w);
                A.index(); P1.fill(SENT); P2.fill(SENT); B1.fill(SENT); B2.fill(SENT);
                R(rp(s.h, s.w, ca, cs, ra, rs, A.p, ld, P1.p, portion, nullptr), "StridedPack");
                H(P<T>::pack2d(s.h, s.w, ca, cs, ra, rs, A.p, ld, P2.p, portion, nullptr), "StridedPack");
                R(ru(s.h, s.w, ca, cs, ra, rs, P1.p, portion, B1.p, ld, nullptr), "StridedUnpack");
                H(P<T>::unpack2d(s.h, s.w, ca, cs, ra, rs, P1.p, portion, B2.p, ld, nullptr), "StridedUnpack");
                ok = ok && same(P1, P2) && same(B1, B2);
            }
        report("StridedPack/Unpack" + tag, ok);
    }
    {
        inter_fn ri = sym<inter_fn>("ref_gpu_interleave_" + sfx);
        bool ok = true;
        for (auto s : {Shape{1, 1}, Shape{33, 17}, Shape{300, 70}, Shape{64, 257}}) {
            const i64 h = s.h, w = s.w;
            const i64 st[][4] = {{1, h + 1, 1, h}, {2, 2 * h + 3, 1, h}, {1, h, 3, 3 * h + 1}, {3, 3 * h, 2, 2 * h + 2},
                                 {w + 2, 1, 1, h + 2}, {w, 1, 2, 2 * h}};
            for (auto& q : st) {
                const i64 nA = (h - 1) * q[0] + (w - 1) * q[1] + 1, nB = (h - 1) * q[2] + (w - 1) * q[3] + 1;
                Dev<T> A(nA), B1(nB), B2(nB);
                A.index(); B1.fill(SENT); B2.fill(SENT);
                R(ri(h, w, A.p, q[0], q[1], B1.p, q[2], q[3], nullptr), "InterleaveMatrix");
                H(P<T>::interleave(h, w, A.p, q[0], q[1], B2.p, q[2], q[3], nullptr), "InterleaveMatrix");
                ok = ok && same(B1, B2);
            }
        }
        report("InterleaveMatrix (general strides, transposing)" + tag, ok);
    }
}

// REF_CHECK_SUMMA=1: the headline workload on ONE GPU, reference-built against product.  On a 1x1 grid the reference's
// SUMMA_NNC (src/blas_like/level3/Gemm/NN.hpp:325-367) degenerates to, per k-panel of width Blocksize():
//   A1[MC,*] <- A1      RowAllGather with stride 1 = a local Copy                  (Copy/RowAllGather.hpp:35-38)
//   B1^T[MR,*] <- B1    Transpose: a local cublasDgeam                              (Transpose.hpp:161-181)
//   C += alpha A1 (B1^T)^T     LocalGemm(NORMAL, TRANSPOSE) -> cublasDgemm          (Gemm.cpp:152-155)
// after one Scale(beta, C) (Gemm.cpp:287; beta = 0 here, i.e. Zero(C)).  Built here from the recompiled reference layer's own entry points and timed
// against hyd_gemm_f64 (SUMMA-C) on the same operands; the results must agree within 1e-13*sqrt(k) relative Frobenius.
static void summa_1x1() {
    typedef int (*copy_fn)(int, i64, i64, const double*, i64, double*, i64, void*);
    typedef int (*scale_fn)(i64, i64, double, double*, i64, void*);
    typedef int (*gemm_fn)(int, int, i64, i64, i64, double, const double*, i64, const double*, i64, double, double*, i64, void*);
    copy_fn rcopy = sym<copy_fn>("ref_gpu_copy_f64");
    scale_fn rfill = sym<scale_fn>("ref_gpu_fill_f64");   // Scale(0, C) is Zero(C) = Fill (level1/Scale.hpp), same argument list
    gemm_fn rgemm = sym<gemm_fn>("ref_gpu_gemm_f64");
    const i64 n = std::getenv("REF_CHECK_N") ? atoll(std::getenv("REF_CHECK_N")) : 16384;
    hyd_local_group* grp = nullptr;
    hyd_world* w = nullptr;
    hyd_grid* g = nullptr;
    H(hyd_local_group_create(1, &grp), "local group"); H(hyd_world_create_local(grp, 0, &w), "world"); H(hyd_grid_create(w, 1, 1, &g), "grid");
    Dev<double> A(n * n), B(n * n), C1(n * n), C2(n * n);
    {
        std::vector<double> h(n * n);
        uint64_t x = 88172645463325252ull;
        auto next = [&]() { x ^= x << 13; x ^= x >> 7; x ^= x << 17; return (double)(x >> 11) / 9007199254740992.0 * 2.0 - 1.0; };
        for (auto& v : h) v = next();
        A.set(h);
        for (auto& v : h) v = next();
        B.set(h);
    }
    cudaStream_t st;
    CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (i64 nb : {(i64)128, (i64)1024}) {
        Dev<double> A1(n * nb), B1t(n * nb);
        auto reference = [&]() {
            R(rfill(n, n, 0.0, C1.p, n, st), "Zero");
            for (i64 k0 = 0; k0 < n; k0 += nb) {
                const i64 kb = std::min(nb, n - k0);
                R(rcopy(0, n, kb, A.p + k0 * n, n, A1.p, n, st), "Copy");                 // A(:, k0:k0+kb)
                R(rcopy(1, n, kb, B.p + k0, n, B1t.p, n, st), "Transpose");              // B(k0:k0+kb, :)^T is n x kb
                R(rgemm(0, 1, n, n, kb, 1.0, A1.p, n, B1t.p, n, 1.0, C1.p, n, st), "Gemm");
            }
        };
        hyd_dm a{A.p, n, n, n, HYD_MC, HYD_MR, 0, 0, 0, 0}, b{B.p, n, n, n, HYD_MC, HYD_MR, 0, 0, 0, 0}, c{C2.p, n, n, n, HYD_MC, HYD_MR, 0, 0, 0, 0};
        auto product = [&]() { H(hyd_gemm_f64(g, HYD_NORMAL, HYD_NORMAL, HYD_GEMM_SUMMA_C, 1.0, &a, &b, 0.0, &c, nb, st), "hyd_gemm_f64"); };
        double ms[2];
        for (int side = 0; side < 2; ++side) {
            auto run = [&]() { if (side == 0) reference(); else product(); };
            run();
            CK(cudaStreamSynchronize(st));
            CK(cudaEventRecord(e0, st));
            for (int i = 0; i < 3; ++i) run();
            CK(cudaEventRecord(e1, st));
            CK(cudaEventSynchronize(e1));
            float t;
            CK(cudaEventElapsedTime(&t, e0, e1));
            ms[side] = t / 3;
        }
        CK(cudaDeviceSynchronize());
        auto x = C1.host(), y = C2.host();
        double num = 0, den = 0;
        for (size_t i = 0; i < x.size(); ++i) { const double d = x[i] - y[i]; num += d * d; den += x[i] * x[i]; }
        const double rel = std::sqrt(num / den), tf = 2.0 * n * n * n / 1e12;
        char d[256];
        std::snprintf(d, sizeof(d), "n=%lld nb=%lld: reference-built 1x1 SUMMA-C %.2f ms = %.2f TFLOP/s, hyd_gemm_f64 %.2f ms = %.2f TFLOP/s, rel. Frobenius diff %.3g",
                      (long long)n, (long long)nb, ms[0], tf / ms[0] * 1e3, ms[1], tf / ms[1] * 1e3, rel);
        report("El::Gemm NN fp64 on a 1x1 grid, reference layer vs product", rel <= 1e-13 * std::sqrt((double)n), d);
    }
    hyd_grid_destroy(g); hyd_world_destroy(w); hyd_local_group_destroy(grp);
}

int main(int argc, char** argv) {
    const char* refpath = std::getenv("REF_GPU_LIB") ? std::getenv("REF_GPU_LIB") : "oracle/_ref/libref_gpu.so";
    g_ref = dlopen(refpath, RTLD_NOW | RTLD_LOCAL);
    if (!g_ref) { std::fprintf(stderr, "cannot load %s: %s\n", refpath, dlerror()); return 2; }
    CK(cudaSetDevice(0));
    ref_err = sym<const char* (*)()>("ref_gpu_last_error");
    if (sym<int (*)(int)>("ref_gpu_init")(0)) { std::fprintf(stderr, "ref_gpu_init: %s\n", ref_err()); return 2; }
    packs<float>();          // fp64 families are in tools/ref_gpu_check.cpp
    level1<double>();
    level1<float>();
    if (std::getenv("REF_CHECK_SUMMA")) summa_1x1();
    const std::string out = "{\"tool\": \"ref_gpu_check_level1\", \"failed_checks\": " + std::to_string(g_fail) + ", \"results\": [\n " + g_json + "\n]}\n";
    std::fputs(out.c_str(), stdout);
    if (argc > 1) { if (FILE* f = std::fopen(argv[1], "w")) { std::fputs(out.c_str(), f); std::fclose(f); } }
    return g_fail ? 1 : 0;
}
