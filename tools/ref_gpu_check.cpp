// ref_gpu_check -- the product (libhydrogen_b200.so, C ABI) against THE REFERENCE'S OWN GPU LOCAL LAYER recompiled for
// sm_100a (oracle/_ref/libref_gpu.so: src/hydrogen/blas/gpu/*.cu, gpu_blas::* -> cuBLAS, the GPU pack/unpack loops of
// Copy/util.hpp; see oracle/ref/ref_gpu.cu), on one GPU, without Python:
//   parity : every copy::util pack/unpack family, InterleaveMatrix with general strides and the transposing copy must
//            produce bit-identical device buffers; InterleaveMatrixUpdate / Axpy within 1 ulp (the reference kernel is
//            `B += alpha*A` under nvcc's default FMA contraction, the product rounds product and sum separately like
//            ?axpy); Gemm within the relative Frobenius tolerance of BASELINE.json;
//   timing : the same calls at BASELINE configs[3]/[1] sizes, CUDA events on the launching stream, median of 5 after
//            2 warm-ups -> GB/s (algorithmic bytes: read once + write once) and TFLOP/s for both sides.
// Prints one JSON object (also to argv[1] if given).  Exit code 0 iff every parity check passed.
// This is checker/measurement code: it is the only place besides tests/ that loads oracle/_ref.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <cstdarg>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../include/hydrogen_b200.h"

#define CK(x)                                                                               \
    do {                                                                                    \
        cudaError_t e_ = (x);                                                               \
        if (e_ != cudaSuccess) {                                                            \
            std::fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
            std::exit(2);                                                                   \
        }                                                                                   \
    } while (0)

typedef int64_t i64;
static void* g_ref = nullptr;
template <typename F> F sym(const char* name) {
    void* p = dlsym(g_ref, name);
    if (!p) { std::fprintf(stderr, "missing symbol %s\n", name); std::exit(2); }
    return (F)p;
}

// signatures shared by hyd_* (product) and ref_gpu_* (reference); stream is void* on both sides
typedef int (*pack_fn)(i64, i64, i64, i64, const double*, i64, double*, i64, void*);
typedef int (*ppack_fn)(i64, i64, i64, i64, i64, i64, i64, i64, const double*, i64, double*, i64, void*);
typedef int (*pack2d_fn)(i64, i64, i64, i64, i64, i64, const double*, i64, double*, i64, void*);
typedef int (*inter_fn)(i64, i64, const double*, i64, i64, double*, i64, i64, void*);
typedef int (*update_fn)(double, i64, i64, const double*, i64, i64, double*, i64, i64, void*);
typedef int (*gemm_fn)(int, int, i64, i64, i64, double, const double*, i64, const double*, i64, double, double*, i64, void*);
typedef int (*tcopy_fn)(int, i64, i64, const double*, i64, double*, i64, void*);

static std::string g_json;
static int g_fail = 0;
static void jadd(const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    std::vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (!g_json.empty()) g_json += ",\n ";
    g_json += buf;
}

struct Dev {
    double* p = nullptr;
    size_t n = 0;
    explicit Dev(size_t n_) : n(n_) { CK(cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(double))); }
    ~Dev() { cudaFree(p); }
    void fill_index() {
        std::vector<double> h(n);
        for (size_t i = 0; i < n; ++i) h[i] = (double)(i % (1u << 24));
        CK(cudaMemcpy(p, h.data(), n * sizeof(double), cudaMemcpyHostToDevice));
    }
    void fill_frac(int a, int b, int c, double d) {
        std::vector<double> h(n);
        for (size_t i = 0; i < n; ++i) h[i] = (double)((i64)((i * a) % b) - c) / d;
        CK(cudaMemcpy(p, h.data(), n * sizeof(double), cudaMemcpyHostToDevice));
    }
    void fill_const(double v) {
        std::vector<double> h(n, v);
        CK(cudaMemcpy(p, h.data(), n * sizeof(double), cudaMemcpyHostToDevice));
    }
    std::vector<double> host() const {
        std::vector<double> h(n);
        CK(cudaMemcpy(h.data(), p, n * sizeof(double), cudaMemcpyDeviceToHost));
        return h;
    }
};

static bool same_bits(const std::vector<double>& a, const std::vector<double>& b) {
    return a.size() == b.size() && std::memcmp(a.data(), b.data(), a.size() * sizeof(double)) == 0;
}
static i64 ceil_div(i64 a, i64 b) { return (a + b - 1) / b; }
static i64 shift_(i64 r, i64 a, i64 s) { return ((r - a) % s + s) % s; }
static i64 length_(i64 n, i64 sh, i64 st) { return n > sh ? (n - sh - 1) / st + 1 : 0; }

template <typename Fn> static double time_ms(Fn&& f, int warm = 2, int reps = 5) {
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    for (int i = 0; i < warm; ++i) f();
    std::vector<float> t;
    for (int i = 0; i < reps; ++i) {
        CK(cudaEventRecord(a, 0));
        f();
        CK(cudaEventRecord(b, 0));
        CK(cudaEventSynchronize(b));
        float ms;
        CK(cudaEventElapsedTime(&ms, a, b));
        t.push_back(ms);
    }
    std::sort(t.begin(), t.end());
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    return t[t.size() / 2];
}

static void check(const char* what, bool ok, const char* detail = "") {
    if (!ok) ++g_fail;
    jadd("{\"check\": \"%s\", \"ok\": %s%s%s}", what, ok ? "true" : "false", *detail ? ", \"detail\": " : "", detail);
    std::fprintf(stderr, "%-58s %s %s\n", what, ok ? "ok" : "FAILED", detail);
}

int main(int argc, char** argv) {
    const bool quick = std::getenv("REF_CHECK_QUICK") != nullptr;  // parity only, small timing sizes
    const char* refpath = std::getenv("REF_GPU_LIB") ? std::getenv("REF_GPU_LIB") : "oracle/_ref/libref_gpu.so";
    g_ref = dlopen(refpath, RTLD_NOW | RTLD_LOCAL);
    if (!g_ref) { std::fprintf(stderr, "cannot load %s: %s\n", refpath, dlerror()); return 2; }
    CK(cudaSetDevice(0));
    if (sym<int (*)(int)>("ref_gpu_init")(0)) {
        std::fprintf(stderr, "ref_gpu_init: %s\n", sym<const char* (*)()>("ref_gpu_last_error")());
        return 2;
    }
    const char* (*ref_err)() = sym<const char* (*)()>("ref_gpu_last_error");
    auto R = [&](int rc, const char* w) { if (rc) { std::fprintf(stderr, "reference %s: %s\n", w, ref_err()); std::exit(2); } };
    auto H = [&](int rc, const char* w) { if (rc) { std::fprintf(stderr, "product %s: %s\n", w, hyd_last_error()); std::exit(2); } };
    const double SENT = -777.0;

    // ------------------------------------------------------------------ parity: pack / unpack families
    struct Shape { i64 h, w; };
    const Shape shapes[] = {{1, 1}, {7, 5}, {50, 33}, {129, 67}, {260, 9}, {3, 300}, {64, 64}, {1000, 131}};
    {
        struct Fam { const char* name; pack_fn rp, ru, hp, hu; bool row; };
        Fam fams[] = {
            {"RowStridedPack/Unpack", sym<pack_fn>("ref_gpu_row_strided_pack_f64"), sym<pack_fn>("ref_gpu_row_strided_unpack_f64"),
             (pack_fn)hyd_row_strided_pack_f64, (pack_fn)hyd_row_strided_unpack_f64, true},
            {"ColStridedPack/Unpack", sym<pack_fn>("ref_gpu_col_strided_pack_f64"), sym<pack_fn>("ref_gpu_col_strided_unpack_f64"),
             (pack_fn)hyd_col_strided_pack_f64, (pack_fn)hyd_col_strided_unpack_f64, false}};
        for (auto& f : fams) {
            bool ok = true;
            int n = 0;
            for (auto s : shapes)
                for (i64 stride : {1, 2, 3, 4, 8}) {
                    const i64 align = (s.h + s.w + stride) % stride, ld = s.h + (s.w % 4);
                    const i64 maxl = f.row ? ceil_div(s.w, stride) : ceil_div(s.h, stride);
                    const i64 portion = std::max<i64>(1, (f.row ? s.h : s.w) * maxl) + (s.h % 3);
                    Dev A(ld * s.w), P1(stride * portion), P2(stride * portion), B1(ld * s.w), B2(ld * s.w);
                    A.fill_index(); P1.fill_const(SENT); P2.fill_const(SENT); B1.fill_const(SENT); B2.fill_const(SENT);
                    R(f.rp(s.h, s.w, align, stride, A.p, ld, P1.p, portion, nullptr), f.name);
                    H(f.hp(s.h, s.w, align, stride, A.p, ld, P2.p, portion, nullptr), f.name);
                    R(f.ru(s.h, s.w, align, stride, P1.p, portion, B1.p, ld, nullptr), f.name);
                    H(f.hu(s.h, s.w, align, stride, P1.p, portion, B2.p, ld, nullptr), f.name);
                    CK(cudaDeviceSynchronize());
                    ok = ok && same_bits(P1.host(), P2.host()) && same_bits(B1.host(), B2.host());
                    ++n;
                }
            char d[64];
            std::snprintf(d, sizeof(d), "\"%d argument sets\"", n);
            check(f.name, ok, d);
        }
    }
    {
        struct Fam { const char* name; ppack_fn rp, ru, hp, hu; bool row; };
        Fam fams[] = {
            {"PartialRowStridedPack/Unpack", sym<ppack_fn>("ref_gpu_partial_row_strided_pack_f64"),
             sym<ppack_fn>("ref_gpu_partial_row_strided_unpack_f64"), (ppack_fn)hyd_partial_row_strided_pack_f64,
             (ppack_fn)hyd_partial_row_strided_unpack_f64, true},
            {"PartialColStridedPack/Unpack", sym<ppack_fn>("ref_gpu_partial_col_strided_pack_f64"),
             sym<ppack_fn>("ref_gpu_partial_col_strided_unpack_f64"), (ppack_fn)hyd_partial_col_strided_pack_f64,
             (ppack_fn)hyd_partial_col_strided_unpack_f64, false}};
        const i64 up[][2] = {{1, 1}, {2, 2}, {4, 2}, {2, 4}, {3, 2}};
        for (auto& f : fams) {
            bool ok = true;
            int n = 0;
            for (auto s : shapes)
                for (auto& q : up) {
                    const i64 uni = q[0], part = q[1], stride = uni * part;
                    const i64 align = (s.h * 3 + s.w) % stride, rankPart = (s.h + s.w) % part;
                    const i64 shiftM = shift_(rankPart, align % part, part);
                    const i64 ndim = f.row ? s.w : s.h, loc = length_(ndim, shiftM, part);
                    const i64 lh = f.row ? s.h : loc, lw = f.row ? loc : s.w;
                    const i64 ld = std::max<i64>(1, lh) + (s.w % 4);
                    const i64 portion = std::max<i64>(1, (f.row ? s.h : s.w) * ceil_div(ndim, stride)) + (s.h % 3);
                    const size_t nm = (size_t)ld * std::max<i64>(lw, 1);
                    Dev A(nm), P1(uni * portion), P2(uni * portion), B1(nm), B2(nm);
                    A.fill_index(); P1.fill_const(SENT); P2.fill_const(SENT); B1.fill_const(SENT); B2.fill_const(SENT);
                    R(f.rp(s.h, s.w, align, stride, uni, part, rankPart, shiftM, A.p, ld, P1.p, portion, nullptr), f.name);
                    H(f.hp(s.h, s.w, align, stride, uni, part, rankPart, shiftM, A.p, ld, P2.p, portion, nullptr), f.name);
                    R(f.ru(s.h, s.w, align, stride, uni, part, rankPart, shiftM, P1.p, portion, B1.p, ld, nullptr), f.name);
                    H(f.hu(s.h, s.w, align, stride, uni, part, rankPart, shiftM, P1.p, portion, B2.p, ld, nullptr), f.name);
                    CK(cudaDeviceSynchronize());
                    ok = ok && same_bits(P1.host(), P2.host()) && same_bits(B1.host(), B2.host());
                    ++n;
                }
            char d[64];
            std::snprintf(d, sizeof(d), "\"%d argument sets\"", n);
            check(f.name, ok, d);
        }
    }
    {
        pack2d_fn rp = sym<pack2d_fn>("ref_gpu_strided_pack_f64"), ru = sym<pack2d_fn>("ref_gpu_strided_unpack_f64");
        bool ok = true;
        int n = 0;
        const i64 cr[][2] = {{1, 1}, {2, 2}, {2, 4}, {8, 4}, {3, 2}};
        for (auto s : shapes)
            for (auto& q : cr) {
                const i64 cs = q[0], rs = q[1], ca = s.h % cs, ra = s.w % rs, ld = s.h + (s.w % 4);
                const i64 portion = std::max<i64>(1, ceil_div(s.h, cs) * ceil_div(s.w, rs)) + 1;
                Dev A(ld * s.w), P1(cs * rs * portion), P2(cs * rs * portion), B1(ld * s.w), B2(ld * s.w);
                A.fill_index(); P1.fill_const(SENT); P2.fill_const(SENT); B1.fill_const(SENT); B2.fill_const(SENT);
                R(rp(s.h, s.w, ca, cs, ra, rs, A.p, ld, P1.p, portion, nullptr), "StridedPack");
                H(hyd_strided_pack_f64(s.h, s.w, ca, cs, ra, rs, A.p, ld, P2.p, portion, nullptr), "StridedPack");
                R(ru(s.h, s.w, ca, cs, ra, rs, P1.p, portion, B1.p, ld, nullptr), "StridedUnpack");
                H(hyd_strided_unpack_f64(s.h, s.w, ca, cs, ra, rs, P1.p, portion, B2.p, ld, nullptr), "StridedUnpack");
                CK(cudaDeviceSynchronize());
                ok = ok && same_bits(P1.host(), P2.host()) && same_bits(B1.host(), B2.host());
                ++n;
            }
        char d[64];
        std::snprintf(d, sizeof(d), "\"%d argument sets\"", n);
        check("StridedPack/Unpack", ok, d);
    }
    {   // InterleaveMatrix with general strides (copy_2d_kernel) and InterleaveMatrixUpdate (axpy kernel)
        inter_fn ri = sym<inter_fn>("ref_gpu_interleave_f64");
        update_fn ru = sym<update_fn>("ref_gpu_interleave_update_f64");
        bool ok = true, upd_ulp = true;
        long upd_equal = 0, upd_total = 0;
        for (auto s : {Shape{1, 1}, Shape{33, 17}, Shape{300, 70}, Shape{64, 257}}) {
            const i64 h = s.h, w = s.w;
            const i64 st[][4] = {{1, h + 1, 1, h}, {2, 2 * h + 3, 1, h}, {1, h, 3, 3 * h + 1}, {3, 3 * h, 2, 2 * h + 2},
                                 {w + 2, 1, 1, h + 2}, {w, 1, 2, 2 * h}};
            for (auto& q : st) {
                const i64 nA = (h - 1) * q[0] + (w - 1) * q[1] + 1, nB = (h - 1) * q[2] + (w - 1) * q[3] + 1;
                Dev A(nA), B1(nB), B2(nB);
                A.fill_index(); B1.fill_const(SENT); B2.fill_const(SENT);
                R(ri(h, w, A.p, q[0], q[1], B1.p, q[2], q[3], nullptr), "InterleaveMatrix");
                H(hyd_interleave_f64(h, w, A.p, q[0], q[1], B2.p, q[2], q[3], nullptr), "InterleaveMatrix");
                CK(cudaDeviceSynchronize());
                ok = ok && same_bits(B1.host(), B2.host());
                A.fill_frac(37, 101, 50, 64.0); B1.fill_frac(53, 97, 48, 32.0); B2.fill_frac(53, 97, 48, 32.0);
                R(ru(0.7, h, w, A.p, q[0], q[1], B1.p, q[2], q[3], nullptr), "InterleaveMatrixUpdate");
                H(hyd_interleave_update_f64(0.7, h, w, A.p, q[0], q[1], B2.p, q[2], q[3], nullptr), "InterleaveMatrixUpdate");
                CK(cudaDeviceSynchronize());
                auto x = B1.host(), y = B2.host();
                for (size_t i = 0; i < x.size(); ++i) {
                    ++upd_total;
                    if (x[i] == y[i]) ++upd_equal;
                    else if (std::fabs(x[i] - y[i]) > 2.3e-16 * std::max(std::fabs(x[i]), 1.0)) upd_ulp = false;
                }
            }
        }
        check("InterleaveMatrix (general strides, transposing)", ok);
        char d[128];
        std::snprintf(d, sizeof(d), "\"%ld of %ld elements bit-identical to the reference's FMA-contracted kernel, rest within 1 ulp\"",
                      upd_equal, upd_total);
        check("InterleaveMatrixUpdate within 1 ulp", upd_ulp, d);
    }
    {   // local Transpose: gpu_blas::Copy(TRANSPOSE, n, m, A, lda, B, ldb) (dims of B) vs hyd_transpose(m, n, ...)
        tcopy_fn rc = sym<tcopy_fn>("ref_gpu_copy_f64");
        bool ok = true;
        for (auto s : shapes) {
            const i64 m = s.h, n = s.w, lda = m + 3, ldb = n + 1;
            Dev A(lda * n), B1(ldb * m), B2(ldb * m);
            A.fill_index(); B1.fill_const(SENT); B2.fill_const(SENT);
            R(rc(1, n, m, A.p, lda, B1.p, ldb, nullptr), "Copy(TRANSPOSE)");
            H(hyd_transpose_f64(m, n, A.p, lda, B2.p, ldb, nullptr), "hyd_transpose");
            CK(cudaDeviceSynchronize());
            ok = ok && same_bits(B1.host(), B2.host());
        }
        check("local Transpose (cublasDgeam in the reference)", ok);
    }
    // ------------------------------------------------------------------ parity + timing: local Gemm
    gemm_fn rg = sym<gemm_fn>("ref_gpu_gemm_f64");
    {
        bool ok = true;
        double worst = 0;
        const i64 mnk[][3] = {{257, 129, 300}, {512, 512, 512}, {1000, 77, 2048}};
        for (auto& q : mnk)
            for (int tA = 0; tA < 2; ++tA)
                for (int tB = 0; tB < 2; ++tB) {
                    const i64 m = q[0], n = q[1], k = q[2];
                    const i64 lda = (tA ? k : m) + 2, ldb = (tB ? n : k) + 2, ldc = m + 2;
                    Dev A(lda * (tA ? m : k)), B(ldb * (tB ? k : n)), C1(ldc * n), C2(ldc * n);
                    A.fill_frac(37, 101, 50, 64.0); B.fill_frac(53, 97, 48, 32.0);
                    C1.fill_frac(11, 89, 44, 16.0); C2.fill_frac(11, 89, 44, 16.0);
                    R(rg(tA, tB, m, n, k, 2.0, A.p, lda, B.p, ldb, 3.0, C1.p, ldc, nullptr), "Gemm");
                    H(hyd_lgemm_f64(tA, tB, m, n, k, 2.0, A.p, lda, B.p, ldb, 3.0, C2.p, ldc, nullptr), "hyd_lgemm_f64");
                    CK(cudaDeviceSynchronize());
                    auto x = C1.host(), y = C2.host();
                    double num = 0, den = 0;
                    for (i64 j = 0; j < n; ++j)
                        for (i64 i = 0; i < m; ++i) {
                            const double d = x[i + j * ldc] - y[i + j * ldc];
                            num += d * d;
                            den += x[i + j * ldc] * x[i + j * ldc];
                        }
                    const double rel = std::sqrt(num / den);
                    worst = std::max(worst, rel / std::sqrt((double)k));
                    ok = ok && rel <= 1e-13 * std::sqrt((double)k);
                    for (i64 j = 0; j < n; ++j)   // ldc padding untouched on both sides
                        for (i64 i = m; i < ldc; ++i) ok = ok && x[i + j * ldc] == y[i + j * ldc];
                }
        char d[96];
        std::snprintf(d, sizeof(d), "\"worst rel. Frobenius error / sqrt(k) = %.3g (bound 1e-13)\"", worst);
        check("local Gemm vs gpu_blas::Gemm (cublasDgemm), 4 orientations", ok, d);
    }
    // ------------------------------------------------------------------ timing
    const i64 N = quick ? 4096 : 16384;   // local block of BASELINE configs[3] on a 2x4 grid is 16384 x 8192
    {
        const i64 h = N, w = N / 2, ld = h;
        for (i64 stride : {2, 4}) {
            const i64 portionR = h * ceil_div(w, stride), portionC = w * ceil_div(h, stride);
            Dev A(ld * w), P(std::max(portionR, portionC) * stride), B(ld * w);
            A.fill_const(1.0);
            const double gb = 2.0 * h * w * 8 / 1e9;
            pack_fn rrp = sym<pack_fn>("ref_gpu_row_strided_pack_f64"), rcp = sym<pack_fn>("ref_gpu_col_strided_pack_f64");
            pack_fn rru = sym<pack_fn>("ref_gpu_row_strided_unpack_f64"), rcu = sym<pack_fn>("ref_gpu_col_strided_unpack_f64");
            const double t_rrp = time_ms([&] { rrp(h, w, 0, stride, A.p, ld, P.p, portionR, nullptr); });
            const double t_hrp = time_ms([&] { hyd_row_strided_pack_f64(h, w, 0, stride, A.p, ld, P.p, portionR, nullptr); });
            const double t_rru = time_ms([&] { rru(h, w, 0, stride, P.p, portionR, B.p, ld, nullptr); });
            const double t_hru = time_ms([&] { hyd_row_strided_unpack_f64(h, w, 0, stride, P.p, portionR, B.p, ld, nullptr); });
            const double t_rcp = time_ms([&] { rcp(h, w, 0, stride, A.p, ld, P.p, portionC, nullptr); });
            const double t_hcp = time_ms([&] { hyd_col_strided_pack_f64(h, w, 0, stride, A.p, ld, P.p, portionC, nullptr); });
            const double t_rcu = time_ms([&] { rcu(h, w, 0, stride, P.p, portionC, B.p, ld, nullptr); });
            const double t_hcu = time_ms([&] { hyd_col_strided_unpack_f64(h, w, 0, stride, P.p, portionC, B.p, ld, nullptr); });
            jadd("{\"timing\": \"pack/unpack fp64 %lldx%lld stride %lld\", \"unit\": \"GB/s (2*h*w*8 bytes)\", "
                 "\"RowStridedPack\": {\"reference\": %.0f, \"product\": %.0f}, \"RowStridedUnpack\": {\"reference\": %.0f, \"product\": %.0f}, "
                 "\"ColStridedPack\": {\"reference\": %.0f, \"product\": %.0f}, \"ColStridedUnpack\": {\"reference\": %.0f, \"product\": %.0f}}",
                 (long long)h, (long long)w, (long long)stride, gb / t_rrp * 1e3, gb / t_hrp * 1e3, gb / t_rru * 1e3, gb / t_hru * 1e3,
                 gb / t_rcp * 1e3, gb / t_hcp * 1e3, gb / t_rcu * 1e3, gb / t_hcu * 1e3);
        }
        {   // local transpose of the same block
            Dev A(h * w), B(h * w);
            A.fill_const(1.0);
            tcopy_fn rc = sym<tcopy_fn>("ref_gpu_copy_f64");
            const double gb = 2.0 * h * w * 8 / 1e9;
            const double t_r = time_ms([&] { rc(1, w, h, A.p, h, B.p, w, nullptr); });
            const double t_h = time_ms([&] { hyd_transpose_f64(h, w, A.p, h, B.p, w, nullptr); });
            jadd("{\"timing\": \"local Transpose fp64 %lldx%lld\", \"unit\": \"GB/s\", \"reference (cublasDgeam)\": %.0f, \"product\": %.0f}",
                 (long long)h, (long long)w, gb / t_r * 1e3, gb / t_h * 1e3);
        }
    }
    {   // local Gemm: the SUMMA-C panel product of BASELINE configs[1] on 2x4 (8192 x 4096 x nb, 'N','T') and a square one
        const i64 cases[][5] = {{8192, 4096, 128, 0, 1}, {8192, 4096, 1024, 0, 1}, {N / 2, N / 2, N / 2, 0, 0}};
        for (auto& q : cases) {
            const i64 m = quick ? q[0] / 4 : q[0], n = quick ? q[1] / 4 : q[1], k = q[2];
            const int tA = (int)q[3], tB = (int)q[4];
            Dev A(m * k), B(n * k), C(m * n);
            A.fill_const(0.5); B.fill_const(0.25); C.fill_const(0.0);
            const i64 lda = tA ? k : m, ldb = tB ? n : k;
            const double tf = 2.0 * m * n * k / 1e12;
            const double t_r = time_ms([&] { rg(tA, tB, m, n, k, 1.0, A.p, lda, B.p, ldb, 1.0, C.p, m, nullptr); });
            const double t_h = time_ms([&] { hyd_lgemm_f64(tA, tB, m, n, k, 1.0, A.p, lda, B.p, ldb, 1.0, C.p, m, nullptr); });
            jadd("{\"timing\": \"local Gemm fp64 %c%c %lldx%lldx%lld beta=1\", \"unit\": \"TFLOP/s\", \"reference (cublasDgemm)\": %.2f, \"product\": %.2f}",
                 tA ? 'T' : 'N', tB ? 'T' : 'N', (long long)m, (long long)n, (long long)k, tf / t_r * 1e3, tf / t_h * 1e3);
        }
    }
    std::string out = "{\"tool\": \"ref_gpu_check\", \"failed_checks\": " + std::to_string(g_fail) + ", \"results\": [\n " + g_json + "\n]}\n";
    std::fputs(out.c_str(), stdout);
    if (argc > 1) {
        if (FILE* f = std::fopen(argv[1], "w")) { std::fputs(out.c_str(), f); std::fclose(f); }
    }
    return g_fail ? 1 : 0;
}
