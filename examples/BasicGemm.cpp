// Drop-in driver in the style of the reference's tests/blas_like/BasicGemm.cpp:96-236, written against
// include/El.hpp: DistMatrix[MC,MR] Gemm with alpha=2, beta=3, compared element-wise with a host
// product, bound 1e-10 (double) / 1e-3 (float) as BasicGemm.cpp:93-94.  P rank-threads share one GPU
// (hyd_local_group); with one process per GPU use mpi::Comm::NCCL instead.
//   usage: BasicGemm [P=4] [gridHeight=2] [m=n=k=300] [blocksize=128]
#include <El.hpp>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>

template <typename T>
static T uni(unsigned long long seed, unsigned long long idx) {  // counter-based uniform [-1,1)
    unsigned long long z = seed * 0x9E3779B97F4A7C15ull + idx + 0x632BE59BD9B4E019ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; z ^= z >> 31;
    return (T)((double)(z >> 11) / 9007199254740992.0 * 2.0 - 1.0);
}

template <typename T>
static int run(int P, int height, El::Int n, El::Int nb, double tol, const char* name) {
    hyd_local_group* grp = nullptr;
    if (hyd_local_group_create(P, &grp)) return 1;
    std::vector<double> worst(P, 0.0);
    std::vector<int> failed(P, 0);
    auto body = [&](int rank) {
        try {
            cudaSetDevice(0);
            El::mpi::Comm comm = El::mpi::Comm::Local(grp, rank);
            {
                El::Grid grid(comm, height);
                El::SetBlocksize(nb);
                // rank-threads that share one GPU must work on distinct non-blocking streams (the legacy null stream
                // would serialise one rank's flag wait against another rank's signal); one process per GPU may use any
                cudaStream_t st;
                cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
                El::DistMatrix<T, El::MC, El::MR> A(n, n, grid), B(n, n, grid), C(n, n, grid);
                A.SetStream(st); B.SetStream(st); C.SetStream(st);
                auto fill = [&](El::AbstractDistMatrix<T>& X, unsigned long long seed) {
                    std::vector<T> h((size_t)X.LocalHeight() * X.LocalWidth());
                    for (El::Int j = 0; j < X.LocalWidth(); ++j)
                        for (El::Int i = 0; i < X.LocalHeight(); ++i)
                            h[i + j * X.LocalHeight()] = uni<T>(seed, X.GlobalRow(i) + X.GlobalCol(j) * n);
                    X.SetLocalFromHost(h.data(), X.LocalHeight());
                };
                fill(A, 1); fill(B, 2); fill(C, 3);
                El::Gemm(El::NORMAL, El::NORMAL, T(2), A, B, T(3), C, El::GEMM_SUMMA_C);
                std::vector<T> got((size_t)C.LocalHeight() * C.LocalWidth());
                C.GetLocalToHost(got.data(), C.LocalHeight());
                for (El::Int j = 0; j < C.LocalWidth(); ++j)
                    for (El::Int i = 0; i < C.LocalHeight(); ++i) {
                        const El::Int gi = C.GlobalRow(i), gj = C.GlobalCol(j);
                        long double s = 0;
                        for (El::Int k = 0; k < n; ++k) s += (long double)uni<T>(1, gi + k * n) * (long double)uni<T>(2, k + gj * n);
                        const double ref = (double)(2 * s + 3 * (long double)uni<T>(3, gi + gj * n));
                        worst[rank] = std::fmax(worst[rank], std::fabs((double)got[i + j * C.LocalHeight()] - ref));
                    }
                // a redistribution through operator=, checked bit-exactly against the ownership formula
                El::DistMatrix<T, El::STAR, El::VR> S(grid);
                S.SetStream(st);
                S = A;
                std::vector<T> sh((size_t)S.LocalHeight() * S.LocalWidth());
                S.GetLocalToHost(sh.data(), S.LocalHeight());
                for (El::Int j = 0; j < S.LocalWidth(); ++j)
                    for (El::Int i = 0; i < S.LocalHeight(); ++i)
                        if (sh[i + j * S.LocalHeight()] != uni<T>(1, S.GlobalRow(i) + S.GlobalCol(j) * n)) failed[rank] = 1;
                cudaStreamSynchronize(st);
                comm.Barrier();
                cudaStreamDestroy(st);
            }
            comm.Free();
        } catch (const std::exception& e) {
            std::fprintf(stderr, "rank %d: %s\n", rank, e.what());
            failed[rank] = 1;
        }
    };
    std::vector<std::thread> th;
    for (int r = 0; r < P; ++r) th.emplace_back(body, r);
    for (auto& t : th) t.join();
    hyd_local_group_destroy(grp);
    double w = 0; int f = 0;
    for (int r = 0; r < P; ++r) { w = std::fmax(w, worst[r]); f |= failed[r]; }
    std::printf("%s: P=%d grid %dx%d n=%lld nb=%lld max|GPU-CPU|=%.3e (bound %.0e) redist %s\n", name, P, height, P / height,
                (long long)n, (long long)nb, w, tol, f ? "FAILED" : "bit-exact");
    if (w > tol) { std::printf("Bad GEMM\n"); return 1; }  // BasicGemm.cpp:147-148
    return f;
}

int main(int argc, char** argv) {
    const int P = argc > 1 ? std::atoi(argv[1]) : 4, h = argc > 2 ? std::atoi(argv[2]) : 2;
    const El::Int n = argc > 3 ? std::atoll(argv[3]) : 300, nb = argc > 4 ? std::atoll(argv[4]) : 128;
    int rc = run<double>(P, h, n, nb, 1e-10, "double");
    rc |= run<float>(P, h, n, nb, 1e-3, "float");
    std::printf(rc ? "FAILED\n" : "PASSED\n");
    return rc;
}
