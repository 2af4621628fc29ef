// Drop-in driver in the style of the reference's tests/blas_like/Syrk.cpp:70-140, written against include/El.hpp:
// DistMatrix[MC,MR] A and C with explicit alignments (A.Align / C.Align), C made symmetric (MakeSymmetric), then
// Syrk(uplo, orientation, alpha, A, beta, C) for both triangles and orientations.  The reference test prints an
// associativity residual; here every entry is compared with a host long-double product: inside the triangle with
// BasicGemm's bound (1e-10 double / 1e-3 float), outside it for bit equality with the original C.
// Also checks the alignment rules of the shim on the way: an unconstrained [MC,*] copy of A adopts A's column alignment
// (Copy/RowAllGather.hpp:23), AlignWith picks the matching dimension (ElementMatrix.cpp:243-300).
//   usage: BasicSyrk [P=4] [gridHeight=2] [n=300] [k=200] [blocksize=64]
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
static int run(int P, int height, El::Int n, El::Int k, El::Int nb, double tol, const char* name) {
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
                cudaStream_t st;
                cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
                const int r = grid.Height(), c = grid.Width();
                for (int variant = 0; variant < 4; ++variant) {
                    const El::UpperOrLower uplo = (variant & 1) ? El::UPPER : El::LOWER;
                    const El::Orientation orient = (variant & 2) ? El::TRANSPOSE : El::NORMAL;
                    El::DistMatrix<T, El::MC, El::MR> A(grid), C(grid);
                    A.SetStream(st); C.SetStream(st);
                    A.Align((variant + 1) % r, variant % c);          // Syrk.cpp:97-98
                    C.Align(variant % r, (variant + 1) % c);
                    const El::Int am = orient == El::NORMAL ? n : k, an = orient == El::NORMAL ? k : n;
                    A.Resize(am, an);
                    C.Resize(n, n);
                    auto fill = [&](El::AbstractDistMatrix<T>& X, unsigned long long seed, El::Int ld) {
                        std::vector<T> h((size_t)X.LocalHeight() * X.LocalWidth());
                        for (El::Int j = 0; j < X.LocalWidth(); ++j)
                            for (El::Int i = 0; i < X.LocalHeight(); ++i)
                                h[i + j * X.LocalHeight()] = uni<T>(seed, X.GlobalRow(i) + X.GlobalCol(j) * ld);
                        X.SetLocalFromHost(h.data(), X.LocalHeight());
                    };
                    fill(A, 1, am); fill(C, 3, n);
                    El::MakeSymmetric(uplo, C);                      // Syrk.cpp:121
                    // symmetric C: entry (i,j) is the filled value of the `uplo`-side position
                    auto c0 = [&](El::Int i, El::Int j) {
                        const bool in = uplo == El::LOWER ? i >= j : i <= j;
                        return in ? uni<T>(3, i + j * n) : uni<T>(3, j + i * n);
                    };
                    // alignment rules on the way (constrained flags, adoption in B = A, AlignWith)
                    El::DistMatrix<T, El::MC, El::STAR> A_MC_STAR(grid);
                    A_MC_STAR.SetStream(st);
                    A_MC_STAR = A;
                    if (A_MC_STAR.ColAlign() != A.ColAlign() || A_MC_STAR.ColConstrained()) failed[rank] = 1;
                    El::DistMatrix<T, El::MR, El::STAR> X_MR_STAR(grid);
                    X_MR_STAR.AlignWith(C);
                    if (X_MR_STAR.ColAlign() != C.RowAlign() || !X_MR_STAR.ColConstrained()) failed[rank] = 1;
                    if (variant == 0) {
                        // Device::CPU matrices (pinned host blocks): cross-device assignment both ways is a copy of the local
                        // block (MC_STAR.cpp:304-390), and Gemm on three of them goes through hyd_gemm_host_*
                        El::DistMatrix<T, El::MC, El::MR, El::ELEMENT, El::Device::CPU> Ah(grid), Gh(grid);
                        El::DistMatrix<T, El::MC, El::MR> A2(grid), Gd(grid);
                        Ah.SetStream(st); Gh.SetStream(st); A2.SetStream(st); Gd.SetStream(st);
                        Ah = A;
                        A2 = Ah;
                        cudaStreamSynchronize(st);
                        if (Ah.ColAlign() != A.ColAlign() || Ah.RowAlign() != A.RowAlign() || A2.ColAlign() != A.ColAlign()) failed[rank] = 1;
                        for (El::Int j = 0; j < Ah.LocalWidth(); ++j)
                            for (El::Int i = 0; i < Ah.LocalHeight(); ++i)
                                if (Ah.LockedBuffer()[i + j * Ah.LDim()] != uni<T>(1, Ah.GlobalRow(i) + Ah.GlobalCol(j) * am)) failed[rank] = 1;
                        El::Gemm(El::NORMAL, El::TRANSPOSE, T(1), A2, A2, Gd);
                        El::Gemm(El::NORMAL, El::TRANSPOSE, T(1), Ah, Ah, Gh);      // returns synchronised, result in host memory
                        std::vector<T> gd((size_t)Gd.LocalHeight() * Gd.LocalWidth());
                        Gd.GetLocalToHost(gd.data(), Gd.LocalHeight());
                        if (Gh.LocalHeight() != Gd.LocalHeight() || Gh.LocalWidth() != Gd.LocalWidth()) failed[rank] = 1;
                        for (El::Int j = 0; j < Gd.LocalWidth(); ++j)
                            for (El::Int i = 0; i < Gd.LocalHeight(); ++i)
                                worst[rank] = std::fmax(worst[rank], std::fabs((double)gd[i + j * Gd.LocalHeight()] -
                                                                               (double)Gh.LockedBuffer()[i + j * Gh.LDim()]));
                    }
                    El::Syrk(uplo, orient, T(2), A, T(3), C);
                    std::vector<T> got((size_t)C.LocalHeight() * C.LocalWidth());
                    C.GetLocalToHost(got.data(), C.LocalHeight());
                    for (El::Int j = 0; j < C.LocalWidth(); ++j)
                        for (El::Int i = 0; i < C.LocalHeight(); ++i) {
                            const El::Int gi = C.GlobalRow(i), gj = C.GlobalCol(j);
                            const T g = got[i + j * C.LocalHeight()];
                            const bool in = uplo == El::LOWER ? gi >= gj : gi <= gj;
                            if (!in) { if (g != c0(gi, gj)) failed[rank] = 1; continue; }
                            long double s = 0;
                            for (El::Int p = 0; p < k; ++p)
                                s += orient == El::NORMAL ? (long double)uni<T>(1, gi + p * am) * (long double)uni<T>(1, gj + p * am)
                                                          : (long double)uni<T>(1, p + gi * am) * (long double)uni<T>(1, p + gj * am);
                            const double ref = (double)(2 * s + 3 * (long double)c0(gi, gj));
                            worst[rank] = std::fmax(worst[rank], std::fabs((double)g - ref));
                        }
                }
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
    std::printf("%s: P=%d grid %dx%d n=%lld k=%lld nb=%lld max|GPU-CPU|=%.3e (bound %.0e) other triangle / alignments %s\n", name, P,
                height, P / height, (long long)n, (long long)k, (long long)nb, w, tol, f ? "FAILED" : "exact");
    if (w > tol) { std::printf("Bad SYRK\n"); return 1; }
    return f;
}

int main(int argc, char** argv) {
    const int P = argc > 1 ? std::atoi(argv[1]) : 4, h = argc > 2 ? std::atoi(argv[2]) : 2;
    const El::Int n = argc > 3 ? std::atoll(argv[3]) : 300, k = argc > 4 ? std::atoll(argv[4]) : 200;
    const El::Int nb = argc > 5 ? std::atoll(argv[5]) : 64;
    int rc = run<double>(P, h, n, k, nb, 1e-10, "double");
    rc |= run<float>(P, h, n, k, nb, 1e-3, "float");
    std::printf(rc ? "FAILED\n" : "PASSED\n");
    return rc;
}
