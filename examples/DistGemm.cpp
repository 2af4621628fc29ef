// One PROCESS per GPU, written the way a Hydrogen client program is (tests/blas_like/BasicGemm.cpp:237-290 of the reference):
//     El::Environment env(argc, argv);  El::mpi::Comm comm = El::mpi::NewWorldComm();
//     El::Grid grid(std::move(comm), gridHeight);  DistMatrix<T,MC,MR> A, B, C;  El::Gemm(NORMAL, NORMAL, alpha, A, B, beta, C);
// against include/El.hpp.  No MPI library: rank / size / device come from the launcher's environment
// (hyd_world_create_env), the data plane is the library's own (NVLink peer-memory push, NCCL as the fallback).
//   launch:  for r in 0 .. N-1:  RANK=r WORLD_SIZE=N LOCAL_RANK=r examples/DistGemm [gridHeight=0 (default)] [n=1000] [nb=128] &
//            (or: python -m torch.distributed.run --no-python --nproc-per-node N examples/DistGemm ...)
// Checks, on every rank's own block: Gemm with alpha = 2, beta = 3 against a host long-double product, bound 1e-10 (double) /
// 1e-3 (float) as BasicGemm.cpp:93-94; B = A into [*,VR] and a distributed Transpose bit for bit against the ownership formula.
#include <El.hpp>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <utility>
#include <vector>

template <typename T>
static T uni(unsigned long long seed, unsigned long long idx) {  // counter-based uniform [-1,1): global entry (i,j) = uni(seed, i + j*n)
    unsigned long long z = seed * 0x9E3779B97F4A7C15ull + idx + 0x632BE59BD9B4E019ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; z ^= z >> 31;
    return (T)((double)(z >> 11) / 9007199254740992.0 * 2.0 - 1.0);
}

template <typename T>
static int TestGemm(El::Int n, const El::Grid& grid, double tol, const char* name) {
    El::DistMatrix<T, El::MC, El::MR> A(n, n, grid), B(n, n, grid), C(n, n, grid);
    auto fill = [&](El::AbstractDistMatrix<T>& X, unsigned long long seed) {
        std::vector<T> h((size_t)X.LocalHeight() * X.LocalWidth());
        for (El::Int j = 0; j < X.LocalWidth(); ++j)
            for (El::Int i = 0; i < X.LocalHeight(); ++i) h[i + j * X.LocalHeight()] = uni<T>(seed, X.GlobalRow(i) + X.GlobalCol(j) * n);
        X.SetLocalFromHost(h.data(), X.LocalHeight());
    };
    fill(A, 1); fill(B, 2); fill(C, 3);
    El::mpi::Barrier(grid.Comm());
    El::Timer timer;
    timer.Start();
    El::Gemm(El::NORMAL, El::NORMAL, T(2), A, B, T(3), C);
    std::vector<T> got((size_t)C.LocalHeight() * C.LocalWidth());
    C.GetLocalToHost(got.data(), C.LocalHeight());      // synchronises C's stream
    El::mpi::Barrier(grid.Comm());
    if (grid.Rank() == 0) El::Output("Distributed Gemm (", name, ", n = ", n, ", ", grid.Height(), " x ", grid.Width(), " grid): ", timer.Stop(), " secs");
    double worst = 0;
    for (El::Int j = 0; j < C.LocalWidth(); ++j)
        for (El::Int i = 0; i < C.LocalHeight(); ++i) {
            const El::Int gi = C.GlobalRow(i), gj = C.GlobalCol(j);
            long double s = 0;
            for (El::Int k = 0; k < n; ++k) s += (long double)uni<T>(1, gi + k * n) * (long double)uni<T>(2, k + gj * n);
            const double ref = (double)(2 * s + 3 * (long double)uni<T>(3, gi + gj * n));
            worst = std::fmax(worst, std::fabs((double)got[i + j * C.LocalHeight()] - ref));
        }
    int bad = worst > tol;
    if (bad) std::fprintf(stderr, "rank %d: Bad GEMM (%s): max |GPU - host| = %.3e > %.0e\n", grid.Rank(), name, worst, tol);
    // redistributions through operator= and Transpose, bit for bit against the ownership formula
    El::DistMatrix<T, El::STAR, El::VR> S(grid);
    S = A;
    std::vector<T> sh((size_t)S.LocalHeight() * S.LocalWidth());
    S.GetLocalToHost(sh.data(), S.LocalHeight());
    for (El::Int j = 0; j < S.LocalWidth(); ++j)
        for (El::Int i = 0; i < S.LocalHeight(); ++i)
            if (sh[i + j * S.LocalHeight()] != uni<T>(1, S.GlobalRow(i) + S.GlobalCol(j) * n)) bad |= 2;
    El::DistMatrix<T, El::MC, El::MR> At(grid);
    El::Transpose(A, At);
    std::vector<T> th((size_t)At.LocalHeight() * At.LocalWidth());
    At.GetLocalToHost(th.data(), At.LocalHeight());
    for (El::Int j = 0; j < At.LocalWidth(); ++j)
        for (El::Int i = 0; i < At.LocalHeight(); ++i)
            if (th[i + j * At.LocalHeight()] != uni<T>(1, At.GlobalCol(j) + At.GlobalRow(i) * n)) bad |= 4;
    if (bad & 6) std::fprintf(stderr, "rank %d: redistribution mismatch (%s, code %d)\n", grid.Rank(), name, bad);
    El::mpi::Barrier(grid.Comm());
    return bad;
}

int main(int argc, char* argv[]) {
    int rc = 1;
    try {
        El::Environment env(argc, argv);
        El::mpi::Comm comm = El::mpi::NewWorldComm();
        const int commRank = El::mpi::Rank(comm), commSize = El::mpi::Size(comm);
        const int gridHeight = argc > 1 ? std::atoi(argv[1]) : 0;
        const El::Int n = argc > 2 ? std::atoll(argv[2]) : 1000, nb = argc > 3 ? std::atoll(argv[3]) : 128;
        {
            El::mpi::Comm keep = comm;   // the handle; the reference moves the communicator into the grid
            El::Grid grid(std::move(comm), gridHeight > 0 ? gridHeight : El::Grid::DefaultHeight(commSize));
            El::SetBlocksize(nb);
            if (commRank == 0) El::Output("DistGemm: ", commSize, " processes, grid ", grid.Height(), " x ", grid.Width(), ", transport ", hyd_world_kind(keep.handle()));
            rc = TestGemm<double>(n, grid, 1e-10, "double");
            rc |= TestGemm<float>(n, grid, 1e-3, "float");
            El::mpi::Barrier(keep);
            comm = keep;
        }
        comm.Free();
        std::printf("rank %d: %s\n", commRank, rc ? "FAILED" : "PASSED");
    } catch (const std::exception& e) {
        std::fprintf(stderr, "DistGemm: %s\n", e.what());
        rc = 1;
    }
    return rc;
}
