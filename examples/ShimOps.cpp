// Shim-level checks of two rows of the hot-path table that BasicGemm / BasicSyrk do not reach:
//   a8  El::LocalGemm (src/blas_like/level3/Gemm.cpp:323-431): the conformity checks of the reference's debug build, then the
//       local product of [MC,*] x [*,MR] (and [MC,*] x [MR,*]^T) blocks -- the stationary-C update written by hand;
//   a24 MultiSync (include/hydrogen/SyncInfo.hpp:136-186, used at Gemm.cpp:117-119): A is produced on stream 1, B on stream 2,
//       C lives on stream 3, no host synchronisation in between; El::Gemm fences the input streams itself, and the result is
//       bit-identical to a fully synchronised run.  Both streams are kept busy first, so an unfenced Gemm would read zeros.
// P rank-threads share one GPU (hyd_local_group).   usage: ShimOps [P=4] [gridHeight=2] [n=300] [blocksize=64]
#include <El.hpp>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

static double uni(unsigned long long seed, unsigned long long idx) {
    unsigned long long z = seed * 0x9E3779B97F4A7C15ull + idx + 0x632BE59BD9B4E019ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; z ^= z >> 31;
    return (double)(z >> 11) / 9007199254740992.0 * 2.0 - 1.0;
}
typedef double T;
typedef El::AbstractDistMatrix<T> ADM;

static void fill(ADM& X, unsigned long long seed, El::Int ld) {
    std::vector<T> h((size_t)X.LocalHeight() * X.LocalWidth());
    for (El::Int j = 0; j < X.LocalWidth(); ++j)
        for (El::Int i = 0; i < X.LocalHeight(); ++i) h[i + j * X.LocalHeight()] = uni(seed, X.GlobalRow(i) + X.GlobalCol(j) * ld);
    X.SetLocalFromHost(h.data(), X.LocalHeight());
}
static std::vector<T> fetch(ADM& X) {
    std::vector<T> h((size_t)X.LocalHeight() * X.LocalWidth());
    X.GetLocalToHost(h.data(), X.LocalHeight());
    return h;
}

int main(int argc, char** argv) {
    const int P = argc > 1 ? std::atoi(argv[1]) : 4, height = argc > 2 ? std::atoi(argv[2]) : 2;
    const El::Int n = argc > 3 ? std::atoll(argv[3]) : 300, nb = argc > 4 ? std::atoll(argv[4]) : 64;
    hyd_local_group* grp = nullptr;
    if (hyd_local_group_create(P, &grp)) return 1;
    std::vector<int> failed(P, 0);
    std::vector<double> worst(P, 0.0);
    auto body = [&](int rank) {
        try {
            cudaSetDevice(0);
            El::mpi::Comm comm = El::mpi::Comm::Local(grp, rank);
            {
                El::Grid grid(comm, height);
                El::SetBlocksize(nb);
                cudaStream_t s1, s2, s3;
                cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking);
                cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking);
                cudaStreamCreateWithFlags(&s3, cudaStreamNonBlocking);
                El::DistMatrix<T, El::MC, El::MR> A0(n, n, grid), B0(n, n, grid), A(n, n, grid), B(n, n, grid), C(n, n, grid), Cref(n, n, grid);
                for (ADM* X : {(ADM*)&A0, (ADM*)&B0, (ADM*)&Cref, (ADM*)&C}) X->SetStream(s3);
                fill(A0, 1, n); fill(B0, 2, n); fill(C, 3, n); fill(Cref, 3, n);
                // ---- reference run: everything on one stream, host-synchronised
                El::Gemm(El::NORMAL, El::NORMAL, T(2), A0, B0, T(3), Cref, El::GEMM_SUMMA_C);
                const std::vector<T> want = fetch(Cref);
                // ---- three streams, no host synchronisation: A appears on s1 and B on s2 only after both streams have
                // worked through a queue of large fills
                A.SetStream(s1); B.SetStream(s2);
                El::Zero(A); El::Zero(B);
                cudaStreamSynchronize(s1); cudaStreamSynchronize(s2); cudaStreamSynchronize(s3);
                El::DistMatrix<T, El::STAR, El::STAR> busy1(4096, 2048, grid), busy2(4096, 2048, grid);
                busy1.SetStream(s1); busy2.SetStream(s2);
                comm.Barrier();
                for (int it = 0; it < 40; ++it) { El::Fill(busy1, T(it)); El::Fill(busy2, T(it)); }
                El::SyncInfoGPU si1, si2;
                si1.stream_ = s1; si2.stream_ = s2;
                El::gpu_blas::Copy(El::NORMAL, A.LocalHeight(), A.LocalWidth(), A0.LockedBuffer(), A0.LDim(), A.Buffer(), A.LDim(), si1);
                El::gpu_blas::Copy(El::NORMAL, B.LocalHeight(), B.LocalWidth(), B0.LockedBuffer(), B0.LDim(), B.Buffer(), B.LDim(), si2);
                El::Gemm(El::NORMAL, El::NORMAL, T(2), A, B, T(3), C, El::GEMM_SUMMA_C);   // fences s1 and s2 itself
                // the exit half of the fence: A may be overwritten on ITS stream right away without disturbing the Gemm
                El::Fill(A, T(-7));
                const std::vector<T> got = fetch(C);
                if (got.size() != want.size() || std::memcmp(got.data(), want.data(), got.size() * sizeof(T)) != 0) {
                    std::fprintf(stderr, "rank %d: three-stream Gemm differs from the synchronised run\n", rank);
                    failed[rank] = 1;
                }
                // ---- LocalGemm: the stationary-C update by hand, C2 = A0 B0 as [MC,*] x [*,MR] and as [MC,*] x [MR,*]^T
                El::DistMatrix<T, El::MC, El::STAR> As(grid);
                El::DistMatrix<T, El::STAR, El::MR> Bs(grid);
                El::DistMatrix<T, El::MR, El::STAR> Bt(grid);
                El::DistMatrix<T, El::MC, El::MR> C2(grid), C3(grid);
                for (ADM* X : {(ADM*)&As, (ADM*)&Bs, (ADM*)&Bt, (ADM*)&C2, (ADM*)&C3}) X->SetStream(s3);
                As.AlignWith(A0); Bs.AlignWith(A0); Bt.AlignWith(A0); C2.AlignWith(A0); C3.AlignWith(A0);
                As = A0;
                Bs = B0;
                El::Transpose(B0, Bt);
                El::LocalGemm(El::NORMAL, El::NORMAL, T(1), As, Bs, C2);
                El::LocalGemm(El::NORMAL, El::TRANSPOSE, T(1), As, Bt, C3);
                const std::vector<T> g2 = fetch(C2), g3 = fetch(C3);
                for (El::Int j = 0; j < C2.LocalWidth(); ++j)
                    for (El::Int i = 0; i < C2.LocalHeight(); ++i) {
                        const El::Int gi = C2.GlobalRow(i), gj = C2.GlobalCol(j);
                        long double s = 0;
                        for (El::Int k = 0; k < n; ++k) s += (long double)uni(1, gi + k * n) * (long double)uni(2, k + gj * n);
                        worst[rank] = std::fmax(worst[rank], std::fabs((double)g2[i + j * C2.LocalHeight()] - (double)s));
                        worst[rank] = std::fmax(worst[rank], std::fabs((double)g3[i + j * C3.LocalHeight()] - (double)s));
                    }
                // the reference's conformity checks (Gemm.cpp:335-431): wrong distributions and wrong shapes throw LogicError
                int thrown = 0;
                try { El::LocalGemm(El::NORMAL, El::NORMAL, T(1), A0, Bs, T(0), C2); } catch (const std::logic_error&) { ++thrown; }
                try { El::LocalGemm(El::TRANSPOSE, El::NORMAL, T(1), As, Bs, T(0), C2); } catch (const std::logic_error&) { ++thrown; }
                El::DistMatrix<T, El::STAR, El::MR> Bbad(n + 1, n, grid);
                Bbad.AlignWith(A0);
                try { El::LocalGemm(El::NORMAL, El::NORMAL, T(1), As, Bbad, T(0), C2); } catch (const std::logic_error&) { ++thrown; }
                if (thrown != 3) { std::fprintf(stderr, "rank %d: LocalGemm accepted non-conformal operands (%d of 3 thrown)\n", rank, thrown); failed[rank] = 1; }
                cudaStreamSynchronize(s1); cudaStreamSynchronize(s2); cudaStreamSynchronize(s3);
                comm.Barrier();
                cudaStreamDestroy(s1); cudaStreamDestroy(s2); cudaStreamDestroy(s3);
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
    std::printf("ShimOps: P=%d grid %dx%d n=%lld: three-stream Gemm %s; LocalGemm max|GPU-CPU|=%.3e (bound 1e-10)\n", P, height, P / height,
                (long long)n, f ? "FAILED" : "bit-identical to the synchronised run", w);
    if (w > 1e-10) f = 1;
    std::printf(f ? "FAILED\n" : "PASSED\n");
    return f;
}
