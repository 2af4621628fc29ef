/* TEST INFRASTRUCTURE ONLY -- see oracle/oracle.h.
 *
 * The reference's own SUMMA loop code, compiled from where it lies, over stand-in matrix classes:
 *   src/blas_like/level3/Gemm/{NN,NT,TN,TT}.hpp        SUMMA_{NN,NT,TN,TT}{A,B,C,Dot} and the SUMMA_{NN,NT,TN,TT} dispatchers with
 *                                                      the GEMM_DEFAULT heuristic: all 16 variants
 *   src/blas_like/level3/Trrk/{Local,NN,NT,TN,TT}.hpp  LocalTrrk (recursion + kernels) and the distributed Trrk loops
 *   src/blas_like/level3/Syrk/{LN,LT,UN,UT}.hpp, Symm/{LL,LU,RL,RU}.hpp   their stationary-C loops
 * (This fork's Gemm/NN.hpp also holds a CPU Cannon_NN and timed / NVSHMEM overloads of SUMMA_NNC with MPI calls, CUDA events
 * and NVSHMEM headers at file scope, NN.hpp:12-13,20-103,370-710: those parts are only parsed -- over empty stand-in headers in
 * oracle/ref/config and a few declared-but-never-called names -- never instantiated.)
 *
 * Those headers are templates over El::DistMatrix / DistMatrixReadProxy / LocalGemm / Transpose / AxpyContract ...; the full
 * class hierarchy needs MPI and cannot be built here.  Below is a minimal stand-in of exactly the interface the loops use,
 * implemented on the oracle's simulated grid (orc_dm: all ranks in one process).  What this pins: the ORDER of operations,
 * the temporaries, their distributions and AlignWith targets, the panel ranges and the local-product forms of
 * oracle/oracle_gemm.c and oracle/oracle_level3.c are the reference's -- the restatement must reproduce the result of the
 * reference's loop code bit for bit (tests/test_ref_pin.py).  What it does not pin: the stand-ins' own pieces (redistribution
 * = the closed-form ownership definition, alignment rules = oracle_core.c, local product = orc_local_gemm), which are pinned
 * separately.  Nothing of the reference is copied here.  Built as oracle/_ref/libref_summa.so (links liboracle.so). */
#include "../oracle.h"
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <utility>

struct CUstream_st;
namespace El {
typedef int Int;
using std::string;
enum class Device : unsigned char { CPU, GPU };
namespace DistNS { enum Dist { MC, MD, MR, VC, VR, STAR, CIRC }; }
using namespace DistNS;
namespace DistWrapNS { enum DistWrap { ELEMENT, BLOCK }; }
using namespace DistWrapNS;
namespace OrientationNS { enum Orientation { NORMAL, TRANSPOSE, ADJOINT }; }
using namespace OrientationNS;
namespace UpperOrLowerNS { enum UpperOrLower { LOWER, UPPER }; }
using namespace UpperOrLowerNS;
namespace GemmAlgorithmNS { enum GemmAlgorithm { GEMM_DEFAULT, GEMM_SUMMA_A, GEMM_SUMMA_B, GEMM_SUMMA_C, GEMM_SUMMA_DOT, GEMM_CANNON }; }
using namespace GemmAlgorithmNS;

template <typename C, class T = void> using EnableIf = typename std::enable_if<C::value, T>::type;
template <typename C, class T = void> using DisableIf = typename std::enable_if<!C::value, T>::type;
template <typename T, Device D> struct IsDeviceValidType : std::true_type {};
template <typename T> struct TypeTraits { static T One() { return T(1); } static T Zero() { return T(0); } };
template <typename... Args> [[noreturn]] inline void LogicError(const Args&...) { throw std::logic_error("reference LogicError"); }
template <typename T> inline const T& Min(const T& a, const T& b) { return a < b ? a : b; }
template <typename T> inline const T& Max(const T& a, const T& b) { return a > b ? a : b; }
template <typename... Args> inline void AssertSameGrids(const Args&...) {}
template <typename... Args> inline void Output(const Args&...) {}

static Int g_blocksize = 128, g_local_trrk_blocksize = 64;
inline Int Blocksize() { return g_blocksize; }
template <typename T> inline Int LocalTrrkBlocksize() { return g_local_trrk_blocksize; }

/* names only the fork-specific parts of Gemm/NN.hpp mention (its Cannon_NN and the timed / NVSHMEM overloads of SUMMA_NNC,
 * NN.hpp:20-103,370-710): declared so that the file parses, never instantiated or called */
template <Device D> struct SyncInfo {};
inline Int Mod(Int a, Int b) { const Int r = a % b; return r < 0 ? r + b : r; }
namespace mpi {
struct Comm { int GetMPIComm() const { return 0; } };
template <typename... Args> inline void SendRecv(Args&&...) {}
}  // namespace mpi

class Grid {
public:
    explicit Grid(orc_grid g) : g_(g) {}
    Int Height() const { return g_.r; }
    Int Width() const { return g_.c; }
    Int Size() const { return g_.r * g_.c; }
    /* every rank is simulated at once: there is no "this rank".  SUMMA_NNA_impl reads these into unused locals (NN.hpp:147-152) */
    Int Row() const { return 0; }
    Int Col() const { return 0; }
    Int Rank() const { return 0; }
    const mpi::Comm& RowComm() const { return comm_; }
    const mpi::Comm& ColComm() const { return comm_; }
    const mpi::Comm& Comm() const { return comm_; }
    orc_grid g_;
    mpi::Comm comm_;
};

/* index ranges: A(ALL, IR(k,k+nb)), C(indOuter, indInner), IR(half,END) */
static const Int END = -100;
template <typename I> struct Range {
    I beg, end;
    Range(I b, I e) : beg(b), end(e) {}
};
inline Range<Int> IR(Int b, Int e) { return Range<Int>(b, e); }
static const Range<Int> ALL(0, END);

/* SUMMA_NNA_impl of this fork fetches a CUDA stream from B's local matrix when B is on the GPU (NN.hpp:130-132); the branch is
 * never taken here (GetLocalDevice() is CPU), the types only have to exist */
template <typename T> class AbstractMatrix {};
template <typename T, Device D = Device::CPU> class Matrix : public AbstractMatrix<T> {
public:
    struct ::CUstream_st* Stream() const { return nullptr; }   /* cudaStream_t */
};

template <typename T>
class AbstractDistMatrix {
public:
    virtual ~AbstractDistMatrix() { release(); }
    const AbstractMatrix<T>& LockedMatrix() const { static const Matrix<T, Device::GPU> none; return none; }
    Int Height() const { return h ? (Int)h->m : 0; }
    Int Width() const { return h ? (Int)h->n : 0; }
    Int ColAlign() const { return ca; }
    Int RowAlign() const { return ra; }
    Dist ColDist() const { return cd; }
    Dist RowDist() const { return rd; }
    const El::Grid& Grid() const { return *grid; }
    Device GetLocalDevice() const { return Device::CPU; }
    int stride(int which) const { return orc_dist_stride(grid->g_, which == 0 ? cd : rd); }
    /* (re)create the storage for an m x n matrix with the current distribution and alignments; contents are not kept */
    void remake(Int m, Int n) {
        if (h && owns && !view && h->m == m && h->n == n && h->colAlign == ca && h->rowAlign == ra) return;
        release();
        h = orc_dm_create(grid->g_, m, n, cd, rd, ca, ra, (int)sizeof(T));
        owns = true;
        view = false;
    }
    void Resize(Int m, Int n) { remake(m, n); }
    void release() {
        if (h && owns) orc_dm_free(h);
        h = nullptr;
        owns = false;
    }
    orc_dm* h = nullptr;
    bool owns = false, view = false;
    const El::Grid* grid = nullptr;
    Dist cd = MC, rd = MR;
    bool colCon = false, rowCon = false;
    int ca = 0, ra = 0;
protected:
    AbstractDistMatrix() {}
    AbstractDistMatrix(const AbstractDistMatrix&) = delete;
};

template <typename T> inline string DimsString(const AbstractDistMatrix<T>& A, const string& name) {
    std::ostringstream os; os << name << " ~ " << A.Height() << " x " << A.Width(); return os.str();
}

template <typename T, Dist U = MC, Dist V = MR, DistWrap W = ELEMENT, Device D = Device::CPU>
class DistMatrix : public AbstractDistMatrix<T> {
public:
    explicit DistMatrix(const El::Grid& g) { this->grid = &g; this->cd = U; this->rd = V; this->remake(0, 0); }
    /* alias of an existing matrix of the same distribution (proxies), or owner of a view handle */
    DistMatrix(const El::Grid& g, orc_dm* handle, bool own, bool is_view) {
        this->grid = &g; this->cd = U; this->rd = V; this->h = handle; this->owns = own; this->view = is_view;
        this->ca = handle->colAlign; this->ra = handle->rowAlign; this->colCon = this->rowCon = true;
    }
    DistMatrix(DistMatrix&& o) noexcept { steal(o); }
    DistMatrix& operator=(DistMatrix&& o) = delete;
    /* ElementalMatrix::AlignWith (ElementMatrix.cpp:222-300) through the oracle's table (orc_align_dim_with) */
    void AlignWith(const AbstractDistMatrix<T>& A) {
        for (int which = 0; which < 2; ++which) {
            int a = 0;
            const int k = orc_align_dim_with(U, V, which, this->stride(which), A.cd, A.ca, A.rd, A.ra, &a);
            if (k < 0) LogicError("Nonsensical alignment");
            if (k > 0) { (which ? this->ra : this->ca) = a; (which ? this->rowCon : this->colCon) = true; }
        }
        this->remake(this->Height(), this->Width());
    }
    /* B = A: an unconstrained dimension adopts the alignment the copy:: primitive passes (orc_copy_adopts), then the
     * closed-form redistribution */
    DistMatrix& operator=(const AbstractDistMatrix<T>& A) {
        int a = 0;
        if (!this->colCon && orc_copy_adopts(A.cd, A.ca, U, this->stride(0), &a)) this->ca = a;
        if (!this->rowCon && orc_copy_adopts(A.rd, A.ra, V, this->stride(1), &a)) this->ra = a;
        this->remake(A.Height(), A.Width());
        orc_redist_closed(A.h, this->h, 0);
        return *this;
    }
    DistMatrix& operator=(const DistMatrix& A) { return *this = static_cast<const AbstractDistMatrix<T>&>(A); }
    /* A(rows, cols): a view sharing storage (El::View / LockedView) */
    DistMatrix operator()(Range<Int> I, Range<Int> J) const {
        const Int i0 = I.beg, i1 = I.end == END ? this->Height() : I.end, j0 = J.beg, j1 = J.end == END ? this->Width() : J.end;
        return DistMatrix(*this->grid, orc_dm_view(this->h, i0, j0, i1 - i0, j1 - j0), true, true);
    }
private:
    void steal(DistMatrix& o) {
        this->grid = o.grid; this->cd = o.cd; this->rd = o.rd; this->h = o.h; this->owns = o.owns; this->view = o.view;
        this->ca = o.ca; this->ra = o.ra; this->colCon = o.colCon; this->rowCon = o.rowCon;
        o.h = nullptr; o.owns = false;
    }
};

/* ---- proxies (include/El/core/Proxy.hpp:160-260,694-795): the matrix itself when it already has the wanted distribution
 * (and constrained alignments), else a redistributed temporary; the read-write form copies back on destruction ---- */
struct ElementalProxyCtrl {
    bool colConstrain = false, rowConstrain = false, rootConstrain = false;
    Int colAlign = 0, rowAlign = 0, root = 0;
};
template <typename S, typename T, Dist U = MC, Dist V = MR, DistWrap W = ELEMENT, Device D = Device::CPU>
class DistMatrixReadProxy {
public:
    typedef DistMatrix<T, U, V, W, D> proxy_type;
    DistMatrixReadProxy(const AbstractDistMatrix<S>& A, const ElementalProxyCtrl& ctrl = ElementalProxyCtrl()) { init(A, ctrl); }
    const proxy_type& GetLocked() const { return *prox_; }
protected:
    void init(const AbstractDistMatrix<S>& A, const ElementalProxyCtrl& ctrl) {
        const bool same = A.ColDist() == U && A.RowDist() == V && (!ctrl.colConstrain || A.ColAlign() == ctrl.colAlign) &&
                          (!ctrl.rowConstrain || A.RowAlign() == ctrl.rowAlign);
        if (same) { prox_.reset(new proxy_type(A.Grid(), A.h, false, A.view)); copied_ = false; return; }
        prox_.reset(new proxy_type(A.Grid()));
        if (ctrl.colConstrain) { prox_->ca = ctrl.colAlign; prox_->colCon = true; }
        if (ctrl.rowConstrain) { prox_->ra = ctrl.rowAlign; prox_->rowCon = true; }
        *prox_ = A;
        copied_ = true;
    }
    std::unique_ptr<proxy_type> prox_;
    bool copied_ = false;
};
template <typename S, typename T, Dist U = MC, Dist V = MR, DistWrap W = ELEMENT, Device D = Device::CPU>
class DistMatrixReadWriteProxy : public DistMatrixReadProxy<S, T, U, V, W, D> {
public:
    DistMatrixReadWriteProxy(AbstractDistMatrix<S>& A, const ElementalProxyCtrl& ctrl = ElementalProxyCtrl())
        : DistMatrixReadProxy<S, T, U, V, W, D>(A, ctrl), orig_(&A) {}
    ~DistMatrixReadWriteProxy() { if (this->copied_) orc_redist_closed(this->prox_->h, orig_->h, 0); }
    typename DistMatrixReadProxy<S, T, U, V, W, D>::proxy_type& Get() { return *this->prox_; }
private:
    AbstractDistMatrix<S>* orig_;
};

/* ---- operations the loops call ---- */
template <typename T>
void LocalGemm(Orientation oA, Orientation oB, T alpha, const AbstractDistMatrix<T>& A, const AbstractDistMatrix<T>& B, T beta,
               AbstractDistMatrix<T>& C) {   /* Gemm.cpp:323-431: blas::Gemm on every rank's blocks */
    const int tA = oA != NORMAL, tB = oB != NORMAL;
    for (int p = 0; p < C.h->P; ++p) {
        const int64_t k = tA ? A.h->lh[p] : A.h->lw[p];
        orc_local_gemm(tA, tB, C.h->lh[p], C.h->lw[p], k, (double)alpha, A.h->loc[p], A.h->ld[p], B.h->loc[p], B.h->ld[p], (double)beta,
                       C.h->loc[p], C.h->ld[p], C.h->es);
    }
}
template <typename T>
void LocalGemm(Orientation oA, Orientation oB, T alpha, const AbstractDistMatrix<T>& A, const AbstractDistMatrix<T>& B,
               AbstractDistMatrix<T>& C) {   /* Gemm.cpp:434-445 */
    C.Resize(oA == NORMAL ? A.Height() : A.Width(), oB == NORMAL ? B.Width() : B.Height());
    LocalGemm(oA, oB, alpha, A, B, T(0), C);
}
template <typename T>
void adopt_for_transpose_or_copy(const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& B, bool transpose) {
    int a = 0;
    if (!B.colCon && orc_copy_adopts(transpose ? A.rd : A.cd, transpose ? A.ra : A.ca, B.cd, B.stride(0), &a)) B.ca = a;
    if (!B.rowCon && orc_copy_adopts(transpose ? A.cd : A.rd, transpose ? A.ca : A.ra, B.rd, B.stride(1), &a)) B.ra = a;
}
template <typename T>
void Transpose(const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& B, bool = false) {   /* Transpose.hpp:191-356 */
    adopt_for_transpose_or_copy(A, B, true);
    B.remake(A.Width(), A.Height());
    orc_redist_closed(A.h, B.h, 1);
}
template <typename T>
void AxpyContract(T alpha, const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& B) {   /* AxpyContract.hpp:483-544 */
    orc_axpy_contract_closed((double)alpha, A.h, B.h, 0);
}
template <typename T>
void TransposeAxpyContract(T alpha, const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& B, bool = false) {
    orc_axpy_contract_closed((double)alpha, A.h, B.h, 1);   /* TransposeAxpyContract.hpp:14-48 */
}
template <typename T>
void Contract(const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& B) {   /* Contract.hpp:14-75 */
    if (A.cd == B.cd && A.rd == B.rd) {
        adopt_for_transpose_or_copy(A, B, false);
        B.remake(A.Height(), A.Width());
        orc_redist_closed(A.h, B.h, 0);
        return;
    }
    /* AlignAndResize / AlignColsAndResize / AlignRowsAndResize(force = false): an unconstrained dimension whose distribution is
     * A's, or whose Partial() is, takes A's alignment */
    if (!B.colCon && (A.cd == B.cd || (int)A.cd == orc_dist_partial(B.cd))) B.ca = A.ca;
    if (!B.rowCon && (A.rd == B.rd || (int)A.rd == orc_dist_partial(B.rd))) B.ra = A.ra;
    B.remake(A.Height(), A.Width());
    orc_dm_fill_const(B.h, 0.0);
    orc_axpy_contract_closed(1.0, A.h, B.h, 0);
}
template <typename T>
void Axpy(T alpha, const AbstractDistMatrix<T>& X, AbstractDistMatrix<T>& Y) {   /* Axpy.hpp:161-187 */
    if (orc_axpy((double)alpha, X.h, Y.h, 0)) LogicError("Axpy");
}
/* LocalAxpyTrapezoid / ScaleTrapezoid on DistMatrices: the oracle's loops, themselves pinned to the reference's
 * AxpyTrapezoid.hpp / ScaleTrapezoid.hpp by oracle/ref/ref_trapezoid.cpp */
template <typename T>
void LocalAxpyTrapezoid(UpperOrLower uplo, T alpha, const AbstractDistMatrix<T>& X, AbstractDistMatrix<T>& Y, Int offset = 0) {
    if (orc_axpy_trapezoid((int)uplo, (double)alpha, X.h, Y.h, offset)) LogicError("LocalAxpyTrapezoid");
}
template <typename T, typename S>
void ScaleTrapezoid(S alpha, UpperOrLower uplo, AbstractDistMatrix<T>& A, Int offset = 0) { orc_scale_trapezoid((double)alpha, (int)uplo, A.h, offset); }
template <typename T>
void MakeTrapezoidal(UpperOrLower uplo, AbstractDistMatrix<T>& A, Int offset = 0) { orc_make_trapezoidal((int)uplo, A.h, offset); }
}  // namespace El

typedef int MPI_Comm;
inline int MPI_Comm_rank(MPI_Comm, int*) { return 0; }
inline int MPI_Barrier(MPI_Comm) { return 0; }

#define EL_DEBUG_CSE
#define EL_DEBUG_ONLY(x)
#define AUTO_PROFILE_REGION(...)   /* the reference's profiling hook (NN.hpp:251,800): nothing to profile here */
#include <cstdio>
#include <vector>
#include <cuda_runtime.h>   /* cudaStream_t / cudaEvent_t in the never-instantiated timed overloads of NN.hpp */
#include <blas_like/level3/Gemm/NN.hpp>
#include <blas_like/level3/Gemm/NT.hpp>
#include <blas_like/level3/Gemm/TN.hpp>
#include <blas_like/level3/Gemm/TT.hpp>
#include <blas_like/level3/Trrk/Local.hpp>
#include <blas_like/level3/Trrk/NN.hpp>
#include <blas_like/level3/Trrk/NT.hpp>
#include <blas_like/level3/Trrk/TN.hpp>
#include <blas_like/level3/Trrk/TT.hpp>
#include <blas_like/level3/Syrk/LN.hpp>
#include <blas_like/level3/Syrk/LT.hpp>
#include <blas_like/level3/Syrk/UN.hpp>
#include <blas_like/level3/Syrk/UT.hpp>
#include <blas_like/level3/Symm/LL.hpp>
#include <blas_like/level3/Symm/LU.hpp>
#include <blas_like/level3/Symm/RL.hpp>
#include <blas_like/level3/Symm/RU.hpp>

namespace {
/* a caller's matrix (any distribution, decided at run time) as the AbstractDistMatrix the loops take */
template <typename T>
struct Wrapped : El::AbstractDistMatrix<T> {
    Wrapped(const El::Grid& g, orc_dm* handle) {
        this->grid = &g; this->h = handle; this->owns = false; this->view = false;
        this->cd = (El::Dist)handle->colDist; this->rd = (El::Dist)handle->rowDist;
        this->ca = handle->colAlign; this->ra = handle->rowAlign; this->colCon = this->rowCon = true;
    }
};
inline El::Orientation orient(int o) { return o == 0 ? El::NORMAL : (o == 1 ? El::TRANSPOSE : El::ADJOINT); }

template <typename T>
int summa(int oA, int oB, int alg, double alpha, orc_dm* A, orc_dm* B, double beta, orc_dm* C) {
    using namespace El;
    El::Grid g(C->g);
    Wrapped<T> a(g, A), b(g, B), c(g, C);
    const Orientation orA = orient(oA), orB = orient(oB);
    orc_scale(beta, C);                                   /* El::Gemm, Gemm.cpp:287: C *= beta, then the orientation switch :288-306 */
    if (orA == NORMAL && orB == NORMAL) gemm::SUMMA_NN(T(alpha), a, b, c, (GemmAlgorithm)alg);
    else if (orA == NORMAL) gemm::SUMMA_NT(orB, T(alpha), a, b, c, (GemmAlgorithm)alg);
    else if (orB == NORMAL) gemm::SUMMA_TN(orA, T(alpha), a, b, c, (GemmAlgorithm)alg);
    else gemm::SUMMA_TT(orA, orB, T(alpha), a, b, c, (GemmAlgorithm)alg);
    return 0;
}
template <typename T>
int trrk(int uplo, int oA, int oB, double alpha, orc_dm* A, orc_dm* B, double beta, orc_dm* C) {
    using namespace El;
    El::Grid g(C->g);
    Wrapped<T> a(g, A), b(g, B), c(g, C);
    const Orientation orA = orient(oA), orB = orient(oB);
    const UpperOrLower ul = uplo ? UPPER : LOWER;
    ScaleTrapezoid(T(beta), ul, c);                       /* El::Trrk on AbstractDistMatrix, Trrk.cpp:98-115 */
    if (orA == NORMAL && orB == NORMAL) trrk::TrrkNN(ul, T(alpha), a, b, c);
    else if (orA == NORMAL) trrk::TrrkNT(ul, orB, T(alpha), a, b, c);
    else if (orB == NORMAL) trrk::TrrkTN(ul, orA, T(alpha), a, b, c);
    else trrk::TrrkTT(ul, orA, orB, T(alpha), a, b, c);
    return 0;
}
}  // namespace

namespace {
template <typename T>
int syrk(int uplo, int orientation, double alpha, orc_dm* A, double beta, orc_dm* C) {
    using namespace El;
    El::Grid g(C->g);
    Wrapped<T> a(g, A), c(g, C);
    const UpperOrLower ul = uplo ? UPPER : LOWER;
    ScaleTrapezoid(T(beta), ul, c);                       /* El::Syrk on AbstractDistMatrix, Syrk.cpp:72-90 */
    /* the stationary-C loops; syrk::LN ... also route k > 10 n to their Dot variants (LN.hpp:152-159), which need the Matrix-level
     * Syrk and are not instantiated here */
    if (ul == LOWER && orientation == 0) syrk::LN_C(T(alpha), a, c, false);
    else if (ul == LOWER) syrk::LT_C(T(alpha), a, c, false);
    else if (orientation == 0) syrk::UN_C(T(alpha), a, c, false);
    else syrk::UT_C(T(alpha), a, c, false);
    return 0;
}
template <typename T>
int symm(int side, int uplo, double alpha, orc_dm* A, orc_dm* B, double beta, orc_dm* C) {
    using namespace El;
    El::Grid g(C->g);
    Wrapped<T> a(g, A), b(g, B), c(g, C);
    orc_scale(beta, C);                                   /* El::Symm on AbstractDistMatrix, Symm.cpp:56-80: C *= beta */
    /* the stationary-C loops; symm::LL ... also route tall problems to their stationary-A variants (LL.hpp:226-229) */
    if (side == 0 && uplo == 0) symm::LLC(T(alpha), a, b, c, false);
    else if (side == 0) symm::LUC(T(alpha), a, b, c, false);
    else if (uplo == 0) symm::RLC(T(alpha), a, b, c, false);
    else symm::RUC(T(alpha), a, b, c, false);
    return 0;
}
}  // namespace

extern "C" {
int ref_summa_syrk(int uplo, int orientation, double alpha, orc_dm* A, double beta, orc_dm* C) {
    try {
        return C->es == 8 ? syrk<double>(uplo, orientation, alpha, A, beta, C) : syrk<float>(uplo, orientation, alpha, A, beta, C);
    } catch (const std::exception&) {
        return -1;
    }
}
int ref_summa_symm(int side, int uplo, double alpha, orc_dm* A, orc_dm* B, double beta, orc_dm* C) {
    try {
        return C->es == 8 ? symm<double>(side, uplo, alpha, A, B, beta, C) : symm<float>(side, uplo, alpha, A, B, beta, C);
    } catch (const std::exception&) {
        return -1;
    }
}
void ref_summa_set_blocksize(int nb) { El::g_blocksize = nb; }
void ref_summa_set_local_trrk_blocksize(int nb) { El::g_local_trrk_blocksize = nb; }
/* El::Gemm(orientA, orientB, alpha, A, B, beta, C, alg) on the simulated grid through the reference's SUMMA_{NT,TN,TT} code.
 * Returns 0, or -1 if the reference raised LogicError. */
int ref_summa_gemm(int oA, int oB, int alg, double alpha, orc_dm* A, orc_dm* B, double beta, orc_dm* C) {
    try {
        return C->es == 8 ? summa<double>(oA, oB, alg, alpha, A, B, beta, C) : summa<float>(oA, oB, alg, alpha, A, B, beta, C);
    } catch (const std::exception&) {
        return -1;
    }
}
/* El::Trrk(uplo, orientA, orientB, alpha, A, B, beta, C) through the reference's Trrk/{NN,NT,TN,TT}.hpp and LocalTrrk */
int ref_summa_trrk(int uplo, int oA, int oB, double alpha, orc_dm* A, orc_dm* B, double beta, orc_dm* C) {
    try {
        return C->es == 8 ? trrk<double>(uplo, oA, oB, alpha, A, B, beta, C) : trrk<float>(uplo, oA, oB, alpha, A, B, beta, C);
    } catch (const std::exception&) {
        return -1;
    }
}
}
