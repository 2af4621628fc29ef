/* TEST INFRASTRUCTURE ONLY -- see oracle/oracle.h.
 *
 * The reference's redistribution primitives, compiled from where they lie and run SPMD -- one thread per simulated rank --
 * over stand-in classes: include/El/blas_like/level1/Copy/{RowAllGather,ColAllGather,...,Translate}.hpp and AxpyContract.hpp.  The per-portion loops they
 * call are the reference's own (Copy/util.hpp, already used by oracle/ref/ref_util.cpp); the stand-ins supply what sits
 * below and around them: an ElementalMatrix view of one rank's block of an orc_dm, and mpi::AllGather / SendRecv /
 * Broadcast / AllToAll over thread barriers with the data order MPI defines (portions in communicator-rank order).
 * This pins oracle_core.c's restated primitive chains (orc_row_allgather ...) -- packing order, portion sizes, the
 * unaligned SendRecv partners -- to the reference's code.  Nothing of the reference is copied here.
 * Built into oracle/_ref/libref_copy.so (links liboracle.so). */
#include "ref_prelude.hpp"

#include <El/core/indexing/decl.hpp>
#include <El/core/indexing/impl.hpp>
#include <El/blas_like/level1/Copy/util.hpp>
#include <El/blas_like/level1/Axpy/util.hpp>

#include "../oracle.h"
#include <condition_variable>
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <thread>
#include <vector>

namespace El {
/* distribution algebra (types.hpp:292-337): the oracle's tables, themselves pinned to types.hpp by oracle/ref/ref_types.cpp */
inline Dist Collect(Dist U) { return (Dist)orc_dist_collect(U); }
inline Dist Partial(Dist U) { return (Dist)orc_dist_partial(U); }
inline Dist PartialUnionRow(Dist U, Dist V) { return (Dist)orc_dist_partial_union_row(U, V); }
inline Dist PartialUnionCol(Dist U, Dist V) { return (Dist)orc_dist_partial_union_col(U, V); }

template <Dist U> constexpr Dist Collect() { return U == CIRC ? CIRC : STAR; }
template <Dist U> constexpr Dist Partial() { return U == VC ? MC : (U == VR ? MR : U); }
template <Dist U, Dist V> constexpr Dist PartialUnionRow() { return U == VC ? MR : (U == VR ? MC : V); }
template <Dist U, Dist V> constexpr Dist PartialUnionCol() { return PartialUnionRow<V, U>(); }
template <Dist U, Dist V> constexpr Dist ProductDist() {   /* types.hpp:343-356 */
    return U == STAR ? V : (V == STAR ? U : ((U == MC && V == MR) ? VC : ((U == MR && V == MC) ? VR : CIRC)));
}
template <Dist U, Dist V> constexpr Dist ProductDistPartner() { return (U == CIRC && V == CIRC) ? CIRC : STAR; }
template <typename T, Device D> struct IsDeviceValidType : std::true_type {};
template <typename T> struct TypeTraits { static T One() { return T(1); } static T Zero() { return T(0); } };

/* ---- SPMD runtime ---- */
struct Barrier {
    explicit Barrier(int n_) : n(n_) {}
    void wait() {
        std::unique_lock<std::mutex> lk(m);
        const int g = gen;
        if (++count == n) { count = 0; ++gen; cv.notify_all(); }
        else cv.wait(lk, [&] { return gen != g; });
    }
    std::mutex m;
    std::condition_variable cv;
    int n, count = 0, gen = 0;
};
struct CommShared {              /* one communicator instance: its members share a barrier, a pointer board and mailboxes */
    explicit CommShared(int n) : bar(n), board(n, nullptr), box((size_t)n * n, nullptr), size(n) {}
    Barrier bar;
    std::vector<const void*> board;
    std::vector<const void*> box;    /* box[from*size + to]: a message in flight (point-to-point) */
    std::mutex m;
    std::condition_variable cv;
    int size;
};
enum class Collective { SENDRECV };

namespace mpi {
const int MIN_COLL_MSG = 1;                                  /* include/El/core/imports/mpi.hpp:181-185 */
inline int Pad(int count) { return count > MIN_COLL_MSG ? count : MIN_COLL_MSG; }
struct Comm {
    CommShared* sh = nullptr;
    int rank = 0;
    int Size() const { return sh ? sh->size : 1; }
    int Rank() const { return rank; }
};
/* portions land in communicator-rank order */
template <typename T, Device D>
void AllGather(const T* sbuf, int sc, T* rbuf, int rc, const Comm& c, SyncInfo<D> const&) {
    c.sh->board[c.rank] = sbuf;
    c.sh->bar.wait();
    for (int q = 0; q < c.sh->size; ++q) std::memcpy(rbuf + (size_t)q * rc, c.sh->board[q], (size_t)sc * sizeof(T));
    c.sh->bar.wait();
}
template <typename T, Device D>
void AllToAll(const T* sbuf, int sc, T* rbuf, int rc, const Comm& c, SyncInfo<D> const&) {
    c.sh->board[c.rank] = sbuf;
    c.sh->bar.wait();
    for (int q = 0; q < c.sh->size; ++q)
        std::memcpy(rbuf + (size_t)q * rc, static_cast<const T*>(c.sh->board[q]) + (size_t)c.rank * sc, (size_t)sc * sizeof(T));
    c.sh->bar.wait();
}
/* point-to-point (not every member of the communicator takes part: Exchange.hpp:36-41): post to the partner's mailbox, take
 * what the other partner posted, and keep the send buffer alive until it has been taken */
template <typename T, Device D>
void SendRecv(const T* sbuf, int sc, int to, T* rbuf, int rc, int from, const Comm& c, SyncInfo<D> const&) {
    (void)sc;
    CommShared& s = *c.sh;
    std::unique_lock<std::mutex> lk(s.m);
    const size_t out = (size_t)c.rank * s.size + to, in = (size_t)from * s.size + c.rank;
    s.cv.wait(lk, [&] { return s.box[out] == nullptr; });
    s.box[out] = sbuf ? static_cast<const void*>(sbuf) : static_cast<const void*>(&s);   /* an empty message still has to be seen */
    s.cv.notify_all();
    s.cv.wait(lk, [&] { return s.box[in] != nullptr; });
    if (rc > 0) std::memcpy(rbuf, s.box[in], (size_t)rc * sizeof(T));
    s.box[in] = nullptr;
    s.cv.notify_all();
    s.cv.wait(lk, [&] { return s.box[out] == nullptr; });
}
/* reductions: member q's portion k summed over q in ASCENDING communicator rank (MPI leaves the order open; the oracle's
 * orc_axpy_contract_closed fixes it the same way).  In-place form (buf holds size*rc entries, the result lands at its front),
 * two-buffer form, and Reduce to a root. */
template <typename T>
inline void reduce_from_board(const Comm& c, T* out, size_t offset, int count) {
    for (int i = 0; i < count; ++i) {
        T sum = static_cast<const T*>(c.sh->board[0])[offset + i];
        for (int q = 1; q < c.sh->size; ++q) sum += static_cast<const T*>(c.sh->board[q])[offset + i];
        out[i] = sum;
    }
}
template <typename T, Device D>
void ReduceScatter(const T* sbuf, T* rbuf, int rc, const Comm& c, SyncInfo<D> const&) {
    c.sh->board[c.rank] = sbuf;
    c.sh->bar.wait();
    reduce_from_board(c, rbuf, (size_t)c.rank * rc, rc);
    c.sh->bar.wait();
}
template <typename T, Device D>
void ReduceScatter(T* buf, int rc, const Comm& c, SyncInfo<D> const& si) {
    std::vector<T> tmp((size_t)(rc > 0 ? rc : 1));
    ReduceScatter(static_cast<const T*>(buf), tmp.data(), rc, c, si);   /* the barrier inside orders the overwrite below */
    std::memcpy(buf, tmp.data(), (size_t)rc * sizeof(T));
}
template <typename T, Device D>
void Reduce(const T* sbuf, T* rbuf, int count, int root, const Comm& c, SyncInfo<D> const&) {
    c.sh->board[c.rank] = sbuf;
    c.sh->bar.wait();
    if (c.rank == root) reduce_from_board(c, rbuf, 0, count);
    c.sh->bar.wait();
}
/* in-place exchange (Copy/Translate.hpp:93-95): send buf to `to`, overwrite it with what `from` sends */
template <typename T, Device D>
void SendRecv(T* buf, int count, int to, int from, const Comm& c, SyncInfo<D> const& si) {
    std::vector<T> tmp(buf, buf + count);
    SendRecv(static_cast<const T*>(tmp.data()), count, to, buf, count, from, c, si);
}
/* only reached when the roots of A and B differ, which never happens here (every matrix has root 0) */
template <typename T, Device D> void Send(const T*, int, int, const Comm&, SyncInfo<D> const&) { throw std::logic_error("mpi::Send: not simulated"); }
template <typename T, Device D> void Recv(T*, int, int, const Comm&, SyncInfo<D> const&) { throw std::logic_error("mpi::Recv: not simulated"); }
inline int Rank(const Comm& c) { return c.rank; }
inline int Size(const Comm& c) { return c.Size(); }
template <typename T, Collective C, Device D> inline void EnsureComm(const Comm&, SyncInfo<D> const&) {}
template <typename T, Device D>
void Broadcast(T* buf, int count, int root, const Comm& c, SyncInfo<D> const&) {
    if (c.rank == root) c.sh->board[root] = buf;
    c.sh->bar.wait();
    if (c.rank != root) std::memcpy(buf, c.sh->board[root], (size_t)count * sizeof(T));
    c.sh->bar.wait();
}
}  // namespace mpi
inline bool CongruentToCommSelf(const mpi::Comm& c) { return c.Size() == 1; }

template <typename T, Device D>
class simple_buffer {            /* include/hydrogen/utils/SimpleBuffer.hpp */
public:
    simple_buffer() {}
    simple_buffer(size_t n, SyncInfo<D> const&) : v_(n) {}
    simple_buffer(size_t n, T const& value, SyncInfo<D> const&) : v_(n, value) {}
    void allocate(size_t n) { v_.assign(n, T()); }
    T* data() { return v_.data(); }
private:
    std::vector<T> v_;
};
template <typename... A> inline int MakeMultiSync(const A&...) { return 0; }

/* ---- one rank's view of a simulated distributed matrix ---- */
struct World {                   /* shared by all rank threads of one call */
    orc_grid g;
    int P;
    Barrier all;
    std::vector<CommShared*> mc, mr, vc, vr, self;   /* communicator instances: MC comm of grid column j = mc[j], ... */
    explicit World(orc_grid g_) : g(g_), P(g_.r * g_.c), all(P) {
        for (int j = 0; j < g.c; ++j) mc.push_back(new CommShared(g.r));
        for (int i = 0; i < g.r; ++i) mr.push_back(new CommShared(g.c));
        vc.push_back(new CommShared(P));
        vr.push_back(new CommShared(P));
        for (int p = 0; p < P; ++p) self.push_back(new CommShared(1));
    }
    ~World() { for (auto* v : {&mc, &mr, &vc, &vr, &self}) for (auto* c : *v) delete c; }
};
struct GridView {
    bool InGrid() const { return true; }
    int Rank() const { return rank; }
    bool operator!=(const GridView&) const { return false; }   /* one grid per call */
    int rank = 0;
};
typedef GridView Grid;
template <typename M, Device D> inline void SetSyncInfo(M&, SyncInfo<D> const&) {}

template <typename T, Device D>
class Matrix {                   /* the local block */
public:
    Matrix(Int h, Int w, T* buf, Int ld) : h_(h), w_(w), ld_(ld), buf_(buf) {}
    Int Height() const { return h_; }
    Int Width() const { return w_; }
    Int LDim() const { return ld_; }
    T* Buffer() { return buf_; }
    const T* LockedBuffer() const { return buf_; }
    Matrix& operator=(const Matrix& A) {   /* B.Matrix() = A.LockedMatrix(): a copy of the entries */
        for (Int j = 0; j < w_; ++j) std::memcpy(buf_ + (size_t)j * ld_, A.buf_ + (size_t)j * A.ld_, (size_t)h_ * sizeof(T));
        return *this;
    }
private:
    Int h_, w_, ld_;
    T* buf_;
};
template <typename T, Device D>
inline SyncInfo<D> SyncInfoFromMatrix(const Matrix<T, D>&) { return SyncInfo<D>{}; }
template <typename T, Device D>
inline void Copy(const Matrix<T, D>& A, Matrix<T, D> B) { B = A; }
template <typename T, Device D>
inline void Zero(Matrix<T, D> A) {
    for (Int j = 0; j < A.Width(); ++j) std::memset(static_cast<void*>(A.Buffer() + (size_t)j * A.LDim()), 0, (size_t)A.Height() * sizeof(T));
}   /* B is a handle onto the block */

template <typename T>
class ElementalMatrix {
public:
    ElementalMatrix(World* w, orc_dm** slot, int rank) : w_(w), slot_(slot), p_(rank) {
        orc_grid_coords(w->g, rank, &mc_, &mr_);
        grid_.rank = rank;
    }
    orc_dm* h() const { return *slot_; }
    Int Height() const { return (Int)h()->m; }
    Int Width() const { return (Int)h()->n; }
    Int ColAlign() const { return h()->colAlign; }
    Int RowAlign() const { return h()->rowAlign; }
    Dist ColDist() const { return (Dist)h()->colDist; }
    Dist RowDist() const { return (Dist)h()->rowDist; }
    Int ColStride() const { return orc_dist_stride(w_->g, h()->colDist); }
    Int RowStride() const { return orc_dist_stride(w_->g, h()->rowDist); }
    Int ColRank() const { return orc_dist_rank(w_->g, h()->colDist, mc_, mr_); }
    Int RowRank() const { return orc_dist_rank(w_->g, h()->rowDist, mc_, mr_); }
    Int ColShift() const { return (Int)h()->colShift[p_]; }
    Int RowShift() const { return (Int)h()->rowShift[p_]; }
    Int LocalHeight() const { return (Int)h()->lh[p_]; }
    Int LocalWidth() const { return (Int)h()->lw[p_]; }
    Int LDim() const { return (Int)h()->ld[p_]; }
    T* Buffer() { return static_cast<T*>(h()->loc[p_]); }
    const T* LockedBuffer() const { return static_cast<const T*>(h()->loc[p_]); }
    const T* LockedBuffer(Int iLoc, Int jLoc) const { return LockedBuffer() + iLoc + (size_t)jLoc * LDim(); }
    T* Buffer(Int iLoc, Int jLoc) { return Buffer() + iLoc + (size_t)jLoc * LDim(); }
    El::Matrix<T, Device::CPU> Matrix() { return El::Matrix<T, Device::CPU>(LocalHeight(), LocalWidth(), Buffer(), LDim()); }
    El::Matrix<T, Device::CPU> LockedMatrix() const {
        return El::Matrix<T, Device::CPU>(LocalHeight(), LocalWidth(), const_cast<T*>(LockedBuffer()), LDim());
    }
    bool Participating() const { return true; }
    bool ColConstrained() const { return colCon; }
    bool RowConstrained() const { return rowCon; }
    bool RootConstrained() const { return true; }
    void SetGrid(const GridView&) {}
    void SetRoot(int) {}
    Int CrossRank() const { return 0; }
    void AlignCols(Int ca, bool = true) { realign(ca, RowAlign(), Height(), Width()); }
    void AlignRows(Int ra, bool = true) { realign(ColAlign(), ra, Height(), Width()); }
    const GridView& Grid() const { return grid_; }
    int Root() const { return 0; }
    Device GetLocalDevice() const { return Device::CPU; }
    mpi::Comm comm_of(int dist) const {
        mpi::Comm c;
        switch (dist) {
            case MC: c.sh = w_->mc[mr_]; c.rank = mc_; break;
            case MR: c.sh = w_->mr[mc_]; c.rank = mr_; break;
            case VC: c.sh = w_->vc[0]; c.rank = mc_ + w_->g.r * mr_; break;
            case VR: c.sh = w_->vr[0]; c.rank = mr_ + w_->g.c * mc_; break;
            default: c.sh = w_->self[p_]; c.rank = 0;
        }
        return c;
    }
    mpi::Comm ColComm() const { return comm_of(h()->colDist); }
    mpi::Comm RowComm() const { return comm_of(h()->rowDist); }
    mpi::Comm CrossComm() const { return comm_of(STAR); }
    /* per-distribution definitions at the bottom of src/core/DistMatrix/ElementMatrix/<dist>.cpp (e.g. VR_STAR.cpp:159-247):
     * Partial(VC) = MC, Partial(VR) = MR; the partial UNION of VC is the MR communicator, of VR the MC one, of anything else
     * COMM_SELF (stride 1, rank 0); DistComm is the communicator of the product distribution */
    static int part(int d) { return d == VC ? MC : (d == VR ? MR : d); }
    static int puni(int d) { return d == VC ? MR : (d == VR ? MC : STAR); }
    int prod() const {
        const int U = h()->colDist, V = h()->rowDist;
        return U == STAR ? V : (V == STAR ? U : ((U == MC && V == MR) ? VC : ((U == MR && V == MC) ? VR : STAR)));
    }
    int rank_in(int d) const { return orc_dist_rank(w_->g, d, mc_, mr_); }
    int size_of(int d) const { return orc_dist_stride(w_->g, d); }
    mpi::Comm PartialColComm() const { return comm_of(part(h()->colDist)); }
    mpi::Comm PartialRowComm() const { return comm_of(part(h()->rowDist)); }
    mpi::Comm PartialUnionColComm() const { return comm_of(puni(h()->colDist)); }
    mpi::Comm PartialUnionRowComm() const { return comm_of(puni(h()->rowDist)); }
    mpi::Comm DistComm() const { return comm_of(prod()); }
    Int PartialColStride() const { return size_of(part(h()->colDist)); }
    Int PartialRowStride() const { return size_of(part(h()->rowDist)); }
    Int PartialUnionColStride() const { return size_of(puni(h()->colDist)); }
    Int PartialUnionRowStride() const { return size_of(puni(h()->rowDist)); }
    Int PartialColRank() const { return rank_in(part(h()->colDist)); }
    Int PartialRowRank() const { return rank_in(part(h()->rowDist)); }
    Int PartialUnionColRank() const { return rank_in(puni(h()->colDist)); }
    Int PartialUnionRowRank() const { return rank_in(puni(h()->rowDist)); }
    Int DistSize() const { return size_of(prod()); }
    Int DistRank() const { return rank_in(prod()); }
    /* collective: every rank calls it with the same arguments; one of them re-creates the shared storage */
    void realign(int ca, int ra, Int m, Int n) {
        w_->all.wait();
        if (p_ == 0) {
            orc_dm* old = *slot_;
            if (old->colAlign != ca || old->rowAlign != ra || old->m != m || old->n != n) {
                *slot_ = orc_dm_create(w_->g, m, n, old->colDist, old->rowDist, ca, ra, old->es);
                orc_dm_free(old);
            }
        }
        w_->all.wait();
    }
    void AlignColsAndResize(Int ca, Int m, Int n, bool = false, bool = false) { realign(colCon ? ColAlign() : ca, RowAlign(), m, n); }
    void AlignRowsAndResize(Int ra, Int m, Int n, bool = false, bool = false) { realign(ColAlign(), rowCon ? RowAlign() : ra, m, n); }
    void AlignAndResize(Int ca, Int ra, Int m, Int n, bool = false, bool = false) {
        realign(colCon ? ColAlign() : ca, rowCon ? RowAlign() : ra, m, n);
    }
    void Resize(Int m, Int n) { realign(ColAlign(), RowAlign(), m, n); }
    bool colCon = false, rowCon = false;
private:
    World* w_;
    orc_dm** slot_;
    int p_, mc_ = 0, mr_ = 0;
    GridView grid_;
};
/* the typed form some primitives take (PartialColAllGather, the AllToAll promotions / demotions, Exchange ...) */
template <typename T, Dist U, Dist V, DistWrap W, Device D>
class DistMatrix : public ElementalMatrix<T> {
public:
    using ElementalMatrix<T>::ElementalMatrix;
};
/* El::Broadcast(A, comm, root) (include/El/blas_like/level1/Broadcast.hpp): the root's local matrix to every member, which all
 * hold a block of the same shape; column by column here */
template <typename T>
inline void Broadcast(ElementalMatrix<T>& B, const mpi::Comm& comm, int root) {
    if (comm.Size() == 1) return;
    for (Int j = 0; j < B.LocalWidth(); ++j) mpi::Broadcast(B.Buffer(0, j), (int)B.LocalHeight(), root, comm, SyncInfo<Device::CPU>{});
}
/* names only the BlockMatrix overloads in the same headers mention: declared so that the files parse, never instantiated */
using std::vector;
template <typename T> class BlockMatrix;
template <typename... A> inline void AssertSameGrids(const A&...) {}
template <typename... A> inline void GeneralPurpose(const A&...) {}
template <typename V> inline void FastResize(V& v, size_t n) { v.resize(n); }
template <typename... A> inline void SendRecv(const A&...) {}
}  // namespace El

#define EL_DEBUG_CSE
#include <El/blas_like/level1/Copy/RowAllGather.hpp>
#include <El/blas_like/level1/Copy/ColAllGather.hpp>
#include <El/blas_like/level1/Copy/PartialColAllGather.hpp>
#include <El/blas_like/level1/Copy/PartialRowAllGather.hpp>
#include <El/blas_like/level1/Copy/ColAllToAllDemote.hpp>
#include <El/blas_like/level1/Copy/RowAllToAllDemote.hpp>
#include <El/blas_like/level1/Copy/ColAllToAllPromote.hpp>
#include <El/blas_like/level1/Copy/RowAllToAllPromote.hpp>
#include <El/blas_like/level1/Copy/Exchange.hpp>
#include <El/blas_like/level1/Copy/ColFilter.hpp>
#include <El/blas_like/level1/Copy/RowFilter.hpp>
#include <El/blas_like/level1/Copy/PartialColFilter.hpp>
#include <El/blas_like/level1/Copy/PartialRowFilter.hpp>
namespace El { namespace copy {
template <typename A, typename B> inline void TranslateBetweenGrids(const A&, B&) { throw std::logic_error("TranslateBetweenGrids: one grid only"); }
} }
#include <El/blas_like/level1/Copy/Translate.hpp>
namespace El {
/* same-distribution Axpy (AxpyContract's first branch): Y += alpha X on the local blocks, through the reference's own loop */
template <typename T>
inline void Axpy(T alpha, const ElementalMatrix<T>& X, ElementalMatrix<T>& Y) {
    axpy::util::InterleaveMatrixUpdate(alpha, X.LocalHeight(), X.LocalWidth(), X.LockedBuffer(), Int(1), X.LDim(), Y.Buffer(), Int(1), Y.LDim(),
                                       SyncInfo<Device::CPU>{});
}
}  // namespace El
/* no explicit-instantiation declarations at the end of the header: the routines are instantiated here, from its definitions */
#define PROTO_INT(T)
#define PROTO_REAL(T)
#define PROTO_COMPLEX(T)
#include <El/blas_like/level1/AxpyContract.hpp>

namespace {
using namespace El;
/* run `body(a, b)` on one thread per rank */
template <typename T, typename F>
void spmd(orc_dm* A, orc_dm** B, int colCon, int rowCon, F body) {
    World w(A->g);
    std::vector<std::thread> th;
    orc_dm* a_slot = A;
    std::vector<int> threw(w.P, 0);
    for (int p = 0; p < w.P; ++p)
        th.emplace_back([&, p] {
            ElementalMatrix<T> a(&w, &a_slot, p), b(&w, B, p);
            b.colCon = colCon != 0;
            b.rowCon = rowCon != 0;
            /* the reference raises its LogicErrors on every rank alike, before any communication: no rank is left waiting */
            try { body(a, b); } catch (const std::exception&) { threw[p] = 1; }
        });
    for (auto& t : th) t.join();
    for (int p = 0; p < w.P; ++p) if (threw[p]) throw std::logic_error("the reference raised LogicError");
}
/* the typed primitives: build DistMatrix<T,U,V> handles for the (source, destination) pair at hand */
#define TYPED(SU, SV, DU, DV, CALL)                                                                                         \
    if (A->colDist == SU && A->rowDist == SV && (*B)->colDist == DU && (*B)->rowDist == DV) {                                \
        World w(A->g);                                                                                                      \
        std::vector<std::thread> th;                                                                                        \
        orc_dm* a_slot = A;                                                                                                 \
        for (int p = 0; p < w.P; ++p)                                                                                       \
            th.emplace_back([&, p] {                                                                                        \
                DistMatrix<T, SU, SV, ELEMENT, Device::CPU> a(&w, &a_slot, p);                                              \
                DistMatrix<T, DU, DV, ELEMENT, Device::CPU> b(&w, B, p);                                                    \
                b.colCon = colCon != 0;                                                                                     \
                b.rowCon = rowCon != 0;                                                                                     \
                CALL;                                                                                                       \
            });                                                                                                             \
        for (auto& t : th) t.join();                                                                                        \
        return 0;                                                                                                           \
    }

template <typename T>
int run(int prim, orc_dm* A, orc_dm** B, int colCon, int rowCon) {
    typedef ElementalMatrix<T> EM;
    constexpr Device CPU = Device::CPU;
    switch (prim) {
        case 0: spmd<T>(A, B, colCon, rowCon, [](EM& a, EM& b) { copy::RowAllGather(a, b); }); return 0;
        case 1: spmd<T>(A, B, colCon, rowCon, [](EM& a, EM& b) { copy::ColAllGather(a, b); }); return 0;
        case 2:
            TYPED(VC, STAR, MC, STAR, copy::PartialColAllGather(a, b))
            TYPED(VR, STAR, MR, STAR, copy::PartialColAllGather(a, b))
            return -2;
        case 3: spmd<T>(A, B, colCon, rowCon, [](EM& a, EM& b) { copy::PartialRowAllGather(a, b); }); return 0;
        case 4:
            TYPED(MC, MR, VC, STAR, (copy::ColAllToAllDemote<T, VC, STAR, CPU>(a, b)))
            TYPED(MR, MC, VR, STAR, (copy::ColAllToAllDemote<T, VR, STAR, CPU>(a, b)))
            return -2;
        case 5:
            TYPED(MC, MR, STAR, VR, (copy::RowAllToAllDemote<T, STAR, VR, CPU>(a, b)))
            TYPED(MR, MC, STAR, VC, (copy::RowAllToAllDemote<T, STAR, VC, CPU>(a, b)))
            return -2;
        case 6:
            TYPED(VC, STAR, MC, MR, copy::ColAllToAllPromote(a, b))
            TYPED(VR, STAR, MR, MC, copy::ColAllToAllPromote(a, b))
            return -2;
        case 7:
            TYPED(STAR, VR, MC, MR, copy::RowAllToAllPromote(a, b))
            TYPED(STAR, VC, MR, MC, copy::RowAllToAllPromote(a, b))
            return -2;
        case 8:
            TYPED(VC, STAR, VR, STAR, (copy::ColwiseVectorExchange<T, MC, MR, CPU>(a, b)))
            TYPED(VR, STAR, VC, STAR, (copy::ColwiseVectorExchange<T, MR, MC, CPU>(a, b)))
            return -2;
        case 9:
            TYPED(STAR, VC, STAR, VR, (copy::RowwiseVectorExchange<T, MC, MR, CPU>(a, b)))
            TYPED(STAR, VR, STAR, VC, (copy::RowwiseVectorExchange<T, MR, MC, CPU>(a, b)))
            return -2;
        case 10: spmd<T>(A, B, colCon, rowCon, [](EM& a, EM& b) { copy::ColFilter(a, b); }); return 0;
        case 11: spmd<T>(A, B, colCon, rowCon, [](EM& a, EM& b) { copy::RowFilter(a, b); }); return 0;
        case 12: spmd<T>(A, B, colCon, rowCon, [](EM& a, EM& b) { copy::PartialColFilter(a, b); }); return 0;
        case 13: spmd<T>(A, B, colCon, rowCon, [](EM& a, EM& b) { copy::PartialRowFilter(a, b); }); return 0;
        case 14:
            TYPED(MC, MR, MC, MR, copy::Translate(a, b))
            TYPED(MR, MC, MR, MC, copy::Translate(a, b))
            TYPED(MC, STAR, MC, STAR, copy::Translate(a, b))
            TYPED(STAR, MR, STAR, MR, copy::Translate(a, b))
            TYPED(MR, STAR, MR, STAR, copy::Translate(a, b))
            TYPED(STAR, MC, STAR, MC, copy::Translate(a, b))
            TYPED(VC, STAR, VC, STAR, copy::Translate(a, b))
            TYPED(STAR, VC, STAR, VC, copy::Translate(a, b))
            TYPED(VR, STAR, VR, STAR, copy::Translate(a, b))
            TYPED(STAR, VR, STAR, VR, copy::Translate(a, b))
            TYPED(STAR, STAR, STAR, STAR, copy::Translate(a, b))
            return -2;
    }
    return -1;
}
}  // namespace

/* B += alpha * (sum of A over the ranks along which it is replicated / partially summed): El::AxpyContract(alpha, A, B),
 * include/El/blas_like/level1/AxpyContract.hpp:484-544, through the reference's own scatter routines */
extern "C" int ref_axpy_contract(double alpha, orc_dm* A, orc_dm** B) {
    try {
        if (A->es == 8) spmd<double>(A, B, 1, 1, [alpha](ElementalMatrix<double>& a, ElementalMatrix<double>& b) { AxpyContract(alpha, a, b); });
        else spmd<float>(A, B, 1, 1, [alpha](ElementalMatrix<float>& a, ElementalMatrix<float>& b) { AxpyContract((float)alpha, a, b); });
    } catch (const std::exception&) {
        return -1;
    }
    return 0;
}

/* B = A through the reference's copy:: primitive `prim` (0 RowAllGather, 1 ColAllGather, 2 PartialColAllGather, 3 PartialRowAllGather, 4 ColAllToAllDemote,
 * 5 RowAllToAllDemote, 6 ColAllToAllPromote, 7 RowAllToAllPromote, 8 ColwiseVectorExchange, 9 RowwiseVectorExchange, 10 ColFilter,
 * 11 RowFilter, 12 PartialColFilter, 13 PartialRowFilter, 14 Translate).  *B is the destination as the caller set it up
 * (distribution, alignments, any size); colCon / rowCon say whether its alignments are constrained (else the primitive may
 * re-align it, in which case *B is replaced).  Returns 0, -1 for an unknown primitive, -2 if the primitive does not serve the pair. */
extern "C" int ref_copy_primitive(int prim, orc_dm* A, orc_dm** B, int colCon, int rowCon) {
    return A->es == 8 ? run<double>(prim, A, B, colCon, rowCon) : run<float>(prim, A, B, colCon, rowCon);
}
