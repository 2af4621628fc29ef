/* TEST INFRASTRUCTURE ONLY -- see oracle/oracle.h.
 *
 * The reference's redistribution primitives, compiled from where they lie and run SPMD -- one thread per simulated rank --
 * over stand-in classes: include/El/blas_like/level1/Copy/{RowAllGather,ColAllGather,...}.hpp.  The per-portion loops they
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

#include "../oracle.h"
#include <condition_variable>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

namespace El {

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
struct CommShared {              /* one communicator instance: its members share a barrier and a pointer board */
    explicit CommShared(int n) : bar(n), board(n, nullptr), size(n) {}
    Barrier bar;
    std::vector<const void*> board;
    int size;
};

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
template <typename T, Device D>
void SendRecv(const T* sbuf, int sc, int to, T* rbuf, int rc, int from, const Comm& c, SyncInfo<D> const&) {
    (void)to;
    c.sh->board[c.rank] = sbuf;
    c.sh->bar.wait();
    std::memcpy(rbuf, c.sh->board[from], (size_t)rc * sizeof(T));
    c.sh->bar.wait();
    (void)sc;
}
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
    simple_buffer(size_t n, SyncInfo<D> const&) : v_(n) {}
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
    int rank = 0;
};

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
inline void Copy(const Matrix<T, D>& A, Matrix<T, D> B) { B = A; }   /* B is a handle onto the block */

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
    El::Matrix<T, Device::CPU> Matrix() { return El::Matrix<T, Device::CPU>(LocalHeight(), LocalWidth(), Buffer(), LDim()); }
    El::Matrix<T, Device::CPU> LockedMatrix() const {
        return El::Matrix<T, Device::CPU>(LocalHeight(), LocalWidth(), const_cast<T*>(LockedBuffer()), LDim());
    }
    bool Participating() const { return true; }
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
    mpi::Comm PartialColComm() const { return comm_of(h()->colDist == VC ? MC : (h()->colDist == VR ? MR : h()->colDist)); }
    mpi::Comm PartialRowComm() const { return comm_of(h()->rowDist == VC ? MC : (h()->rowDist == VR ? MR : h()->rowDist)); }
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
template <typename T>
inline void Broadcast(ElementalMatrix<T>&, const mpi::Comm&, int) {}   /* cross-communicator broadcast: [o,o] / MD only */
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

namespace {
using namespace El;
/* run `body(a, b)` on one thread per rank */
template <typename T, typename F>
void spmd(orc_dm* A, orc_dm** B, int colCon, int rowCon, F body) {
    World w(A->g);
    std::vector<std::thread> th;
    orc_dm* a_slot = A;
    for (int p = 0; p < w.P; ++p)
        th.emplace_back([&, p] {
            ElementalMatrix<T> a(&w, &a_slot, p), b(&w, B, p);
            b.colCon = colCon != 0;
            b.rowCon = rowCon != 0;
            body(a, b);
        });
    for (auto& t : th) t.join();
}
template <typename T>
int run(int prim, orc_dm* A, orc_dm** B, int colCon, int rowCon) {
    switch (prim) {
        case 0: spmd<T>(A, B, colCon, rowCon, [](ElementalMatrix<T>& a, ElementalMatrix<T>& b) { copy::RowAllGather(a, b); }); return 0;
    }
    return -1;
}
}  // namespace

/* B = A through the reference's copy:: primitive `prim` (0 RowAllGather ...).  *B is the destination as the caller set it up
 * (distribution, alignments, any size); colCon / rowCon say whether its alignments are constrained (else the primitive may
 * re-align it, in which case *B is replaced).  Returns 0, or -1 for an unknown primitive. */
extern "C" int ref_copy_primitive(int prim, orc_dm* A, orc_dm** B, int colCon, int rowCon) {
    return A->es == 8 ? run<double>(prim, A, B, colCon, rowCon) : run<float>(prim, A, B, colCon, rowCon);
}
