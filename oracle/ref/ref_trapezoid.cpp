/* TEST INFRASTRUCTURE ONLY -- see oracle/oracle.h.
 * The reference's trapezoid level-1 routines compiled from where they lie:
 *   include/El/blas_like/level1/ScaleTrapezoid.hpp   (Matrix :14-43, AbstractDistMatrix :45-92)
 *   include/El/blas_like/level1/AxpyTrapezoid.hpp    (Matrix :14-50, LocalAxpyTrapezoid on AbstractDistMatrix :52-98)
 *   include/El/blas_like/level1/MakeTrapezoidal.hpp  (Matrix :14-44, AbstractDistMatrix :46-86)
 * The headers are templates over Matrix<T> / AbstractDistMatrix<T>; they only use a handful of accessors, so minimal
 * stand-ins of those two classes (an external buffer plus the index map of one rank's block) are enough to instantiate the
 * reference's own loops.  Nothing of the reference is copied here.  Built into oracle/_ref/libref_util.so. */
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>

namespace El {
typedef int Int;
using std::unique_ptr;
enum class Device : unsigned char { CPU, GPU };
namespace UpperOrLowerNS { enum UpperOrLower { LOWER, UPPER }; }
using namespace UpperOrLowerNS;

template <typename... Args> inline void LogicError(const Args&...) { throw std::logic_error("reference LogicError"); }
template <typename T> inline const T& Min(const T& a, const T& b) { return a < b ? a : b; }
template <typename T> inline const T& Max(const T& a, const T& b) { return a > b ? a : b; }
template <typename T> inline void MemZero(T* p, size_t n) { std::memset(static_cast<void*>(p), 0, n * sizeof(T)); }

/* blas::Axpy is the vendor BLAS' ?axpy_ (src/core/imports/blas/Axpy.hpp), third-party code outside /root/reference whose rounding
 * of alpha*x + y (FMA or not) is unpinned: the fixtures use alpha = 2^k, for which both give the same bits */
namespace blas {
template <typename T> inline void Axpy(Int n, const T& alpha, const T* x, Int incx, T* y, Int incy) {
    for (Int i = 0; i < n; ++i) { volatile T t = alpha * x[i * incx]; y[i * incy] += t; }
}
}  // namespace blas

/* what the three headers ask of a local matrix */
template <typename T, Device D = Device::CPU>
class Matrix {
public:
    Matrix(Int h, Int w, T* buf, Int ld) : h_(h), w_(w), ld_(ld), buf_(buf) {}
    Int Height() const { return h_; }
    Int Width() const { return w_; }
    Int LDim() const { return ld_; }
    T* Buffer() { return buf_; }
    const T* LockedBuffer() const { return buf_; }
private:
    Int h_, w_, ld_;
    T* buf_;
};

/* ... and of one rank's block of a distributed matrix: global sizes, local sizes, the index map
 * (GlobalCol: AbstractDistMatrix.cpp GlobalCol = rowShift + jLoc*rowStride; LocalRowOffset(i) = Length_(i, colShift, colStride),
 * include/El/core/DistMatrix/Abstract.hpp) */
template <typename T>
class AbstractDistMatrix {
public:
    AbstractDistMatrix(Int m, Int n, Int colShift, Int colStride, Int rowShift, Int rowStride, T* buf, Int ld)
        : m_(m), n_(n), cs_(colShift), cst_(colStride), rs_(rowShift), rst_(rowStride), ld_(ld), buf_(buf) {}
    Int Height() const { return m_; }
    Int Width() const { return n_; }
    Int LocalHeight() const { return len(m_, cs_, cst_); }
    Int LocalWidth() const { return len(n_, rs_, rst_); }
    Int LDim() const { return ld_; }
    T* Buffer() { return buf_; }
    const T* LockedBuffer() const { return buf_; }
    Int GlobalCol(Int jLoc) const { return rs_ + jLoc * rst_; }
    Int LocalRowOffset(Int i) const { return len(i, cs_, cst_); }
private:
    static Int len(Int n, Int shift, Int stride) { return n > shift ? (n - shift - 1) / stride + 1 : 0; }
    Int m_, n_, cs_, cst_, rs_, rst_, ld_;
    T* buf_;
};
template <typename T> class ElementalMatrix;
template <typename T> class BlockMatrix;
}  // namespace El

#define EL_DEBUG_CSE
#define EL_DEBUG_ONLY(x)
#define EL_PARALLEL_FOR
/* no explicit-instantiation declarations: the routines are instantiated here, from the header's own definitions */
#define PROTO_INT(T)
#define PROTO_REAL(T)
#define PROTO_COMPLEX(T)
#include <El/blas_like/level1/ScaleTrapezoid.hpp>
#define PROTO_INT(T)
#define PROTO_REAL(T)
#define PROTO_COMPLEX(T)
#include <El/blas_like/level1/AxpyTrapezoid.hpp>
#define PROTO_INT(T)
#define PROTO_REAL(T)
#define PROTO_COMPLEX(T)
#include <El/blas_like/level1/MakeTrapezoidal.hpp>

namespace {
template <typename T>
El::AbstractDistMatrix<T> dm(const int* g, T* buf) { return El::AbstractDistMatrix<T>(g[0], g[1], g[2], g[3], g[4], g[5], buf, g[6]); }
}

/* geom = {height, width, colShift, colStride, rowShift, rowStride, ldim}; a local Matrix is {m, n, 0, 1, 0, 1, ldim} */
#define REF_TRAP(SFX, T)                                                                                                   \
    extern "C" void ref_scale_trapezoid_##SFX(double alpha, int uplo, const int* geom, T* A, int offset, int local) {     \
        if (local) { El::Matrix<T> M(geom[0], geom[1], A, geom[6]); El::ScaleTrapezoid(T(alpha), (El::UpperOrLower)uplo, M, offset); } \
        else { auto D = dm<T>(geom, A); El::ScaleTrapezoid(T(alpha), (El::UpperOrLower)uplo, D, offset); }                 \
    }                                                                                                                      \
    extern "C" void ref_axpy_trapezoid_##SFX(int uplo, double alpha, const int* geom, const T* X, int ldx, T* Y, int offset, int local) { \
        if (local) {                                                                                                       \
            El::Matrix<T> MX(geom[0], geom[1], const_cast<T*>(X), ldx), MY(geom[0], geom[1], Y, geom[6]);                   \
            El::AxpyTrapezoid((El::UpperOrLower)uplo, T(alpha), MX, MY, offset);                                            \
        } else {                                                                                                           \
            int gx[7]; for (int i = 0; i < 7; ++i) gx[i] = geom[i]; gx[6] = ldx;                                           \
            auto DX = dm<T>(gx, const_cast<T*>(X)); auto DY = dm<T>(geom, Y);                                               \
            El::LocalAxpyTrapezoid((El::UpperOrLower)uplo, T(alpha), DX, DY, offset);                                       \
        }                                                                                                                  \
    }                                                                                                                      \
    extern "C" void ref_make_trapezoidal_##SFX(int uplo, const int* geom, T* A, int offset, int local) {                   \
        if (local) { El::Matrix<T> M(geom[0], geom[1], A, geom[6]); El::MakeTrapezoidal((El::UpperOrLower)uplo, M, offset); } \
        else { auto D = dm<T>(geom, A); El::MakeTrapezoidal((El::UpperOrLower)uplo, D, offset); }                           \
    }
REF_TRAP(f64, double)
REF_TRAP(f32, float)
