// ref_prelude.hpp -- the minimum environment the REFERENCE's own headers need to compile stand-alone
// (TEST INFRASTRUCTURE ONLY; nothing here is reference source).
//
// The reference cannot be built as a whole here (MPI + LAPACK + Aluminum + NVSHMEM),
// but three of its headers on the hot path are self-contained template/inline code once the handful of
// names below exist:
//   include/El/core/indexing/{decl,impl}.hpp          Length_/MaxLength_/Mod_/Shift_/GCD_ ...
//   include/El/blas_like/level1/Copy/util.hpp         every CPU pack/unpack loop of copy::util
//   include/El/blas_like/level1/Axpy/util.hpp         axpy::util::InterleaveMatrixUpdate (CPU)
// They are #included from where they lie under /root/reference (never copied) by ref_util.cpp.
// What this file supplies stands in for: El::Int (include/El/core/types.hpp), the EL_* macros
// (include/El/core/environment/decl.hpp), hydrogen::Device / SyncInfo (include/hydrogen/Device.hpp,
// SyncInfo.hpp), the EnableIf family (include/hydrogen/meta/MetaUtilities.hpp), LogicError, Min,
// StridedMemCopy (include/El/core/Memory), lapack::Copy = ?lacpy('F') and blas::Axpy = ?axpy
// (src/core/imports/{lapack,blas}.cpp -> vendor LAPACK/BLAS, absent from this image: written as the
// plain loops those routines are documented to perform, compiled with -ffp-contract=off).
#pragma once
#include <cstring>
#include <stdexcept>
#include <string>
#include <sstream>
#include <type_traits>
#include <cmath>

#define EL_NO_EXCEPT noexcept
#define EL_NO_RELEASE_EXCEPT noexcept
#define EL_DEBUG_CSE
#define EL_DEBUG_ONLY(...)

namespace El {

#ifdef REF_USE_64BIT_INTS
typedef long long int Int;   // EL_USE_64BIT_INTS (CMakeLists.txt:196-208)
typedef long long unsigned Unsigned;
#else
typedef int Int;             // the reference's default
typedef unsigned Unsigned;
#endif
typedef Int BlasInt;

#ifndef REF_GPU
enum class Device : unsigned char { CPU, GPU };
template <Device D> struct SyncInfo {};
// float and double are storage types on the CPU (hydrogen/Device.hpp: IsStorageType<T,Device::CPU>)
template <typename T, Device D> struct IsStorageType : std::integral_constant<bool, D == Device::CPU && std::is_arithmetic<T>::value> {};
#else
// GPU build (ref_gpu.cu): Device, SyncInfo<Device::GPU>{stream_,event_}, IsStorageType, gpu_blas::* and the
// *_GPU_impl kernels are the reference's own (include/hydrogen/**), pulled in before this file.
using namespace hydrogen;
#endif
// names axpy::util::UpdateWithLocalData (Axpy/util.hpp:54-75, never instantiated here) must be able to parse
enum Dist { MC, MD, MR, VC, VR, STAR, CIRC };       // include/El/core/types.hpp:208-216
enum DistWrap { ELEMENT, BLOCK };
template <typename T> class ElementalMatrix;
template <typename T, Device D = Device::CPU> class Matrix;
template <typename T, Dist U = MC, Dist V = MR, DistWrap W = ELEMENT, Device D = Device::CPU> class DistMatrix;

template <typename C> using EnableIf = typename std::enable_if<C::value>::type;
template <typename C> using DisableIf = typename std::enable_if<!C::value>::type;
template <typename T> using IsIntegral = std::is_integral<T>;

template <typename... Args> inline void LogicError(const Args&... args) {
    std::ostringstream os;
    int dummy[] = {0, ((void)(os << args), 0)...};
    (void)dummy;
    throw std::logic_error(os.str());
}
template <typename T> inline const T& Min(const T& a, const T& b) noexcept { return a < b ? a : b; }
template <typename T> inline const T& Max(const T& a, const T& b) noexcept { return a > b ? a : b; }
template <typename T> inline T Sqrt(const T& a) { return (T)std::sqrt((double)a); }

template <typename T> inline void StridedMemCopy(T* dest, Int destStride, const T* source, Int sourceStride, Int numEntries) {
    for (Int i = 0; i < numEntries; ++i) dest[(std::ptrdiff_t)i * destStride] = source[(std::ptrdiff_t)i * sourceStride];
}

namespace lapack {
// ?lacpy with uplo = 'F': B(0:m,0:n) := A(0:m,0:n), column-major
template <typename T> inline void Copy(char, BlasInt m, BlasInt n, const T* A, BlasInt lda, T* B, BlasInt ldb) {
    for (BlasInt j = 0; j < n; ++j)
        if (m > 0) std::memcpy(B + (std::ptrdiff_t)j * ldb, A + (std::ptrdiff_t)j * lda, sizeof(T) * (size_t)m);
}
}  // namespace lapack
namespace blas {
// ?axpy: y := alpha x + y (one product rounding, one sum rounding; no FMA contraction)
template <typename T> inline void Axpy(BlasInt n, const T& alpha, const T* x, BlasInt incx, T* y, BlasInt incy) {
    for (BlasInt i = 0; i < n; ++i) {
        volatile T p = alpha * x[(std::ptrdiff_t)i * incx];
        y[(std::ptrdiff_t)i * incy] = y[(std::ptrdiff_t)i * incy] + p;
    }
}
}  // namespace blas

// Declarations the reference keeps in include/El/blas_like/level1/decl.hpp:96-112,424-700 (that header
// declares all of level 1 and needs the whole class hierarchy, so only the default template arguments of
// the copy::util CPU overloads are re-declared here; the definitions come from the reference's util.hpp).
namespace axpy { namespace util {
template <typename Ring> void InterleaveMatrixUpdate(Ring, Int, Int, Ring const*, Int, Int, Ring*, Int, Int, SyncInfo<Device::CPU>);
#ifdef REF_GPU
template <typename Ring> void InterleaveMatrixUpdate(Ring, Int, Int, Ring const*, Int, Int, Ring*, Int, Int, SyncInfo<Device::GPU>);
#endif
} }
namespace copy { namespace util {
#define REF_CPU_ONLY typename = EnableIf<IsStorageType<T, Device::CPU>>
template <typename T, REF_CPU_ONLY> void InterleaveMatrix(Int, Int, T const*, Int, Int, T*, Int, Int, SyncInfo<Device::CPU>);
template <typename T, REF_CPU_ONLY> void RowStridedPack(Int, Int, Int, Int, const T*, Int, T*, Int, SyncInfo<Device::CPU>);
template <typename T, REF_CPU_ONLY> void RowStridedUnpack(Int, Int, Int, Int, T const*, Int, T*, Int, SyncInfo<Device::CPU>);
template <typename T, REF_CPU_ONLY> void PartialRowStridedPack(Int, Int, Int, Int, Int, Int, Int, Int, T const*, Int, T*, Int, SyncInfo<Device::CPU>);
template <typename T, REF_CPU_ONLY> void PartialRowStridedUnpack(Int, Int, Int, Int, Int, Int, Int, Int, const T*, Int, T*, Int, SyncInfo<Device::CPU>);
#undef REF_CPU_ONLY
#ifdef REF_GPU   // decl.hpp:470-515
#define REF_GPU_ONLY typename = EnableIf<IsStorageType<T, Device::GPU>>
template <typename T, REF_GPU_ONLY> void InterleaveMatrix(Int, Int, T const*, Int, Int, T*, Int, Int, SyncInfo<Device::GPU>);
template <typename T, REF_GPU_ONLY> void RowStridedPack(Int, Int, Int, Int, const T*, Int, T*, Int, SyncInfo<Device::GPU>);
template <typename T, REF_GPU_ONLY> void RowStridedUnpack(Int, Int, Int, Int, T const*, Int, T*, Int, SyncInfo<Device::GPU>);
template <typename T, REF_GPU_ONLY> void PartialRowStridedPack(Int, Int, Int, Int, Int, Int, Int, Int, T const*, Int, T*, Int, SyncInfo<Device::GPU>);
template <typename T, REF_GPU_ONLY> void PartialRowStridedUnpack(Int, Int, Int, Int, Int, Int, Int, Int, const T*, Int, T*, Int, SyncInfo<Device::GPU>);
#undef REF_GPU_ONLY
#endif
#define REF_ON_D typename = EnableIf<IsStorageType<T, D>>
template <typename T, Device D, REF_ON_D> void ColStridedPack(Int, Int, Int, Int, T const*, Int, T*, Int, SyncInfo<D>);
template <typename T, Device D, REF_ON_D> void ColStridedColumnPack(Int, Int, Int, T const*, T*, Int, SyncInfo<D>);
template <typename T, Device D = Device::CPU, REF_ON_D> void ColStridedUnpack(Int, Int, Int, Int, T const*, Int, T*, Int, SyncInfo<D>);
template <typename T, Device D = Device::CPU, REF_ON_D> void PartialColStridedPack(Int, Int, Int, Int, Int, Int, Int, Int, T const*, Int, T*, Int, SyncInfo<D>);
template <typename T, Device D = Device::CPU, REF_ON_D> void PartialColStridedUnpack(Int, Int, Int, Int, Int, Int, Int, Int, T const*, Int, T*, Int, SyncInfo<D>);
template <typename T, Device D, REF_ON_D> void PartialColStridedColumnPack(Int, Int, Int, Int, Int, Int, Int, T const*, T*, Int, SyncInfo<D>);
template <typename T, Device D, REF_ON_D> void PartialColStridedColumnUnpack(Int, Int, Int, Int, Int, Int, Int, T const*, Int, T*, SyncInfo<D>);
template <typename T, Device D, REF_ON_D> void StridedPack(Int, Int, Int, Int, Int, Int, T const*, Int, T*, Int, SyncInfo<D>);
template <typename T, Device D, REF_ON_D> void StridedUnpack(Int, Int, Int, Int, Int, Int, T const*, Int, T*, Int, SyncInfo<D>);
#undef REF_ON_D
} }
}  // namespace El
