// ref_util.cpp -- extern "C" handles onto the REFERENCE's own code, compiled from where it lies under
// /root/reference (REF_ROOT): indexing (include/El/core/indexing/impl.hpp), the copy::util CPU
// pack/unpack loops (include/El/blas_like/level1/Copy/util.hpp) and axpy::util::InterleaveMatrixUpdate
// (include/El/blas_like/level1/Axpy/util.hpp).  Built by oracle/Makefile into oracle/_ref/libref_util.so.
// TEST INFRASTRUCTURE ONLY: used by tests/ to pin oracle/oracle_{core,util}.c and to generate the
// golden vectors under tests/golden/ (tests/golden/make_golden.py).
#include "ref_prelude.hpp"

#include <El/core/indexing/decl.hpp>
#include <El/core/indexing/impl.hpp>
#include <El/blas_like/level1/Copy/util.hpp>
#include <El/blas_like/level1/Axpy/util.hpp>

#include <cstdint>

using namespace El;
typedef SyncInfo<Device::CPU> SI;

extern "C" {

int ref_int_bytes() { return (int)sizeof(Int); }

// indexing/impl.hpp:33-66,190-194,244-245,276-283
int64_t ref_length(int64_t n, int64_t shift, int64_t stride) { return Length_((Int)n, (Int)shift, (Int)stride); }
int64_t ref_length4(int64_t n, int64_t rank, int64_t align, int64_t stride) { return Length_((Int)n, (Int)rank, (Int)align, (Int)stride); }
int64_t ref_max_length(int64_t n, int64_t stride) { return MaxLength_((Int)n, (Int)stride); }
int64_t ref_mod(int64_t a, int64_t b) { return Mod_((Int)a, (Int)b); }
int64_t ref_shift(int64_t rank, int64_t align, int64_t stride) { return Shift_((Int)rank, (Int)align, (Int)stride); }
int64_t ref_gcd(int64_t a, int64_t b) { return GCD_((Int)a, (Int)b); }
int64_t ref_global_index(int64_t iLoc, int64_t shift, int64_t stride) { return GlobalIndex((Int)iLoc, (Int)shift, (Int)stride); }

#define REF_FOR_TYPES(X) X(f64, double) X(f32, float)

#define REF_DEFINE(sfx, T)                                                                                          \
    void ref_interleave_##sfx(int64_t h, int64_t w, const T* A, int64_t csA, int64_t rsA, T* B, int64_t csB,       \
                              int64_t rsB) {                                                                        \
        copy::util::InterleaveMatrix((Int)h, (Int)w, A, (Int)csA, (Int)rsA, B, (Int)csB, (Int)rsB, SI{});          \
    }                                                                                                               \
    void ref_row_strided_pack_##sfx(int64_t h, int64_t w, int64_t al, int64_t st, const T* A, int64_t ld, T* Bp,   \
                                    int64_t portion) {                                                              \
        copy::util::RowStridedPack((Int)h, (Int)w, (Int)al, (Int)st, A, (Int)ld, Bp, (Int)portion, SI{});          \
    }                                                                                                               \
    void ref_row_strided_unpack_##sfx(int64_t h, int64_t w, int64_t al, int64_t st, const T* Ap, int64_t portion,  \
                                      T* B, int64_t ld) {                                                           \
        copy::util::RowStridedUnpack((Int)h, (Int)w, (Int)al, (Int)st, Ap, (Int)portion, B, (Int)ld, SI{});        \
    }                                                                                                               \
    void ref_partial_row_strided_pack_##sfx(int64_t h, int64_t w, int64_t al, int64_t st, int64_t uni,             \
                                            int64_t part, int64_t rankPart, int64_t shiftA, const T* A, int64_t ld, \
                                            T* Bp, int64_t portion) {                                               \
        copy::util::PartialRowStridedPack((Int)h, (Int)w, (Int)al, (Int)st, (Int)uni, (Int)part, (Int)rankPart,    \
                                          (Int)shiftA, A, (Int)ld, Bp, (Int)portion, SI{});                         \
    }                                                                                                               \
    void ref_partial_row_strided_unpack_##sfx(int64_t h, int64_t w, int64_t al, int64_t st, int64_t uni,           \
                                              int64_t part, int64_t rankPart, int64_t shiftB, const T* Ap,          \
                                              int64_t portion, T* B, int64_t ld) {                                  \
        copy::util::PartialRowStridedUnpack((Int)h, (Int)w, (Int)al, (Int)st, (Int)uni, (Int)part, (Int)rankPart,  \
                                            (Int)shiftB, Ap, (Int)portion, B, (Int)ld, SI{});                       \
    }                                                                                                               \
    void ref_col_strided_pack_##sfx(int64_t h, int64_t w, int64_t al, int64_t st, const T* A, int64_t ld, T* Bp,   \
                                    int64_t portion) {                                                              \
        copy::util::ColStridedPack<T, Device::CPU>((Int)h, (Int)w, (Int)al, (Int)st, A, (Int)ld, Bp, (Int)portion, \
                                                   SI{});                                                           \
    }                                                                                                               \
    void ref_col_strided_unpack_##sfx(int64_t h, int64_t w, int64_t al, int64_t st, const T* Ap, int64_t portion,  \
                                      T* B, int64_t ld) {                                                           \
        copy::util::ColStridedUnpack<T, Device::CPU>((Int)h, (Int)w, (Int)al, (Int)st, Ap, (Int)portion, B,        \
                                                     (Int)ld, SI{});                                                \
    }                                                                                                               \
    void ref_partial_col_strided_pack_##sfx(int64_t h, int64_t w, int64_t al, int64_t st, int64_t uni,             \
                                            int64_t part, int64_t rankPart, int64_t shiftA, const T* A, int64_t ld, \
                                            T* Bp, int64_t portion) {                                               \
        copy::util::PartialColStridedPack<T, Device::CPU>((Int)h, (Int)w, (Int)al, (Int)st, (Int)uni, (Int)part,   \
                                                          (Int)rankPart, (Int)shiftA, A, (Int)ld, Bp, (Int)portion, \
                                                          SI{});                                                    \
    }                                                                                                               \
    void ref_partial_col_strided_unpack_##sfx(int64_t h, int64_t w, int64_t al, int64_t st, int64_t uni,           \
                                              int64_t part, int64_t rankPart, int64_t shiftB, const T* Ap,          \
                                              int64_t portion, T* B, int64_t ld) {                                  \
        copy::util::PartialColStridedUnpack<T, Device::CPU>((Int)h, (Int)w, (Int)al, (Int)st, (Int)uni, (Int)part, \
                                                            (Int)rankPart, (Int)shiftB, Ap, (Int)portion, B,        \
                                                            (Int)ld, SI{});                                         \
    }                                                                                                               \
    void ref_strided_pack_##sfx(int64_t h, int64_t w, int64_t cal, int64_t cst, int64_t ral, int64_t rst,          \
                                const T* A, int64_t ld, T* Bp, int64_t portion) {                                   \
        copy::util::StridedPack<T, Device::CPU>((Int)h, (Int)w, (Int)cal, (Int)cst, (Int)ral, (Int)rst, A, (Int)ld, \
                                                Bp, (Int)portion, SI{});                                            \
    }                                                                                                               \
    void ref_strided_unpack_##sfx(int64_t h, int64_t w, int64_t cal, int64_t cst, int64_t ral, int64_t rst,        \
                                  const T* Ap, int64_t portion, T* B, int64_t ld) {                                 \
        copy::util::StridedUnpack<T, Device::CPU>((Int)h, (Int)w, (Int)cal, (Int)cst, (Int)ral, (Int)rst, Ap,      \
                                                  (Int)portion, B, (Int)ld, SI{});                                  \
    }                                                                                                               \
    void ref_interleave_update_##sfx(double alpha, int64_t h, int64_t w, const T* A, int64_t csA, int64_t rsA,     \
                                     T* B, int64_t csB, int64_t rsB) {                                              \
        axpy::util::InterleaveMatrixUpdate((T)alpha, (Int)h, (Int)w, A, (Int)csA, (Int)rsA, B, (Int)csB, (Int)rsB, \
                                           SI{});                                                                   \
    }

REF_FOR_TYPES(REF_DEFINE)

}  // extern "C"
