// Seam S2: the reference's pack / unpack utilities on raw device buffers, with the reference's own
// argument lists, so that code written against copy::util / axpy::util can call this library portion by portion:
//   copy::util::InterleaveMatrix, {Row,Col}Strided{Pack,Unpack}, Partial{Row,Col}Strided{Pack,Unpack},
//   Strided{Pack,Unpack}            include/El/blas_like/level1/Copy/util.hpp:122-230 (CPU), :244-359 (GPU), :363-722
//   axpy::util::InterleaveMatrixUpdate   include/El/blas_like/level1/Axpy/util.hpp:23-51
//
// The reference issues one cudaMemcpy2DAsync / copy_2d_kernel per portion k.  Every portion is a lattice
// (lattice.cuh), so here ALL portions of a call go out in ONE lattice_copy_kernel launch.  The distributed engine
// (redist.cu) does not call these: it plans whole redistributions as per-peer lattices directly.  These entry
// points exist for callers that keep the reference's own collective choreography.
//
// Naming follows the reference: "colStride" is the distance between consecutive ROWS of a column (the stride of
// the column distribution), "rowStride" the distance between consecutive columns.
#include "../../include/hydrogen_b200.h"
#include "grid.h"
#include "hyd_internal.h"
#include "lattice.cuh"
#include <algorithm>
#include <vector>

namespace hyd {
namespace {

template <typename T>
CopyDesc lat(const T* src, i64 s_rs, i64 s_cs, T* dst, i64 d_rs, i64 d_cs, i64 nr, i64 nc) {
    CopyDesc d;
    d.src = src; d.dst = dst;
    d.s_rs = s_rs; d.s_cs = s_cs;
    d.d_rs = d_rs; d.d_cs = d_cs;
    d.nr = nr; d.nc = nc;
    return d;
}

template <typename T>
int launch(const std::vector<CopyDesc>& v, cudaStream_t s, const char* who) {
    for (size_t i0 = 0; i0 < v.size(); i0 += kMaxDesc) {
        CopyBatch b;
        b.n = (int)std::min((size_t)kMaxDesc, v.size() - i0);
        i64 mx = 0;
        for (int i = 0; i < b.n; ++i) { b.d[i] = v[i0 + i]; mx = std::max(mx, b.d[i].nr * b.d[i].nc); }
        if (mx == 0) continue;
        const i64 want = (mx + 4095) / 4096;
        const i64 cap = std::max<i64>(1, (i64)num_sms() * 8 / b.n);
        const unsigned grid = (unsigned)(std::max<i64>(1, std::min(want, cap)) * b.n);
        lattice_copy_kernel<T><<<grid, 256, 0, s>>>(b);
        ++g_launches;
        if (check_cuda(cudaGetLastError(), who)) return 1;
    }
    return 0;
}

// Row-interleaved unpack (Col / PartialCol / 2-D Strided unpack): the S portions of a call each hold every S-th row of
// the destination block.  Unpacked one lattice at a time (as the reference does, one copy_2d_kernel per portion), every
// 32-byte sector of the block is written S times, an S-th at a time; lattice_merge_kernel writes each destination
// vector once, fetching its elements from the portions that own them.  src_of_residue[r] = portion holding rows
// r, r+S, ...; cs_of_residue[r] = its leading dimension.
template <typename T>
int launch_merge(int S, T* const* src_of_residue, const i64* cs_of_residue, T* dst, i64 d_cs, i64 nr, i64 nc, cudaStream_t s,
                 const char* who) {
    if (nr == 0 || nc == 0) return 0;
    MergeDesc md;
    md.S = S;
    for (int r = 0; r < S; ++r) { md.src[r] = src_of_residue[r]; md.s_rs[r] = 1; md.s_cs[r] = cs_of_residue[r]; }
    md.dst = dst;
    md.d_cs = d_cs;
    md.nr = nr;
    md.nc = nc;
    const i64 V = 16 / (i64)sizeof(T);
    const i64 want = ((nr + V - 1) / V + 255) / 256 * ((nc + 7) / 8);
    const unsigned grid = (unsigned)std::max<i64>(1, std::min<i64>(want, (i64)num_sms() * 8));
    lattice_merge_kernel<T><<<grid, 256, 0, s>>>(md);
    ++g_launches;
    return check_cuda(cudaGetLastError(), who);
}

inline int bad_dims(i64 h, i64 w, const char* who) { return (h < 0 || w < 0) ? set_error("%s: negative dimension", who) : 0; }
inline int bad_stride(i64 align, i64 stride, const char* who) {
    if (stride <= 0) return set_error("%s: stride must be positive", who);
    if (align < 0 || align >= stride) return set_error("%s: alignment out of range", who);
    return 0;
}

// B[i*csB + j*rsB] = A[i*csA + j*rsA]                                              (util.hpp:122-146, :244-267)
template <typename T>
int interleave(i64 h, i64 w, const T* A, i64 csA, i64 rsA, T* B, i64 csB, i64 rsB, cudaStream_t s) {
    const char* who = "hyd_interleave";
    if (bad_dims(h, w, who)) return 1;
    if (h == 0 || w == 0) return 0;
    if (csA <= 0 || rsA <= 0 || csB <= 0 || rsB <= 0) return set_error("%s: strides must be positive", who);
    if (!A || !B) return set_error("%s: null buffer", who);
    std::vector<CopyDesc> v{lat(A, csA, rsA, B, csB, rsB, h, w)};
    return launch<T>(v, s, who);
}

// portion k = columns rowShift_k + j*rowStride of A, stored h x localWidth_k, ld h       (util.hpp:148-165, :269-288)
template <typename T>
int row_strided(bool pack, i64 h, i64 w, i64 rowAlign, i64 rowStride, T* M, i64 ld, T* P, i64 portion, cudaStream_t s) {
    const char* who = pack ? "hyd_row_strided_pack" : "hyd_row_strided_unpack";
    if (bad_dims(h, w, who) || bad_stride(rowAlign, rowStride, who)) return 1;
    if (h == 0 || w == 0) return 0;
    if (ld < h) return set_error("%s: leading dimension too small", who);
    if (portion < h * max_length_(w, rowStride)) return set_error("%s: portionSize too small", who);
    std::vector<CopyDesc> v;
    for (i64 k = 0; k < rowStride; ++k) {
        const i64 rowShift = shift_(k, rowAlign, rowStride);
        const i64 lw = length_(w, rowShift, rowStride);
        T* m = M + rowShift * ld;
        T* p = P + k * portion;
        v.push_back(pack ? lat<T>(m, 1, rowStride * ld, p, 1, h, h, lw) : lat<T>(p, 1, h, m, 1, rowStride * ld, h, lw));
    }
    return launch<T>(v, s, who);
}

// portion k = columns of the union rank rowRankPart + k*rowStridePart; A (or B) itself holds the columns of
// rowShiftA + j*rowStridePart                                                              (util.hpp:186-230, :311-359)
template <typename T>
int partial_row_strided(bool pack, i64 h, i64 w, i64 rowAlign, i64 rowStride, i64 uni, i64 part, i64 rankPart, i64 shiftM,
                        T* M, i64 ld, T* P, i64 portion, cudaStream_t s) {
    const char* who = pack ? "hyd_partial_row_strided_pack" : "hyd_partial_row_strided_unpack";
    if (bad_dims(h, w, who) || bad_stride(rowAlign, rowStride, who)) return 1;
    if (uni <= 0 || part <= 0 || uni * part != rowStride) return set_error("%s: union x part must equal the stride", who);
    if (rankPart < 0 || rankPart >= part || shiftM < 0 || shiftM >= part) return set_error("%s: bad partial rank / shift", who);
    if (h == 0 || w == 0) return 0;
    if (ld < h) return set_error("%s: leading dimension too small", who);
    if (portion < h * max_length_(w, rowStride)) return set_error("%s: portionSize too small", who);
    std::vector<CopyDesc> v;
    for (i64 k = 0; k < uni; ++k) {
        const i64 rowShift = shift_(rankPart + k * part, rowAlign, rowStride);
        if (mod_(rowShift - shiftM, part) != 0) return set_error("%s: shift is not congruent to the partial rank", who);
        const i64 off = (rowShift - shiftM) / part;
        if (off < 0) return set_error("%s: shift is not congruent to the partial rank", who);
        const i64 lw = length_(w, rowShift, rowStride);
        T* m = M + off * ld;
        T* p = P + k * portion;
        v.push_back(pack ? lat<T>(m, 1, uni * ld, p, 1, h, h, lw) : lat<T>(p, 1, h, m, 1, uni * ld, h, lw));
    }
    return launch<T>(v, s, who);
}

// portion k = rows colShift_k + i*colStride of A, stored localHeight_k x w, ld localHeight_k   (util.hpp:363-381, :403-421)
template <typename T>
int col_strided(bool pack, i64 h, i64 w, i64 colAlign, i64 colStride, T* M, i64 ld, T* P, i64 portion, cudaStream_t s) {
    const char* who = pack ? "hyd_col_strided_pack" : "hyd_col_strided_unpack";
    if (bad_dims(h, w, who) || bad_stride(colAlign, colStride, who)) return 1;
    if (h == 0 || w == 0) return 0;
    if (ld < h) return set_error("%s: leading dimension too small", who);
    if (portion < w * max_length_(h, colStride)) return set_error("%s: portionSize too small", who);
    if (!pack && colStride >= 2 && colStride <= kMaxMerge) {
        T* src[kMaxMerge];
        i64 cs[kMaxMerge];
        for (i64 k = 0; k < colStride; ++k) {
            const i64 colShift = shift_(k, colAlign, colStride);  // portion k owns rows colShift, colShift + colStride, ...
            src[colShift] = P + k * portion;
            cs[colShift] = length_(h, colShift, colStride);
        }
        return launch_merge<T>((int)colStride, src, cs, M, ld, h, w, s, who);
    }
    std::vector<CopyDesc> v;
    for (i64 k = 0; k < colStride; ++k) {
        const i64 colShift = shift_(k, colAlign, colStride);
        const i64 lh = length_(h, colShift, colStride);
        T* m = M + colShift;
        T* p = P + k * portion;
        v.push_back(pack ? lat<T>(m, colStride, ld, p, 1, lh, lh, w) : lat<T>(p, 1, lh, m, colStride, ld, lh, w));
    }
    return launch<T>(v, s, who);
}

template <typename T>
int partial_col_strided(bool pack, i64 h, i64 w, i64 colAlign, i64 colStride, i64 uni, i64 part, i64 rankPart, i64 shiftM,
                        T* M, i64 ld, T* P, i64 portion, cudaStream_t s) {  // util.hpp:464-486, :510-533
    const char* who = pack ? "hyd_partial_col_strided_pack" : "hyd_partial_col_strided_unpack";
    if (bad_dims(h, w, who) || bad_stride(colAlign, colStride, who)) return 1;
    if (uni <= 0 || part <= 0 || uni * part != colStride) return set_error("%s: union x part must equal the stride", who);
    if (rankPart < 0 || rankPart >= part || shiftM < 0 || shiftM >= part) return set_error("%s: bad partial rank / shift", who);
    if (h == 0 || w == 0) return 0;
    const i64 lhM = length_(h, shiftM, part);  // rows the matrix itself holds
    if (ld < lhM) return set_error("%s: leading dimension too small", who);
    if (portion < w * max_length_(h, colStride)) return set_error("%s: portionSize too small", who);
    if (!pack && uni >= 2 && uni <= kMaxMerge) {
        T* src[kMaxMerge];
        i64 cs[kMaxMerge];
        for (i64 k = 0; k < uni; ++k) {
            const i64 colShift = shift_(rankPart + k * part, colAlign, colStride);
            if (mod_(colShift - shiftM, part) != 0) return set_error("%s: shift is not congruent to the partial rank", who);
            const i64 off = (colShift - shiftM) / part;  // portion k owns local rows off, off + uni, ... (off < uni, distinct)
            if (off < 0 || off >= uni) return set_error("%s: shift is not congruent to the partial rank", who);
            src[off] = P + k * portion;
            cs[off] = length_(h, colShift, colStride);
        }
        return launch_merge<T>((int)uni, src, cs, M, ld, lhM, w, s, who);
    }
    std::vector<CopyDesc> v;
    for (i64 k = 0; k < uni; ++k) {
        const i64 colShift = shift_(rankPart + k * part, colAlign, colStride);
        if (mod_(colShift - shiftM, part) != 0) return set_error("%s: shift is not congruent to the partial rank", who);
        const i64 off = (colShift - shiftM) / part;
        if (off < 0) return set_error("%s: shift is not congruent to the partial rank", who);
        const i64 lh = length_(h, colShift, colStride);
        T* m = M + off;
        T* p = P + k * portion;
        v.push_back(pack ? lat<T>(m, uni, ld, p, 1, lh, lh, w) : lat<T>(p, 1, lh, m, uni, ld, lh, w));
    }
    return launch<T>(v, s, who);
}

// portion k + l*colStride = rows colShift_k + i*colStride, columns rowShift_l + j*rowStride   (util.hpp:671-722)
template <typename T>
int strided(bool pack, i64 h, i64 w, i64 colAlign, i64 colStride, i64 rowAlign, i64 rowStride, T* M, i64 ld, T* P,
            i64 portion, cudaStream_t s) {
    const char* who = pack ? "hyd_strided_pack" : "hyd_strided_unpack";
    if (bad_dims(h, w, who) || bad_stride(colAlign, colStride, who) || bad_stride(rowAlign, rowStride, who)) return 1;
    if (h == 0 || w == 0) return 0;
    if (ld < h) return set_error("%s: leading dimension too small", who);
    if (portion < max_length_(h, colStride) * max_length_(w, rowStride)) return set_error("%s: portionSize too small", who);
    if (!pack && colStride >= 2 && colStride <= kMaxMerge) {
        for (i64 l = 0; l < rowStride; ++l) {  // one merged pass per column class
            const i64 rowShift = shift_(l, rowAlign, rowStride);
            const i64 lw = length_(w, rowShift, rowStride);
            T* src[kMaxMerge];
            i64 cs[kMaxMerge];
            for (i64 k = 0; k < colStride; ++k) {
                const i64 colShift = shift_(k, colAlign, colStride);
                src[colShift] = P + (k + l * colStride) * portion;
                cs[colShift] = length_(h, colShift, colStride);
            }
            if (launch_merge<T>((int)colStride, src, cs, M + rowShift * ld, rowStride * ld, h, lw, s, who)) return 1;
        }
        return 0;
    }
    std::vector<CopyDesc> v;
    for (i64 l = 0; l < rowStride; ++l) {
        const i64 rowShift = shift_(l, rowAlign, rowStride);
        const i64 lw = length_(w, rowShift, rowStride);
        for (i64 k = 0; k < colStride; ++k) {
            const i64 colShift = shift_(k, colAlign, colStride);
            const i64 lh = length_(h, colShift, colStride);
            T* m = M + colShift + rowShift * ld;
            T* p = P + (k + l * colStride) * portion;
            v.push_back(pack ? lat<T>(m, colStride, rowStride * ld, p, 1, lh, lh, lw)
                             : lat<T>(p, 1, lh, m, colStride, rowStride * ld, lh, lw));
        }
    }
    return launch<T>(v, s, who);
}

// B[i*csB + j*rsB] += alpha * A[i*csA + j*rsA]: one rounding for the product, one for the sum (the same arithmetic
// as the reduce-side unpack, lattice_reduce_kernel).                                           Axpy/util.hpp:23-51
template <typename T>
__global__ void __launch_bounds__(256) lattice_axpy_kernel(T alpha, i64 nr, i64 nc, const T* __restrict__ A, i64 a_rs, i64 a_cs,
                                                           T* __restrict__ B, i64 b_rs, i64 b_cs) {
    const i64 chunks_a = (nr + 255) / 256;
    const i64 col_groups = (nc + 3) / 4;
    for (i64 tile = blockIdx.x; tile < chunks_a * col_groups; tile += gridDim.x) {
        const i64 a = (tile % chunks_a) * 256 + threadIdx.x;
        const i64 b0 = (tile / chunks_a) * 4;
        if (a >= nr) continue;
        T x[4], y[4];
#pragma unroll
        for (int q = 0; q < 4; ++q)
            if (b0 + q < nc) {
                x[q] = __ldg(A + a * a_rs + (b0 + q) * a_cs);
                y[q] = B[a * b_rs + (b0 + q) * b_cs];
            }
#pragma unroll
        for (int q = 0; q < 4; ++q)
            if (b0 + q < nc) B[a * b_rs + (b0 + q) * b_cs] = add_rn(y[q], mul_rn(alpha, x[q]));
    }
}

template <typename T>
int interleave_update(T alpha, i64 h, i64 w, const T* A, i64 csA, i64 rsA, T* B, i64 csB, i64 rsB, cudaStream_t s) {
    const char* who = "hyd_interleave_update";
    if (bad_dims(h, w, who)) return 1;
    if (h == 0 || w == 0) return 0;
    if (csA <= 0 || rsA <= 0 || csB <= 0 || rsB <= 0) return set_error("%s: strides must be positive", who);
    if (!A || !B) return set_error("%s: null buffer", who);
    const i64 tiles = ((h + 255) / 256) * ((w + 3) / 4);
    const unsigned grid = (unsigned)std::max<i64>(1, std::min<i64>(tiles, (i64)num_sms() * 8));
    lattice_axpy_kernel<T><<<grid, 256, 0, s>>>(alpha, h, w, A, csA, rsA, B, csB, rsB);
    ++g_launches;
    return check_cuda(cudaGetLastError(), who);
}

}  // namespace
}  // namespace hyd

using hyd::i64;
#define S(x) ((cudaStream_t)(x))

extern "C" {

#define HYD_PACKS(T, sfx)                                                                                                    \
    int hyd_interleave_##sfx(int64_t h, int64_t w, const T* A, int64_t csA, int64_t rsA, T* B, int64_t csB, int64_t rsB,     \
                             hyd_stream_t s) {                                                                               \
        return hyd::interleave<T>(h, w, A, csA, rsA, B, csB, rsB, S(s));                                                     \
    }                                                                                                                        \
    int hyd_interleave_update_##sfx(T alpha, int64_t h, int64_t w, const T* A, int64_t csA, int64_t rsA, T* B, int64_t csB,  \
                                    int64_t rsB, hyd_stream_t s) {                                                           \
        return hyd::interleave_update<T>(alpha, h, w, A, csA, rsA, B, csB, rsB, S(s));                                       \
    }                                                                                                                        \
    int hyd_row_strided_pack_##sfx(int64_t h, int64_t w, int64_t rowAlign, int64_t rowStride, const T* A, int64_t ldA,       \
                                   T* BPortions, int64_t portionSize, hyd_stream_t s) {                                      \
        return hyd::row_strided<T>(true, h, w, rowAlign, rowStride, const_cast<T*>(A), ldA, BPortions, portionSize, S(s));   \
    }                                                                                                                        \
    int hyd_row_strided_unpack_##sfx(int64_t h, int64_t w, int64_t rowAlign, int64_t rowStride, const T* APortions,          \
                                     int64_t portionSize, T* B, int64_t ldB, hyd_stream_t s) {                               \
        return hyd::row_strided<T>(false, h, w, rowAlign, rowStride, B, ldB, const_cast<T*>(APortions), portionSize, S(s));  \
    }                                                                                                                        \
    int hyd_partial_row_strided_pack_##sfx(int64_t h, int64_t w, int64_t rowAlign, int64_t rowStride, int64_t rowStrideUnion,\
                                           int64_t rowStridePart, int64_t rowRankPart, int64_t rowShiftA, const T* A,        \
                                           int64_t ldA, T* BPortions, int64_t portionSize, hyd_stream_t s) {                 \
        return hyd::partial_row_strided<T>(true, h, w, rowAlign, rowStride, rowStrideUnion, rowStridePart, rowRankPart,      \
                                           rowShiftA, const_cast<T*>(A), ldA, BPortions, portionSize, S(s));                 \
    }                                                                                                                        \
    int hyd_partial_row_strided_unpack_##sfx(int64_t h, int64_t w, int64_t rowAlign, int64_t rowStride,                      \
                                             int64_t rowStrideUnion, int64_t rowStridePart, int64_t rowRankPart,             \
                                             int64_t rowShiftB, const T* APortions, int64_t portionSize, T* B, int64_t ldB,  \
                                             hyd_stream_t s) {                                                               \
        return hyd::partial_row_strided<T>(false, h, w, rowAlign, rowStride, rowStrideUnion, rowStridePart, rowRankPart,     \
                                           rowShiftB, B, ldB, const_cast<T*>(APortions), portionSize, S(s));                 \
    }                                                                                                                        \
    int hyd_col_strided_pack_##sfx(int64_t h, int64_t w, int64_t colAlign, int64_t colStride, const T* A, int64_t ldA,       \
                                   T* BPortions, int64_t portionSize, hyd_stream_t s) {                                      \
        return hyd::col_strided<T>(true, h, w, colAlign, colStride, const_cast<T*>(A), ldA, BPortions, portionSize, S(s));   \
    }                                                                                                                        \
    int hyd_col_strided_unpack_##sfx(int64_t h, int64_t w, int64_t colAlign, int64_t colStride, const T* APortions,          \
                                     int64_t portionSize, T* B, int64_t ldB, hyd_stream_t s) {                               \
        return hyd::col_strided<T>(false, h, w, colAlign, colStride, B, ldB, const_cast<T*>(APortions), portionSize, S(s));  \
    }                                                                                                                        \
    int hyd_partial_col_strided_pack_##sfx(int64_t h, int64_t w, int64_t colAlign, int64_t colStride, int64_t colStrideUnion,\
                                           int64_t colStridePart, int64_t colRankPart, int64_t colShiftA, const T* A,        \
                                           int64_t ldA, T* BPortions, int64_t portionSize, hyd_stream_t s) {                 \
        return hyd::partial_col_strided<T>(true, h, w, colAlign, colStride, colStrideUnion, colStridePart, colRankPart,      \
                                           colShiftA, const_cast<T*>(A), ldA, BPortions, portionSize, S(s));                 \
    }                                                                                                                        \
    int hyd_partial_col_strided_unpack_##sfx(int64_t h, int64_t w, int64_t colAlign, int64_t colStride,                      \
                                             int64_t colStrideUnion, int64_t colStridePart, int64_t colRankPart,             \
                                             int64_t colShiftB, const T* APortions, int64_t portionSize, T* B, int64_t ldB,  \
                                             hyd_stream_t s) {                                                               \
        return hyd::partial_col_strided<T>(false, h, w, colAlign, colStride, colStrideUnion, colStridePart, colRankPart,     \
                                           colShiftB, B, ldB, const_cast<T*>(APortions), portionSize, S(s));                 \
    }                                                                                                                        \
    int hyd_strided_pack_##sfx(int64_t h, int64_t w, int64_t colAlign, int64_t colStride, int64_t rowAlign,                  \
                               int64_t rowStride, const T* A, int64_t ldA, T* BPortions, int64_t portionSize,                \
                               hyd_stream_t s) {                                                                             \
        return hyd::strided<T>(true, h, w, colAlign, colStride, rowAlign, rowStride, const_cast<T*>(A), ldA, BPortions,      \
                               portionSize, S(s));                                                                           \
    }                                                                                                                        \
    int hyd_strided_unpack_##sfx(int64_t h, int64_t w, int64_t colAlign, int64_t colStride, int64_t rowAlign,                \
                                 int64_t rowStride, const T* APortions, int64_t portionSize, T* B, int64_t ldB,              \
                                 hyd_stream_t s) {                                                                           \
        return hyd::strided<T>(false, h, w, colAlign, colStride, rowAlign, rowStride, B, ldB, const_cast<T*>(APortions),     \
                               portionSize, S(s));                                                                           \
    }

HYD_PACKS(double, f64)
HYD_PACKS(float, f32)
#undef HYD_PACKS

}  // extern "C"
