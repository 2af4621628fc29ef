/* TEST INFRASTRUCTURE ONLY -- see oracle/oracle.h.
 * The reference's distribution algebra compiled from where it lies: include/El/core/types.hpp (Collect :292-297,
 * Partial :299-307, PartialUnionRow/Col :309-337, the Dist / Orientation / GridOrder enumerators).  types.hpp only needs the
 * few names below; nothing of the reference is copied here.  Built into oracle/_ref/libref_util.so by oracle/Makefile.   */
#include <climits>
#include <complex>
#include <limits>
#include <string>
#include <type_traits>
#include <vector>
namespace hydrogen { enum class TransposeMode { NORMAL, TRANSPOSE, CONJ_TRANSPOSE }; }
namespace El {
typedef long long Int;
typedef unsigned long long Unsigned;
using hydrogen::TransposeMode;
template <typename T> struct Complex;
template <typename T> struct BaseHelper { typedef T type; };
template <typename T> using Base = typename BaseHelper<T>::type;
template <typename C, class T = void> using EnableIf = typename std::enable_if<C::value, T>::type;
template <typename C, class T = void> using DisableIf = typename std::enable_if<!C::value, T>::type;
}  // namespace El
#define EL_NO_EXCEPT noexcept
#define EL_NO_RELEASE_EXCEPT noexcept
#include <El/core/types.hpp>

extern "C" {
int ref_dist_collect(int d) { return (int)El::Collect((El::Dist)d); }
int ref_dist_partial(int d) { return (int)El::Partial((El::Dist)d); }
int ref_dist_partial_union_row(int u, int v) { return (int)El::PartialUnionRow((El::Dist)u, (El::Dist)v); }
int ref_dist_partial_union_col(int u, int v) { return (int)El::PartialUnionCol((El::Dist)u, (El::Dist)v); }
/* enumerator values by name: 0 MC, 1 MD, 2 MR, 3 VC, 4 VR, 5 STAR, 6 CIRC; 10 NORMAL, 11 TRANSPOSE, 12 ADJOINT; 20 ROW_MAJOR,
 * 21 COLUMN_MAJOR; 30 ELEMENT, 31 BLOCK */
int ref_enum_value(int which) {
    switch (which) {
        case 0: return El::MC; case 1: return El::MD; case 2: return El::MR; case 3: return El::VC; case 4: return El::VR;
        case 5: return El::STAR; case 6: return El::CIRC;
        case 10: return El::NORMAL; case 11: return El::TRANSPOSE; case 12: return El::ADJOINT;
        case 20: return El::ROW_MAJOR; case 21: return El::COLUMN_MAJOR;
        case 30: return El::ELEMENT; case 31: return El::BLOCK;
    }
    return -1;
}
}
