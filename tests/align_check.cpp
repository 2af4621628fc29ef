// CPU-only dump of the shim's distribution algebra and alignment rules (include/El.hpp: Collect / Partial / PartialUnion*,
// detail::align_dim_with, detail::copy_adopts) for tests/test_shim_align.py.  No device call is made: only the pure inline
// functions are used, so this links without libhydrogen_b200.so.
#include <cstdio>
#include "El.hpp"

int main() {
    using namespace El;
    const Dist all[7] = {MC, MD, MR, VC, VR, STAR, CIRC};
    std::printf("{\"enum\": {\"MC\": %d, \"MD\": %d, \"MR\": %d, \"VC\": %d, \"VR\": %d, \"STAR\": %d, \"CIRC\": %d, \"NORMAL\": %d, "
                "\"TRANSPOSE\": %d, \"ADJOINT\": %d, \"ROW_MAJOR\": %d, \"COLUMN_MAJOR\": %d, \"ELEMENT\": %d, \"BLOCK\": %d},\n",
                MC, MD, MR, VC, VR, STAR, CIRC, NORMAL, TRANSPOSE, ADJOINT, ROW_MAJOR, COLUMN_MAJOR, ELEMENT, BLOCK);
    std::printf("\"collect\": [");
    for (int i = 0; i < 7; ++i) std::printf("%s%d", i ? ", " : "", (int)Collect(all[i]));
    std::printf("],\n\"partial\": [");
    for (int i = 0; i < 7; ++i) std::printf("%s%d", i ? ", " : "", (int)Partial(all[i]));
    std::printf("],\n\"partial_union_row\": [");
    for (int i = 0; i < 7; ++i) for (int j = 0; j < 7; ++j) std::printf("%s%d", i + j ? ", " : "", (int)PartialUnionRow(all[i], all[j]));
    std::printf("],\n\"partial_union_col\": [");
    for (int i = 0; i < 7; ++i) for (int j = 0; j < 7; ++j) std::printf("%s%d", i + j ? ", " : "", (int)PartialUnionCol(all[i], all[j]));
    // align_with: matrix [U,V] on an r x c grid aligned with DistData {dCol, ca, dRow, ra}; result per dimension
    // [code, align] with code as detail::align_dim_with returns it.  Strides: MC r, MR c, VC/VR r*c, STAR/CIRC 1, MD lcm.
    const int r = 2, c = 3;
    auto stride = [&](Dist d) { return d == MC ? r : d == MR ? c : (d == VC || d == VR) ? r * c : d == MD ? 6 : 1; };
    std::printf("],\n\"align_with\": [");
    bool first = true;
    const int ca = 5, ra = 4;   // larger than some strides on purpose: the PartialUnion branches reduce them
    for (int u = 0; u < 7; ++u) for (int v = 0; v < 7; ++v) for (int dc = 0; dc < 7; ++dc) for (int dr = 0; dr < 7; ++dr) {
        int a0 = -1, a1 = -1;
        const int k0 = detail::align_dim_with(all[u], PartialUnionCol(all[u], all[v]), stride(all[u]), all[dc], ca, all[dr], ra, &a0);
        const int k1 = detail::align_dim_with(all[v], PartialUnionRow(all[u], all[v]), stride(all[v]), all[dc], ca, all[dr], ra, &a1);
        std::printf("%s[%d, %d, %d, %d]", first ? "" : ", ", k0, a0, k1, a1);
        first = false;
    }
    std::printf("],\n\"copy_adopts\": [");
    first = true;
    for (int s = 0; s < 7; ++s) for (int d = 0; d < 7; ++d) {
        int a = -1;
        const bool k = detail::copy_adopts(all[s], 5, all[d], stride(all[d]), &a);
        std::printf("%s[%d, %d]", first ? "" : ", ", (int)k, a);
        first = false;
    }
    std::printf("]}\n");
    return 0;
}
