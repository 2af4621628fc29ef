"""Regenerate tests/golden/copy_util.npz from the REFERENCE's own code.

Runs only where /root/reference exists: `make -C oracle ref` compiles oracle/_ref/libref_util.so from the reference's
headers in place (include/El/core/indexing/impl.hpp, include/El/blas_like/level1/Copy/util.hpp,
include/El/blas_like/level1/Axpy/util.hpp; shim: oracle/ref/ref_util.cpp), then every case of tests/golden_cases.py is
run through it and the outputs are stored.  Usage:  python tests/golden/make_golden.py
"""
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import golden_cases as G  # noqa: E402


def main():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "ref", "-s"])
    ref = G.RefBackend()
    assert ref.lib.ref_int_bytes() == 4, "fixtures are generated with the reference's default 32-bit Int"
    cases = G.build_cases()
    outs = [ref.run(fn, dn, p) for (fn, dn, p) in cases]
    idx_in = G.indexing_inputs()
    idx_out = np.zeros((len(idx_in), 6), dtype=np.int64)
    for i, (n, rank, align, stride) in enumerate(idx_in.tolist()):
        shift = ref.lib.ref_shift(rank, align, stride)
        idx_out[i] = (shift, ref.lib.ref_length(n, shift, stride), ref.lib.ref_length4(n, rank, align, stride),
                      ref.lib.ref_max_length(n, stride), ref.lib.ref_mod(rank - align - stride, stride),
                      ref.lib.ref_gcd(max(n, 1), stride))
    G.save_golden(G.GOLDEN, cases, outs, idx_in, idx_out)
    write_dist_algebra()
    write_trapezoid()
    write_summa()
    print(f"{len(cases)} cases, {sum(o.size for o in outs)} output elements, {os.path.getsize(G.GOLDEN)} bytes -> {G.GOLDEN}")


def write_trapezoid():
    """tests/golden/trapezoid.json: SHA-256 of what the reference's ScaleTrapezoid / AxpyTrapezoid / MakeTrapezoidal loops
    (compiled in place by oracle/ref/ref_trapezoid.cpp) produce for every case of tests/trapezoid_cases.py."""
    import trapezoid_cases as T
    ref = T.RefBackend()
    T.save_golden([T.digest(ref.run(c)) for c in T.build_cases()])


def write_summa():
    """tests/golden/summa.json: SHA-256 of C after the reference's own SUMMA_{NT,TN,TT} / Trrk loop code (compiled in place over
    stand-in matrix classes, oracle/ref/ref_summa.cpp) for every case of tests/summa_cases.py."""
    import summa_cases as S
    ref = S.RefBackend()
    S.save_golden([ref.run(c) for c in S.build_cases()])


def write_dist_algebra():
    """tests/golden/dist_algebra.json: the reference's Collect / Partial / PartialUnion tables and enumerator values
    (include/El/core/types.hpp, compiled in place by oracle/ref/ref_types.cpp) for tests/test_shim_align.py."""
    import ctypes
    import json
    ref = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_util.so"))
    names = {"MC": 0, "MD": 1, "MR": 2, "VC": 3, "VR": 4, "STAR": 5, "CIRC": 6, "NORMAL": 10, "TRANSPOSE": 11, "ADJOINT": 12,
             "ROW_MAJOR": 20, "COLUMN_MAJOR": 21, "ELEMENT": 30, "BLOCK": 31}
    table = {
        "enum": {k: ref.ref_enum_value(v) for k, v in names.items()},
        "collect": [ref.ref_dist_collect(d) for d in range(7)],
        "partial": [ref.ref_dist_partial(d) for d in range(7)],
        "partial_union_row": [ref.ref_dist_partial_union_row(u, v) for u in range(7) for v in range(7)],
        "partial_union_col": [ref.ref_dist_partial_union_col(u, v) for u in range(7) for v in range(7)],
    }
    with open(os.path.join(HERE, "dist_algebra.json"), "w") as f:
        json.dump(table, f, indent=0)
        f.write("\n")


if __name__ == "__main__":
    main()
