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
    print(f"{len(cases)} cases, {sum(o.size for o in outs)} output elements, {os.path.getsize(G.GOLDEN)} bytes -> {G.GOLDEN}")


if __name__ == "__main__":
    main()
