"""Case list and backends that pin the oracle's SUMMA / Trrk restatement (oracle/oracle_gemm.c, oracle_level3.c) to the
REFERENCE's own loop code: src/blas_like/level3/Gemm/{NN,NT,TN,TT}.hpp and Trrk/{Local,NN,NT,TN,TT}.hpp compiled in place over
stand-in matrix classes on the simulated grid (oracle/ref/ref_summa.cpp -> oracle/_ref/libref_summa.so); likewise the stationary-C
loops of Syrk/{LN,LT,UN,UT}.hpp and Symm/{LL,LU,RL,RU}.hpp.

tests/golden/make_golden.py runs every case through the reference's loops and stores the SHA-256 of the resulting global C
in tests/golden/summa.json; tests/test_ref_pin.py requires the oracle to reproduce every digest (always), and the reference
to still produce them where oracle/_ref exists.  Local products use the oracle's plain triple loop (no vendor BLAS), so the
bits depend only on the order of operations -- which is what is being pinned."""
import ctypes as C
import hashlib
import itertools
import json
import os

import numpy as np

import oracle_lib as O
from oracle_lib import MC, MR, VC, VR, STAR

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden", "summa.json")
REF_LIB = os.path.join(os.path.dirname(HERE), "oracle", "_ref", "libref_summa.so")
SEED = 20261106
DISTS = [((MC, MR), (MC, MR), (MC, MR)), ((VC, STAR), (STAR, VR), (MR, MC)), ((STAR, VC), (VC, STAR), (MC, MR)),
         ((STAR, STAR), (MR, MC), (STAR, MR))]
GRIDS = [(1, 1), (2, 2), (2, 3), (4, 2)]


def build_cases():
    rng = np.random.default_rng(SEED)
    cases = []

    def aligns(g, ds):
        return [int(rng.integers(0, O.lib.orc_dist_stride(g, d))) for pair in ds for d in pair]

    for (r, c) in GRIDS:
        g = O.Grid(r, c, 1)
        # Gemm: NN / NT / TN / TT x SUMMA A / B / C / Dot / GEMM_DEFAULT
        for (oA, oB), alg, ds, (m, n, k, nb), dt in itertools.product([(0, 0), (0, 1), (1, 0), (1, 1)], (1, 2, 3, 4, 0), DISTS,
                                                                       [(45, 38, 52, 16), (7, 60, 130, 32)], ("f64",)):
            cases.append(dict(op="gemm", r=r, c=c, oA=oA, oB=oB, alg=alg, ds=ds, m=m, n=n, k=k, nb=nb, nbl=64, dt=dt, al=aligns(g, ds), uplo=0))
        for (oA, oB), alg in itertools.product([(0, 0), (0, 1), (1, 0), (1, 1)], (3, 4)):
            cases.append(dict(op="gemm", r=r, c=c, oA=oA, oB=oB, alg=alg, ds=DISTS[0], m=33, n=29, k=41, nb=8, nbl=64, dt="f32",
                              al=aligns(g, DISTS[0]), uplo=0))
        # Trrk: all four orientations, both triangles, both branches of LocalTrrk (kernel / recursion)
        for uplo, oA, oB, ds, (n, k, nb, nbl) in itertools.product((0, 1), (0, 1), (0, 1), DISTS[:3], [(150, 45, 16, 8), (97, 70, 32, 64)]):
            cases.append(dict(op="trrk", r=r, c=c, oA=oA, oB=oB, alg=3, ds=ds, m=n, n=n, k=k, nb=nb, nbl=nbl, dt="f64", al=aligns(g, ds), uplo=uplo))
        # Syrk (stationary-C loops Syrk/{LN,LT,UN,UT}.hpp; oA = orientation) and Symm (Symm/{LL,LU,RL,RU}.hpp *C; oA = side)
        for uplo, orient, ds, (n, k, nb, nbl) in itertools.product((0, 1), (0, 1), DISTS[:3], [(150, 45, 16, 8), (97, 70, 32, 64)]):
            cases.append(dict(op="syrk", r=r, c=c, oA=orient, oB=1 - orient, alg=3, ds=ds, m=n, n=n, k=k, nb=nb, nbl=nbl, dt="f64",
                              al=aligns(g, ds), uplo=uplo))
        for side, uplo, ds, (m, n, nb) in itertools.product((0, 1), (0, 1), DISTS[:3], [(57, 43, 16), (30, 71, 32)]):
            cases.append(dict(op="symm", r=r, c=c, oA=side, oB=0, alg=3, ds=ds, m=m, n=n, k=(m if side == 0 else n), nb=nb, nbl=64,
                              dt="f64", al=aligns(g, ds), uplo=uplo))
    return cases


def _operands(case):
    g = O.Grid(case["r"], case["c"], 1)
    dt = np.float64 if case["dt"] == "f64" else np.float32
    m, n, k, oA, oB = case["m"], case["n"], case["k"], case["oA"], case["oB"]
    (dA, dB, dC), al = case["ds"], case["al"]
    am, an = (k, m) if oA else (m, k)
    bm, bn = (n, k) if oB else (k, n)
    if case["op"] == "symm":      # A is the symmetric k x k matrix (k = m for LEFT, n for RIGHT), B is m x n
        am, an, bm, bn = k, k, m, n
    return (O.DM(g, am, an, dA[0], dA[1], al[0], al[1], dt).fill(1), O.DM(g, bm, bn, dB[0], dB[1], al[2], al[3], dt).fill(2),
            O.DM(g, m, n, dC[0], dC[1], al[4], al[5], dt).fill(3))


def digest(Cm):
    G, mism = Cm.to_global()
    assert mism == 0
    return hashlib.sha256(np.ascontiguousarray(G).tobytes()).hexdigest()


class OracleBackend:
    def run(self, case):
        O.use_blas(False)
        A, B, Cm = _operands(case)
        O.lib.orc_set_local_trrk_blocksize(case["nbl"])
        try:
            if case["op"] == "gemm":
                rc = O.lib.orc_gemm(case["oA"], case["oB"], case["alg"], 2.0, A.h, B.h, 3.0, Cm.h, case["nb"])
            elif case["op"] == "trrk":
                rc = O.lib.orc_trrk(case["uplo"], case["oA"], case["oB"], 2.0, A.h, B.h, 3.0, Cm.h, case["nb"])
            elif case["op"] == "syrk":
                rc = O.lib.orc_syrk(case["uplo"], case["oA"], 2.0, A.h, 3.0, Cm.h, case["nb"])
            else:
                rc = O.lib.orc_symm(case["oA"], case["uplo"], 2.0, A.h, B.h, 3.0, Cm.h, case["nb"])
        finally:
            O.lib.orc_set_local_trrk_blocksize(64)
        assert rc == 0, rc
        return digest(Cm)


class RefBackend:
    """The reference's own loop code (oracle/_ref/libref_summa.so)."""
    def __init__(self):
        self.lib = C.CDLL(REF_LIB)
        for fn in (self.lib.ref_summa_gemm, self.lib.ref_summa_trrk):
            fn.restype = C.c_int
            fn.argtypes = [C.c_int] * 3 + [C.c_double, C.c_void_p, C.c_void_p, C.c_double, C.c_void_p]
        self.lib.ref_summa_syrk.restype = C.c_int
        self.lib.ref_summa_syrk.argtypes = [C.c_int] * 2 + [C.c_double, C.c_void_p, C.c_double, C.c_void_p]
        self.lib.ref_summa_symm.restype = C.c_int
        self.lib.ref_summa_symm.argtypes = [C.c_int] * 2 + [C.c_double, C.c_void_p, C.c_void_p, C.c_double, C.c_void_p]

    def run(self, case):
        O.use_blas(False)
        A, B, Cm = _operands(case)
        self.lib.ref_summa_set_blocksize(case["nb"])
        self.lib.ref_summa_set_local_trrk_blocksize(case["nbl"])
        if case["op"] == "gemm":
            rc = self.lib.ref_summa_gemm(case["oA"], case["oB"], case["alg"], 2.0, A.h, B.h, 3.0, Cm.h)
        elif case["op"] == "trrk":
            rc = self.lib.ref_summa_trrk(case["uplo"], case["oA"], case["oB"], 2.0, A.h, B.h, 3.0, Cm.h)
        elif case["op"] == "syrk":
            rc = self.lib.ref_summa_syrk(case["uplo"], case["oA"], 2.0, A.h, 3.0, Cm.h)
        else:
            rc = self.lib.ref_summa_symm(case["oA"], case["uplo"], 2.0, A.h, B.h, 3.0, Cm.h)
        assert rc == 0, rc
        return digest(Cm)


def save_golden(digests):
    with open(GOLDEN, "w") as f:
        json.dump({"seed": SEED, "cases": len(digests), "sha256": digests}, f, indent=0)
        f.write("\n")


def load_golden():
    with open(GOLDEN) as f:
        return json.load(f)


def check_against_golden(backend, select=None):
    gold = load_golden()
    cases = build_cases()
    assert gold["seed"] == SEED and gold["cases"] == len(cases)
    n = 0
    for i, case in enumerate(cases):
        if select is not None and not select(i, case):
            continue
        assert backend.run(case) == gold["sha256"][i], (i, case)
        n += 1
    return n
