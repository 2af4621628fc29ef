"""Case list and backends for the trapezoid level-1 routines (ScaleTrapezoid / AxpyTrapezoid / MakeTrapezoidal), shared by
tests/golden/make_golden.py (which runs them through the REFERENCE's own headers compiled in place, oracle/ref/ref_trapezoid.cpp,
and stores SHA-256 digests in tests/golden/trapezoid.json), tests/test_ref_pin.py (oracle vs fixtures, and vs the reference
live where oracle/_ref exists) and tests/test_gpu_golden.py (the CUDA kernels vs the same fixtures).

A case is one distributed matrix on a simulated r x c grid; every rank's local block is pushed through the routine and the
digest covers all ranks' blocks in rank order.  `local` cases are 1x1 grids run through the reference's Matrix<T> overloads."""
import ctypes as C
import hashlib
import json
import os

import numpy as np

import oracle_lib as O
from oracle_lib import MC, MR, VC, VR, STAR

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden", "trapezoid.json")
REF_LIB = os.path.join(os.path.dirname(HERE), "oracle", "_ref", "libref_util.so")
SEED = 20261105
OPS = ("scale", "axpy", "make")


def build_cases():
    rng = np.random.default_rng(SEED)
    cases = []
    grids = [(1, 1), (2, 2), (2, 3), (4, 2)]
    dists = [(MC, MR), (MR, MC), (VC, STAR), (STAR, VR), (MC, STAR), (STAR, STAR)]
    for op in OPS:
        for dt in ("f64", "f32"):
            for r, c in grids:
                for k in range(10):
                    cd, rd = dists[int(rng.integers(0, len(dists)))]
                    m, n = int(rng.integers(0, 40)), int(rng.integers(0, 40))
                    if op == "axpy" and k % 2 == 0:
                        n = m                        # the Matrix overload of AxpyTrapezoid is only meaningful for square matrices
                    g = O.Grid(r, c, 1)
                    ca = int(rng.integers(0, O.lib.orc_dist_stride(g, cd)))
                    ra = int(rng.integers(0, O.lib.orc_dist_stride(g, rd)))
                    uplo, offset = int(rng.integers(0, 2)), int(rng.integers(-6, 7))
                    local = r * c == 1 and k % 2 == 0 and (op != "axpy" or m == n)
                    # The reference's Matrix overload of ScaleTrapezoid(UPPER) scales rows 0..j-offset of column j WITHOUT clamping
                    # to the height (ScaleTrapezoid.hpp:27-33): it writes out of bounds unless width - offset <= height.  Such
                    # shapes go through its AbstractDistMatrix overload (which clamps, :66-68) instead.
                    if op == "scale" and uplo == 1 and n - offset > m:
                        local = False
                    # LocalAxpyTrapezoid(UPPER) has the same hazard (rows 0..j-offset unclamped, AxpyTrapezoid.hpp:75-82; its
                    # square-matrix precondition is only checked in debug builds and does not cover negative offsets)
                    if op == "axpy" and uplo == 1 and n - offset > m:
                        uplo = 0
                    cases.append(dict(op=op, dt=dt, r=r, c=c, cd=cd, rd=rd, ca=ca, ra=ra, m=m, n=n, uplo=uplo, offset=offset,
                                      alpha=float(2.0 ** int(rng.integers(-2, 3)) * (-1) ** k), local=int(local),
                                      pad=int(rng.integers(0, 3)), seed=int(rng.integers(0, 2 ** 31))))
    return cases


def _np(dt):
    return np.float64 if dt == "f64" else np.float32


def geometry(case, p):
    """{height, width, colShift, colStride, rowShift, rowStride} of rank p's block, from the oracle's index functions."""
    g = O.Grid(case["r"], case["c"], 1)
    mc, mr = C.c_int(), C.c_int()
    O.lib.orc_grid_coords(g, p, C.byref(mc), C.byref(mr))
    cst, rst = O.lib.orc_dist_stride(g, case["cd"]), O.lib.orc_dist_stride(g, case["rd"])
    cs = O.lib.orc_shift(O.lib.orc_dist_rank(g, case["cd"], mc.value, mr.value), case["ca"], cst)
    rs = O.lib.orc_shift(O.lib.orc_dist_rank(g, case["rd"], mc.value, mr.value), case["ra"], rst)
    return case["m"], case["n"], cs, cst, rs, rst


def inputs(case):
    """Global Y and X (seeded); blocks are cut out with the geometry."""
    rng = np.random.default_rng(case["seed"])
    dt = _np(case["dt"])
    return rng.uniform(-1, 1, size=(case["m"], case["n"])).astype(dt), rng.uniform(-1, 1, size=(case["m"], case["n"])).astype(dt)


def block(G, geo):
    m, n, cs, cst, rs, rst = geo
    return np.asfortranarray(G[cs::cst, rs::rst])


def digest(blocks):
    h = hashlib.sha256()
    for b in blocks:
        h.update(np.ascontiguousarray(b.T).tobytes())
    return h.hexdigest()


class RefBackend:
    """The reference's own loops (oracle/_ref/libref_util.so), one rank's block at a time."""
    def __init__(self):
        self.lib = C.CDLL(REF_LIB)

    def run(self, case):
        Y, X = inputs(case)
        out = []
        for p in range(case["r"] * case["c"]):
            geo = geometry(case, p)
            y, x = block(Y, geo), block(X, geo)
            lh, lw = y.shape
            ld = max(lh, 1) + case["pad"]
            buf = np.full((ld, max(lw, 1)), 7.0, dtype=y.dtype, order="F")
            buf[:lh, :lw] = y
            xb = np.zeros((max(lh, 1), max(lw, 1)), dtype=y.dtype, order="F")
            xb[:lh, :lw] = x
            geom = (C.c_int * 7)(*geo, ld)
            sfx = case["dt"]
            if case["op"] == "scale":
                getattr(self.lib, "ref_scale_trapezoid_" + sfx)(C.c_double(case["alpha"]), case["uplo"], geom, C.c_void_p(buf.ctypes.data), case["offset"], case["local"])
            elif case["op"] == "axpy":
                getattr(self.lib, "ref_axpy_trapezoid_" + sfx)(case["uplo"], C.c_double(case["alpha"]), geom, C.c_void_p(xb.ctypes.data), max(lh, 1),
                                                              C.c_void_p(buf.ctypes.data), case["offset"], case["local"])
            else:
                getattr(self.lib, "ref_make_trapezoidal_" + sfx)(case["uplo"], geom, C.c_void_p(buf.ctypes.data), case["offset"], case["local"])
            assert (buf[lh:, :] == 7.0).all()          # the ldim padding is never touched
            out.append(buf[:lh, :lw].copy())
        return out


class OracleBackend:
    """oracle/oracle_level3.c on the simulated grid, all ranks at once."""
    def run(self, case):
        Y, X = inputs(case)
        g = O.Grid(case["r"], case["c"], 1)
        dt = _np(case["dt"])
        y = O.DM(g, case["m"], case["n"], case["cd"], case["rd"], case["ca"], case["ra"], dt).from_global(Y)
        if case["op"] == "scale":
            O.lib.orc_scale_trapezoid(case["alpha"], case["uplo"], y.h, case["offset"])
        elif case["op"] == "axpy":
            x = O.DM(g, case["m"], case["n"], case["cd"], case["rd"], case["ca"], case["ra"], dt).from_global(X)
            assert O.lib.orc_axpy_trapezoid(case["uplo"], case["alpha"], x.h, y.h, case["offset"]) == 0
        else:
            O.lib.orc_make_trapezoidal(case["uplo"], y.h, case["offset"])
        return [y.block(p).copy() for p in range(case["r"] * case["c"])]   # block() is a view into storage the DM frees


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
        if select is not None and not select(case):
            continue
        assert digest(backend.run(case)) == gold["sha256"][i], (i, case)
        n += 1
    return n
