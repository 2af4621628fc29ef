"""Deterministic case list for the copy::util / axpy::util / indexing golden vectors (tests/golden/).

One description of every case, three back ends that run it:
  * RefBackend    -- the REFERENCE's own CPU code, oracle/_ref/libref_util.so, compiled by oracle/Makefile from
                     /root/reference/include/El/{core/indexing/impl.hpp, blas_like/level1/Copy/util.hpp,
                     blas_like/level1/Axpy/util.hpp} (only where /root/reference exists; generates the fixtures);
  * OracleBackend -- the C restatement oracle/liboracle.so (checked against the fixtures in -m "not gpu");
  * CudaBackend   -- the product, libhydrogen_b200.so through the C ABI (checked against the fixtures in -m gpu).
Inputs are closed-form (never stored): index patterns exact in fp32 and dyadic rationals; only parameters and
expected outputs are committed (tests/golden/copy_util.npz, written by tests/golden/make_golden.py).
"""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden", "copy_util.npz")
REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libref_util.so")
SENT = -777.0
DTYPES = {"f64": np.float64, "f32": np.float32}

FNS = ["row_pack", "row_unpack", "col_pack", "col_unpack", "prow_pack", "prow_unpack", "pcol_pack", "pcol_unpack",
       "pack2d", "unpack2d", "interleave", "update"]
# C entry point stems shared by the three libraries (prefix ref_/orc_/hyd_)
STEM = {"row_pack": "row_strided_pack", "row_unpack": "row_strided_unpack", "col_pack": "col_strided_pack",
        "col_unpack": "col_strided_unpack", "prow_pack": "partial_row_strided_pack",
        "prow_unpack": "partial_row_strided_unpack", "pcol_pack": "partial_col_strided_pack",
        "pcol_unpack": "partial_col_strided_unpack", "pack2d": "strided_pack", "unpack2d": "strided_unpack",
        "interleave": "interleave", "update": "interleave_update"}
ALPHAS = [0.7, -1.3, 2.0, 1.0 / 3.0]


def index_vec(n, dt):
    """Every element distinct (below 2^24) and exact in fp32."""
    return (np.arange(n, dtype=np.int64) % (1 << 24)).astype(dt)


def frac_vec(n, dt, a, b, c, d):
    """Dyadic rationals ((i*a) mod b - c) / d: exact in fp32 and fp64, products with ALPHAS are not."""
    i = np.arange(n, dtype=np.int64)
    return (((i * a) % b - c) / float(d)).astype(dt)


def _shift(rank, align, stride):
    return (rank - align) % stride


def _length(n, shift, stride):
    return (n - shift - 1) // stride + 1 if n > shift else 0


def _ceil(a, b):
    return -(-a // b)


def build_cases():
    """List of (fn, dtype-name, params tuple).  Fixed seed: the list is part of the fixture's definition."""
    rng = np.random.default_rng(20261104)
    shapes = [(1, 1), (7, 5), (50, 33), (129, 67), (260, 9), (3, 300), (0, 4), (5, 0), (64, 64)]
    cases = []
    for dn in ("f64", "f32"):
        for kind in ("row", "col"):
            for (h, w) in shapes:
                for stride in (1, 2, 3, 4, 8):
                    align = int(rng.integers(0, stride))
                    ld = max(1, h) + int(rng.integers(0, 4))
                    maxl = _ceil(w, stride) if kind == "row" else _ceil(h, stride)
                    portion = max(1, (h if kind == "row" else w) * maxl) + int(rng.integers(0, 3))
                    p = (h, w, align, stride, ld, portion)
                    cases.append((kind + "_pack", dn, p))
                    cases.append((kind + "_unpack", dn, p))
                for (uni, part) in [(1, 1), (2, 2), (4, 2), (2, 4), (3, 2)]:
                    stride = uni * part
                    align = int(rng.integers(0, stride))
                    rank_part = int(rng.integers(0, part))
                    shift_m = _shift(rank_part, align % part, part)
                    n_dim = w if kind == "row" else h
                    loc = _length(n_dim, shift_m, part)
                    lh, lw = (h, loc) if kind == "row" else (loc, w)
                    ld = max(1, lh) + int(rng.integers(0, 4))
                    portion = max(1, (h if kind == "row" else w) * _ceil(n_dim, stride)) + int(rng.integers(0, 3))
                    p = (h, w, align, stride, uni, part, rank_part, shift_m, ld, portion, lh, lw)
                    cases.append(("p" + kind + "_pack", dn, p))
                    cases.append(("p" + kind + "_unpack", dn, p))
        for (h, w) in shapes:
            for (cs, rs) in [(1, 1), (2, 2), (2, 4), (8, 4), (3, 2)]:
                ca, ra = int(rng.integers(0, cs)), int(rng.integers(0, rs))
                ld = max(1, h) + int(rng.integers(0, 4))
                portion = max(1, _ceil(h, cs) * _ceil(w, rs)) + int(rng.integers(0, 3))
                p = (h, w, ca, cs, ra, rs, ld, portion)
                cases.append(("pack2d", dn, p))
                cases.append(("unpack2d", dn, p))
        for (h, w) in [(1, 1), (33, 17), (300, 70), (64, 257)]:
            for ai, (csA, rsA, csB, rsB) in enumerate([(1, h + 1, 1, h), (2, 2 * h + 3, 1, h), (1, h, 3, 3 * h + 1),
                                                       (3, 3 * h, 2, 2 * h + 2), (w + 2, 1, 1, h + 2), (w, 1, 2, 2 * h)]):
                cases.append(("interleave", dn, (h, w, csA, rsA, csB, rsB)))
                cases.append(("update", dn, (h, w, csA, rsA, csB, rsB, ai % len(ALPHAS))))
    return cases


def io_of(fn, dn, p):
    """(source vector, destination vector before the call) of a case."""
    dt = DTYPES[dn]
    if fn in ("row_pack", "col_pack"):
        h, w, align, stride, ld, portion = p
        return index_vec(ld * max(w, 1), dt), np.full(stride * portion, SENT, dt)
    if fn in ("row_unpack", "col_unpack"):
        h, w, align, stride, ld, portion = p
        return index_vec(stride * portion, dt), np.full(ld * max(w, 1), SENT, dt)
    if fn in ("prow_pack", "pcol_pack"):
        ld, portion, lh, lw, uni = p[8], p[9], p[10], p[11], p[4]
        return index_vec(ld * max(lw, 1), dt), np.full(uni * portion, SENT, dt)
    if fn in ("prow_unpack", "pcol_unpack"):
        ld, portion, lh, lw, uni = p[8], p[9], p[10], p[11], p[4]
        return index_vec(uni * portion, dt), np.full(ld * max(lw, 1), SENT, dt)
    if fn == "pack2d":
        h, w, ca, cs, ra, rs, ld, portion = p
        return index_vec(ld * max(w, 1), dt), np.full(cs * rs * portion, SENT, dt)
    if fn == "unpack2d":
        h, w, ca, cs, ra, rs, ld, portion = p
        return index_vec(cs * rs * portion, dt), np.full(ld * max(w, 1), SENT, dt)
    h, w, csA, rsA, csB, rsB = p[:6]
    nA = (h - 1) * csA + (w - 1) * rsA + 1
    nB = (h - 1) * csB + (w - 1) * rsB + 1
    if fn == "interleave":
        return index_vec(nA, dt), np.full(nB, SENT, dt)
    return frac_vec(nA, dt, 37, 101, 50, 64), frac_vec(nB, dt, 53, 97, 48, 32)


def call_args(fn, p, src, dst, ptr):
    """Argument list (without prefix-specific extras) in the order every library uses; ptr(x) wraps a buffer."""
    if fn in ("row_pack", "col_pack"):
        h, w, align, stride, ld, portion = p
        return [h, w, align, stride, ptr(src), ld, ptr(dst), portion]
    if fn in ("row_unpack", "col_unpack"):
        h, w, align, stride, ld, portion = p
        return [h, w, align, stride, ptr(src), portion, ptr(dst), ld]
    if fn in ("prow_pack", "pcol_pack"):
        return list(p[:8]) + [ptr(src), p[8], ptr(dst), p[9]]
    if fn in ("prow_unpack", "pcol_unpack"):
        return list(p[:8]) + [ptr(src), p[9], ptr(dst), p[8]]
    if fn == "pack2d":
        return list(p[:6]) + [ptr(src), p[6], ptr(dst), p[7]]
    if fn == "unpack2d":
        return list(p[:6]) + [ptr(src), p[7], ptr(dst), p[6]]
    h, w, csA, rsA, csB, rsB = p[:6]
    return [h, w, ptr(src), csA, rsA, ptr(dst), csB, rsB]


def _hp(a):
    return C.c_void_p(a.ctypes.data)


class RefBackend:
    """The reference's own loops (oracle/_ref/libref_util.so)."""
    name = "reference"

    def __init__(self, path=REF_LIB):
        self.lib = C.CDLL(path)
        i64, vp = C.c_int64, C.c_void_p
        for sf in ("f64", "f32"):
            for fn in FNS:
                f = getattr(self.lib, f"ref_{STEM[fn]}_{sf}")
                f.restype = None
                n_int = {"row": 4, "col": 4, "prow": 8, "pcol": 8}.get(fn.split("_")[0], 6)
                if fn == "interleave":
                    f.argtypes = [i64, i64, vp, i64, i64, vp, i64, i64]
                elif fn == "update":
                    f.argtypes = [C.c_double, i64, i64, vp, i64, i64, vp, i64, i64]
                else:
                    f.argtypes = [i64] * n_int + [vp, i64, vp, i64]
        for n, k in (("ref_length", 3), ("ref_length4", 4), ("ref_max_length", 2), ("ref_mod", 2), ("ref_shift", 3),
                     ("ref_gcd", 2), ("ref_global_index", 3)):
            f = getattr(self.lib, n)
            f.restype, f.argtypes = i64, [i64] * k

    def run(self, fn, dn, p):
        src, dst = io_of(fn, dn, p)
        a = call_args(fn, p, src, dst, _hp)
        if fn == "update":
            a = [float(DTYPES[dn](ALPHAS[p[6]]))] + a
        getattr(self.lib, f"ref_{STEM[fn]}_{dn}")(*a)
        return dst


class OracleBackend:
    """The C restatement (oracle/liboracle.so)."""
    name = "oracle"

    def __init__(self):
        import oracle_lib as O
        self.lib = O.lib

    def run(self, fn, dn, p):
        src, dst = io_of(fn, dn, p)
        a = call_args(fn, p, src, dst, _hp) + [np.dtype(DTYPES[dn]).itemsize]
        if fn == "update":
            a = [float(DTYPES[dn](ALPHAS[p[6]]))] + a
        getattr(self.lib, f"orc_{STEM[fn]}")(*a)
        return dst


class CudaBackend:
    """The product: libhydrogen_b200.so through the C ABI (seam S2), device buffers via torch."""
    name = "cuda"

    def __init__(self):
        import torch
        from elemental_b200 import lib, check
        self.torch, self.lib, self.check = torch, lib, check

    def run(self, fn, dn, p):
        torch = self.torch
        src, dst = io_of(fn, dn, p)
        dsrc = torch.from_numpy(np.ascontiguousarray(src)).cuda()
        ddst = torch.from_numpy(np.ascontiguousarray(dst)).cuda()
        a = call_args(fn, p, dsrc, ddst, lambda t: C.c_void_p(t.data_ptr()))
        if fn == "update":
            al = DTYPES[dn](ALPHAS[p[6]])
            a = [C.c_double(float(al)) if dn == "f64" else C.c_float(float(al))] + a
        a.append(C.c_void_p(torch.cuda.current_stream().cuda_stream))
        self.check(getattr(self.lib, f"hyd_{STEM[fn]}_{dn}")(*a))
        return ddst.cpu().numpy()


# ---- indexing table (include/El/core/indexing/impl.hpp) ----
def indexing_inputs():
    rng = np.random.default_rng(7)
    rows = []
    for stride in (1, 2, 3, 4, 5, 7, 8, 16, 24):
        for n in (0, 1, 2, stride - 1, stride, stride + 1, 50, 129, 1000, 32768, 2 ** 31 - 1):
            if n < 0:
                continue
            for _ in range(3):
                rows.append((n, int(rng.integers(0, stride)), int(rng.integers(0, stride)), stride))
    return np.array(rows, dtype=np.int64)   # columns: n, rank, align, stride


FULL_LIMIT = 1536   # outputs up to this many elements are stored in full; every output also has a SHA-256 digest


def digest(a):
    import hashlib
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), dtype=np.uint8)


def save_golden(path, cases, outs, idx_in, idx_out):
    fn_code = np.array([FNS.index(c[0]) for c in cases], dtype=np.int16)
    dt_code = np.array([0 if c[1] == "f64" else 1 for c in cases], dtype=np.int8)
    params = np.full((len(cases), 12), -1, dtype=np.int64)
    for i, c in enumerate(cases):
        params[i, :len(c[2])] = c[2]
    sha = np.stack([digest(o) for o in outs])            # of the output in its own dtype (f64 / f32 bytes)
    kept = [o if o.size <= FULL_LIMIT else o[:0] for o in outs]
    off = np.zeros(len(cases) + 1, dtype=np.int64)
    off[1:] = np.cumsum([o.size for o in kept])
    flat = np.concatenate([o.astype(np.float64) for o in kept])   # exact for both dtypes
    np.savez_compressed(path, fn=fn_code, dt=dt_code, params=params, sha=sha, off=off, out=flat,
                        idx_in=idx_in, idx_out=idx_out)


NPAR = {"row_pack": 6, "row_unpack": 6, "col_pack": 6, "col_unpack": 6, "prow_pack": 12, "prow_unpack": 12,
        "pcol_pack": 12, "pcol_unpack": 12, "pack2d": 8, "unpack2d": 8, "interleave": 6, "update": 7}


def load_golden(path=GOLDEN):
    """-> cases, digests (uint8[32] each), full outputs (None where only the digest is stored), idx_in, idx_out."""
    z = np.load(path)
    cases, outs = [], []
    for i in range(len(z["fn"])):
        fn = FNS[int(z["fn"][i])]
        dn = "f64" if int(z["dt"][i]) == 0 else "f32"
        cases.append((fn, dn, tuple(int(x) for x in z["params"][i, :NPAR[fn]])))
        a, b = int(z["off"][i]), int(z["off"][i + 1])
        outs.append(z["out"][a:b].astype(DTYPES[dn]) if b > a else None)
    return cases, z["sha"], outs, z["idx_in"], z["idx_out"]


def check_against_golden(backend, select=None):
    """Run every (selected) golden case through `backend`; returns the number of cases checked."""
    cases, sha, outs, _, _ = load_golden()
    n = 0
    for i, (fn, dn, p) in enumerate(cases):
        if select is not None and not select(fn, dn, p):
            continue
        got = backend.run(fn, dn, p)
        if outs[i] is not None:
            assert np.array_equal(got, outs[i]), (backend.name, fn, dn, p)
        assert np.array_equal(digest(got), sha[i]), (backend.name, fn, dn, p)
        n += 1
    return n
