"""Helpers shared by the -m gpu parity tests: move oracle-side distributed matrices onto the GPU
ranks and back.  The oracle (tests/oracle_lib.py) is the checker; the product path is the C ABI."""
import numpy as np
import torch

import oracle_lib as O
from elemental_b200 import host as H


def to_rank(grid, oracle_dm, ldim_extra=0):
    """This rank's device DistMatrix holding the same block as the oracle's copy of rank grid.Rank()."""
    p = grid.Rank()
    lh = oracle_dm.lh(p)
    q = 16 // oracle_dm.dtype.itemsize
    ld = max(q, -(-lh // q) * q) + ldim_extra
    d = H.DistMatrix(grid, oracle_dm.m, oracle_dm.n, oracle_dm.cd, oracle_dm.rd, oracle_dm.dtype, oracle_dm.ca, oracle_dm.ra,
                     ldim=ld)
    assert (d.lh, d.lw) == (lh, oracle_dm.lw(p)), ((d.lh, d.lw), (lh, oracle_dm.lw(p)))
    d.store.fill_(-12345.0)  # sentinel in the ldim padding
    if d.lh and d.lw:
        blk = np.ascontiguousarray(oracle_dm.block(p).T)
        d.store[: d.lw, : d.lh].copy_(torch.from_numpy(blk))
    return d


def block_of(d):
    return d.local_numpy()


def padding_untouched(d):
    if d.ld == d.lh or d.lw == 0:
        return True
    return bool((d.store[: d.lw, d.lh:] == -12345.0).all().item())


def rand_align(rng, og, dist):
    return int(rng.integers(0, O.lib.orc_dist_stride(og, dist)))
