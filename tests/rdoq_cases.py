"""Seeded RDOQ inputs shared by the oracle pinning test, the golden generator and the GPU parity test."""
import numpy as np


def coefficient_block(rs, width, kind):
    """Transform-coefficient-like int16 blocks: energy decaying away from DC, a few shapes of sparsity."""
    n = width
    yy, xx = np.mgrid[0:n, 0:n]
    if kind == "decay":
        scale = 600.0 / (1.0 + 0.9 * (xx + yy))
    elif kind == "flat":
        scale = np.full((n, n), 90.0)
    elif kind == "sparse":
        scale = 400.0 * (rs.random((n, n)) < 0.08)
    elif kind == "big":
        scale = 9000.0 / (1.0 + 0.3 * (xx + yy))
    elif kind == "tiny":
        scale = np.full((n, n), 14.0)
    elif kind == "zero":
        scale = np.zeros((n, n))
    else:
        raise ValueError(kind)
    v = np.rint(rs.normal(0.0, 1.0, (n, n)) * scale)
    return np.clip(v, -32768, 32767).astype(np.int16).reshape(-1)


KINDS = ("decay", "flat", "sparse", "big", "tiny", "zero")


def cases(seed, per_shape=2, widths=(4, 8, 16, 32)):
    """Yields dicts: width, type (0 luma / 2 chroma), scan_idx, cu_is_intra, tr_depth, qp, signhide, lambda, coef."""
    rs = np.random.default_rng(seed)
    for width in widths:
        for type_ in (0, 2):
            if type_ == 2 and width == 32:
                continue                       # chroma transform blocks are at most 16x16
            for scan_idx in ((0, 1, 2) if width <= 8 else (0,)):
                for kind in KINDS:
                    for _ in range(per_shape):
                        qp = int(rs.integers(12, 46))
                        yield dict(width=width, type=type_, scan_idx=scan_idx,
                                   cu_is_intra=int(rs.integers(0, 2)), tr_depth=int(rs.integers(0, 3)),
                                   qp=qp, signhide=int(rs.integers(0, 2)), slice_is_intra=int(rs.integers(0, 2)),
                                   lam=float(0.57 * 2.0 ** ((qp - 12) / 3.0) * rs.uniform(0.5, 1.5)),
                                   kind=kind, coef=coefficient_block(rs, width, kind))
