"""Generates tests/golden/rdoq_vectors.npz: kvz_rdoq (rdo.c:548-885) outputs of the UNMODIFIED reference compiled
into oracle/_ref/libkvazaar_ref.so, for the seeded cases of tests/rdoq_cases.py with random CABAC context states
(needs /root/reference: build container only).  tests/test_golden.py checks the oracle (CPU) and the CUDA path
(GPU box) against these vectors without the reference being present.

    python tests/golden/make_golden_rdoq.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import kvzlibs  # noqa: E402
import rdoq_cases  # noqa: E402
from test_oracle_rdoq import run_ref  # noqa: E402

PARAMS = ("width", "type", "scan_idx", "cu_is_intra", "tr_depth", "qp", "signhide", "slice_is_intra")


def main():
    rs = np.random.default_rng(2024)
    params, lams, ctxs, coefs, outs = [], [], [], [], []
    for case in rdoq_cases.cases(77, per_shape=1):
        ctx = kvzlibs.random_rdoq_ctx(rs)
        out = run_ref(case, ctx)
        params.append([case[k] for k in PARAMS])
        lams.append(case["lam"])
        ctxs.append(np.frombuffer(bytes(ctx), np.uint8))
        coefs.append(case["coef"])
        outs.append(out)
    np.savez_compressed(os.path.join(HERE, "rdoq_vectors.npz"), params=np.array(params, np.int32),
                        lam=np.array(lams, np.float64), ctx=np.stack(ctxs), coef=np.concatenate(coefs),
                        levels=np.concatenate(outs))
    print(len(params), "cases,", sum(int(np.any(o != 0)) for o in outs), "with non-zero levels")


if __name__ == "__main__":
    main()
