"""Pins the oracle's intra prediction restatement (planar, DC with edge
filter, the 33 angular modes, reference smoothing, boundary post-filter) on the compiled
reference's kvz_intra_predict (intra.c:281-332), generic and AVX2 strategies."""
import numpy as np
import pytest

import kvzlibs
from kvzlibs import ptr


@pytest.mark.parametrize("cpuid", [0, 1], ids=["generic", "avx2"])
def test_intra_predict_vs_reference(oracle_lib, ref_lib, cpuid):
    L = kvzlibs.ref(cpuid)
    rs = np.random.default_rng(60)
    for lg in (2, 3, 4, 5):
        w = 1 << lg
        for it in range(6):
            if it == 0:
                top = np.full(2 * w + 1, 255, np.uint8); left = np.zeros(2 * w + 1, np.uint8)
            else:
                top = rs.integers(0, 256, 2 * w + 1, dtype=np.uint8)
                left = rs.integers(0, 256, 2 * w + 1, dtype=np.uint8)
            left[0] = top[0]
            for mode in range(35):
                for color in (0, 1):
                    if color == 1 and lg == 5:
                        continue
                    for fb in (0, 1):
                        want = kvzlibs.aligned((w * w,), np.uint8); got = np.zeros(w * w, np.uint8)
                        L.ref_shim_intra_predict(ptr(top), ptr(left), lg, mode, color, fb, ptr(want))
                        oracle_lib.kvzo_intra_predict_8(ptr(top), ptr(left), lg, mode, color, fb, ptr(got))
                        assert np.array_equal(got, want), (lg, it, mode, color, fb)
