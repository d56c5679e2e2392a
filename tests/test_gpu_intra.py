"""GPU parity: all 35 intra predictions per PU and their SATD / SAD costs (the arithmetic of
search_intra_rough) against the oracle, through the C ABI.  Bit-exact."""
import numpy as np
import pytest

import kvzlibs
from kvzlibs import ptr

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("bitdepth", [8, 10])
def test_intra_predictions_and_costs(bitdepth):
    import torch
    from kvazaar_b200 import blocks
    O = kvzlibs.oracle()
    rs = np.random.default_rng(70)
    W, H = 256, 128
    pix, top_v, sfx = (np.uint8, 255, "8") if bitdepth == 8 else (np.uint16, 1023, "10")
    orig = kvzlibs.synth_luma(31, W, H, 1)[0]
    if bitdepth > 8:
        orig = np.ascontiguousarray((orig.astype(np.uint16) << 2) | rs.integers(0, 4, orig.shape, dtype=np.uint16))
    dev = lambda a: torch.from_numpy(a.view(np.int16) if a.dtype == np.uint16 else a).cuda()
    rows, poff = [], 0
    ref_chunks = []
    for it in range(96):
        lg = int(rs.integers(2, 6))
        w = 1 << lg
        chroma = int(lg < 5 and rs.random() < 0.3)
        if it % 7 == 0:
            top = np.full(2 * w + 1, top_v, pix); left = np.zeros(2 * w + 1, pix)
        else:
            base = rs.integers(0, top_v + 1)
            step = 40 * (top_v + 1) // 256
            top = np.clip(base + rs.integers(-step, step + 1, 2 * w + 1).cumsum(), 0, top_v).astype(pix)
            left = np.clip(base + rs.integers(-step, step + 1, 2 * w + 1).cumsum(), 0, top_v).astype(pix)
        left[0] = top[0]
        x, y = int(rs.integers(0, (W - w) // 4 + 1)) * 4, int(rs.integers(0, (H - h_) // 4 + 1)) * 4 if (h_ := w) else 0
        roff = sum(len(c) for c in ref_chunks)
        ref_chunks += [top, left]
        rows.append((roff, y * W + x, W, poff, lg, chroma, int(rs.integers(0, 2)), 1))
        poff += 35 * w * w
    jobs = np.array(rows, blocks.INTRA_JOB_DT)
    refs = np.concatenate(ref_chunks)
    costs, preds = blocks.intra_mode_costs(dev(refs), dev(orig), jobs, want_preds=True, bitdepth=bitdepth)
    costs, preds = costs.cpu().numpy(), preds.cpu().numpy().view(pix)
    for i, jb in enumerate(jobs):
        lg = int(jb["log2_width"]); w = 1 << lg
        top = refs[int(jb["ref_offset"]):][:2 * w + 1].copy(); left = refs[int(jb["ref_offset"]) + 2 * w + 1:][:2 * w + 1].copy()
        y, x = int(jb["orig_offset"]) // W, int(jb["orig_offset"]) % W
        blk = np.ascontiguousarray(orig[y:y + w, x:x + w])
        for mode in range(35):
            want = np.zeros(w * w, pix)
            getattr(O, "kvzo_intra_predict_" + sfx)(ptr(top), ptr(left), lg, mode, int(jb["is_chroma"]), int(jb["filter_boundary"]), ptr(want))
            got = preds[int(jb["pred_offset"]) + mode * w * w:][:w * w]
            assert np.array_equal(got, want), ("pred", i, lg, mode, jb)
            assert costs[i, mode, 0] == getattr(O, "kvzo_satd_nxn_" + sfx)(w, ptr(want), ptr(blk)), ("satd", i, lg, mode)
            assert costs[i, mode, 1] == getattr(O, "kvzo_sad_nxn_" + sfx)(w, ptr(want), ptr(blk)), ("sad", i, lg, mode)
