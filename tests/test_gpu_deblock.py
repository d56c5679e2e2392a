"""GPU parity: kvz_cuda_deblock_picture (kvz_filter_deblock_lcu over every LCU, filter.c:770-779) against the
oracle, through the C ABI.  Bit-exact; the device runs two picture-wide passes where the oracle follows the
reference's LCU walk."""
import numpy as np
import pytest

import deblock_oracle as do
from test_oracle_deblock import oracle_deblock, CASES

pytestmark = pytest.mark.gpu


def _dev(a):
    import torch
    return torch.from_numpy(a.view(np.int16) if a.dtype == np.uint16 else a).cuda()


def _run(w, h, bitdepth=8, qp=32, cu_qp=0, beta=0, tc=0, slice_b=0, chroma=1, p_intra=0.3, qp_spread=0, blockiness=12, seed=0):
    from kvazaar_b200 import blocks
    rs = np.random.default_rng(seed + w + 7 * h)
    ref_LX = [[i for i in range(16)], [(i + 1) % 4 for i in range(16)]]
    cfg = do.make_cfg(w, h, qp=qp, cu_qp=cu_qp, beta=beta, tc=tc, slice_b=slice_b, chroma=chroma, ref_LX=ref_LX)
    scus = do.random_scus(rs, w, h, slice_b=bool(slice_b), p_intra=p_intra, qp=qp, qp_spread=qp_spread)
    planes = do.picture(rs, w, h, bitdepth, blockiness)
    want = oracle_deblock(cfg, scus, planes, bitdepth)
    d = [_dev(p) for p in (planes if chroma else planes[:1])]
    blocks.deblock_picture(d, scus.view(blocks.SCU_DT), qp, cu_qp, beta, tc, slice_b, ref_LX, bitdepth)
    for k, t in enumerate(d):
        got = t.cpu().numpy()
        got = got.view(np.uint16) if bitdepth > 8 else got
        assert np.array_equal(got, want[k]), (k, int((got != want[k]).sum()))
    assert not np.array_equal(want[0], planes[0])


@pytest.mark.parametrize("case", CASES, ids=[str(i) for i in range(len(CASES))])
def test_cases_of_the_reference_pinning(case):
    kw = {k: v for k, v in case.items() if k != "size"}
    _run(case["size"][0], case["size"][1], **kw)


@pytest.mark.parametrize("bitdepth", [8, 10])
@pytest.mark.parametrize("size", [(8, 8), (72, 8), (8, 72), (200, 136), (416, 240), (1920, 1080)])
def test_sizes_and_bit_depths(size, bitdepth):
    if size in ((8, 8),):
        # a single 8x8 CU: no edge inside the picture, nothing may change
        import torch
        from kvazaar_b200 import blocks
        rs = np.random.default_rng(1)
        scus = do.random_scus(rs, 8, 8)
        planes = do.picture(rs, 8, 8, bitdepth)
        d = [_dev(p) for p in planes]
        blocks.deblock_picture(d, scus.view(blocks.SCU_DT), 32, bitdepth=bitdepth)
        for t, p in zip(d, planes):
            g = t.cpu().numpy()
            assert np.array_equal(g.view(np.uint16) if bitdepth > 8 else g, p)
        return
    _run(size[0], size[1], bitdepth=bitdepth, slice_b=size[0] % 16 == 0, cu_qp=1, qp_spread=4, seed=bitdepth)


def test_strided_planes_and_idempotent_rejects():
    import torch
    from kvazaar_b200 import blocks, KvzCudaError
    rs = np.random.default_rng(2)
    w, h = 200, 136
    scus = do.random_scus(rs, w, h)
    planes = do.picture(rs, w, h)
    want = oracle_deblock(do.make_cfg(w, h), scus, planes)
    big = [torch.zeros((p.shape[0], p.shape[1] + 56), dtype=torch.uint8, device="cuda") for p in planes]
    views = [b[:, 24:24 + p.shape[1]] for b, p in zip(big, planes)]
    for v, p in zip(views, planes):
        v.copy_(_dev(p))
    blocks.deblock_picture(views, scus.view(blocks.SCU_DT), 32, ref_LX=[[i for i in range(16)], [(i + 1) % 4 for i in range(16)]])
    for v, wnt, b in zip(views, want, big):
        assert np.array_equal(v.cpu().numpy(), wnt)
        assert int(b[:, :24].sum()) == 0 and int(b[:, 24 + v.shape[1]:].sum()) == 0
    with pytest.raises(KvzCudaError):
        blocks.deblock_picture([torch.zeros((12, 16), dtype=torch.uint8, device="cuda")], scus.view(blocks.SCU_DT), 32)
