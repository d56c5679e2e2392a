"""GPU parity: kvz_cuda_picture_checksum / kvz_cuda_picture_sse (picture hash SEI checksum, nal-generic.c:58-82;
compute_psnr's sums, encmain.c:101-131) against the oracle, through the C ABI.  Bit-exact."""
import numpy as np
import pytest

from test_oracle_picture_stats import oracle_checksum, oracle_sse

pytestmark = pytest.mark.gpu


def _planes(rs, w, h, bitdepth):
    dt, top = (np.uint8, 256) if bitdepth == 8 else (np.uint16, 1 << bitdepth)
    return [rs.integers(0, top, s, dtype=dt) for s in ((h, w), (h // 2, w // 2), (h // 2, w // 2))]


def _dev(a):
    import torch
    return torch.from_numpy(a.view(np.int16) if a.dtype == np.uint16 else a).cuda()


@pytest.mark.parametrize("bitdepth", [8, 10])
@pytest.mark.parametrize("size", [(2, 2), (26, 14), (200, 136), (416, 240), (1920, 1080), (3840, 2160)])
def test_checksum_420_picture(size, bitdepth):
    from kvazaar_b200.blocks import picture_checksum
    rs = np.random.default_rng(size[0] + bitdepth)
    planes = _planes(rs, size[0], size[1], bitdepth)
    got = picture_checksum([_dev(p) for p in planes], bitdepth).cpu().numpy().view(np.uint32)
    want = [oracle_checksum(p, bitdepth) for p in planes]
    assert got.tolist() == want


@pytest.mark.parametrize("bitdepth", [8, 10])
def test_checksum_strided_unaligned_views_and_single_plane(bitdepth):
    from kvazaar_b200.blocks import picture_checksum
    rs = np.random.default_rng(5)
    big = _planes(rs, 700, 300, bitdepth)[0]
    for x0, w in ((0, 700), (1, 333), (3, 517), (16, 64), (5, 1)):
        v = big[7:291, x0:x0 + w]
        got = picture_checksum([_dev(big)[7:291, x0:x0 + w]], bitdepth).cpu().numpy().view(np.uint32)
        assert got.tolist() == [oracle_checksum(v, bitdepth)]


@pytest.mark.parametrize("bitdepth", [8, 10])
@pytest.mark.parametrize("size", [(2, 2), (26, 14), (200, 136), (1920, 1080), (3840, 2160)])
def test_sse_420_picture(size, bitdepth):
    from kvazaar_b200.blocks import picture_sse
    rs = np.random.default_rng(size[1] + bitdepth)
    src, rec = _planes(rs, size[0], size[1], bitdepth), _planes(rs, size[0], size[1], bitdepth)
    got = picture_sse([_dev(p) for p in src], [_dev(p) for p in rec], bitdepth).cpu().numpy()
    assert got.tolist() == [oracle_sse(a, b, bitdepth) for a, b in zip(src, rec)]
    same = picture_sse([_dev(p) for p in src], [_dev(p) for p in src], bitdepth).cpu().numpy()
    assert same.tolist() == [0, 0, 0]


def test_sse_strided_views_and_extremes():
    from kvazaar_b200.blocks import picture_sse
    rs = np.random.default_rng(6)
    big = rs.integers(0, 256, (300, 724), dtype=np.uint8)
    d = _dev(big)
    for xa, xb, w in ((0, 12, 700), (1, 2, 333), (3, 19, 517)):
        got = picture_sse([d[:, xa:xa + w]], [d[:, xb:xb + w]]).cpu().numpy()
        assert got.tolist() == [oracle_sse(big[:, xa:xa + w], big[:, xb:xb + w])]
    # the largest sum a 4K 10-bit plane can reach exceeds 2^32 by far
    a = np.zeros((2160, 3840), np.uint16)
    b = np.full((2160, 3840), 1023, np.uint16)
    got = picture_sse([_dev(a)], [_dev(b)], 10).cpu().numpy()
    assert got.tolist() == [2160 * 3840 * 1023 * 1023]
