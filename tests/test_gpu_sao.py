"""GPU parity: kvz_cuda_sao_ctu_stats / kvz_cuda_sao_reconstruct (SAO statistics of sao.c:337-455 with the strategies
of sao-generic.c:50-182; kvz_sao_reconstruct sao.c:277-337 over a whole picture) against the oracle, through the
C ABI.  Bit-exact."""
import ctypes as C

import numpy as np
import pytest

import kvzlibs
import sao_oracle as so
from sao_oracle import _pp, ints

pytestmark = pytest.mark.gpu


def _picture(rs, w, h, bitdepth, noise=8):
    """source planes and a 'deblocked reconstruction' = source + smooth-ish noise, 4:2:0"""
    top = 1 << bitdepth
    dt = np.uint8 if bitdepth == 8 else np.uint16
    orig, rec = [], []
    for pw, ph in ((w, h), (w // 2, h // 2), (w // 2, h // 2)):
        base = rs.integers(0, top, (ph, pw)).astype(np.int32)
        base = (base + np.roll(base, 1, 0) + np.roll(base, 1, 1) + np.roll(base, 3, 1)) // 4
        orig.append(np.clip(base, 0, top - 1).astype(dt))
        rec.append(np.clip(base + rs.integers(-noise, noise + 1, base.shape) * (top // 256), 0, top - 1).astype(dt))
    return orig, rec


def _dev(a):
    import torch
    return torch.from_numpy(a.view(np.int16) if a.dtype == np.uint16 else a).cuda()


@pytest.mark.parametrize("bitdepth", [8, 10])
@pytest.mark.parametrize("size", [(64, 64), (152, 104), (416, 240), (1920, 1080)])
def test_ctu_stats(size, bitdepth):
    from kvazaar_b200 import blocks
    rs = np.random.default_rng(size[0] + bitdepth)
    orig, rec = _picture(rs, size[0], size[1], bitdepth)
    got = blocks.sao_ctu_stats([_dev(p) for p in orig], [_dev(p) for p in rec], bitdepth).cpu().numpy().view(blocks.SAO_STATS_DT)
    want = np.concatenate(so.oracle_ctu_stats(orig, rec, bitdepth))
    assert len(got) == len(want)
    assert np.array_equal(got["edge"], want["edge"])
    assert np.array_equal(got["band"], want["band"])


def test_ddistortion_from_stats_equals_the_strategies():
    """the host glue's formulas on the device statistics against kvzo_sao_edge_ddistortion / kvzo_sao_band_ddistortion"""
    from kvazaar_b200 import blocks
    O = kvzlibs.oracle()
    rs = np.random.default_rng(21)
    orig, rec = _picture(rs, 200, 136, 8, noise=12)
    st = blocks.sao_ctu_stats([_dev(p) for p in orig], [_dev(p) for p in rec]).cpu().numpy().view(blocks.SAO_STATS_DT)
    base = 0
    for k in range(3):
        ctu = 64 if k == 0 else 32
        for i, y0, x0, h, w in so.ctu_blocks(orig[k], ctu):
            ob = np.ascontiguousarray(orig[k][y0:y0 + h, x0:x0 + w])
            rb = np.ascontiguousarray(rec[k][y0:y0 + h, x0:x0 + w])
            s = st[base + i]
            for e in range(4):
                offs = [0] + [int(v) for v in rs.integers(-7, 8, 4)]
                want = O.kvzo_sao_edge_ddistortion_8(_pp(ob), _pp(rb), w, h, e, ints(offs))
                assert want == sum(int(s["edge"][e][1][c]) * offs[c] ** 2 - 2 * offs[c] * int(s["edge"][e][0][c]) for c in range(5))
            pos = int(rs.integers(0, 29))
            bands = [int(v) for v in rs.integers(-7, 8, 4)]
            want = O.kvzo_sao_band_ddistortion_8(_pp(ob), _pp(rb), w, h, pos, ints(bands))
            assert want == sum(int(s["band"][1][pos + b]) * bands[b] ** 2 - 2 * bands[b] * int(s["band"][0][pos + b]) for b in range(4))
        base += len(list(so.ctu_blocks(orig[k], ctu)))


@pytest.mark.parametrize("bitdepth", [8, 10])
@pytest.mark.parametrize("size", [(64, 64), (2, 2), (152, 104), (416, 240), (1920, 1080)])
def test_reconstruct_picture(size, bitdepth):
    import torch
    from kvazaar_b200 import blocks
    rs = np.random.default_rng(size[1] + bitdepth)
    _, pre = _picture(rs, size[0], size[1], bitdepth, noise=30)
    counts = [len(list(so.ctu_blocks(p, c))) for p, c in zip(pre, (64, 32, 32))]
    luma, chroma = so.random_params(rs, counts[0], bitdepth), so.random_params(rs, counts[1], bitdepth)
    d_in = [_dev(p) for p in pre]
    d_out = [torch.zeros_like(t) for t in d_in]
    blocks.sao_reconstruct(d_in, d_out, luma.view(blocks.SAO_PARAMS_DT), chroma.view(blocks.SAO_PARAMS_DT), bitdepth)
    want = so.oracle_reconstruct(pre, luma, chroma, bitdepth)
    for k in range(3):
        got = d_out[k].cpu().numpy()
        got = got.view(np.uint16) if bitdepth > 8 else got
        assert np.array_equal(got, want[k]), k
    # type none everywhere: a copy
    none = np.zeros_like(luma)
    blocks.sao_reconstruct(d_in, d_out, none.view(blocks.SAO_PARAMS_DT), np.zeros_like(chroma).view(blocks.SAO_PARAMS_DT), bitdepth)
    assert all(torch.equal(a, b) for a, b in zip(d_in, d_out))


def test_reconstruct_luma_only_strided_and_rejects_aliasing():
    import torch
    from kvazaar_b200 import blocks, KvzCudaError
    rs = np.random.default_rng(33)
    _, pre = _picture(rs, 200, 136, 8, noise=30)
    big = torch.zeros((136, 256), dtype=torch.uint8, device="cuda")
    big[:, 8:208] = _dev(pre[0])
    out = torch.zeros((136, 320), dtype=torch.uint8, device="cuda")
    luma = so.random_params(rs, 12, p_none=0.0)
    blocks.sao_reconstruct([big[:, 8:208]], [out[:, 16:216]], luma.view(blocks.SAO_PARAMS_DT), None)
    want = so.oracle_reconstruct(pre[:1], luma, None)
    assert np.array_equal(out[:, 16:216].cpu().numpy(), want[0])
    assert int(out[:, :16].sum()) == 0 and int(out[:, 216:].sum()) == 0
    with pytest.raises(KvzCudaError):
        blocks.sao_reconstruct([big[:, 8:208]], [big[:, 8:208]], luma.view(blocks.SAO_PARAMS_DT), None)


def test_in_loop_filter_chain_on_resident_planes():
    """deblock -> SAO statistics -> SAO reconstruction -> picture checksum / SSE, every step on the planes the step
    before left on the device, against the same chain of oracle calls"""
    import torch
    import deblock_oracle as do
    from kvazaar_b200 import blocks
    from test_oracle_deblock import oracle_deblock
    from test_oracle_picture_stats import oracle_checksum, oracle_sse
    rs = np.random.default_rng(55)
    w, h, qp = 264, 200, 34
    src = do.picture(rs, w, h, blockiness=0)
    rec = [np.clip(p.astype(np.int32) + np.kron(rs.integers(-10, 11, ((p.shape[0] + 7) // 8, (p.shape[1] + 7) // 8)),
                                                np.ones((8, 8), np.int32))[:p.shape[0], :p.shape[1]], 0, 255).astype(np.uint8) for p in src]
    scus = do.random_scus(rs, w, h, p_intra=0.4, qp=qp)
    ref_LX = [[i for i in range(16)], [(i + 1) % 4 for i in range(16)]]
    # oracle chain
    o_db = oracle_deblock(do.make_cfg(w, h, qp=qp, ref_LX=ref_LX), scus, rec)
    o_stats = np.concatenate(so.oracle_ctu_stats(src, o_db))
    counts = [len(list(so.ctu_blocks(p, c))) for p, c in zip(src, (64, 32, 32))]
    luma, chroma = so.random_params(rs, counts[0]), so.random_params(rs, counts[1])
    o_out = so.oracle_reconstruct(o_db, luma, chroma)
    # device chain
    d_src, d_rec = [_dev(p) for p in src], [_dev(p) for p in rec]
    blocks.deblock_picture(d_rec, scus.view(blocks.SCU_DT), qp, ref_LX=ref_LX)
    st = blocks.sao_ctu_stats(d_src, d_rec).cpu().numpy().view(blocks.SAO_STATS_DT)
    assert np.array_equal(st["edge"], o_stats["edge"]) and np.array_equal(st["band"], o_stats["band"])
    d_out = [torch.empty_like(t) for t in d_rec]
    blocks.sao_reconstruct(d_rec, d_out, luma.view(blocks.SAO_PARAMS_DT), chroma.view(blocks.SAO_PARAMS_DT))
    ck = blocks.picture_checksum(d_out).cpu().numpy().view(np.uint32)
    sse = blocks.picture_sse(d_src, d_out).cpu().numpy()
    assert ck.tolist() == [oracle_checksum(p) for p in o_out]
    assert sse.tolist() == [oracle_sse(a, b) for a, b in zip(src, o_out)]
    assert not np.array_equal(o_out[0], rec[0])
