"""End-to-end parity on the GPU: the reference encoder with the CUDA path bound in (kvazaar_b200/host/, driven through
its own CLI -> kvz_api) produces the bitstream -- md5 -- and the per-frame bits / PSNR lines of the unmodified reference
for identical options (tests/golden/encode_golden.json; where oracle/_ref travelled, also a live run of it).

batch mode:   every search_pu_inter reference loop runs on the device in the host's CTU wavefront order
              (kvz_cuda_enc_search_pu); all IME algorithms, FME levels, bipred, tiles, odd picture sizes.
percall mode: kvz_strategyselector_init picks the "cuda" strategies (priority 50) for every type of the seven families;
              KVAZAAR_OVERRIDE_<type> still flips a type back (strategyselector.c:258-324)."""
import json
import os
import re

import pytest

import encode_cases as ec

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not os.path.exists(ec.CUDA_BIN), reason="kvazaar_cuda not built")]
THREADS = min(16, os.cpu_count() or 4)


@pytest.fixture(scope="module")
def golden():
    return json.load(open(ec.GOLDEN))


def _counters(stderr):
    out = {}
    for m in re.finditer(r"(\w+)=(\d+)", " ".join(l for l in stderr.splitlines() if l.startswith("kvz_glue:"))):
        out[m.group(1)] = out.get(m.group(1), 0) + int(m.group(2))
    m = re.search(r"kernels launched (\d+)", stderr)
    out["launches"] = int(m.group(1)) if m else 0
    return out


@pytest.mark.parametrize("name", list(ec.CASES))
def test_batch_mode_bitstream_equals_reference(name, golden, tmp_path):
    case = ec.CASES[name]
    r = ec.encode(ec.CUDA_BIN, str(tmp_path), name, case, THREADS, env={"KVZ_CUDA_MODE": "batch", "KVZ_CUDA_STATS": "1"})
    assert r["rc"] == 0, r["stderr"][-600:]
    assert r["md5"] == golden[name]["md5"], name
    assert r["frames"] == [tuple(f) for f in golden[name]["frames"]]
    c = _counters(r["stderr"])
    # the device did the searching: every search_pu_inter answered by the device (one launch carries up to four PUs,
    # most of them sent ahead of the walk: kvz_glue_search_inter.c), every inter frame begun
    assert c["me_pu_searches"] > 50 * (case[2] - 1) and c["launches"] >= c["me_pu_searches"] // 4
    assert c["me_frames"] >= case[2] - 1 - (case[2] - 1) // 64
    if os.path.exists(ec.REF_BIN) and name in ("c2_small_full32", "tiles_2x2_full16"):
        live = ec.encode(ec.REF_BIN, str(tmp_path), name, case, THREADS)
        assert live["md5"] == r["md5"] and live["frames"] == r["frames"]


@pytest.mark.parametrize("name", list(ec.PERCALL_CASES))
def test_percall_mode_selects_cuda_strategies_and_matches(name, golden, tmp_path):
    case = ec.PERCALL_CASES[name]
    r = ec.encode(ec.CUDA_BIN, str(tmp_path), name, case, THREADS, env={"KVZ_CUDA_MODE": "percall", "KVZ_CUDA_STATS": "1"})
    assert r["rc"] == 0, r["stderr"][-600:]
    assert r["md5"] == golden[name]["md5"], name
    assert r["frames"] == [tuple(f) for f in golden[name]["frames"]]
    m = re.search(r"per-call strategies served: picture=(\d+) dct=(\d+) quant=(\d+) intra=(\d+) ipol=(\d+) sao=(\d+) nal=(\d+)",
                  r["stderr"])
    served = dict(zip(("picture", "dct", "quant", "intra", "ipol", "sao", "nal"), map(int, m.groups())))
    assert served["picture"] > 1000 and served["quant"] > 100 and served["intra"] > 100 and served["nal"] >= 3
    if "ultrafast" not in name:                 # medium: FME, SAO and RDOQ (on the device: no CPU fallback) are on
        assert served["ipol"] > 10 and served["sao"] > 10
    assert "me_pu_searches=0" in r["stderr"]    # the search itself ran the host code over the cuda strategies


def test_percall_override_flips_a_type_back(golden, tmp_path):
    name = "c1_ultrafast_4"
    case = ec.PERCALL_CASES[name]
    env = {"KVZ_CUDA_MODE": "percall", "KVZ_CUDA_STATS": "1"}
    base = ec.encode(ec.CUDA_BIN, str(tmp_path), name, case, THREADS, env=env)
    env.update({"KVAZAAR_OVERRIDE_angular_pred": "generic", "KVAZAAR_OVERRIDE_intra_pred_planar": "avx2"})
    r = ec.encode(ec.CUDA_BIN, str(tmp_path), name, case, THREADS, env=env)
    assert r["rc"] == 0 and r["md5"] == golden[name]["md5"] == base["md5"]
    intra = lambda s: int(re.search(r"served: .* intra=(\d+)", s).group(1))
    assert intra(base["stderr"]) > 100 and intra(r["stderr"]) == 0
