"""CPU checks of the end-to-end encode goldens (tests/golden/encode_golden.json, made by make_golden_encode.py from the
unmodified reference CLI): the compiled reference still reproduces them here, the reference-with-glue binary
(kvazaar_b200/host/_build/kvazaar_cuda) is the same encoder when the CUDA path is switched off, and it refuses to
encode -- loudly, without a CPU fallback -- when the CUDA path is on and there is no GPU."""
import json
import os

import pytest

import encode_cases as ec

pytestmark = pytest.mark.skipif(not os.path.exists(ec.REF_BIN), reason="oracle/_ref not built")


@pytest.fixture(scope="module")
def golden():
    return json.load(open(ec.GOLDEN))


def _frames(g):
    return [tuple(f) for f in g["frames"]]


@pytest.mark.parametrize("name", ["c1_ultrafast", "c2_small_full8"])
def test_reference_cli_reproduces_golden(name, golden, tmp_path):
    r = ec.encode(ec.REF_BIN, str(tmp_path), name, ec.CASES[name], 4)
    assert r["rc"] == 0 and r["md5"] == golden[name]["md5"]
    assert r["frames"] == _frames(golden[name])


@pytest.mark.skipif(not os.path.exists(ec.CUDA_BIN), reason="kvazaar_cuda not built")
@pytest.mark.parametrize("name", ["c1_ultrafast", "c2_small_hexbs"])
def test_glue_binary_with_cuda_off_is_the_reference(name, golden, tmp_path):
    r = ec.encode(ec.CUDA_BIN, str(tmp_path), name, ec.CASES[name], 4, env={"KVZ_CUDA_MODE": "off"})
    assert r["rc"] == 0 and r["md5"] == golden[name]["md5"]
    assert r["frames"] == _frames(golden[name])


@pytest.mark.skipif(not os.path.exists(ec.CUDA_BIN), reason="kvazaar_cuda not built")
@pytest.mark.parametrize("mode", ["batch", "percall"])
def test_glue_binary_fails_loudly_without_a_gpu(mode, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    r = ec.encode(ec.CUDA_BIN, str(tmp_path), "c1_ultrafast", ec.CASES["c1_ultrafast"], 2, env={"KVZ_CUDA_MODE": mode})
    assert r["rc"] != 0 and r["md5"] is None
    assert "no CUDA device" in r["stderr"] and "no CPU fallback" in r["stderr"]
