"""CPU check of the host glue's launching of PU searches ahead of the CU walk (kvazaar_b200/host/kvz_glue_search_inter.c).

tests/stub/enc_stub.c stands in for the device behind include/kvz_cuda_enc.h (LD_PRELOAD): its "results" are a hash of
exactly the records a launch carries -- a pure function of what the device is given, like the real search, but NOT a
search.  If the glue ever took the results of a launch made with other records (another hypothesis, another frame, a
stale slot), the encode would differ from the same encode with launching ahead switched off.  They must be the same
bitstream; and launches made ahead must actually be used.  No GPU involved; the product path never loads the stub."""
import os
import re
import subprocess

import pytest

import encode_cases as ec

ROOT = ec.ROOT
pytestmark = pytest.mark.skipif(not os.path.exists(ec.CUDA_BIN), reason="kvazaar_cuda not built")

CASES = {
    "medium_full32": (416, 240, 9, 2, "--preset medium --me full32 --gop 8 -r 4 --owf 0"),
    "owf2_lowdelay": (416, 240, 9, 3, "--preset medium --me full16 --gop lp-g4d3t1 -r 2 --owf 2"),
    "odd_size_bipred": (200, 136, 9, 5, "--preset medium --me full16 --gop 8 -r 3 --bipred --owf 0"),
    "tiles": (416, 240, 6, 2, "--preset medium --me full16 --gop 8 -r 2 --tiles 2x2 --owf 0"),
}


@pytest.fixture(scope="module")
def stub(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("stub") / "enc_stub.so")
    subprocess.run(["gcc", "-O2", "-shared", "-fPIC", "-I" + os.path.join(ROOT, "include"), "-o", so,
                    os.path.join(ROOT, "tests", "stub", "enc_stub.c")], check=True)
    return so


def _run(stub, tmp, name, threads, **env):
    e = {"LD_PRELOAD": stub, "KVZ_CUDA_MODE": "batch", "KVZ_CUDA_STATS": "1"}
    e.update(env)
    r = ec.encode(ec.CUDA_BIN, tmp, name, CASES[name], threads, env=e)
    assert r["rc"] == 0, r["stderr"][-800:]
    m = re.search(r"(\d+) answered by a launch made ahead, (\d+) searched on the spot; PUs launched (\d+), of them ahead of time (\d+)",
                  r["stderr"])
    assert m, r["stderr"][-800:]
    return r["md5"], [int(v) for v in m.groups()]


@pytest.mark.parametrize("name", list(CASES))
def test_launching_ahead_does_not_change_the_bitstream(name, stub, tmp_path):
    tmp = str(tmp_path)
    plain, c0 = _run(stub, tmp, name, 4, KVZ_CUDA_SPECULATE="0")
    assert c0[0] == 0 and c0[3] == 0 and c0[1] == c0[2] > 100          # nothing ahead: one PU per search
    ahead, c1 = _run(stub, tmp, name, 4)
    assert ahead == plain
    assert c1[0] + c1[1] == c0[1]                                       # the host asked for the same searches
    assert c1[0] > c0[1] // 10 and c1[3] > c1[0]                        # and a good part was answered ahead
    no_children, c2 = _run(stub, tmp, name, 3, KVZ_CUDA_SPEC_CHILDREN="0")
    assert no_children == plain and 0 < c2[3] < c1[3]
    if "--owf 0" in CASES[name][4]:                                     # with --owf 0 the bitstream does not depend on --threads
        one_thread, _ = _run(stub, tmp, name, 1)
        assert one_thread == plain
