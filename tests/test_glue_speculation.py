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


def test_kvz_api_driver_gives_the_cli_bitstream(stub, tmp_path):
    """kvazaar_b200/encoder.py (kvz_api through ctypes, the library dlopen'ed: what bench.py drives) and the reference's CLI
    linked against the same library produce the same bitstream -- with the stand-in device, on the CPU."""
    import sys
    name = "medium_full32"
    w, h, n, seed, opts = CASES[name]
    cli, _ = _run(stub, str(tmp_path), name, 4)
    code = f"""
import sys, hashlib
sys.path.insert(0, {os.path.join(ROOT, 'tools')!r}); sys.path.insert(0, {ROOT!r})
import make_clip
from kvazaar_b200 import encoder
frames = list(make_clip.clip_planes({seed}, {w}, {h}, {n}))
md5, out = hashlib.md5(), 0
with encoder.Encoder(encoder.CUDA_LIB, {w}, {h}, encoder.options_from_cli({opts!r} + ' --threads 4')) as e:
    for y, u, v in frames:
        d, _ = e.encode(e.picture(y, u, v))
        if d: md5.update(d); out += 1
    while out < {n}:
        d, _ = e.encode(None)
        if d is None: break
        md5.update(d); out += 1
print(out, md5.hexdigest())
"""
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, LD_PRELOAD=stub, KVZ_CUDA_MODE="batch"))
    assert p.returncode == 0, p.stderr[-600:]
    out, md5 = p.stdout.strip().splitlines()[-1].split()
    assert int(out) == n and md5 == cli
