"""percall-mode debugging: md5 of a case for the reference, glue-off, percall with ALL types forced generic, and percall
with option variations"""
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import encode_cases as ec  # noqa: E402
from percall_bisect import FAMILIES  # noqa: E402

w, h, n, seed = 200, 136, 2, 5
base = "--preset medium --gop 8 -r 2 --owf 0"
with tempfile.TemporaryDirectory() as tmp:
    def run(label, binary, opts, env):
        r = ec.encode(binary, tmp, "dbg", (w, h, n, seed, opts), 8, env=env)
        print(f"{label:50s} rc={r['rc']} md5={r['md5']}", flush=True)
        return r
    all_generic = {"KVZ_CUDA_MODE": "percall"}
    for fam, types in FAMILIES.items():
        for t in types:
            all_generic["KVAZAAR_OVERRIDE_" + t] = "generic"
    for extra in ("", " --rdoq 0", " --sao off", " --subme 0", " --no-deblock", " --no-signhide", " --rdoq 0 --sao off --subme 0",
                  " --me-early-termination off", " --pu-depth-inter 1-3 --pu-depth-intra 2-3"):
        o = base + extra
        a = run("ref" + extra, ec.REF_BIN, o, None)
        b = run("percall" + extra, ec.CUDA_BIN, o, {"KVZ_CUDA_MODE": "percall"})
        c = run("percall all-generic" + extra, ec.CUDA_BIN, o, all_generic)
        print("   ", "OK" if a["md5"] == b["md5"] else "MISMATCH", "| all-generic", "OK" if a["md5"] == c["md5"] else "MISMATCH")
