"""Generates tests/golden/encode_golden.json: md5 of the bitstream and the per-frame bits / PSNR lines the UNMODIFIED
reference CLI (oracle/_ref/kvazaar, built by oracle/Makefile from /root/reference) prints for every case of
tests/encode_cases.py.  Run in the build container:  python tests/golden/make_golden_encode.py
Each case is encoded with 1 and with 8 threads and the two bitstreams must agree (with --owf 0 the stream does not
depend on the thread count), which is what lets the GPU box use its own core count."""
import json
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import encode_cases as ec  # noqa: E402


def main():
    out = {}
    with tempfile.TemporaryDirectory() as tmp:
        for group in (ec.CASES, ec.PERCALL_CASES):
            for name, case in group.items():
                a = ec.encode(ec.REF_BIN, tmp, name, case, 8)
                assert a["rc"] == 0 and a["md5"], (name, a["stderr"][-400:])
                if case[0] * case[1] <= 416 * 240:
                    b = ec.encode(ec.REF_BIN, tmp, name, case, 1)
                    assert b["md5"] == a["md5"], f"{name}: the bitstream depends on --threads"
                out[name] = {"md5": a["md5"], "frames": a["frames"], "options": case[4],
                             "clip": list(case[:4])}
                print(name, a["md5"], len(a["frames"]), "frames")
    with open(ec.GOLDEN, "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
