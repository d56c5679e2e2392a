"""The end-to-end encode cases shared by tests/golden/make_golden_encode.py and the encode parity tests: clips of
SURVEY.md 8(d) (tools/make_clip.py) and the option strings of BASELINE.json's configs, threads and owf pinned."""
import hashlib
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import make_clip  # noqa: E402

REF_BIN = os.path.join(ROOT, "oracle", "_ref", "kvazaar")
CUDA_BIN = os.path.join(ROOT, "kvazaar_b200", "host", "_build", "kvazaar_cuda")
GOLDEN = os.path.join(ROOT, "tests", "golden", "encode_golden.json")

# name -> (width, height, frames, clip seed, options); --threads is pinned because the bitstream does not depend on
# it with --owf 0 (checked by the generator with 1 and 8 threads) but the wall time does
CASES = {
    # BASELINE config 1: the reference's own CPU-runnable case
    "c1_ultrafast": (416, 240, 16, 1, "--preset ultrafast -q 32 --gop lp-g4d3t1 --owf 0"),
    # config 2's options (medium, full search +-32, GOP 8, 4 references: RDOQ, SAO, deblocking, FME level 4) at 416x240
    "c2_small_full32": (416, 240, 17, 2, "--preset medium --me full32 --gop 8 -r 4 --owf 0"),
    # the same search at full8 / full16 / hexbs / dia / tz (the other IME algorithms of search_inter.c:1208-1235)
    "c2_small_full8": (416, 240, 9, 2, "--preset medium --me full8 --gop 8 -r 4 --owf 0"),
    "c2_small_full16_r2": (416, 240, 9, 2, "--preset medium --me full16 --gop 8 -r 2 --owf 0"),
    "c2_small_hexbs": (416, 240, 9, 2, "--preset medium --gop 8 -r 4 --owf 0"),
    "c2_small_tz_subme2": (416, 240, 9, 2, "--preset medium --me tz --subme 2 --gop 8 -r 3 --owf 0"),
    "c2_small_dia_lowdelay": (416, 240, 9, 2, "--preset fast --me dia --gop lp-g4d3t1 --owf 0"),
    # bi-prediction on top of the device search (search_pu_inter_bipred stays the reference's function)
    "c2_small_bipred": (416, 240, 9, 2, "--preset medium --me full16 --gop 8 -r 4 --bipred --owf 0"),
    # tiles: one session context per tile, tile-relative PU origins (encoderstate.c:869-903)
    "tiles_2x2_full16": (416, 240, 9, 2, "--preset medium --me full16 --gop 8 -r 2 --tiles 2x2 --owf 0"),
    # a picture that is not a multiple of 64 in either direction and white noise (worst case for early-outs)
    "odd_size_full16": (200, 136, 9, 5, "--preset medium --me full16 --gop 8 -r 4 --owf 0"),
    # overlapped wavefront (frames in flight, encoderstate.c:751-804): pictures still being reconstructed are brought to
    # the device block by block and their SAD maps made per CTU; vectors restricted by fracmv_within_tile
    "owf2_full32": (416, 240, 17, 2, "--preset medium --me full32 --gop 8 -r 4 --owf 2"),
    "owf1_full16_bipred": (416, 240, 17, 3, "--preset medium --me full16 --gop 8 -r 3 --bipred --owf 1"),
    "owf3_lowdelay_hexbs": (416, 240, 16, 1, "--preset fast --gop lp-g4d3t1 --owf 3"),
    "owf4_odd_size_full8": (200, 136, 17, 5, "--preset medium --me full8 --gop 8 -r 4 --owf 4"),
    "owf2_1080p_full32": (1920, 1080, 6, 2, "--preset medium --me full32 --gop 8 -r 4 --owf 2"),
    # config 2 itself, first frames
    "c2_1080p_3frames": (1920, 1080, 3, 2, "--preset medium --me full32 --gop 8 -r 4 --owf 0"),
    # the bench's configuration for a whole GOP: frames in flight and PU searches launched ahead of the CU walk
    "c2_1080p_owf3_gop": (1920, 1080, 9, 2, "--preset medium --me full32 --gop 8 -r 4 --owf 3"),
    # search range 64 (windows of 129 x 129: the boxes are at their caps and many windows are evaluated on demand)
    "full64_small": (416, 240, 9, 2, "--preset medium --me full64 --gop 8 -r 2 --owf 0"),
    # the hostile clip: white noise, arbitrary vectors, nothing predicts anything (launches ahead mostly miss)
    "noise_full16_owf1": (416, 240, 6, 99, "--preset medium --me full16 --gop 8 -r 2 --owf 1"),
}
# cases cheap enough for the per-call face (one launch per strategy call)
PERCALL_CASES = {
    "c1_ultrafast_4": (416, 240, 4, 1, "--preset ultrafast -q 32 --gop lp-g4d3t1 --owf 0"),
    # width a multiple of 16: at widths = 8 (mod 16) the reference's own AVX2 and generic SAO strategies disagree
    #  and the cuda strategies follow generic, the definition
    "medium_rdoq_2": (208, 144, 2, 5, "--preset medium --gop 8 -r 2 --owf 0"),
}


def clip_path(tmpdir, name, case):
    w, h, n, seed, _ = case
    path = os.path.join(tmpdir, f"{name}_{w}x{h}.yuv")
    if not os.path.exists(path):
        make_clip.write_clip(path, seed, w, h, n)
    return path


def encode(binary, tmpdir, name, case, threads, env=None, timeout=900):
    """-> dict(md5, frames=[(poc, bits, psnr_y, psnr_u, psnr_v)], rc, stderr)"""
    w, h, n, seed, opts = case
    out = os.path.join(tmpdir, f"{name}_{os.path.basename(binary)}_{threads}.hevc")
    cmd = [binary, "-i", clip_path(tmpdir, name, case), "--input-res", f"{w}x{h}", "-o", out, "-n", str(n),
           "--threads", str(threads)] + opts.split()
    e = dict(os.environ)
    e.update(env or {})
    p = subprocess.run(cmd, env=e, capture_output=True, text=True, timeout=timeout)
    frames = [(int(m.group(1)), int(m.group(2)), m.group(3), m.group(4), m.group(5)) for m in
              re.finditer(r"POC\s+(\d+) QP\s+\d+ \(\S+\)\s+(\d+) bits PSNR Y ([\d.]+) U ([\d.]+) V ([\d.]+)", p.stderr)]
    md5 = None
    if p.returncode == 0 and os.path.exists(out):
        md5 = hashlib.md5(open(out, "rb").read()).hexdigest()
    if os.path.exists(out):
        os.remove(out)
    return {"md5": md5, "frames": frames, "rc": p.returncode, "stderr": p.stderr}
