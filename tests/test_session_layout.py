"""CPU check: the ctypes mirrors in kvazaar_b200/session.py have the layout gcc gives the structs of
include/kvz_cuda_enc.h (size and the offset of every field a driver fills)."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

PROBE = r"""
#include <stddef.h>
#include <stdio.h>
#include "kvz_cuda_enc.h"
#define F(S, M) printf(#S "." #M " %zu\n", offsetof(S, M))
int main(void)
{
  printf("kvz_cuda_enc_cfg %zu\nkvz_cuda_enc_ref %zu\nkvz_cuda_enc_frame %zu\n", sizeof(kvz_cuda_enc_cfg), sizeof(kvz_cuda_enc_ref),
         sizeof(kvz_cuda_enc_frame));
  F(kvz_cuda_enc_cfg, device); F(kvz_cuda_enc_cfg, range); F(kvz_cuda_enc_cfg, fme_level); F(kvz_cuda_enc_cfg, owf);
  F(kvz_cuda_enc_cfg, max_ref_lcu_down); F(kvz_cuda_enc_cfg, max_map_mbytes);
  F(kvz_cuda_enc_ref, luma); F(kvz_cuda_enc_ref, stride); F(kvz_cuda_enc_ref, poc); F(kvz_cuda_enc_ref, key);
  F(kvz_cuda_enc_ref, mv_hints); F(kvz_cuda_enc_ref, n_mv_hints); F(kvz_cuda_enc_ref, in_flight);
  F(kvz_cuda_enc_frame, frame_id); F(kvz_cuda_enc_frame, src_luma); F(kvz_cuda_enc_frame, src_stride); F(kvz_cuda_enc_frame, src_key);
  F(kvz_cuda_enc_frame, src_pts); F(kvz_cuda_enc_frame, tile_x); F(kvz_cuda_enc_frame, tile_h); F(kvz_cuda_enc_frame, ref_LX);
  F(kvz_cuda_enc_frame, n_refs); F(kvz_cuda_enc_frame, refs);
  return 0;
}
"""


def test_ctypes_mirrors_match_the_header(tmp_path):
    from kvazaar_b200 import session
    src, exe = tmp_path / "probe.c", tmp_path / "probe"
    src.write_text(PROBE)
    subprocess.run(["gcc", "-std=c99", "-I" + os.path.join(ROOT, "include"), "-o", str(exe), str(src)], check=True)
    want = dict(line.split() for line in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.splitlines())
    mirrors = {"kvz_cuda_enc_cfg": session.EncCfg, "kvz_cuda_enc_ref": session.EncRef, "kvz_cuda_enc_frame": session.EncFrame}
    import ctypes as C
    for key, value in want.items():
        if "." in key:
            struct, field = key.split(".")
            assert getattr(mirrors[struct], field).offset == int(value), key
        else:
            assert C.sizeof(mirrors[key]) == int(value), key
