"""Finds which strategy family makes a percall-mode encode differ from the golden: re-runs the case with every type of
one family forced back to the reference's generic strategy (KVAZAAR_OVERRIDE_<type>=generic, strategyselector.c:258-324).
usage: python tools/percall_bisect.py CASE"""
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import encode_cases as ec  # noqa: E402

FAMILIES = {
    "picture": ["reg_sad", "satd_any_size", "satd_any_size_quad", "pixels_calc_ssd"] +
               [f"{k}_{n}x{n}{d}" for k in ("sad", "satd") for n in (4, 8, 16, 32, 64) for d in ("", "_dual")],
    "dct": ["fast_forward_dst_4x4", "fast_inverse_dst_4x4"] + [f"{k}_{n}x{n}" for k in ("dct", "idct") for n in (4, 8, 16, 32)],
    "quant": ["quant", "quantize_residual", "dequant", "coeff_abs_sum"],
    "quantize_residual_only": ["quantize_residual"],
    "intra": ["angular_pred", "intra_pred_planar"],
    "ipol": ["filter_hpel_blocks_hor_ver_luma", "filter_hpel_blocks_diag_luma", "filter_qpel_blocks_hor_ver_luma",
             "filter_qpel_blocks_diag_luma", "sample_quarterpel_luma", "sample_octpel_chroma",
             "sample_14bit_quarterpel_luma", "sample_14bit_octpel_chroma"],
    "sao": ["sao_edge_ddistortion", "calc_sao_edge_dir", "sao_reconstruct_color", "sao_band_ddistortion"],
    "sao_reconstruct_only": ["sao_reconstruct_color"],
    "sao_band_only": ["sao_band_ddistortion"],
    "sao_edge_only": ["sao_edge_ddistortion", "calc_sao_edge_dir"],
    "nal": ["array_checksum"],
}


def main():
    name = sys.argv[1]
    case = ec.PERCALL_CASES.get(name) or ec.CASES[name]
    want = json.load(open(ec.GOLDEN))[name]["md5"]
    with tempfile.TemporaryDirectory() as tmp:
        for fam, types in [("(none)", [])] + list(FAMILIES.items()):
            env = {"KVZ_CUDA_MODE": "percall", "KVZ_CUDA_STATS": "1"}
            for t in types:
                env["KVAZAAR_OVERRIDE_" + t] = "generic"
            r = ec.encode(ec.CUDA_BIN, tmp, name, case, 8, env=env)
            print(f"{fam:24s} rc={r['rc']} md5={r['md5']} {'== golden' if r['md5'] == want else '!= golden'}", flush=True)
            if r["rc"] != 0:
                print(r["stderr"][-400:])


if __name__ == "__main__":
    main()
