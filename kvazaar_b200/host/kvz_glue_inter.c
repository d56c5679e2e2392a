/*
 * kvazaar_b200/host/kvz_glue_inter.c -- compiled INSTEAD of src/inter.c: it includes that file in place (from the
 * reference tree, at build time; nothing is copied) and adds ONE function next to kvz_inter_get_mv_cand.
 *
 * search_pu_inter_ref asks for the AMVP candidates of a PU once per reference picture (search_inter.c:1164-1181), and
 * every kvz_inter_get_mv_cand call derives the PU's spatial and temporal neighbours again (inter.c:1209-1228) although
 * they do not depend on the picture -- only the choice among them does (get_mv_cand_from_candidates, inter.c:1102).
 * The glue derives the records of about three PUs per search (the real one and the ones it launches ahead,
 * kvz_glue_search_inter.c), so it asks for all pictures of a PU at once: the neighbours are looked up once.
 */
#include "inter.c"

#include "kvz_glue.h"

/* mv_cand[i] = what kvz_inter_get_mv_cand(state, x, y, width, height, mv_cand, cur_cu, lcu, reflist[i]) gives for a
 * cur_cu whose inter.mv_ref[reflist[i]] is lx_idx[i] */
void kvz_glue_inter_get_mv_cands(const encoder_state_t *const state, int32_t x, int32_t y, int32_t width, int32_t height,
                                 lcu_t *lcu, int n, const int8_t *reflist, const int8_t *lx_idx, int16_t (*mv_cand)[2][2])
{
  merge_candidates_t merge_cand = { {0, 0}, {0, 0, 0}, 0, 0 };
  get_spatial_merge_candidates(x, y, width, height, state->tile->frame->width, state->tile->frame->height, lcu, &merge_cand);
  get_temporal_merge_candidates(state, x, y, width, height, 1, 0, &merge_cand);
  cu_info_t probe;
  memset(&probe, 0, sizeof(probe));
  for (int i = 0; i < n; i++) {
    probe.inter.mv_ref[reflist[i]] = lx_idx[i];
    get_mv_cand_from_candidates(state, x, y, width, height, &merge_cand, &probe, reflist[i], mv_cand[i]);
  }
}
