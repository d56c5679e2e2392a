/*
 * kvazaar_b200/host/kvz_glue_api.c -- the public API table with the CUDA session's lifetime tied to the encoder's.
 *
 * Compiled INSTEAD of src/kvazaar.c: it includes that file in place (from the reference tree, at build time) with
 * kvz_api_get renamed, and defines kvz_api_get (src/kvazaar.h:707) to return the same twelve entry points
 * (src/kvazaar.c:378-395) except that encoder_close first closes the encoder-session of the motion search
 * (include/kvz_cuda_enc.h) -- device planes, SAD maps, streams -- which the glue of search_inter.c opens lazily at the
 * first inter frame.  Signatures, structures and semantics of kvz_api are unchanged.
 */
#define kvz_api_get kvz_api_get_host
#include "kvazaar.c"
#undef kvz_api_get

#include "kvz_glue.h"

static void glue_encoder_close(kvz_encoder *encoder)
{
  kvazaar_close(encoder);            /* stops the thread queue first: frames still in flight (--owf) finish their searches */
  kvz_glue_session_close();
}

static kvz_encoder *glue_encoder_open(const kvz_config *cfg)
{
  kvz_glue_session_close();          /* "Only one encoder may be open at a time" (src/kvazaar.h:619) */
  return kvazaar_open(cfg);
}

const kvz_api *kvz_api_get(int bit_depth)
{
  static kvz_api api;
  static int made = 0;
  if (!made) {
    api = *kvz_api_get_host(bit_depth);
    api.encoder_open = glue_encoder_open;
    api.encoder_close = glue_encoder_close;
    made = 1;
  }
  return &api;
}
