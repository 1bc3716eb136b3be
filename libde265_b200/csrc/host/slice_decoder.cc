// slice_decoder.cc — the decoder's instantiation of the slice-data syntax walk (slice_syntax.h) with the
// CABAC decoding engine, one picture at a time.
#include <stdio.h>
#include "slice_syntax.h"

#ifdef B265_PROF
unsigned long long g_prof[16];   // developer aid (-DB265_PROF): TSC ticks per parser section, single-threaded runs only
extern "C" __attribute__((visibility("default"))) void b265_prof_get(unsigned long long* out) {
  for (int i = 0; i < 16; i++) out[i] = g_prof[i];
}
#endif

namespace b265h {


int decode_picture_slices(Picture& pic) {
  PROF(5);
  PicDecodeState st;
  const Sps& sps = *pic.sps;
  st.wpp.assign((size_t)sps.ctbs_h * CTX_COUNT, 0);
  st.wpp_valid.assign(sps.ctbs_h, 0);
  int err = 0;
  for (const SliceUnit& u : pic.slice_units) {
    int r = decode_slice_unit<Cabac>(pic, u, st, Cabac::Setup());
    if (r && !err) err = r;
    if (r == E_NOT_IMPLEMENTED) break;
  }
  // what later pictures need for temporal motion vector prediction
  pic.self->slice_refs = pic.slice_refs;
  if (err) pic.self->failed.store(1, std::memory_order_relaxed);
  pic.self->parsed.store(1, std::memory_order_release);
  return err;
}

}  // namespace b265h
