// postfilter.cu — host-side sequencing of the in-loop filters of a batch (S5: decctx.cc:757-766,
// run_postprocessing_filters_sequential decctx.cc:1859): deblock V -> deblock H -> SAO (boundary strengths are derived inside the edge kernels).
#include <algorithm>
#include "backend.h"
#include "kernels.h"

int launch_postfilters(b265_ctx* ctx, const FrameDev* d_frames, const FrameDev* h, int n, cudaEvent_t ev,
                       const uint32_t* mc_items, int n_mc) {
  bool any_dbk = false, any_sao = false, chroma = false;
  int mw = 0, mh = 0;
  for (int j = 0; j < n; j++) {
    if ((h[j].flags & B265_FRAME_DEBLOCK) && h[j].unit_info) {
      if (!h[j].qp_y) return B265_ERR_INVALID_ARG;
      any_dbk = true;
    }
    if ((h[j].flags & B265_FRAME_SAO) && h[j].sao) any_sao = true;
    chroma |= h[j].cw > 0;
    mw = std::max(mw, h[j].width);
    mh = std::max(mh, h[j].height);
  }
  if (any_dbk) {
    launch_deblock(ctx->stream, d_frames, n, mw, mh, mc_items, n_mc, chroma);
    ctx->launches += 2 + (chroma ? 2 : 0) + (n_mc > 0 ? 1 : 0);
    CK_LAUNCH("deblock");
  }
  if (ev) cudaEventRecord(ev, ctx->stream);
  if (any_sao) {
    launch_sao(ctx->stream, d_frames, n, mw, mh);
    ctx->launches++;
    CK_LAUNCH("k_sao");
  }
  return B265_OK;
}
