// api_mc.cu — kernel-level C ABI for motion compensation (b265_mc_batch, include/b265.h part B).
#include <vector>
#include "backend.h"
#include "kernels.h"

extern "C" int b265_mc_batch(b265_ctx* ctx, int mem, const b265_pu* pus, int n_pus, const b265_slice* slices,
                             int n_slices, const uint8_t* const* ref_planes, int n_refs,
                             uint8_t* const* dst_planes, int width, int height, int pitch, int pitch_c) {
  if (!ctx || !pus || !slices || !ref_planes || !dst_planes || n_pus < 0 || n_refs < 1 || n_refs > 16 ||
      n_slices < 1 || width <= 0 || height <= 0)
    return B265_ERR_INVALID_ARG;
  if (n_pus == 0) return B265_OK;
  std::vector<b265_pu> pieces;   // <= 16 x 8 pieces (pu_expand)
  pieces.reserve((size_t)n_pus * 2);
  for (int i = 0; i < n_pus; i++) {
    const size_t at = pieces.size();
    pieces.resize(at + pu_pieces(pus[i]));
    pu_expand(pus[i], pieces.data() + at);
  }
  pus = pieces.data();
  n_pus = (int)pieces.size();
  CK(cudaSetDevice(ctx->device));
  const bool chroma = dst_planes[1] != nullptr;
  const int cw = (width + 1) / 2, ch = (height + 1) / 2;
  const size_t ysz = (size_t)pitch * height, csz = chroma ? (size_t)pitch_c * ch : 0;
  const size_t pic = (ysz + 2 * csz + 255) & ~(size_t)255;

  // device layout: [FrameDev][items][pus][slices][pictures: dst, ref0..]
  size_t off_items = 256 * ((sizeof(FrameDev) + 255) / 256);
  size_t off_pus = off_items + (((size_t)n_pus * 4 + 255) & ~(size_t)255);
  size_t off_sl = off_pus + (((size_t)n_pus * sizeof(b265_pu) + 255) & ~(size_t)255);
  size_t off_pic = off_sl + (((size_t)n_slices * sizeof(b265_slice) + 255) & ~(size_t)255);
  size_t total = off_pic + (mem == B265_MEM_HOST ? pic * (1 + n_refs) : 0);
  int r = devbuf_reserve(ctx, &ctx->d_stage, total, false);
  if (r) return r;
  uint8_t* base = (uint8_t*)ctx->d_stage.p;

  FrameDev f = {};
  f.width = width; f.height = height;
  f.cw = chroma ? cw : 0; f.ch = chroma ? ch : 0;
  f.pitch[0] = pitch; f.pitch[1] = f.pitch[2] = pitch_c;
  f.pus = (const b265_pu*)(base + off_pus);
  f.slices = (const b265_slice*)(base + off_sl);
  f.n_refs = n_refs;
  if (mem == B265_MEM_HOST) {
    for (int p = 0; p <= n_refs; p++) {
      uint8_t* d = base + off_pic + pic * p;
      uint8_t* pl[3] = {d, chroma ? d + ysz : nullptr, chroma ? d + ysz + csz : nullptr};
      for (int c = 0; c < 3; c++) {
        if (!pl[c]) continue;
        const uint8_t* src = (p == 0) ? dst_planes[c] : ref_planes[3 * (p - 1) + c];
        CK(cudaMemcpyAsync(pl[c], src, c ? csz : ysz, cudaMemcpyHostToDevice, ctx->stream));
        if (p == 0) f.plane[c] = pl[c]; else f.ref_plane[p - 1][c] = pl[c];
      }
    }
  } else {
    for (int c = 0; c < 3; c++) {
      f.plane[c] = dst_planes[c];
      for (int p = 0; p < n_refs; p++) f.ref_plane[p][c] = ref_planes[3 * p + c];
    }
  }
  std::vector<uint32_t> items(n_pus);
  for (int i = 0; i < n_pus; i++) items[i] = (uint32_t)i;
  CK(cudaMemcpyAsync(base, &f, sizeof(f), cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(base + off_items, items.data(), (size_t)n_pus * 4, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(base + off_pus, pus, (size_t)n_pus * sizeof(b265_pu), cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(base + off_sl, slices, (size_t)n_slices * sizeof(b265_slice), cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));  // items/f are stack/heap temporaries
  launch_mc(ctx->stream, (const FrameDev*)base, (const uint32_t*)(base + off_items), n_pus, ctx->sm_count);
  ctx->launches++;
  CK_LAUNCH("k_mc");
  if (mem == B265_MEM_HOST) {
    for (int c = 0; c < 3; c++)
      if (f.plane[c]) CK(cudaMemcpyAsync(dst_planes[c], f.plane[c], c ? csz : ysz, cudaMemcpyDeviceToHost, ctx->stream));
  }
  CK(cudaStreamSynchronize(ctx->stream));
  return B265_OK;
}
