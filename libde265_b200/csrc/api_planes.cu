// api_planes.cu — kernel-level C ABI on raw planes (include/b265.h part B): sparse residual add,
// whole-plane deblocking and SAO with explicit metadata.  These run the SAME device code as the
// frame path (k_residual.cu, k_deblock.cu, k_sao.cu) on caller-supplied host or device planes; they
// exist so each stage can be parity-tested and timed in isolation (SURVEY §8c function-level oracle).
#include <string.h>
#include <algorithm>
#include <vector>
#include "backend.h"
#include "kernels.h"

namespace {

inline size_t al256(size_t v) { return (v + 255) & ~(size_t)255; }

// Uploads the metadata of `J` (lists + maps) behind a FrameDev into ctx->d_stage and reserves
// `extra` more bytes; returns device base and offset of the extra area.  Synchronous (small data).
struct Staged {
  uint8_t* base = nullptr;
  size_t extra_off = 0;
  FrameDev f = {};
  std::vector<uint2> sao_rec;   // host copy of the SAO control records, alive until push_framedev has synchronised
};

int stage_meta(b265_ctx* ctx, const b265_frame_job& J, size_t extra, bool need_maps, Staged& S) {
  const int cw_ = (J.width + (1 << J.log2_ctb) - 1) >> J.log2_ctb, ch_ = (J.height + (1 << J.log2_ctb) - 1) >> J.log2_ctb;
  const size_t nctb = (size_t)cw_ * ch_;
  const size_t n_units = (size_t)(J.width >> 2) * (J.height >> 2);
  const size_t n_qp = (size_t)(J.width >> J.log2_min_cb) * (J.height >> J.log2_min_cb);
  size_t off = al256(sizeof(FrameDev));
  const size_t o_tus = off; off += al256(sizeof(b265_tu) * (size_t)J.n_tus);
  const size_t o_co = off; off += al256(sizeof(b265_coeff) * (size_t)J.n_coeffs);
  const size_t o_sl = off; off += al256(sizeof(b265_slice) * (size_t)(J.n_slices > 0 ? J.n_slices : 1));
  const size_t o_ctb = off; off += J.ctbs ? al256(sizeof(b265_ctb) * nctb) : 0;
  const size_t o_sao = off; off += J.sao ? al256(sizeof(b265_sao) * nctb) : 0;
  const size_t o_srec = off; off += J.sao ? al256(sizeof(uint2) * 3 * nctb) : 0;
  const size_t o_ui = off; off += J.unit_info ? al256(n_units) : 0;
  const size_t o_qp = off; off += J.qp_y ? al256(n_qp) : 0;
  const size_t o_scal = off; off += J.scaling ? al256(sizeof(b265_scaling)) : 0;
  S.extra_off = off;
  off += extra;
  int r = devbuf_reserve(ctx, &ctx->d_stage, off, false);
  if (r) return r;
  uint8_t* D = (uint8_t*)ctx->d_stage.p;
  S.base = D;
  cudaStream_t st = ctx->stream;
  auto up = [&](size_t o, const void* src, size_t bytes) -> cudaError_t {
    if (!src || !bytes) return cudaSuccess;
    return cudaMemcpyAsync(D + o, src, bytes, cudaMemcpyHostToDevice, st);
  };
  static const b265_slice default_slice = {0, 2, 0, 0, 0, 1, 1, 1, 0, 0, 0, 0, 0, {}, {}};
  CK(up(o_tus, J.tus, sizeof(b265_tu) * (size_t)J.n_tus));
  CK(up(o_co, J.coeffs, sizeof(b265_coeff) * (size_t)J.n_coeffs));
  if (J.n_slices > 0) CK(up(o_sl, J.slices, sizeof(b265_slice) * (size_t)J.n_slices));
  else CK(up(o_sl, &default_slice, sizeof(b265_slice)));
  CK(up(o_ctb, J.ctbs, sizeof(b265_ctb) * nctb));
  CK(up(o_sao, J.sao, sizeof(b265_sao) * nctb));
  if (J.sao) {
    S.sao_rec.resize(3 * nctb);
    sao_records(J, cw_, ch_, true, S.sao_rec.data());
    CK(up(o_srec, S.sao_rec.data(), sizeof(uint2) * 3 * nctb));
  }
  CK(up(o_ui, J.unit_info, n_units));
  CK(up(o_qp, J.qp_y, n_qp));
  CK(up(o_scal, J.scaling, sizeof(b265_scaling)));
  FrameDev& f = S.f;
  f.width = J.width; f.height = J.height;
  f.cw = J.chroma_format_idc ? (J.width + 1) >> 1 : 0;
  f.ch = J.chroma_format_idc ? (J.height + 1) >> 1 : 0;
  f.log2_ctb = J.log2_ctb; f.log2_min_cb = J.log2_min_cb; f.log2_min_tb = J.log2_min_tb;
  f.ctbs_w = cw_; f.ctbs_h = ch_;
  f.flags = J.flags;
  f.cb_qp_offset = J.cb_qp_offset; f.cr_qp_offset = J.cr_qp_offset;
  f.tus = (const b265_tu*)(D + o_tus);
  f.coeffs = (const b265_coeff*)(D + o_co);
  f.slices = (const b265_slice*)(D + o_sl);
  f.ctbs = J.ctbs ? (const b265_ctb*)(D + o_ctb) : nullptr;
  f.sao = J.sao ? (const b265_sao*)(D + o_sao) : nullptr;
  f.sao_rec = J.sao ? (const uint2*)(D + o_srec) : nullptr;
  f.unit_info = J.unit_info ? D + o_ui : nullptr;
  f.qp_y = J.qp_y ? (const int8_t*)(D + o_qp) : nullptr;
  f.scaling = J.scaling ? (const b265_scaling*)(D + o_scal) : nullptr;
  f.bs = nullptr;
  if (need_maps && (!f.unit_info || !f.qp_y)) return B265_ERR_INVALID_ARG;
  return B265_OK;
}

int push_framedev(b265_ctx* ctx, Staged& S) {
  CK(cudaMemcpyAsync(S.base, &S.f, sizeof(FrameDev), cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));  // S.f lives on the caller's stack
  return B265_OK;
}

}  // namespace

extern "C" {

int b265_residual_sparse(b265_ctx* ctx, int mem, const b265_tu* tus, int n_tus, const b265_coeff* coeffs, int n_coeffs,
                         uint8_t* dst, int pitch, int width, int height) {
  if (!ctx || !tus || n_tus < 0 || (!coeffs && n_coeffs > 0) || !dst || pitch < width || width <= 0 || height <= 0)
    return B265_ERR_INVALID_ARG;
  if (n_tus == 0) return B265_OK;
  CK(cudaSetDevice(ctx->device));
  b265_frame_job J;
  memset(&J, 0, sizeof(J));
  J.width = width; J.height = height; J.chroma_format_idc = 0; J.bit_depth = 8;
  J.log2_ctb = 6; J.log2_min_cb = 3; J.log2_min_tb = 2;
  J.tus = tus; J.n_tus = n_tus; J.coeffs = coeffs; J.n_coeffs = n_coeffs;
  std::vector<uint32_t> items[4];
  for (int i = 0; i < n_tus; i++) {
    const int lg = (tus[i].cidx_log2 >> 2) & 7;
    if (lg < 2 || lg > 5 || tus[i].x + (1 << lg) > width || tus[i].y + (1 << lg) > height) return B265_ERR_INVALID_ARG;
    if (tus[i].ncoeff) items[lg - 2].push_back((uint32_t)i);
  }
  const size_t plane_bytes = al256((size_t)pitch * height);
  const size_t items_bytes = al256(4 * (size_t)n_tus) * 4;
  Staged S;
  int r = stage_meta(ctx, J, items_bytes + (mem == B265_MEM_HOST ? plane_bytes : 0), false, S);
  if (r) return r;
  uint8_t* d_items = S.base + S.extra_off;
  uint8_t* d_plane = mem == B265_MEM_HOST ? d_items + items_bytes : dst;
  if (mem == B265_MEM_HOST) CK(cudaMemcpyAsync(d_plane, dst, (size_t)pitch * height, cudaMemcpyHostToDevice, ctx->stream));
  for (int c = 0; c < 3; c++) { S.f.plane[c] = d_plane; S.f.pitch[c] = pitch; }  // every TU addresses `dst`
  r = push_framedev(ctx, S);
  if (r) return r;
  // the packed kernels add to the prediction with 128-bit accesses: 16-byte aligned plane, pitch and TU column
  int n_plain[4] = {0, 0, 0, 0};
  if (pitch % 16 == 0 && (reinterpret_cast<uintptr_t>(d_plane) & 15) == 0)
    for (int s = 2; s < 4; s++) {
      auto plain = [&](uint32_t i) { return !(tus[i].flags & (B265_TU_SKIP | B265_TU_BYPASS)) && (tus[i].x & 15) == 0; };
      n_plain[s] = (int)(std::stable_partition(items[s].begin(), items[s].end(), plain) - items[s].begin());
    }
  for (int s = 0; s < 4; s++) {
    if (items[s].empty()) continue;
    uint32_t* d = (uint32_t*)(d_items + (size_t)s * al256(4 * (size_t)n_tus));
    CK(cudaMemcpyAsync(d, items[s].data(), 4 * items[s].size(), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    const int n = (int)items[s].size(), np = n_plain[s];
    if (np) {   // plain DCT TUs of 16x16 / 32x32: packed dp2a kernels (as in the frame path)
      launch_residual_sparse_fast(ctx->stream, 2 + s, (const FrameDev*)S.base, d, np, ctx->sm_count, false);
      ctx->launches++;
      CK_LAUNCH("k_residual_sparse_fast");
    }
    if (n > np) {
      launch_residual_sparse(ctx->stream, 2 + s, (const FrameDev*)S.base, d + np, n - np, ctx->sm_count);
      ctx->launches++;
      CK_LAUNCH("k_residual_sparse");
    }
  }
  if (mem == B265_MEM_HOST) CK(cudaMemcpyAsync(dst, d_plane, (size_t)pitch * height, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return B265_OK;
}

static int plane_geometry(const b265_frame_job& J, int pitch, int pitch_c, size_t sz[3], int np) {
  const int ch = (J.height + 1) >> 1;
  sz[0] = al256((size_t)pitch * J.height);
  sz[1] = sz[2] = np == 3 ? al256((size_t)pitch_c * ch) : 0;
  return 0;
}

int b265_deblock_planes(b265_ctx* ctx, int mem, const b265_frame_job* meta, uint8_t* const* planes, int pitch, int pitch_c) {
  if (!ctx || !meta || !planes || !planes[0] || meta->bit_depth != 8 || meta->chroma_format_idc > 1) return B265_ERR_INVALID_ARG;
  CK(cudaSetDevice(ctx->device));
  const int np = (meta->chroma_format_idc && planes[1] && planes[2]) ? 3 : 1;
  size_t sz[3];
  plane_geometry(*meta, pitch, pitch_c, sz, np);
  Staged S;
  b265_frame_job J = *meta;
  J.n_tus = 0; J.n_coeffs = 0; J.tus = nullptr; J.coeffs = nullptr;
  J.flags |= B265_FRAME_DEBLOCK;
  int r = stage_meta(ctx, J, mem == B265_MEM_HOST ? sz[0] + sz[1] + sz[2] : 0, true, S);
  if (r) return r;
  uint8_t* d = S.base + S.extra_off;
  for (int c = 0; c < 3; c++) {
    S.f.pitch[c] = c ? pitch_c : pitch;
    if (c >= np) { S.f.plane[c] = nullptr; continue; }
    if (mem == B265_MEM_HOST) {
      S.f.plane[c] = d;
      CK(cudaMemcpyAsync(d, planes[c], (size_t)S.f.pitch[c] * (c ? (J.height + 1) >> 1 : J.height), cudaMemcpyHostToDevice, ctx->stream));
      d += sz[c];
    } else {
      S.f.plane[c] = planes[c];
    }
    S.f.out[c] = S.f.plane[c];
  }
  if (np == 1) { S.f.cw = S.f.ch = 0; }
  r = push_framedev(ctx, S);
  if (r) return r;
  launch_deblock(ctx->stream, (const FrameDev*)S.base, 1, J.width, J.height, nullptr, 0, np == 3);
  ctx->launches += 2 + (np == 3 ? 2 : 0);
  CK_LAUNCH("deblock");
  if (mem == B265_MEM_HOST)
    for (int c = 0; c < np; c++)
      CK(cudaMemcpyAsync(planes[c], S.f.plane[c], (size_t)S.f.pitch[c] * (c ? (J.height + 1) >> 1 : J.height), cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return B265_OK;
}

int b265_sao_planes(b265_ctx* ctx, int mem, const b265_frame_job* meta, const uint8_t* const* in_planes,
                    uint8_t* const* out_planes, int pitch, int pitch_c) {
  if (!ctx || !meta || !in_planes || !out_planes || !in_planes[0] || !out_planes[0] || !meta->sao || meta->bit_depth != 8 ||
      meta->chroma_format_idc > 1)
    return B265_ERR_INVALID_ARG;
  CK(cudaSetDevice(ctx->device));
  const int np = (meta->chroma_format_idc && in_planes[1] && in_planes[2] && out_planes[1] && out_planes[2]) ? 3 : 1;
  size_t sz[3];
  plane_geometry(*meta, pitch, pitch_c, sz, np);
  Staged S;
  b265_frame_job J = *meta;
  J.n_tus = 0; J.n_coeffs = 0; J.tus = nullptr; J.coeffs = nullptr;
  J.flags |= B265_FRAME_SAO;
  int r = stage_meta(ctx, J, mem == B265_MEM_HOST ? 2 * (sz[0] + sz[1] + sz[2]) : 0, false, S);
  if (r) return r;
  uint8_t* d = S.base + S.extra_off;
  for (int c = 0; c < 3; c++) {
    S.f.pitch[c] = c ? pitch_c : pitch;
    if (c >= np) { S.f.plane[c] = S.f.out[c] = nullptr; continue; }
    const size_t bytes = (size_t)S.f.pitch[c] * (c ? (J.height + 1) >> 1 : J.height);
    if (mem == B265_MEM_HOST) {
      S.f.plane[c] = d;
      S.f.out[c] = d + sz[c];
      CK(cudaMemcpyAsync(d, in_planes[c], bytes, cudaMemcpyHostToDevice, ctx->stream));
      d += 2 * sz[c];
    } else {
      S.f.plane[c] = const_cast<uint8_t*>(in_planes[c]);
      S.f.out[c] = out_planes[c];
    }
    (void)bytes;  // k_sao writes every sample of the output plane (unchanged ones are copied through)
  }
  if (np == 1) { S.f.cw = S.f.ch = 0; }
  r = push_framedev(ctx, S);
  if (r) return r;
  launch_sao(ctx->stream, (const FrameDev*)S.base, 1, J.width, J.height);
  ctx->launches++;
  CK_LAUNCH("k_sao");
  if (mem == B265_MEM_HOST)
    for (int c = 0; c < np; c++)
      CK(cudaMemcpyAsync(out_planes[c], S.f.out[c], (size_t)S.f.pitch[c] * (c ? (J.height + 1) >> 1 : J.height),
                         cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return B265_OK;
}

int b265_intra_batch(b265_ctx* ctx, int mem, int log2, int cidx, int n_blocks, const uint8_t* modes, const uint8_t* borders,
                     int strong_filter, int no_edge_filter, uint8_t* out) {
  if (!ctx || log2 < 2 || log2 > 5 || cidx < 0 || cidx > 2 || n_blocks < 0 || !modes || !borders || !out)
    return B265_ERR_INVALID_ARG;
  if (n_blocks == 0) return B265_OK;
  if (mem == B265_MEM_HOST)
    for (int i = 0; i < n_blocks; i++)
      if (modes[i] > 34) return B265_ERR_INVALID_ARG;
  CK(cudaSetDevice(ctx->device));
  const size_t nb = (size_t)n_blocks, bsz = (4u << log2) + 1, px = nb << (2 * log2);
  if (mem == B265_MEM_DEVICE) {
    launch_intra_blocks(ctx->stream, log2, cidx, n_blocks, modes, borders, strong_filter, no_edge_filter, out);
    ctx->launches++;
    CK_LAUNCH("k_intra_blocks");
    return B265_OK;
  }
  const size_t o_b = al256(nb), o_out = o_b + al256(nb * bsz);
  int r = devbuf_reserve(ctx, &ctx->d_stage, o_out + al256(px), false);
  if (r) return r;
  uint8_t* D = (uint8_t*)ctx->d_stage.p;
  CK(cudaMemcpyAsync(D, modes, nb, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(D + o_b, borders, nb * bsz, cudaMemcpyHostToDevice, ctx->stream));
  launch_intra_blocks(ctx->stream, log2, cidx, n_blocks, D, D + o_b, strong_filter, no_edge_filter, D + o_out);
  ctx->launches++;
  CK_LAUNCH("k_intra_blocks");
  CK(cudaMemcpyAsync(out, D + o_out, px, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return B265_OK;
}

int b265_interp_batch(b265_ctx* ctx, int mem, int luma, int n_blocks, int w, int h, const uint8_t* src, ptrdiff_t src_stride,
                      ptrdiff_t src_block_stride, const uint8_t* frac_xy, int16_t* out) {
  if (!ctx || n_blocks < 0 || w <= 0 || h <= 0 || w > 64 || h > 64 || !src || !frac_xy || !out || src_stride < w ||
      src_block_stride < 0)
    return B265_ERR_INVALID_ARG;
  if (n_blocks == 0) return B265_OK;
  CK(cudaSetDevice(ctx->device));
  const size_t nb = (size_t)n_blocks, px = nb * w * h;
  if (mem == B265_MEM_DEVICE) {
    launch_interp_blocks(ctx->stream, luma, n_blocks, w, h, src, src_stride, src_block_stride, frac_xy, out);
    ctx->launches++;
    CK_LAUNCH("k_interp_blocks");
    return B265_OK;
  }
  const int before = luma ? 3 : 1, after = luma ? 4 : 2;  // halo of the 8-tap / 4-tap filters
  for (size_t i = 0; i < 2 * nb; i++)
    if (frac_xy[i] > (luma ? 3 : 7)) return B265_ERR_INVALID_ARG;
  // host source window: from the first block's top-left halo sample to the last block's bottom-right one
  const size_t lead = (size_t)before * src_stride + before;
  const size_t span = lead + (nb - 1) * src_block_stride + (size_t)(h - 1 + after) * src_stride + w + after;
  const size_t o_f = al256(span), o_out = o_f + al256(2 * nb);
  int r = devbuf_reserve(ctx, &ctx->d_stage, o_out + al256(px * 2), false);
  if (r) return r;
  uint8_t* D = (uint8_t*)ctx->d_stage.p;
  CK(cudaMemcpyAsync(D, src - lead, span, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(D + o_f, frac_xy, 2 * nb, cudaMemcpyHostToDevice, ctx->stream));
  launch_interp_blocks(ctx->stream, luma, n_blocks, w, h, D + lead, src_stride, src_block_stride, D + o_f,
                       (int16_t*)(D + o_out));
  ctx->launches++;
  CK_LAUNCH("k_interp_blocks");
  CK(cudaMemcpyAsync(out, D + o_out, px * 2, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return B265_OK;
}

}  // extern "C"
