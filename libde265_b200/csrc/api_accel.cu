// api_accel.cu — the reference's `acceleration_functions` slot table (libde265/acceleration.h:29-201) filled with
// CUDA entry points: init_acceleration_functions_cuda(), the counterpart of init_acceleration_functions_sse()
// (x86/sse.cc:37-103).  Every slot takes HOST pointers and one block, like the CPU slot it replaces: the block is
// packed into a dense staging buffer, handed to the batched kernel-level entry points of b265.h (the device code
// the frame path uses) with n_blocks = 1, and unpacked.  See include/b265_acceleration.h for what this is for.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <mutex>
#include <vector>
#include "../../include/b265_acceleration.h"
#include "backend.h"
#include "kernels.h"

// ---------------------------------------------------------------------------------------------
// the four prediction writers on dense blocks: out = clip8(((a*w0 + b*w1 + rnd) >> sh) + post)
// (put_unweighted_pred_8 / put_weighted_pred_avg_8 / put_weighted_pred_8 / put_weighted_bipred_8,
//  fallback-motion.cc:33-158; the same single formula k_mc.cu's writer uses)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_pred_write(long long n, const int16_t* __restrict__ a, const int16_t* __restrict__ b,
                                                    int w0, int w1, int rnd, int sh, int post, uint8_t* __restrict__ dst) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int va = a[i], vb = b ? b[i] : 0;
    dst[i] = (uint8_t)clip8(((va * w0 + vb * w1 + rnd) >> sh) + post);
  }
}

extern "C" int b265_pred_write_batch(b265_ctx* ctx, int mem, int mode, int n_blocks, int w, int h, const int16_t* src1,
                                     const int16_t* src2, int w1, int o1, int w2, int o2, int log2WD, uint8_t* dst) {
  if (!ctx || mode < 0 || mode > 3 || n_blocks < 0 || w <= 0 || h <= 0 || !src1 || !dst) return B265_ERR_INVALID_ARG;
  const bool two = mode == 1 || mode == 3;
  if (two && !src2) return B265_ERR_INVALID_ARG;
  if (mode >= 2 && (log2WD < 1 || log2WD > 30)) return B265_ERR_INVALID_ARG;   // the slots assert log2WD >= 1 (fallback-motion.cc:58,80)
  if (n_blocks == 0) return B265_OK;
  int k0, k1, rnd, sh, post;
  switch (mode) {
    case 0: k0 = 1; k1 = 0; rnd = 32; sh = 6; post = 0; break;
    case 1: k0 = 1; k1 = 1; rnd = 64; sh = 7; post = 0; break;
    case 2: k0 = w1; k1 = 0; rnd = 1 << (log2WD - 1); sh = log2WD; post = o1; break;
    default: k0 = w1; k1 = w2; rnd = (o1 + o2 + 1) << log2WD; sh = log2WD + 1; post = 0; break;
  }
  CK(cudaSetDevice(ctx->device));
  const long long n = (long long)n_blocks * w * h;
  long long grid = (n + 255) / 256;
  if (grid > 4LL * ctx->sm_count * 8) grid = 4LL * ctx->sm_count * 8;
  if (mem == B265_MEM_DEVICE) {
    k_pred_write<<<(int)grid, 256, 0, ctx->stream>>>(n, src1, two ? src2 : nullptr, k0, k1, rnd, sh, post, dst);
    ctx->launches++;
    CK_LAUNCH("k_pred_write");
    return B265_OK;
  }
  const size_t bytes16 = ((size_t)n * 2 + 255) & ~(size_t)255;
  int r = devbuf_reserve(ctx, &ctx->d_stage, 2 * bytes16 + (size_t)n + 256, false);
  if (r) return r;
  uint8_t* D = (uint8_t*)ctx->d_stage.p;
  CK(cudaMemcpyAsync(D, src1, (size_t)n * 2, cudaMemcpyHostToDevice, ctx->stream));
  if (two) CK(cudaMemcpyAsync(D + bytes16, src2, (size_t)n * 2, cudaMemcpyHostToDevice, ctx->stream));
  k_pred_write<<<(int)grid, 256, 0, ctx->stream>>>(n, (const int16_t*)D, two ? (const int16_t*)(D + bytes16) : nullptr, k0, k1, rnd,
                                                   sh, post, D + 2 * bytes16);
  ctx->launches++;
  CK_LAUNCH("k_pred_write");
  CK(cudaMemcpyAsync(dst, D + 2 * bytes16, (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return B265_OK;
}

// ---------------------------------------------------------------------------------------------
// the slots
// ---------------------------------------------------------------------------------------------
namespace {

std::mutex g_mu;              // the slots may be called from several decoder threads: one block at a time
b265_ctx* g_ctx = nullptr;
int g_device = -1;
std::vector<int16_t> g_a, g_b;
std::vector<uint8_t> g_px, g_tile;

b265_ctx* slot_ctx() {   // g_mu held
  if (!g_ctx) {
    int dev = g_device;
    if (dev < 0) {
      const char* e = getenv("B265_DEVICE");
      dev = e ? atoi(e) : 0;
    }
    if (b265_create(dev, &g_ctx) != B265_OK) g_ctx = nullptr;
  }
  return g_ctx;
}

// A slot has no way to report failure (void functions).  A CUDA error after a successful initialisation is fatal for
// the decode: say so and stop rather than return wrong samples.
void slot_check(int status, const char* what) {
  if (status == B265_OK) return;
  fprintf(stderr, "b265: acceleration slot %s failed: %s (%s)\n", what, b265_status_string(status), g_ctx ? b265_last_error(g_ctx) : "no context");
  abort();
}

void pack16(std::vector<int16_t>& v, const int16_t* src, ptrdiff_t stride, int w, int h) {
  v.resize((size_t)w * h);
  for (int y = 0; y < h; y++) memcpy(&v[(size_t)y * w], src + y * stride, (size_t)w * 2);
}

void writer(int mode, uint8_t* dst, ptrdiff_t dststride, const int16_t* s1, const int16_t* s2, ptrdiff_t srcstride, int w, int h,
            int w1, int o1, int w2, int o2, int log2WD) {
  std::lock_guard<std::mutex> l(g_mu);
  pack16(g_a, s1, srcstride, w, h);
  if (s2) pack16(g_b, s2, srcstride, w, h);
  g_px.resize((size_t)w * h);
  slot_check(b265_pred_write_batch(slot_ctx(), B265_MEM_HOST, mode, 1, w, h, g_a.data(), s2 ? g_b.data() : nullptr, w1, o1, w2, o2,
                                   log2WD, g_px.data()), "put_*_pred_8");
  for (int y = 0; y < h; y++) memcpy(dst + y * dststride, &g_px[(size_t)y * w], (size_t)w);
}

void s_put_unweighted_pred_8(uint8_t* dst, ptrdiff_t ds, const int16_t* src, ptrdiff_t ss, int w, int h) {
  writer(0, dst, ds, src, nullptr, ss, w, h, 0, 0, 0, 0, 0);
}
void s_put_weighted_pred_avg_8(uint8_t* dst, ptrdiff_t ds, const int16_t* s1, const int16_t* s2, ptrdiff_t ss, int w, int h) {
  writer(1, dst, ds, s1, s2, ss, w, h, 0, 0, 0, 0, 0);
}
void s_put_weighted_pred_8(uint8_t* dst, ptrdiff_t ds, const int16_t* src, ptrdiff_t ss, int w, int h, int wt, int o, int log2WD) {
  writer(2, dst, ds, src, nullptr, ss, w, h, wt, o, 0, 0, log2WD);
}
void s_put_weighted_bipred_8(uint8_t* dst, ptrdiff_t ds, const int16_t* s1, const int16_t* s2, ptrdiff_t ss, int w, int h, int w1,
                             int o1, int w2, int o2, int log2WD) {
  writer(3, dst, ds, s1, s2, ss, w, h, w1, o1, w2, o2, log2WD);
}

// Interpolation slots: the caller guarantees the halo only on the sides the fraction needs (3/4 samples for luma, 1/2 for
// chroma); the block is copied into a tile with the full halo (zero where the contract gives nothing - those taps are zero).
void interp(bool luma, int16_t* out, ptrdiff_t outstride, const uint8_t* src, ptrdiff_t srcstride, int w, int h, int fx, int fy) {
  const int before = luma ? 3 : 1, after = luma ? 4 : 2;
  if (w > 64 || h > 64) {   // the slots are called with blocks up to 64x64; larger requests are cut
    for (int y = 0; y < h; y += 64)
      for (int x = 0; x < w; x += 64)
        interp(luma, out + y * outstride + x, outstride, src + y * srcstride + x, srcstride, w - x < 64 ? w - x : 64,
               h - y < 64 ? h - y : 64, fx, fy);
    return;
  }
  std::lock_guard<std::mutex> l(g_mu);
  const int tw = w + before + after, th = h + before + after;
  g_tile.assign((size_t)tw * th, 0);
  const int x0 = fx ? -before : 0, x1 = fx ? w + after : w, y0 = fy ? -before : 0, y1 = fy ? h + after : h;
  for (int y = y0; y < y1; y++) memcpy(&g_tile[(size_t)(y + before) * tw + before + x0], src + y * srcstride + x0, (size_t)(x1 - x0));
  const uint8_t frac[2] = {(uint8_t)fx, (uint8_t)fy};
  g_a.resize((size_t)w * h);
  slot_check(b265_interp_batch(slot_ctx(), B265_MEM_HOST, luma ? 1 : 0, 1, w, h, &g_tile[(size_t)before * tw + before], tw, 0, frac,
                               g_a.data()), "put_hevc_qpel/epel_8");
  for (int y = 0; y < h; y++) memcpy(out + y * outstride, &g_a[(size_t)y * w], (size_t)w * 2);
}

template <int X, int Y>
void s_qpel(int16_t* dst, ptrdiff_t ds, const uint8_t* src, ptrdiff_t ss, int w, int h, int16_t*) { interp(true, dst, ds, src, ss, w, h, X, Y); }
void s_epel(int16_t* dst, ptrdiff_t ds, const uint8_t* src, ptrdiff_t ss, int w, int h, int, int, int16_t*) {
  interp(false, dst, ds, src, ss, w, h, 0, 0);   // the integer-position slot ignores mx / my (fallback-motion.cc:257-274)
}
void s_epel_frac(int16_t* dst, ptrdiff_t ds, const uint8_t* src, ptrdiff_t ss, int w, int h, int mx, int my, int16_t*, int) {
  interp(false, dst, ds, src, ss, w, h, mx & 7, my & 7);
}

void residual(int kind, int log2, uint8_t* dst, const int16_t* coeffs, ptrdiff_t stride) {
  std::lock_guard<std::mutex> l(g_mu);
  const int n = 1 << log2;
  g_px.resize((size_t)n * n);
  for (int y = 0; y < n; y++) memcpy(&g_px[(size_t)y * n], dst + y * stride, (size_t)n);
  slot_check(b265_residual_batch(slot_ctx(), B265_MEM_HOST, kind, log2, 0, 1, coeffs, nullptr, g_px.data()), "transform_*_8");
  for (int y = 0; y < n; y++) memcpy(dst + y * stride, &g_px[(size_t)y * n], (size_t)n);
}
void s_transform_skip_8(uint8_t* dst, const int16_t* c, ptrdiff_t stride) { residual(B265_TR_SKIP, 2, dst, c, stride); }
void s_transform_dst_8(uint8_t* dst, const int16_t* c, ptrdiff_t stride) { residual(B265_TR_DST, 2, dst, c, stride); }
template <int LOG2>
void s_transform_add_8(uint8_t* dst, const int16_t* c, ptrdiff_t stride) { residual(B265_TR_DCT, LOG2, dst, c, stride); }

}  // namespace

extern "C" {

void b265_acceleration_set_device(int device) {
  std::lock_guard<std::mutex> l(g_mu);
  g_device = device;
}

uint64_t b265_acceleration_launch_count(void) {
  std::lock_guard<std::mutex> l(g_mu);
  return g_ctx ? b265_launch_count(g_ctx) : 0;
}

int init_acceleration_functions_cuda(struct b265_acceleration_functions* a) {
  if (!a) return B265_ERR_INVALID_ARG;
  {
    std::lock_guard<std::mutex> l(g_mu);
    if (!slot_ctx()) return B265_ERR_NO_DEVICE;
  }
  a->put_unweighted_pred_8 = s_put_unweighted_pred_8;
  a->put_weighted_pred_avg_8 = s_put_weighted_pred_avg_8;
  a->put_weighted_pred_8 = s_put_weighted_pred_8;
  a->put_weighted_bipred_8 = s_put_weighted_bipred_8;
  a->put_hevc_epel_8 = s_epel;
  a->put_hevc_epel_h_8 = s_epel_frac;
  a->put_hevc_epel_v_8 = s_epel_frac;
  a->put_hevc_epel_hv_8 = s_epel_frac;
#define Q(X, Y) a->put_hevc_qpel_8[X][Y] = s_qpel<X, Y>
  Q(0, 0); Q(0, 1); Q(0, 2); Q(0, 3); Q(1, 0); Q(1, 1); Q(1, 2); Q(1, 3);
  Q(2, 0); Q(2, 1); Q(2, 2); Q(2, 3); Q(3, 0); Q(3, 1); Q(3, 2); Q(3, 3);
#undef Q
  a->transform_skip_8 = s_transform_skip_8;
  a->transform_4x4_dst_add_8 = s_transform_dst_8;
  a->transform_add_8[0] = s_transform_add_8<2>;
  a->transform_add_8[1] = s_transform_add_8<3>;
  a->transform_add_8[2] = s_transform_add_8<4>;
  a->transform_add_8[3] = s_transform_add_8<5>;
  return B265_OK;
}

}  // extern "C"
