// backend.cu — C ABI (include/b265.h): context, device memory, surfaces, kernel-level batch calls.
// There is deliberately no CPU fallback anywhere in this file: without a usable CUDA device
// b265_create() fails and nothing else can be called.
#include "backend.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <chrono>
#include <mutex>
#include <vector>
#include "kernels.h"

extern "C" {

const char* b265_version(void) { return "b265 0.1 (sm_100a)"; }

const char* b265_status_string(int s) {
  switch (s) {
    case B265_OK: return "ok";
    case B265_ERR_NO_DEVICE: return "no usable CUDA device";
    case B265_ERR_OUT_OF_MEMORY: return "out of memory";
    case B265_ERR_INVALID_ARG: return "invalid argument";
    case B265_ERR_CUDA: return "CUDA error";
    case B265_ERR_UNSUPPORTED: return "unsupported";
    default: return "unknown";
  }
}
}  // extern "C"

int b265_fail(b265_ctx* ctx, int status, const char* what, cudaError_t e) {
  if (ctx) {
    ctx->last_error = std::string(what) + ": " + cudaGetErrorString(e);
  }
  if (getenv("B265_DEBUG")) fprintf(stderr, "b265: %s: %s\n", what, cudaGetErrorString(e));
  if (e == cudaErrorMemoryAllocation) return B265_ERR_OUT_OF_MEMORY;
  return status;
}

int devbuf_reserve(b265_ctx* ctx, DevBuf* b, size_t bytes, bool host) {
  if (bytes <= b->cap) return B265_OK;
  size_t cap = bytes + bytes / 4 + 4096;
  if (b->p) {
    if (host) cudaFreeHost(b->p); else cudaFree(b->p);
    b->p = nullptr;
    b->cap = 0;
  }
  if (host) CK(cudaMallocHost(&b->p, cap));
  else CK(cudaMalloc(&b->p, cap));
  b->cap = cap;
  return B265_OK;
}

// every live context of the process (b265_busy_peers)
static std::mutex g_ctx_mu;
static std::vector<b265_ctx*> g_ctxs;

int64_t b265_now_ns() {
  return std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// A peer counts when its last batch is still on the device, or was handed over within the last 200 ms: contexts that take
// turns (the decoder's reconstruction threads, a server's streams) keep their shares while one of them is between batches
// (one full-size grid among shared ones stalls every peer for the length of its wavefront).
int b265_busy_peers(b265_ctx* ctx) {
  std::lock_guard<std::mutex> lk(g_ctx_mu);
  const int64_t now = b265_now_ns();
  int n = 0;
  for (b265_ctx* c : g_ctxs) {
    if (c == ctx || c->device != ctx->device || !c->busy_recorded.load(std::memory_order_acquire)) continue;
    if (now - c->last_batch_ns.load(std::memory_order_relaxed) < 200000000 || cudaEventQuery(c->ev_busy) == cudaErrorNotReady) n++;
  }
  cudaGetLastError();   // cudaErrorNotReady is not sticky, but leave no trace for the next CK_LAUNCH
  return n;
}

extern "C" {

const char* b265_last_error(const b265_ctx* ctx) { return ctx ? ctx->last_error.c_str() : ""; }

int b265_create(int device, b265_ctx** out) {
  if (!out) return B265_ERR_INVALID_ARG;
  *out = nullptr;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n) {
    cudaGetLastError();
    return B265_ERR_NO_DEVICE;
  }
  if (cudaSetDevice(device) != cudaSuccess) return B265_ERR_NO_DEVICE;
  b265_ctx* ctx = new b265_ctx();
  ctx->device = device;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { delete ctx; return B265_ERR_NO_DEVICE; }
  ctx->sm_count = prop.multiProcessorCount;
  if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
    delete ctx;
    return B265_ERR_NO_DEVICE;
  }
  cudaEventCreate(&ctx->ev0);
  cudaEventCreate(&ctx->ev1);
  cudaEventCreateWithFlags(&ctx->ev_sync, cudaEventBlockingSync | cudaEventDisableTiming);
  cudaEventCreateWithFlags(&ctx->ev_busy, cudaEventDisableTiming);
  cudaEventCreateWithFlags(&ctx->ev_wait, cudaEventDisableTiming);
  for (auto& e : ctx->stage_ev) cudaEventCreate(&e);
  {
    std::lock_guard<std::mutex> lk(g_ctx_mu);
    g_ctxs.push_back(ctx);
  }
  *out = ctx;
  return B265_OK;
}

void b265_free_batch(b265_ctx* ctx, b265_batch* batch);

void b265_destroy(b265_ctx* ctx) {
  if (!ctx) return;
  {
    std::lock_guard<std::mutex> lk(g_ctx_mu);
    g_ctxs.erase(std::remove(g_ctxs.begin(), g_ctxs.end(), ctx), g_ctxs.end());
  }
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  if (ctx->submit_batch) b265_free_batch(ctx, ctx->submit_batch);
  for (auto& s : ctx->surfaces)
    if (s.used && s.base) cudaFree(s.base);
  if (ctx->flush_buf) cudaFree(ctx->flush_buf);
  if (ctx->d_stage.p) cudaFree(ctx->d_stage.p);
  if (ctx->d_hash.p) cudaFree(ctx->d_hash.p);
  if (ctx->d_submit.p) cudaFree(ctx->d_submit.p);
  if (ctx->h_submit.p) cudaFreeHost(ctx->h_submit.p);
  if (ctx->h_stage.p) cudaFreeHost(ctx->h_stage.p);
  cudaEventDestroy(ctx->ev0);
  cudaEventDestroy(ctx->ev1);
  if (ctx->ev_sync) cudaEventDestroy(ctx->ev_sync);
  if (ctx->ev_busy) cudaEventDestroy(ctx->ev_busy);
  if (ctx->ev_wait) cudaEventDestroy(ctx->ev_wait);
  for (auto& e : ctx->stage_ev) cudaEventDestroy(e);
  cudaStreamDestroy(ctx->stream);
  delete ctx;
}

int b265_sync(b265_ctx* ctx) {
  if (!ctx) return B265_ERR_INVALID_ARG;
  // the host cores belong to the CABAC parse threads: wait on a blocking-sync event (the calling thread sleeps)
  // rather than cudaStreamSynchronize, which spins
  if (ctx->ev_sync) {
    CK(cudaEventRecord(ctx->ev_sync, ctx->stream));
    CK(cudaEventSynchronize(ctx->ev_sync));
  } else {
    CK(cudaStreamSynchronize(ctx->stream));
  }
  return B265_OK;
}

int b265_stream_wait(b265_ctx* ctx, b265_ctx* other) {
  if (!ctx || !other || ctx == other) return B265_ERR_INVALID_ARG;
  if (ctx->device != other->device) return b265_fail(ctx, B265_ERR_INVALID_ARG, "b265_stream_wait: contexts on different devices", cudaSuccess);
  CK(cudaSetDevice(ctx->device));
  CK(cudaEventRecord(ctx->ev_wait, other->stream));   // the waiter's event: calls on one context are serialised by its owner anyway
  CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_wait, 0));
  return B265_OK;
}

uint64_t b265_launch_count(const b265_ctx* ctx) { return ctx ? ctx->launches : 0; }

int b265_host_alloc(b265_ctx* ctx, size_t bytes, void** out) {
  if (!ctx || !out) return B265_ERR_INVALID_ARG;
  CK(cudaSetDevice(ctx->device));
  CK(cudaMallocHost(out, bytes ? bytes : 1));
  return B265_OK;
}
void b265_host_free(b265_ctx* ctx, void* p) {
  (void)ctx;
  if (p) cudaFreeHost(p);
}
int b265_dev_alloc(b265_ctx* ctx, size_t bytes, void** out) {
  if (!ctx || !out) return B265_ERR_INVALID_ARG;
  CK(cudaSetDevice(ctx->device));
  CK(cudaMalloc(out, bytes ? bytes : 1));
  return B265_OK;
}
void b265_dev_free(b265_ctx* ctx, void* p) {
  (void)ctx;
  if (p) cudaFree(p);
}
int b265_memcpy_h2d(b265_ctx* ctx, void* dst, const void* src, size_t bytes) {
  if (!ctx) return B265_ERR_INVALID_ARG;
  CK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return B265_OK;
}
int b265_memcpy_d2h(b265_ctx* ctx, void* dst, const void* src, size_t bytes) {
  if (!ctx) return B265_ERR_INVALID_ARG;
  CK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return B265_OK;
}
int b265_timer_start(b265_ctx* ctx) {
  if (!ctx) return B265_ERR_INVALID_ARG;
  CK(cudaEventRecord(ctx->ev0, ctx->stream));
  return B265_OK;
}
int b265_timer_stop(b265_ctx* ctx, float* ms) {
  if (!ctx || !ms) return B265_ERR_INVALID_ARG;
  CK(cudaEventRecord(ctx->ev1, ctx->stream));
  CK(cudaEventSynchronize(ctx->ev1));
  CK(cudaEventElapsedTime(ms, ctx->ev0, ctx->ev1));
  return B265_OK;
}
int b265_flush_l2(b265_ctx* ctx) {
  if (!ctx) return B265_ERR_INVALID_ARG;
  if (!ctx->flush_buf) {
    ctx->flush_bytes = (size_t)256 << 20;  // 2x the 126 MB L2
    CK(cudaMalloc(&ctx->flush_buf, ctx->flush_bytes));
  }
  CK(cudaMemsetAsync(ctx->flush_buf, 0x5a, ctx->flush_bytes, ctx->stream));
  return B265_OK;
}

// ------------------------------------------------------------------------------------------
// surfaces
// ------------------------------------------------------------------------------------------
static Surface* get_surface(b265_ctx* ctx, int id) {
  if (!ctx || id < 0 || id >= (int)ctx->surfaces.size() || !ctx->surfaces[id].used) return nullptr;
  return &ctx->surfaces[id];
}

int b265_surface_alloc(b265_ctx* ctx, int width, int height, int chroma, int* id_out) {
  if (!ctx || !id_out || width <= 0 || height <= 0 || chroma < 0 || chroma > 1) return B265_ERR_INVALID_ARG;
  CK(cudaSetDevice(ctx->device));
  int id = -1;
  for (size_t i = 0; i < ctx->surfaces.size(); i++)
    if (!ctx->surfaces[i].used) { id = (int)i; break; }
  if (id < 0) { ctx->surfaces.emplace_back(); id = (int)ctx->surfaces.size() - 1; }
  Surface& s = ctx->surfaces[id];
  // keep the allocation if the geometry matches a previously released one
  bool reuse = s.base && s.width == width && s.height == height && s.chroma == chroma;
  if (!reuse) {
    if (s.base) cudaFree(s.base);
    s = Surface();
    s.width = width; s.height = height; s.chroma = chroma;
    s.pw[0] = width; s.ph[0] = height;
    s.pitch[0] = (width + 255) & ~255;
    size_t luma = (size_t)s.pitch[0] * ((height + 63) & ~63);
    size_t cbytes = 0;
    if (chroma) {
      s.pw[1] = s.pw[2] = (width + 1) / 2;
      s.ph[1] = s.ph[2] = (height + 1) / 2;
      s.pitch[1] = s.pitch[2] = (s.pw[1] + 127) & ~127;
      cbytes = (size_t)s.pitch[1] * ((s.ph[1] + 31) & ~31);
    }
    s.bytes = luma + 2 * cbytes + 256;
    CK(cudaMalloc((void**)&s.base, s.bytes));
    s.plane[0] = s.base;
    s.plane[1] = chroma ? s.base + luma : nullptr;
    s.plane[2] = chroma ? s.base + luma + cbytes : nullptr;
  }
  s.used = true;
  *id_out = id;
  return B265_OK;
}

int b265_surface_release(b265_ctx* ctx, int id) {
  Surface* s = get_surface(ctx, id);
  if (!s) return B265_ERR_INVALID_ARG;
  s->used = false;  // memory is kept for reuse by the next alloc of the same geometry
  return B265_OK;
}

int b265_surface_fill(b265_ctx* ctx, int id, int luma, int chroma) {
  Surface* s = get_surface(ctx, id);
  if (!s) return B265_ERR_INVALID_ARG;
  CK(cudaMemset2DAsync(s->plane[0], s->pitch[0], luma, s->pw[0], s->ph[0], ctx->stream));
  for (int c = 1; c < 3; c++)
    if (s->plane[c]) CK(cudaMemset2DAsync(s->plane[c], s->pitch[c], chroma, s->pw[c], s->ph[c], ctx->stream));
  return B265_OK;
}

int b265_surface_upload(b265_ctx* ctx, int id, int plane, const uint8_t* src, ptrdiff_t stride) {
  Surface* s = get_surface(ctx, id);
  if (!s || plane < 0 || plane > 2 || !s->plane[plane] || !src) return B265_ERR_INVALID_ARG;
  CK(cudaMemcpy2DAsync(s->plane[plane], s->pitch[plane], src, stride, s->pw[plane], s->ph[plane],
                       cudaMemcpyHostToDevice, ctx->stream));
  return B265_OK;
}

int b265_surface_download(b265_ctx* ctx, int id, int plane, uint8_t* dst, ptrdiff_t stride, int sync) {
  Surface* s = get_surface(ctx, id);
  if (!s || plane < 0 || plane > 2 || !s->plane[plane] || !dst) return B265_ERR_INVALID_ARG;
  CK(cudaMemcpy2DAsync(dst, stride, s->plane[plane], s->pitch[plane], s->pw[plane], s->ph[plane],
                       cudaMemcpyDeviceToHost, ctx->stream));
  if (sync) CK(cudaStreamSynchronize(ctx->stream));
  return B265_OK;
}

int b265_surface_plane(b265_ctx* ctx, int id, int plane, void** dev_ptr, int* pitch, int* width, int* height) {
  Surface* s = get_surface(ctx, id);
  if (!s || plane < 0 || plane > 2) return B265_ERR_INVALID_ARG;
  if (dev_ptr) *dev_ptr = s->plane[plane];
  if (pitch) *pitch = s->pitch[plane];
  if (width) *width = s->pw[plane];
  if (height) *height = s->ph[plane];
  return B265_OK;
}

int b265_surface_hash(b265_ctx* ctx, const int* ids, int n, int type, uint8_t* out) {
  if (!ctx || !ids || n <= 0 || n > 4096 || type < 0 || type > 2 || !out) return B265_ERR_INVALID_ARG;
  CK(cudaSetDevice(ctx->device));
  std::vector<HashJob> jobs(n);
  int max_h = 1;
  for (int i = 0; i < n; i++) {
    if (ids[i] < 0 || ids[i] >= (int)ctx->surfaces.size() || !ctx->surfaces[ids[i]].used) return B265_ERR_INVALID_ARG;
    const Surface& s = ctx->surfaces[ids[i]];
    if (s.width & 7) return B265_ERR_UNSUPPORTED;   // rows are hashed in 4-byte words (luma and 4:2:0 chroma)
    for (int c = 0; c < 3; c++) {
      const bool present = c == 0 || s.chroma;
      jobs[i].plane[c] = present ? s.plane[c] : nullptr;
      jobs[i].w[c] = c ? (s.width + 1) / 2 : s.width;
      jobs[i].h[c] = c ? (s.height + 1) / 2 : s.height;
      jobs[i].pitch[c] = s.pitch[c];
      if (jobs[i].h[c] > max_h) max_h = jobs[i].h[c];
    }
  }
  const size_t o_out = (sizeof(HashJob) * (size_t)n + 255) & ~(size_t)255, o_rows = o_out + (((size_t)n * 48 + 255) & ~(size_t)255);
  int r = devbuf_reserve(ctx, &ctx->d_hash, o_rows + (size_t)n * 3 * max_h * 2 + 256, false);
  if (r) return r;
  uint8_t* D = (uint8_t*)ctx->d_hash.p;
  CK(cudaMemcpyAsync(D, jobs.data(), sizeof(HashJob) * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));   // pageable: staged before return
  ctx->launches += launch_hash(ctx->stream, type, (const HashJob*)D, n, max_h, (uint16_t*)(D + o_rows), D + o_out);
  CK_LAUNCH("k_hash");
  CK(cudaMemcpyAsync(out, D + o_out, (size_t)n * 48, cudaMemcpyDeviceToHost, ctx->stream));
  return B265_OK;
}

// ------------------------------------------------------------------------------------------
// kernel-level: dense residual batch
// ------------------------------------------------------------------------------------------
int b265_residual_batch(b265_ctx* ctx, int mem, int kind, int log2, int rdpcm, int n_blocks,
                        const int16_t* coeffs, const uint8_t* qp, uint8_t* dst) {
  if (!ctx || log2 < 2 || log2 > 5 || n_blocks < 0 || !coeffs || !dst) return B265_ERR_INVALID_ARG;
  if (kind < 0 || kind > 3 || (kind == B265_TR_DST && log2 != 2)) return B265_ERR_INVALID_ARG;
  if (rdpcm && kind != B265_TR_SKIP && kind != B265_TR_BYPASS) return B265_ERR_INVALID_ARG;
  if (n_blocks == 0) return B265_OK;
  CK(cudaSetDevice(ctx->device));
  const size_t px = (size_t)n_blocks << (2 * log2);
  if (mem == B265_MEM_DEVICE) {
    launch_residual_dense(ctx->stream, log2, kind, rdpcm, n_blocks, coeffs, qp, dst, ctx->sm_count);
    ctx->launches++;
    CK_LAUNCH("k_residual_dense");
    return B265_OK;
  }
  size_t need = px * 2 + px + (qp ? (size_t)n_blocks : 0) + 1024;
  int r = devbuf_reserve(ctx, &ctx->d_stage, need, false);
  if (r) return r;
  uint8_t* base = (uint8_t*)ctx->d_stage.p;
  int16_t* d_coef = (int16_t*)base;
  uint8_t* d_dst = base + px * 2;
  uint8_t* d_qp = qp ? d_dst + ((px + 255) & ~(size_t)255) : nullptr;
  CK(cudaMemcpyAsync(d_coef, coeffs, px * 2, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(d_dst, dst, px, cudaMemcpyHostToDevice, ctx->stream));
  if (qp) CK(cudaMemcpyAsync(d_qp, qp, n_blocks, cudaMemcpyHostToDevice, ctx->stream));
  launch_residual_dense(ctx->stream, log2, kind, rdpcm, n_blocks, d_coef, d_qp, d_dst, ctx->sm_count);
  ctx->launches++;
  CK_LAUNCH("k_residual_dense");
  CK(cudaMemcpyAsync(dst, d_dst, px, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return B265_OK;
}

}  // extern "C"
