// backend.h — internal definition of b265_ctx (one CUDA device, one stream) shared by the C-ABI files.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <atomic>
#include <string>
#include <vector>
#include "../../include/b265.h"
#include "common.cuh"

struct Surface {
  bool used = false;
  int width = 0, height = 0, chroma = 1;
  uint8_t* base = nullptr;
  uint8_t* plane[3] = {nullptr, nullptr, nullptr};
  int pitch[3] = {0, 0, 0};
  int pw[3] = {0, 0, 0}, ph[3] = {0, 0, 0};
  size_t bytes = 0;
};

// growable device / pinned-host scratch
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
};

struct b265_ctx {
  int device = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  cudaEvent_t ev_sync = nullptr;   // blocking-sync event: b265_sync() sleeps instead of spinning on a host core
  cudaEvent_t ev_busy = nullptr;   // recorded after the last kernel of every batch: not ready = this context has work on the device
  std::atomic<bool> busy_recorded{false};
  std::atomic<int64_t> last_batch_ns{0};   // steady-clock time of the last b265_run_batch / b265_submit_frames
  cudaEvent_t ev_wait = nullptr;   // b265_stream_wait
  cudaEvent_t stage_ev[8] = {};
  bool stage_timing = false;
  float stage_ms[6] = {0, 0, 0, 0, 0, 0};
  uint64_t launches = 0;
  std::string last_error;
  std::vector<Surface> surfaces;
  void* flush_buf = nullptr;
  size_t flush_bytes = 0;
  // staging for b265_submit_frames (reused between calls)
  DevBuf d_stage, h_stage;
  DevBuf d_submit, h_submit;   // grow-only staging of b265_submit_frames
  DevBuf d_hash;               // b265_surface_hash: jobs, digests, CRC row scratch
  b265_batch* submit_batch = nullptr;
};

int b265_fail(b265_ctx* ctx, int status, const char* what, cudaError_t e);
#define CK(call)                                                              \
  do {                                                                        \
    cudaError_t e__ = (call);                                                 \
    if (e__ != cudaSuccess) return b265_fail(ctx, B265_ERR_CUDA, #call, e__); \
  } while (0)
#define CK_LAUNCH(name)                                                              \
  do {                                                                               \
    cudaError_t e__ = cudaGetLastError();                                            \
    if (e__ != cudaSuccess) return b265_fail(ctx, B265_ERR_CUDA, "launch " name, e__); \
  } while (0)

int devbuf_reserve(b265_ctx* ctx, DevBuf* b, size_t bytes, bool host);
// Contexts of this process on ctx's device, other than ctx, whose last batch has not finished (backend.cu).  The intra
// wavefront sizes its persistent grid by it: concurrent batches share the SMs instead of queueing behind a full grid.
int b265_busy_peers(b265_ctx* ctx);
int64_t b265_now_ns();

// Prediction blocks larger than 16 x 8 are cut into pieces of at most 16 x 8 when the lists are staged for the device:
// motion compensation and the motion raster are per-sample functions of (position, motion), so the pieces give the same
// result, every work item of the MC kernel costs at most one half-warp pass (k_mc.cu), and the scheduling is balanced.
#define B265_MC_PIECE_W 16
#define B265_MC_PIECE_H 8
static inline int pu_pieces(const b265_pu& pu) {
  return ((pu.w + B265_MC_PIECE_W - 1) / B265_MC_PIECE_W) * ((pu.h + B265_MC_PIECE_H - 1) / B265_MC_PIECE_H);
}
static inline int pu_expand(const b265_pu& pu, b265_pu* out) {
  int n = 0;
  for (int y = 0; y < pu.h; y += B265_MC_PIECE_H)
    for (int x = 0; x < pu.w; x += B265_MC_PIECE_W) {
      b265_pu q = pu;
      q.x = (uint16_t)(pu.x + x); q.y = (uint16_t)(pu.y + y);
      q.w = (uint8_t)(pu.w - x < B265_MC_PIECE_W ? pu.w - x : B265_MC_PIECE_W);
      q.h = (uint8_t)(pu.h - y < B265_MC_PIECE_H ? pu.h - y : B265_MC_PIECE_H);
      out[n++] = q;
    }
  return n;
}
