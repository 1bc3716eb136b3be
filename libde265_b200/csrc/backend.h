// backend.h — internal definition of b265_ctx (one CUDA device, one stream) shared by the C-ABI files.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
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
