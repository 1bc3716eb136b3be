// common.cuh — shared device helpers for the b265 reconstruction kernels (sm_100a).
// Arithmetic macros follow libde265/util.h:109-116 (Clip3 / Clip_BitDepth / Sign).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/b265.h"

#define B265_WARPS_PER_CTA 8
#define B265_FULL_MASK 0xffffffffu

__device__ __forceinline__ int clip3(int lo, int hi, int v) { return min(max(v, lo), hi); }
__device__ __forceinline__ int clip16(int v) { return min(max(v, -32768), 32767); }
__device__ __forceinline__ int clip8(int v) { return min(max(v, 0), 255); }
__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

// loads that must observe stores made by other SMs of the same launch (bypass the non-coherent L1)
__device__ __forceinline__ uint8_t ld_cg_u8(const uint8_t* p) {
  unsigned v;
  asm volatile("ld.global.cg.u8 %0, [%1];" : "=r"(v) : "l"(p));
  return (uint8_t)v;
}
__device__ __forceinline__ uint32_t ld_cg_u32(const void* p) {
  uint32_t v;
  asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint8_t ld_acquire_u8(const uint8_t* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u8 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return (uint8_t)v;
}
__device__ __forceinline__ uint8_t ld_relaxed_u8(const uint8_t* p) {
  unsigned v;
  asm volatile("ld.relaxed.gpu.global.u8 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return (uint8_t)v;
}
__device__ __forceinline__ uint32_t ld_relaxed_u32(const void* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u32(void* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const void* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u64(void* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void fence_acq_rel_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
__device__ __forceinline__ void st_release_u8(uint8_t* p, uint8_t v) {
  asm volatile("st.release.gpu.global.u8 [%0], %1;" ::"l"(p), "r"((unsigned)v) : "memory");
}

// Device-side view of one picture being reconstructed (built by backend.cu from b265_frame_job).
struct FrameDev {
  uint8_t* plane[3];       // recon target (pre-deblock); == out[] when SAO is off
  uint8_t* out[3];         // final (post-SAO) surface planes
  int pitch[3];
  int width, height;       // luma
  int cw, ch;              // chroma plane size (0 for 4:0:0)
  int log2_ctb, log2_min_cb, log2_min_tb;
  int ctbs_w, ctbs_h;
  uint32_t flags;          // B265_FRAME_*
  int cb_qp_offset, cr_qp_offset;
  const b265_tu* tus;
  const b265_coeff* coeffs;
  const b265_pu* pus;
  const b265_slice* slices;
  const b265_ctb* ctbs;       // may be NULL
  const b265_sao* sao;        // may be NULL
  const uint2* sao_rec;       // [3][ctbs_w*ctbs_h] per plane and CTB {ctl, 4 offset bytes}, built on the host (sao_records)
  const uint8_t* unit_info;   // may be NULL
  const int8_t* qp_y;         // may be NULL
  const b265_scaling* scaling;
  const b265_pcm* pcms;       // PCM coding blocks (S4) and their samples (Y, Cb, Cr per block)
  const uint8_t* pcm_samples;
  int n_pcms;
  int n_tus;
  uint64_t* tb_mask;          // per TU: availability mask of the intra TB's neighbouring units (k_intra_prep)
  int16_t* resid[3];          // residuals of the coded intra TBs (same geometry/pitch as plane[], in samples)
  const struct CtbRec* ctb_rec; // per CTB (raster): its TU range and which planes hold intra TBs (built on the host)
  uint8_t* nbinfo;            // intra wavefront, [3][ctbs_w*ctbs_h]: neighbouring CTBs usable for prediction (k_intra_sched_prep)
  int slot_base;              // first (plane, CTB) slot of this picture in batch-wide numbering (trace only)
  uint8_t* mbox[3];           // intra wavefront border mailboxes: per CTB [bottom row S][right column S] as 8-byte words
                              // {4 samples, tag}; zero before the launch, written once by the CTB's warp
  uint8_t* bs;                // boundary strength: per 4x4 luma unit, bits 0-1 vertical edge, 2-3 horizontal
  uint32_t* motion;           // per 4x4 luma unit: 4 words {ref ids, mv0, mv1, flags} rasterised from PUs
  const uint8_t* ref_plane[16][3];
  int ref_poc[16];
  int n_refs;
};

// sao_rec ctl word: bits 0-1 SaoTypeIdx (0 when the slice switches SAO off for the component), bits 2-6 band position /
// edge class, bit 7 = a CTB around this one belongs to another slice or tile (edge offset then follows the reference's
// boundary rules sample by sample on the CTB rim, sao.cc:119-164)
#define B265_SAOREC_FOREIGN 0x80

// one picture to hash (k_hash.cu): planes of an 8-bit surface; plane[c] == NULL: no such plane (4:0:0)
// Interleave the low 4 bits of x (even positions) and y (odd positions).
__host__ __device__ __forceinline__ unsigned morton4(unsigned x, unsigned y) {
  x = (x | (x << 2)) & 0x33; x = (x | (x << 1)) & 0x55;
  y = (y | (y << 2)) & 0x33; y = (y | (y << 1)) & 0x55;
  return x | (y << 1);
}
// int16 residual scratch of the coded intra TBs (FrameDev::resid): BLOCK-LINEAR.  A TB's N x N residual is one contiguous
// run of N*N values, line l at +l*N (lines = rows, or columns for the column-wise predicted modes 2..17), at the z-order
// position of the TB inside its CTB (TBs are aligned quadtree blocks, so such runs never overlap): the row-per-lane stores of
// the residual pre-pass and loads of the wavefront then touch 1/4 .. 1/16 of the cache lines a pitched plane costs.
// (x, y) in samples of the component, lg_s = log2 of the CTB size in that component.
__host__ __device__ __forceinline__ size_t resid_block(int x, int y, int lg_s, int ctbs_w) {
  const int m = (1 << lg_s) - 1;
  return ((size_t)((y >> lg_s) * ctbs_w + (x >> lg_s)) << (2 * lg_s)) + ((size_t)morton4((x & m) >> 2, (y & m) >> 2) << 4);
}

struct HashJob {
  const uint8_t* plane[3];
  int w[3], h[3], pitch[3];
};

// Intra wavefront work unit = one colour plane of one CTB ("ticket").  The host only states, per CTB, where its TUs are
// (decoding order, all planes interleaved) and which planes hold intra TBs; dependency counters and the ready queue are
// built and driven on the device (k_intra.cu).
struct CtbRec {
  uint32_t tu_first;
  uint16_t tu_count;
  uint8_t has;     // bit c: plane c of this CTB holds intra TBs (= a ticket exists)
  uint8_t mixed;   // bit c: ... and also samples produced by earlier launches (inter / PCM)
};
// control block of one launch of the wavefront (device memory, zeroed before k_intra_sched_prep)
#define B265_MAX_CTB_ROWS 1024
struct IntraSched {
  unsigned head;     // next queue index to hand to a warp
  unsigned tail;     // next queue index to fill
  unsigned done;     // finished runs
  unsigned n_runs;   // runs = maximal sequences of horizontally adjacent CTBs with intra TBs (per plane and picture)
  // runs per CTB-row level (all pictures and planes of the launch), counted by k_intra_prep; and how many of a level's
  // queue positions are taken (k_intra_sched_prep): the runs of row r sit after all runs of row r - 1
  unsigned level_cnt[B265_MAX_CTB_ROWS];
  unsigned level_fill[B265_MAX_CTB_ROWS];
};
#define B265_QUEUE_EXIT 0xffffffffu   // queue payload: nothing left, the warp returns
#define B265_QUEUE_SKIP 0xfffffffeu   // queue payload: a row without intra TBs
