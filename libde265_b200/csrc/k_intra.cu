// k_intra.cu — intra prediction (+ residual add) of all intra TBs of a batch of pictures as ONE
// persistent launch: a CTB-level wavefront.
//
// Intra TBs form dependency chains in z-scan order inside a CTB (every TB reads the reconstructed
// samples left of / above it, including below-left and above-right), so a CTB plane is walked
// SEQUENTIALLY by one warp and parallelism comes from (a) the three colour planes, (b) the CTB
// anti-diagonal wave = ctbX + 2*ctbY (a CTB needs its left, upper-left, upper and upper-right neighbours,
// intrapred.cc:597-687), and (c) the pictures of the batch, which are mutually independent.
// The kernel is therefore bound by the LATENCY of that one warp; everything that does not depend on
// neighbouring samples is kept off its path:
//  * the residuals of the intra TBs (dequant + inverse transform) are computed beforehand by the fully
//    parallel k_residual_sparse<.., true> launches into an int16 residual plane (frames.cu);
//  * the availability mask of every TB is computed beforehand by k_intra_prep (one thread per TB);
//  * the samples a TB needs from earlier TBs of the same CTB come from the CTB's edge store in shared
//    memory (intra.cuh), so global memory is read once per CTB (bordering row/column of the neighbouring
//    CTBs, one L2 round trip after the flag wait) and never per TB.
//
// Scheduling (all on the device): the unit of work is a RUN - a maximal sequence of horizontally adjacent CTBs that hold
// intra TBs, in one colour plane of one picture - walked left to right by one warp.  In an all-intra picture a run is a
// whole CTB row (the shape of HEVC's own wavefront parallel processing); in P/B pictures the scattered intra CTBs become
// many short, mostly independent runs.  Inside a run nothing is exchanged through memory: the right column, the corner
// and the upper-right row piece of the CTB just finished ARE the left column, the corner and the upper row piece of the
// next one, so they are moved inside the warp's shared-memory edge store.  Between rows the bottom row of every CTB
// travels through a flag-in-data mailbox (below); a run needs the CTB above-right of its position, i.e. it trails the row
// above by two CTBs, and the words it needs next are requested one CTB ahead.
// Runs sit in a ring-buffer queue ordered by CTB row (all runs of row r after all runs of row r - 1, over all pictures
// and planes of the launch): a run only ever waits for runs of the row above, which were taken earlier by warps that are
// running or done - so polling cannot deadlock for any grid size.  A warp takes a run and walks it to its end; when the
// row above is not far enough AND other runs are waiting in the queue, the warp may put its run (with the column it
// reached) back at the end of the queue and take another one; with nothing else to do it just keeps polling (lowest
// latency for a lone picture).  (The give-back threshold is a launch parameter, B265_INTRA_YIELD_POLLS; measured best on
// B200: never.)  No dependency counters, no atomics and no fences on the path between two dependent CTBs.
//
// HBM traffic per TB: N*N bytes written (+ 2*N*N bytes of residual read when the TB is coded).
#include <stdio.h>
#include <stdlib.h>
#include <algorithm>
#include <vector>
#include "transform.cuh"
#include "intra.cuh"
#include "intra_seg.cuh"
#include "kernels.h"


// log2 sizes whose TBs take the size-independent segment code (intra_seg.cuh) instead of their own unrolled instantiation.
// Measured on B200 (4K / 1080p enc265 streams, generator I pictures; batches of 1..128; gpurun_out r2h-r2j): every size on
// its own instantiation has the shortest chain per TB (a lone 4K picture 1.35 ms; all sizes on the segment code 1.88 ms: the
// 4-pass loop of a 32x32 TB loses the instruction-level parallelism of the unrolled form, and the run-time size costs address
// arithmetic on the chain); only the 16x16 block gains from the segment form - one pass with all 32 lanes busy and half the
// chain per lane (1.28 ms alone, -3..5 % at every batch size).  The smaller code (35 KB instead of 61 KB) only paid off in
// batches of >= 16 4K pictures, not on the 1080p stream.
#ifndef INTRA_SEG_SIZES
#define INTRA_SEG_SIZES 0x10
#endif
#define INTRA_WARPS 4   // one warp per SM sub-partition; several CTAs per SM when the batch is large
#ifndef INTRA_MIN_CTAS
#define INTRA_MIN_CTAS 4  // resident CTAs per SM the register budget is sized for
#endif

// ---------------------------------------------------------------------------------------------
// k_intra_prep: availability mask of every intra TB (one thread per TU), FrameDev::tb_mask[tu];
// and, one thread per (plane, CTB): the CTBs that start a run are counted per CTB-row level
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ bool run_start(const CtbRec* __restrict__ recs, int ctbs_w, int ctb, int c) {
  if (!((recs[ctb].has >> c) & 1)) return false;
  return (ctb % ctbs_w) == 0 || !((recs[ctb - 1].has >> c) & 1);
}

__global__ void __launch_bounds__(256) k_intra_prep(const FrameDev* __restrict__ frames, IntraSched* sched) {
  const FrameDev& f = frames[blockIdx.y];
  {
    const int nctb = f.ctbs_w * f.ctbs_h, planes = f.cw ? 3 : 1;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < planes * nctb; i += gridDim.x * blockDim.x) {
      const int c = i / nctb, ctb = i - c * nctb;
      if (run_start(f.ctb_rec, f.ctbs_w, ctb, c)) atomicAdd(&sched->level_cnt[ctb / f.ctbs_w], 1u);
    }
  }
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < f.n_tus; i += gridDim.x * blockDim.x) {
    const b265_tu tu = f.tus[i];
    if (tu.intra_mode == B265_INTRA_NONE) continue;
    const int c = tu.cidx_log2 & 3, lg = (tu.cidx_log2 >> 2) & 7;
    IntraGeom g;
    g.sh = (c && f.cw) ? 1 : 0;
    g.width = f.width; g.height = f.height;
    g.log2_ctb = f.log2_ctb; g.log2_min_tb = f.log2_min_tb;
    g.cx = (tu.x << g.sh) >> f.log2_ctb; g.cy = (tu.y << g.sh) >> f.log2_ctb;
    g.unit_info = (f.flags & B265_FRAME_CONSTRAINED_INTRA) ? f.unit_info : nullptr;
    g.unit_pitch = f.width >> 2;
    g.nb = 0;
    const b265_ctb* ctbs = f.ctbs;
    const b265_ctb me = ctbs ? ctbs[g.cy * f.ctbs_w + g.cx] : b265_ctb{0, 0, 0};
#pragma unroll
    for (int k = 0; k < 4; k++) {  // left, upper-left, upper, upper-right
      const int nx = g.cx + (k == 3 ? 1 : (k == 2 ? 0 : -1)), ny = g.cy - (k == 0 ? 0 : 1);
      if (nx < 0 || ny < 0 || nx >= f.ctbs_w) continue;
      bool ok = true;
      if (ctbs) {
        const b265_ctb r = ctbs[ny * f.ctbs_w + nx];
        ok = r.slice == me.slice && r.tile == me.tile && r.addr_ts < me.addr_ts;
      }
      if (ok) g.nb |= 1u << k;
    }
    // bits 0..32: availability of the neighbour units; bits 40..63: where the TB's residual lies in the block-linear scratch
    // (resid_block, in units of 16 values), so that the wavefront has no address arithmetic on its chain
    f.tb_mask[i] = intra_avail_mask(g, tu.x, tu.y, lg) |
                   ((uint64_t)(resid_block(tu.x, tu.y, f.log2_ctb - g.sh, f.ctbs_w) >> 4) << 40);
  }
}

void launch_intra_prep(cudaStream_t st, const FrameDev* frames, int n_frames, int max_tus, int max_slots, IntraSched* sched) {
  if (n_frames <= 0 || (max_tus <= 0 && max_slots <= 0)) return;
  int gx = (std::max(max_tus, max_slots) + 255) / 256;
  if (gx > 592) gx = 592;
  k_intra_prep<<<dim3(gx, n_frames), 256, 0, st>>>(frames, sched);
}

// ---------------------------------------------------------------------------------------------
// the wavefront
// ---------------------------------------------------------------------------------------------
// residual line `l` of a coded TB (written by k_residual_sparse<LOG2, true>, transposed for modes 2..17)
template <int N>
__device__ __forceinline__ void load_resid_line(const int16_t* __restrict__ p, uint32_t (&w)[N / 2]) {
  if (N == 4) {
    const uint2 v = __ldg(reinterpret_cast<const uint2*>(p));
    w[0] = v.x; w[1] = v.y;
  } else {
#pragma unroll
    for (int q = 0; q < N / 8; q++) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(p) + q);
      w[4 * q] = v.x; w[4 * q + 1] = v.y; w[4 * q + 2] = v.z; w[4 * q + 3] = v.w;
    }
  }
}

template <int N>
__device__ __forceinline__ void store_words(uint8_t* p, const uint32_t (&px)[N / 4]) {
  if (N == 4) *reinterpret_cast<uint32_t*>(p) = px[0];
  else if (N == 8) *reinterpret_cast<uint2*>(p) = make_uint2(px[0], px[1]);
  else {
#pragma unroll
    for (int q = 0; q < N / 16; q++)
      reinterpret_cast<uint4*>(p)[q] = make_uint4(px[4 * q], px[4 * q + 1], px[4 * q + 2], px[4 * q + 3]);
  }
}
// the edge store is only 4-byte aligned in general
template <int N>
__device__ __forceinline__ void store_words_smem(uint8_t* p, const uint32_t (&px)[N / 4]) {
#pragma unroll
  for (int q = 0; q < N / 4; q++) reinterpret_cast<uint32_t*>(p)[q] = px[q];
}

template <int LOG2, bool TRACE>
__device__ __forceinline__ void intra_tb(const CtbEdges& E, const IntraGeom& g, int cidx, int x, int y, int mode, int flags,
                                         bool coded, uint64_t M, const int16_t* rblk, uint8_t* wborder, long long (&acc)[4]) {
  constexpr int N = 1 << LOG2, STRIDE = IntraBuf<N>::STRIDE;
  long long c0 = 0, c1 = 0, c2 = 0, c3 = 0;
  if (TRACE) c0 = clock64();
  const int lane = lane_id();
  const bool active = lane < N;
  const int l = active ? lane : 0;
  // the residual line is independent of the neighbours: issue its load first
  uint32_t r[N / 2];
  if (coded && active) load_resid_line<N>(rblk + l * N, r);
  const int xr = x - E.x0, yr = y - E.y0;
  uint8_t* buf0 = wborder;
  uint8_t* buf1 = buf0 + STRIDE;
  gather_border<LOG2>(E, xr, yr, M, lane, buf0);
  if (TRACE) c1 = clock64();
  uint8_t* pb = filter_border<LOG2>(buf0, buf1, lane, mode, cidx, g.strong);
  if (TRACE) c2 = clock64();
  uint32_t h[N / 2];
  const bool by_column = predict_line<LOG2>(pb, lane, mode, cidx, !(flags & B265_TU_NO_EDGE_FILT), h);
  if (TRACE) c3 = clock64();
  if (active) {
    uint32_t px[N / 4];
    finish_line<N>(h, r, coded, px);
    uint8_t* d = g.plane + (size_t)y * g.pitch + x;
    uint8_t* erow = E.rows + ((yr + N) >> 2) * E.rstride + 16 + xr;   // row below the TB's last row ... i.e. its bottom row
    uint8_t* ecol = E.cols + ((xr + N) >> 2) * E.S + yr;              // its right column
    // The corner sample of a later TB lies on an edge row AND an edge column but only needs to be this TB's
    // bottom row OR right column: the right column is therefore also entered into the edge rows it crosses.
    if (!by_column) {   // lane = row
      store_words<N>(d + (size_t)l * g.pitch, px);
      ecol[l] = (uint8_t)(px[N / 4 - 1] >> 24);
      if ((l & 3) == 3) E.rows[((yr + l + 1) >> 2) * E.rstride + 16 + xr + N - 1] = (uint8_t)(px[N / 4 - 1] >> 24);
      if (l == N - 1) store_words_smem<N>(erow, px);
    } else {            // lane = column
#pragma unroll
      for (int j = 0; j < N; j++) d[(size_t)j * g.pitch + l] = (uint8_t)(px[j >> 2] >> (8 * (j & 3)));
      erow[l] = (uint8_t)(px[N / 4 - 1] >> 24);
      if (l == N - 1) {
        store_words_smem<N>(ecol, px);
#pragma unroll
        for (int q = 0; q < N / 4; q++) E.rows[((yr + 4 * q + 4) >> 2) * E.rstride + 16 + xr + N - 1] = (uint8_t)(px[q] >> 24);
      }
    }
  }
  __syncwarp();  // the next TB of this CTB reads the edge store
  if (TRACE) {
    acc[0] += c1 - c0; acc[1] += c2 - c1; acc[2] += c3 - c2; acc[3] += clock64() - c3;
  }
}


// 4x4 blocks without the border buffer: a third of the TBs of a textured picture are 4x4 and sit in long serial chains (one
// CTB can hold 64 of them), where the cost of a TB is the LATENCY of its instruction chain, not its 16 samples.  Every lane
// reads the five neighbour units (lower-left, left, corner, top, upper-right) straight from the edge store (same addresses
// in all lanes: broadcasts), substitutes the missing ones in registers and predicts from registers; one barrier per TB (the
// edge store update) instead of three, no shared-memory round trip between gather and prediction.  Same arithmetic as
// gather_border / predict_line (intra.cuh), bit-identical.
//   selectors that build ref[-4..-1] of the main line from the other line's samples 1..4 (inverse-angle projection,
//   intrapred.cc:937-951), per mode 11..25
__constant__ uint16_t c_neg_sel4[15] = {0x0, 0x0, 0x3000, 0x1000, 0x1300, 0x1200, 0x130, 0x120, 0x130, 0x1200, 0x1300, 0x1000, 0x3000, 0x0, 0x0};

template <bool TRACE>
__device__ __forceinline__ void intra_tb4(const CtbEdges& E, const IntraGeom& g, int cidx, int x, int y, int mode, int flags,
                                          bool coded, uint32_t M, const int16_t* rblk, long long (&acc)[4]) {
  long long c0 = 0;
  if (TRACE) c0 = clock64();
  const int lane = lane_id();
  const bool active = lane < 4;
  const int l = lane & 3;
  uint2 res = make_uint2(0, 0);
  if (coded && active) res = __ldg(reinterpret_cast<const uint2*>(rblk + 4 * l));
  const int xr = x - E.x0, yr = y - E.y0;
  const uint8_t* colp = E.cols + (xr >> 2) * E.S + yr;              // column xr-1, rows yr ..
  const uint8_t* rowp = E.rows + (yr >> 2) * E.rstride + 16 + xr;   // row yr-1, columns xr ..
  // units in scan order: 0 lower-left (rows yr+4..7), 1 left, 2 corner, 3 top, 4 upper-right
  uint32_t U[5];
  U[0] = *reinterpret_cast<const uint32_t*>(colp + 4);
  U[1] = *reinterpret_cast<const uint32_t*>(colp);
  U[2] = __byte_perm(*reinterpret_cast<const uint32_t*>(rowp - 4), 0, 0x3333);
  U[3] = *reinterpret_cast<const uint32_t*>(rowp);
  U[4] = *reinterpret_cast<const uint32_t*>(rowp + 4);
  // reference sample substitution (intrapred.cc:395-431): the scan runs UP the left column (first sample of a left unit =
  // byte 3, last = byte 0) and RIGHT along the top row (first = byte 0, last = byte 3)
  if (M == 0) {
#pragma unroll
    for (int i = 0; i < 5; i++) U[i] = 0x80808080u;
  } else {
    const int first = __ffs((int)M) - 1;
    uint32_t lead = 0;
#pragma unroll
    for (int i = 0; i < 5; i++)
      if (i == first) lead = __byte_perm(U[i], 0, i < 2 ? 0x3333 : 0x0000);
#pragma unroll
    for (int i = 0; i < 5; i++) {
      if (i < first) U[i] = lead;
      else if (i > 0 && !((M >> i) & 1)) U[i] = __byte_perm(U[i - 1], 0, (i - 1) < 2 ? 0x0000 : 0x3333);
    }
  }
  const uint32_t L0 = U[1], L1 = U[0], T0 = U[3], T1 = U[4], corner = U[2] & 0xff;   // L[1..4], L[5..8], T[1..4], T[5..8]
  uint32_t h0, h1;
  bool by_column = false;
  if (mode >= 2) {   // angular (intrapred.cc:903-1005)
    const bool horiz = mode < 18;
    const uint32_t m0 = horiz ? L0 : T0, m1 = horiz ? L1 : T1, o0 = horiz ? T0 : L0;
    const int ang = c_intra_angle[mode];
    // ref[-4..11] as four words; ref[0] = corner, ref[k] = main line sample k
    const uint32_t R0 = (ang < 0) ? __byte_perm(o0, 0, c_neg_sel4[mode - 11]) : 0u;
    const uint32_t R1 = corner | (m0 << 8), R2 = __funnelshift_r(m0, m1, 24), R3 = m1 >> 24;
    const int t = (l + 1) * ang, idx = t >> 5, f = t & 31;
    const int p = idx + 5, w = p >> 2, b = 8 * (p & 3);   // ref[idx+1] is byte p of the sixteen
    const uint32_t A = w == 0 ? R0 : (w == 1 ? R1 : R2), B = w == 0 ? R1 : (w == 1 ? R2 : R3), Cw = w == 0 ? R2 : (w == 1 ? R3 : 0u);
    const uint32_t lo0 = __funnelshift_r(A, B, b), lo1 = __funnelshift_r(B, Cw, b);
    const uint32_t W0 = (uint32_t)(32 - f) | ((uint32_t)f << 8);
    const uint32_t b34 = __funnelshift_r(lo0, lo1, 24);
    const uint32_t v0 = __dp4a(lo0, W0, 16u), v1 = __dp4a(lo0, W0 << 8, 16u), v2 = __dp4a(lo0, W0 << 16, 16u), v3 = __dp4a(b34, W0, 16u);
    h0 = (__byte_perm(v0, v1, 0x5410) >> 5) & 0x00ff00ffu;
    h1 = (__byte_perm(v2, v3, 0x5410) >> 5) & 0x00ff00ffu;
    if ((mode == 26 || mode == 10) && cidx == 0 && !(flags & B265_TU_NO_EDGE_FILT))
      h0 = (h0 & 0xffff0000u) | (uint32_t)clip8((int)(m0 & 0xff) + (((int)((o0 >> (8 * l)) & 0xff) - (int)corner) >> 1));
    by_column = horiz;
  } else if (mode == 0) {   // planar (intrapred.cc:1009-1032), lane = row
    const uint32_t abd = ((L0 >> (8 * l)) & 0xff) | ((T1 & 0xff) << 8) | ((L1 & 0xff) << 24);
    const uint32_t Wy = ((uint32_t)(3 - l) << 16) | ((uint32_t)(l + 1) << 24);
    uint32_t v[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const uint32_t op = __byte_perm(abd, T0, 0x3010u | ((4u + i) << 8));   // [L[1+y], T[5], T[1+x], L[5]]
      v[i] = __dp4a(op, Wy | (uint32_t)(3 - i) | ((uint32_t)(i + 1) << 8), 4u) >> 3;
    }
    h0 = v[0] | (v[1] << 16);
    h1 = v[2] | (v[3] << 16);
  } else {   // DC (intrapred.cc:1036-1069), lane = row
    const int dc = (int)(__dp4a(T0, 0x01010101u, 0u) + __dp4a(L0, 0x01010101u, 4u)) >> 3;
    h0 = h1 = (uint32_t)dc * 0x00010001u;
    if (cidx == 0) {
      if (l == 0) {
        h0 = (uint32_t)(((int)(T0 & 0xff) + 3 * dc + 2) >> 2) | ((uint32_t)(((int)((T0 >> 8) & 0xff) + 3 * dc + 2) >> 2) << 16);
        h1 = (uint32_t)(((int)((T0 >> 16) & 0xff) + 3 * dc + 2) >> 2) | ((uint32_t)(((int)(T0 >> 24) + 3 * dc + 2) >> 2) << 16);
        h0 = (h0 & 0xffff0000u) | (uint32_t)(((int)(L0 & 0xff) + 2 * dc + (int)(T0 & 0xff) + 2) >> 2);
      } else {
        h0 = (h0 & 0xffff0000u) | (uint32_t)(((int)((L0 >> (8 * l)) & 0xff) + 3 * dc + 2) >> 2);
      }
    }
  }
  if (coded) {
    h0 = __viaddmin_s16x2_relu(h0, res.x, 0x00ff00ffu);
    h1 = __viaddmin_s16x2_relu(h1, res.y, 0x00ff00ffu);
  }
  const uint32_t px = __byte_perm(h0, h1, 0x6420);
  if (active) {
    uint8_t* d = g.plane + (size_t)y * g.pitch + x;
    uint8_t* erow = E.rows + ((yr + 4) >> 2) * E.rstride + 16 + xr;   // the TB's bottom row
    uint8_t* ecol = E.cols + ((xr + 4) >> 2) * E.S + yr;              // its right column
    if (!by_column) {   // lane = row
      *reinterpret_cast<uint32_t*>(d + (size_t)l * g.pitch) = px;
      ecol[l] = (uint8_t)(px >> 24);
      if (l == 3) *reinterpret_cast<uint32_t*>(erow) = px;
    } else {            // lane = column
#pragma unroll
      for (int j = 0; j < 4; j++) d[(size_t)j * g.pitch + l] = (uint8_t)(px >> (8 * j));
      erow[l] = (uint8_t)(px >> 24);
      if (l == 3) *reinterpret_cast<uint32_t*>(ecol) = px;
    }
  }
  __syncwarp();  // the next TB of this CTB reads the edge store
  if (TRACE) acc[0] += clock64() - c0;
}

// Size-independent form (intra_seg.cuh): a lane's unit of work is one 8-sample segment of a line; a 32x32 block is four
// passes of the warp, a 16x16 block one pass with all lanes busy.  One body for every size.
template <bool TRACE>
__device__ __forceinline__ void intra_tb_seg(const CtbEdges& E, const IntraGeom& g, int cidx, int x, int y, int log2, int mode, int flags,
                                             bool coded, uint64_t M, const int16_t* rblk, uint8_t* wborder, long long (&acc)[4]) {
  long long c0 = 0, c1 = 0, c2 = 0;
  if (TRACE) c0 = clock64();
  const int lane = lane_id();
  const int N = 1 << log2;
  const int lseg = log2 < 3 ? 0 : log2 - 3;        // segments per line = 1 << lseg (1, 1, 2, 4)
  const int items = N << lseg;                     // 4, 8, 32, 128
  const bool small = log2 == 2;                    // half segments: 4 samples per line
  // the residual of the first pass is independent of the neighbours: issue its load first
  int item = lane;
  uint32_t r[4] = {0, 0, 0, 0};
  const int16_t* rbase = rblk;
  auto load_resid = [&](int it, uint32_t (&rr)[4]) {
    if (!coded || it >= items) return;
    const int l = it >> lseg, sg = it & ((1 << lseg) - 1);
    const int16_t* p = rbase + l * N + 8 * sg;
    if (small) {
      const uint2 v = __ldg(reinterpret_cast<const uint2*>(p));
      rr[0] = v.x; rr[1] = v.y;
    } else {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
      rr[0] = v.x; rr[1] = v.y; rr[2] = v.z; rr[3] = v.w;
    }
  };
  load_resid(item, r);
  const int xr = x - E.x0, yr = y - E.y0;
  uint8_t* buf0 = wborder;
  uint8_t* buf1 = buf0 + ibuf_stride(N);
  if (log2 == 5) gather_border_rt<true>(E, N, xr, yr, M, lane, buf0);
  else gather_border_rt<false>(E, N, xr, yr, M, lane, buf0);
  if (TRACE) c1 = clock64();
  uint8_t* pb = filter_border_rt(buf0, buf1, N, lane, mode, cidx, g.strong);
  if (TRACE) c2 = clock64();
  IntraPred P;
  intra_pred_setup(P, pb, N, log2, lane, mode, cidx, !(flags & B265_TU_NO_EDGE_FILT));
  uint8_t* d = g.plane + (size_t)y * g.pitch + x;
  uint8_t* erow = E.rows + ((yr + N) >> 2) * E.rstride + 16 + xr;   // the TB's bottom row in the edge store
  uint8_t* ecol = E.cols + ((xr + N) >> 2) * E.S + yr;              // its right column
  const int last_seg = (1 << lseg) - 1;
  for (; item < items; item += 32) {   // warp-uniform trip count is not needed: nothing inside synchronises
    uint32_t r_next[4] = {0, 0, 0, 0};
    load_resid(item + 32, r_next);
    const int l = item >> lseg, sg = item & last_seg;
    uint32_t h[4], px[2];
    intra_pred_seg(P, l, sg, h);
    finish_seg(h, r, coded, px);
    const uint32_t last_word = small ? px[0] : px[1];
    // The corner sample of a later TB lies on an edge row AND an edge column but only needs to be this TB's
    // bottom row OR right column: the right column is therefore also entered into the edge rows it crosses.
    if (!P.by_column) {   // line = row
      uint8_t* o = d + (size_t)l * g.pitch + 8 * sg;
      if (small) *reinterpret_cast<uint32_t*>(o) = px[0];
      else *reinterpret_cast<uint2*>(o) = make_uint2(px[0], px[1]);
      if (sg == last_seg) {
        ecol[l] = (uint8_t)(last_word >> 24);
        if ((l & 3) == 3) E.rows[((yr + l + 1) >> 2) * E.rstride + 16 + xr + N - 1] = (uint8_t)(last_word >> 24);
      }
      if (l == N - 1) {
        reinterpret_cast<uint32_t*>(erow + 8 * sg)[0] = px[0];
        if (!small) reinterpret_cast<uint32_t*>(erow + 8 * sg)[1] = px[1];
      }
    } else {              // line = column, segment = 8 rows
      uint8_t* o = d + (size_t)(8 * sg) * g.pitch + l;
#pragma unroll
      for (int j = 0; j < 8; j++)
        if (j < 4 || !small) o[(size_t)j * g.pitch] = (uint8_t)(px[j >> 2] >> (8 * (j & 3)));
      if (sg == last_seg) erow[l] = (uint8_t)(last_word >> 24);
      if (l == N - 1) {
        reinterpret_cast<uint32_t*>(ecol + 8 * sg)[0] = px[0];
        E.rows[((yr + 8 * sg + 4) >> 2) * E.rstride + 16 + xr + N - 1] = (uint8_t)(px[0] >> 24);
        if (!small) {
          reinterpret_cast<uint32_t*>(ecol + 8 * sg)[1] = px[1];
          E.rows[((yr + 8 * sg + 8) >> 2) * E.rstride + 16 + xr + N - 1] = (uint8_t)(px[1] >> 24);
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 4; i++) r[i] = r_next[i];
  }
  __syncwarp();  // the next TB of this CTB reads the edge store
  if (TRACE) {
    acc[0] += c1 - c0; acc[1] += c2 - c1; acc[3] += clock64() - c2;
  }
}

// ---- border exchange between CTB rows: flag-in-data mailboxes --------------------------------------------
// A CTB's warp posts the CTB's bottom row and right column (the only samples later CTBs can read) as
// 8-byte words {4 samples, tag != 0} written with ONE 64-bit store each: a reader that sees the tag has
// the samples, so neither side needs a fence (the scheme of low-latency collective protocols).  The
// mailboxes are zeroed before the launch.  Layout per CTB: S/4 words bottom row, S/4 words right column.
// Neighbours without intra TBs (inter/PCM CTBs) were finished by earlier launches: their samples are read
// from the plane.  The right column is only read back when a row is resumed by another warp.
//
// The polling loop is executed by the WHOLE warp with a warp-uniform exit: a per-lane loop leaves the
// warp split into independently scheduled groups, which then walk the TBs separately and pay the slow
// path of every warp collective (measured: 3x).

// mailbox word `lane - q` of the upper-right neighbour's bottom row, requested ahead of time by the lanes that will consume it
__device__ __forceinline__ unsigned long long prefetch_upper_right(const uint8_t* mbox, int S, int ctbs_w, int cx, int cy) {
  const int lane = lane_id(), q = S / 4;
  unsigned long long v = 0;
  if (cy > 0 && cx + 1 < ctbs_w && lane >= q && lane < 2 * q)
    v = ld_relaxed_u64(mbox + (size_t)((cy - 1) * ctbs_w + cx + 1) * (4 * S) + 8 * (lane - q));
  return v;
}

// Inside a row: the finished CTB's right column, the corner above it and the upper-right row piece become the next CTB's
// left column, corner and upper row piece (all inside the edge store).
__device__ __forceinline__ void shift_edges(const CtbEdges& E) {
  const int lane = lane_id();
  const int S = E.S, q = S / 4;
  uint32_t top = 0, left = 0;
  const uint8_t corner = E.rows[16 + S - 1];
  if (lane < q) {
    top = *reinterpret_cast<const uint32_t*>(E.rows + 16 + S + 4 * lane);
    left = *reinterpret_cast<const uint32_t*>(E.cols + q * S + 4 * lane);
  }
  __syncwarp();
  if (lane < q) {
    *reinterpret_cast<uint32_t*>(E.rows + 16 + 4 * lane) = top;
    *reinterpret_cast<uint32_t*>(E.cols + 4 * lane) = left;
    E.rows[(lane + 1) * E.rstride + 15] = (uint8_t)(left >> 24);   // sample (-1, 4w+3) of the inner edge rows
  }
  if (lane == 0) E.rows[15] = corner;
}

// Borders of the CTB at (g.cx, g.cy) into the edge store.  reuse: everything but the upper-right piece is already there
// (shift_edges).  pre: prefetch_upper_right() of this CTB.  Returns false when the neighbours have not posted after
// `yield_polls` polls while other rows wait in the queue: the caller gives the row back.
__device__ __forceinline__ bool fetch_borders(const IntraGeom& g, const CtbEdges& E, const CtbRec* recs, int plane, const uint8_t* mbox,
                                              int ctbs_w, bool interior, bool reuse, unsigned long long pre, const IntraSched* sched,
                                              unsigned sleep_ns, unsigned yield_polls) {
  const int lane = lane_id();
  const int S = E.S, q = S / 4, pitch = g.pitch;
  const int W = (g.width + g.sh) >> g.sh, H = (g.height + g.sh) >> g.sh;
  const uint8_t* org = g.plane + (size_t)E.y0 * pitch + E.x0;
  (void)recs; (void)plane;
  if (reuse && !interior) {
    // The common step inside a row: only the upper-right CTB's bottom row is missing, its words were requested one CTB
    // ahead by the lanes [q, 2q) that store them (slot == lane).
    bool wr = false, have = true;
    uint32_t val = 0;
    const uint8_t* src = nullptr;
    if (lane >= q && lane < 2 * q && ((g.nb >> 3) & 1)) {
      const int w = lane - q;
      if ((g.nb >> 7) & 1) {
        wr = true;
        if ((uint32_t)(pre >> 32) != 0) val = (uint32_t)pre;
        else { have = false; src = mbox + (size_t)((g.cy - 1) * ctbs_w + g.cx + 1) * (4 * S) + 8 * w; }
      } else if (E.x0 + S + 4 * w < W) {
        wr = true;
        val = ld_cg_u32(org - pitch + S + 4 * w);
      }
    }
    unsigned polls = 0;
    while (!__all_sync(B265_FULL_MASK, have)) {
      if (polls) {
        if (polls >= yield_polls) {   // warp-uniform: is another row waiting for a warp?
          const unsigned hd = ld_relaxed_u32(&sched->head), tl = ld_relaxed_u32(&sched->tail);
          if ((int)(tl - hd) > 0) return false;
          polls = 1;
        }
        if (sleep_ns) __nanosleep(sleep_ns);
      }
      polls++;
      if (!have) {
        const unsigned long long v = ld_relaxed_u64(src);
        if ((uint32_t)(v >> 32) != 0) { val = (uint32_t)v; have = true; }
      }
    }
    if (wr) *reinterpret_cast<uint32_t*>(E.rows + 16 + 4 * lane) = val;
    __syncwarp();
    return true;
  }
  // slots: [0,q) upper CTB's bottom row, [q,2q) upper-right CTB's bottom row, [2q,3q) left CTB's right column,
  // 3q: corner (last word of the upper-left CTB's bottom row); two slots per lane (3q+1 <= 49)
  const uint8_t* src[2];
  int kind[2];   // 0 nothing, 1 mailbox word, 2 plane (row word), 3 plane (column word, 4 strided bytes), 4 plane corner byte
  int slotv[2];
  const int lq = g.log2_ctb - g.sh - 2;   // q = 1 << lq
  const bool two = !reuse && 3 * q >= 32; // more than 32 slots: only for 64-sample CTB planes
#pragma unroll
  for (int k = 0; k < 2; k++) {
    const int slot = lane + 32 * k;
    slotv[k] = slot;
    kind[k] = 0;
    src[k] = nullptr;
    if (k == 1 && !two) break;   // warp-uniform
    if (slot > 3 * q) continue;
    const int piece = slot >> lq, w = slot == 3 * q ? q - 1 : slot & (q - 1);
    if (reuse && piece != 1) continue;
    // neighbour CTB of the piece
    const int nx = g.cx + (piece == 1 ? 1 : (piece == 0 ? 0 : -1)), ny = g.cy - (piece == 2 ? 0 : 1);
    // not usable for prediction (outside the picture, other slice/tile): never read, never waited for
    const int nbit = piece == 0 ? 2 : piece == 1 ? 3 : piece == 2 ? 0 : 1;
    if (!((g.nb >> nbit) & 1)) continue;
    const int nctb = ny * ctbs_w + nx;
    if ((g.nb >> (4 + nbit)) & 1) {   // the neighbour posts its borders in this launch (k_intra_sched_prep)
      kind[k] = 1;
      src[k] = mbox + (size_t)nctb * (4 * S) + (piece == 2 ? 2 * S : 0) + 8 * w;
    } else if (piece == 2) {
      if (E.y0 + 4 * w < H) { kind[k] = 3; src[k] = org + (size_t)(4 * w) * pitch - 1; }
    } else if (piece == 3) {
      kind[k] = 4; src[k] = org - pitch - 1;
    } else {
      const int x = (piece == 1 ? S : 0) + 4 * w;
      if (E.x0 + x < W) { kind[k] = 2; src[k] = org - pitch + x; }
    }
  }
  uint32_t val[2] = {0, 0};
  bool have[2] = {kind[0] != 1, kind[1] != 1};
  // the word requested one CTB ahead (same lane, slot [q,2q) is always slot 0 of its lane)
  if (kind[0] == 1 && lane >= q && lane < 2 * q && (uint32_t)(pre >> 32) != 0) { val[0] = (uint32_t)pre; have[0] = true; }
  // plane reads (rare: mixed pictures)
#pragma unroll
  for (int k = 0; k < 2; k++) {
    if (k == 1 && !two) break;
    if (kind[k] == 2) val[k] = ld_cg_u32(src[k]);
    else if (kind[k] == 3) {
      for (int i = 0; i < 4; i++) val[k] |= (uint32_t)ld_cg_u8(src[k] + (size_t)i * pitch) << (8 * i);
    } else if (kind[k] == 4) val[k] = (uint32_t)ld_cg_u8(src[k]) << 24;
  }
  // every lane polls its own word(s): restricting the polling to one lane per neighbour, or pausing longer
  // between polls, was measured slower at every batch size (the wavefront is bound by detection latency)
  unsigned polls = 0;
  while (!__all_sync(B265_FULL_MASK, have[0] && have[1])) {
    if (polls) {
      if (polls >= yield_polls) {   // warp-uniform: is another row waiting for a warp?
        const unsigned hd = ld_relaxed_u32(&sched->head), tl = ld_relaxed_u32(&sched->tail);
        if ((int)(tl - hd) > 0) return false;
        polls = 1;
      }
      if (sleep_ns) __nanosleep(sleep_ns);
    }
    polls++;
#pragma unroll
    for (int k = 0; k < 2; k++)
      if (!have[k] && (k == 0 || two)) {
        const unsigned long long v = ld_relaxed_u64(src[k]);
        if ((uint32_t)(v >> 32) != 0) { val[k] = (uint32_t)v; have[k] = true; }
      }
  }
  // into the edge store
#pragma unroll
  for (int k = 0; k < 2; k++) {
    if (k == 1 && !two) break;
    if (kind[k] == 0) continue;
    const int slot = slotv[k];
    if (slot < 2 * q) *reinterpret_cast<uint32_t*>(E.rows + 16 + 4 * slot) = val[k];
    else if (slot < 3 * q) {
      const int w = slot - 2 * q;
      *reinterpret_cast<uint32_t*>(E.cols + 4 * w) = val[k];
      E.rows[(w + 1) * E.rstride + 15] = (uint8_t)(val[k] >> 24);   // sample (-1, 4w+3) of the inner edge rows
    } else E.rows[15] = (uint8_t)(val[k] >> 24);
  }
  if (interior) {
    // CTBs that also hold inter/PCM samples: their edge rows 4q-1 (q >= 1) and columns 4q-1 come from the plane
    const int wpr = S / 4;
    for (int i = lane; i < (S / 4) * wpr; i += 32) {
      const int qq = 1 + i / wpr, w = i % wpr, r = 4 * qq - 1;
      if (E.y0 + r < H && E.x0 + 4 * w < W)
        *reinterpret_cast<uint32_t*>(E.rows + qq * E.rstride + 16 + 4 * w) = ld_cg_u32(org + (size_t)r * pitch + 4 * w);
    }
    for (int i = lane; i < (S / 4) * S; i += 32) {
      const int qq = 1 + i / S, r = i % S, cc = 4 * qq - 1;
      if (E.y0 + r < H && E.x0 + cc < W) E.cols[qq * S + r] = ld_cg_u8(org + (size_t)r * pitch + cc);
    }
  }
  __syncwarp();
  return true;
}

// post this CTB's bottom row and right column (edge store rows[S/4], cols[S/4]) to its mailbox
__device__ __forceinline__ void post_borders(const CtbEdges& E, uint8_t* mbox, int ctb) {
  const int lane = lane_id();
  const int S = E.S, q = S / 4;
  if (lane < 2 * q) {
    const uint8_t* s = lane < q ? E.rows + q * E.rstride + 16 + 4 * lane : E.cols + q * S + 4 * (lane - q);
    const unsigned long long v = (unsigned long long)*reinterpret_cast<const uint32_t*>(s) | (1ull << 32);
    st_relaxed_u64(mbox + (size_t)ctb * (4 * S) + 8 * lane, v);
  }
}

// ---- scheduling ------------------------------------------------------------------------------------------
// Ring-buffer queue of CTB rows.  Entry i lives in queue[i & (entries - 1)] as {payload, i + 1}: a warp that took index i
// sleeps until the tag is i + 1.  payload = frame << 22 | plane << 20 | row << 10 | first column still to do.
__device__ __forceinline__ uint32_t row_payload(int frame, int plane, int row, int col) {
  return ((uint32_t)frame << 22) | ((uint32_t)plane << 20) | ((uint32_t)row << 10) | (uint32_t)col;
}
__device__ __forceinline__ void queue_put(unsigned long long* queue, unsigned mask, unsigned idx, uint32_t payload) {
  st_relaxed_u64(queue + (idx & mask), ((unsigned long long)(idx + 1u) << 32) | payload);
}

// Per (plane, CTB) slot: which neighbouring CTBs may be used for prediction (same slice, same tile, earlier in decoding
// order; intrapred.cc:486-508): bit0 left, bit1 upper-left, bit2 upper, bit3 upper-right.  The thread of a CTB that starts
// a run also enters the run into the queue, inside the index range of its CTB-row level: level_base[row] (exclusive scan of
// the per-level counts of k_intra_prep, recomputed by every block) + the level's fill counter.  A run only waits for runs
// of the row above, which therefore all have smaller queue indices - whatever the order inside a level.
__global__ void __launch_bounds__(256) k_intra_sched_prep(const FrameDev* __restrict__ frames, IntraSched* sched, unsigned long long* queue,
                                                          unsigned qmask, int max_rows) {
  __shared__ unsigned s_base[B265_MAX_CTB_ROWS];
  __shared__ unsigned s_part[256];
  {  // exclusive scan of level_cnt[0 .. max_rows): 4 levels per thread
    unsigned v[4], sum = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int l = 4 * threadIdx.x + k;
      v[k] = l < max_rows ? ld_cg_u32(&sched->level_cnt[l]) : 0u;
      sum += v[k];
    }
    s_part[threadIdx.x] = sum;
    __syncthreads();
    for (int o = 1; o < 256; o <<= 1) {
      const unsigned add = threadIdx.x >= (unsigned)o ? s_part[threadIdx.x - o] : 0u;
      __syncthreads();
      s_part[threadIdx.x] += add;
      __syncthreads();
    }
    unsigned run = s_part[threadIdx.x] - sum;
#pragma unroll
    for (int k = 0; k < 4; k++) { s_base[4 * threadIdx.x + k] = run; run += v[k]; }
    if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 255) { sched->tail = s_part[255]; sched->n_runs = s_part[255]; }
    __syncthreads();
  }
  const FrameDev& f = frames[blockIdx.y];
  const int nctb = f.ctbs_w * f.ctbs_h, planes = f.cw ? 3 : 1;
  const CtbRec* __restrict__ recs = f.ctb_rec;
  const b265_ctb* __restrict__ ctbs = f.ctbs;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < planes * nctb; i += gridDim.x * blockDim.x) {
    const int c = i / nctb, ctb = i - c * nctb;
    const int cx = ctb % f.ctbs_w, cy = ctb / f.ctbs_w;
    uint32_t nb = 0;
    if ((recs[ctb].has >> c) & 1) {
      const b265_ctb me = ctbs ? ctbs[ctb] : b265_ctb{0, 0, 0};
#pragma unroll
      for (int k = 0; k < 4; k++) {  // left, upper-left, upper, upper-right
        const int nx = cx + (k == 3 ? 1 : (k == 2 ? 0 : -1)), ny = cy - (k == 0 ? 0 : 1);
        if (nx < 0 || ny < 0 || nx >= f.ctbs_w) continue;
        bool ok = true;
        if (ctbs) {
          const b265_ctb r = ctbs[ny * f.ctbs_w + nx];
          ok = r.slice == me.slice && r.tile == me.tile && r.addr_ts < me.addr_ts;
        }
        // bit k: usable for prediction; bit 4 + k: and it posts its borders in this launch (it has intra TBs in this plane),
        // so the wavefront never reads a neighbour's CTB record on its chain
        if (ok) nb |= (1u << k) | (((recs[ny * f.ctbs_w + nx].has >> c) & 1u) << (4 + k));
      }
      if (run_start(recs, f.ctbs_w, ctb, c))
        queue_put(queue, qmask, s_base[cy] + atomicAdd(&sched->level_fill[cy], 1u), row_payload(blockIdx.y, c, cy, cx));
    }
    f.nbinfo[i] = (uint8_t)nb;
  }
}

// Words 0, 1 and 3 of a 16-byte TU record.  Word 2 (the coefficient offset) is not needed here, and a 128-bit load would leave
// its destination register dead: the compiler reuses such a register at once, and the first write to it then waits for the
// whole load (write-after-write on an in-flight load; it was 7.7 % of all stall samples, on the chain between two CTBs).
__device__ __forceinline__ uint4 load_tu_xyw(const uint4* p) {
  const uint2 a = __ldg(reinterpret_cast<const uint2*>(p));
  const uint32_t w = __ldg(reinterpret_cast<const uint32_t*>(p) + 3);
  return make_uint4(a.x, a.y, 0u, w);
}

// TRACE (developer aid, B265_INTRA_TRACE=<file>): per (plane, CTB) slot 12 x uint64 {t_start, t_setup, t_borders, t_tbs, t_end,
// n_tbs, gather, filter, predict, add+store, globaltimer, (how << 32) | sm id}; how: 0 = first CTB after taking the row from the
// queue, 1 = next CTB of the row with the edge store reused, 2 = next CTB with a full border fetch
template <bool TRACE, int SEGMAX>
__global__ void __launch_bounds__(INTRA_WARPS * 32, INTRA_MIN_CTAS)
k_intra_wavefront(const FrameDev* __restrict__ frames, IntraSched* sched, unsigned long long* queue, unsigned qmask, unsigned total_warps,
                  unsigned sleep_ns, unsigned yield_polls, int tile_bytes, unsigned long long* trace) {
  extern __shared__ __align__(16) uint8_t dyn_smem[];
  __shared__ __align__(16) uint8_t s_border[INTRA_WARPS][B265_BORDER_BYTES];
  const int warp = threadIdx.x >> 5, lane = lane_id();
  uint8_t* my_edges = dyn_smem + (size_t)warp * tile_bytes;
  const unsigned n_runs = ld_cg_u32(&sched->n_runs);
  // picture / plane constants of the current row, reloaded only when a queue entry names another picture or plane
  IntraGeom g;
  int ctbs_w = 1, cur_frame = -1, cur_plane = -1, slot_base = 0;
  const CtbRec* __restrict__ recs = nullptr;
  const uint8_t* __restrict__ nbinfo = nullptr;
  uint8_t* mbox = nullptr;
  const uint4* tus4 = nullptr;
  const uint2* masks = nullptr;
  auto load_constants = [&](int frame, int c) {
    const FrameDev* __restrict__ f = frames + frame;
    const uint32_t fflags = f->flags;
    ctbs_w = f->ctbs_w;
    recs = f->ctb_rec;
    nbinfo = f->nbinfo + (size_t)c * ctbs_w * f->ctbs_h;
    mbox = f->mbox[c];
    slot_base = f->slot_base + c * ctbs_w * f->ctbs_h;
    tus4 = reinterpret_cast<const uint4*>(f->tus);  // 16-byte records, 256-byte aligned list
    masks = reinterpret_cast<const uint2*>(f->tb_mask);
    g.sh = (c && f->cw) ? 1 : 0;
    g.width = f->width; g.height = f->height;
    g.log2_ctb = f->log2_ctb; g.log2_min_tb = f->log2_min_tb; g.ctbs_w = ctbs_w;
    g.plane = f->plane[c]; g.resid = f->resid[c]; g.pitch = f->pitch[c];
    g.unit_info = (fflags & B265_FRAME_CONSTRAINED_INTRA) ? f->unit_info : nullptr;
    g.unit_pitch = g.width >> 2;
    g.strong = (fflags & B265_FRAME_STRONG_INTRA_SMOOTHING) != 0;
    cur_frame = frame; cur_plane = c;
  };
  for (;;) {
    // ---- take a row ----
    unsigned idx = 0;
    if (lane == 0) idx = atomicAdd(&sched->head, 1u);
    idx = __shfl_sync(B265_FULL_MASK, idx, 0);
    uint32_t payload;
    {
      unsigned ns = 32;
      for (;;) {   // sleep on the entry: entries are filled in index order, so the first row that is put back wakes the first sleeper
        const unsigned long long e = ld_relaxed_u64(queue + (idx & qmask));
        if ((uint32_t)(e >> 32) == idx + 1u) { payload = (uint32_t)e; break; }
        __nanosleep(ns);
        if (ns < 1024) ns <<= 1;
      }
    }
    if (payload == B265_QUEUE_EXIT) return;
    if (payload == B265_QUEUE_SKIP) continue;
    const int frame = (int)(payload >> 22), c = (int)((payload >> 20) & 3), row = (int)((payload >> 10) & 0x3ff);
    int cx = (int)(payload & 0x3ff);
    if (frame != cur_frame || c != cur_plane) load_constants(frame, c);
    CtbEdges E;
    E.S = (1 << g.log2_ctb) >> g.sh;
    E.rstride = 2 * E.S + 16;
    E.rows = my_edges;
    E.cols = my_edges + (E.S / 4 + 1) * E.rstride;
    E.y0 = row * E.S;
    g.cy = row;
    const CtbRec* __restrict__ row_recs = recs + row * ctbs_w;
    const uint8_t* __restrict__ row_nb = nbinfo + row * ctbs_w;
    // record + neighbour mask of the CTB about to be processed (requested one CTB ahead inside the loop)
    uint2 rr = __ldg(reinterpret_cast<const uint2*>(row_recs + cx));
    uint32_t nb = __ldg(row_nb + cx);
    unsigned long long pre = prefetch_upper_right(mbox, E.S, ctbs_w, cx, row);
    bool chained = false;      // the previous CTB of the row was processed by this warp just now (edge store still holds it)
    uint32_t nb_prev = 0;
    bool yielded = false;
    while (cx < ctbs_w) {
      const uint32_t has = (rr.y >> 16) & 0xff;
      // next CTB's record, neighbour mask and upper-right mailbox words travel while this one is processed
      uint2 rr_next = make_uint2(0, 0);
      uint32_t nb_next = 0;
      unsigned long long pre_next = 0;
      if (cx + 1 < ctbs_w) {
        rr_next = __ldg(reinterpret_cast<const uint2*>(row_recs + cx + 1));
        nb_next = __ldg(row_nb + cx + 1);
        pre_next = prefetch_upper_right(mbox, E.S, ctbs_w, cx + 1, row);
      }
      if ((has >> c) & 1) {
        long long tr0 = 0, tr1 = 0, tr2 = 0, tr3 = 0;
        unsigned long long gt0 = 0;
        if (TRACE) {
          tr0 = clock64();
          asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt0));
        }
        const uint32_t tu_first = rr.x, tu_count = rr.y & 0xffff;
        const bool mixed = ((rr.y >> 24) >> c) & 1;
        // prefetch the first 32 TU records of the CTB while the borders are fetched
        uint4 mine = make_uint4(0, 0xff000000u, 0, 0);
        uint2 mymask = make_uint2(0, 0);
        if ((uint32_t)lane < tu_count) {
          mine = load_tu_xyw(tus4 + tu_first + lane);
          mymask = __ldg(masks + tu_first + lane);
        }
        g.cx = cx;
        g.nb = nb;
        E.x0 = cx * E.S;
        // the upper row piece and the corner in the edge store were fetched for the previous CTB only if they were usable for it
        const bool reuse = chained && (!((nb >> 2) & 1) || ((nb_prev >> 3) & 1)) && (!((nb >> 1) & 1) || ((nb_prev >> 2) & 1));
        if (reuse) shift_edges(E);
        if (TRACE) tr1 = clock64();
        if (!fetch_borders(g, E, recs, c, mbox, ctbs_w, mixed, reuse, pre, sched, sleep_ns, yield_polls)) {
          // blocked while other rows wait: give the row back (its left neighbour's column is in the mailbox)
          if (lane == 0) queue_put(queue, qmask, atomicAdd(&sched->tail, 1u), row_payload(frame, c, row, cx));
          yielded = true;
          break;
        }
        if (TRACE) tr2 = clock64();
        int n_tbs = 0;
        long long acc[4] = {0, 0, 0, 0};
        for (uint32_t base = 0; base < tu_count; base += 32) {
          if (base) {
            mine = make_uint4(0, 0xff000000u, 0, 0);
            if (base + lane < tu_count) {
              mine = load_tu_xyw(tus4 + tu_first + base + lane);
              mymask = __ldg(masks + tu_first + base + lane);
            }
          }
          const int n_here = min(32u, tu_count - base);
          // lanes whose TU is an intra TB of this plane
          const bool want = lane < n_here && (mine.y >> 24) != B265_INTRA_NONE && (int)(mine.y & 3) == c;
          unsigned todo = __ballot_sync(B265_FULL_MASK, want);
          while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const uint32_t r0 = __shfl_sync(B265_FULL_MASK, mine.x, src), r1 = __shfl_sync(B265_FULL_MASK, mine.y, src);
            const uint32_t r3 = __shfl_sync(B265_FULL_MASK, mine.w, src);
            const int x = (int)(r0 & 0xffff), y = (int)(r0 >> 16);
            const int flags = (int)((r1 >> 8) & 0xff), mode = (int)(r1 >> 24);
            const bool coded = (r3 & 0xffff) != 0;
            const uint32_t mhi = __shfl_sync(B265_FULL_MASK, mymask.y, src);
            const uint64_t M = (uint64_t)__shfl_sync(B265_FULL_MASK, mymask.x, src) | ((uint64_t)(mhi & 0xffu) << 32);
            const int16_t* rblk = g.resid + ((size_t)(mhi >> 8) << 4);   // the TB's residual block (k_intra_prep)
            if (TRACE) n_tbs++;
            // two forms of the same arithmetic, bit-identical; which sizes take which: INTRA_SEG_SIZES above
            {
              const int lg = (int)((r1 >> 2) & 7);   // SEGMAX is a mask of the log2 sizes that take the segment code
              if ((SEGMAX >> lg) & 1) intra_tb_seg<TRACE>(E, g, c, x, y, lg, mode, flags, coded, M, rblk, s_border[warp], acc);
              else if (!((SEGMAX >> 5) & 1) && lg == 5) intra_tb<5, TRACE>(E, g, c, x, y, mode, flags, coded, M, rblk, s_border[warp], acc);
              else if (!((SEGMAX >> 4) & 1) && lg == 4) intra_tb<4, TRACE>(E, g, c, x, y, mode, flags, coded, M, rblk, s_border[warp], acc);
              else if (!((SEGMAX >> 3) & 1) && lg == 3) intra_tb<3, TRACE>(E, g, c, x, y, mode, flags, coded, M, rblk, s_border[warp], acc);
              else if (!((SEGMAX >> 2) & 1)) intra_tb4<TRACE>(E, g, c, x, y, mode, flags, coded, (uint32_t)M, rblk, acc);
            }
          }
        }
        if (TRACE) tr3 = clock64();
        // post the borders the row below (and a warp that resumes this row) reads; the picture itself is complete at the end of the launch
        post_borders(E, mbox, row * ctbs_w + cx);
        if (TRACE && lane == 0) {
          unsigned smid;
          asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
          unsigned long long* o = trace + (size_t)(slot_base + row * ctbs_w + cx) * 12;
          o[0] = tr0; o[1] = tr1; o[2] = tr2; o[3] = tr3; o[4] = clock64(); o[5] = n_tbs; o[6] = acc[0]; o[7] = acc[1];
          o[8] = acc[2]; o[9] = acc[3]; o[10] = gt0; o[11] = ((unsigned long long)(reuse ? 1 : (chained ? 2 : 0)) << 32) | smid;
        }
        chained = true;
        nb_prev = nb;
      } else {
        break;   // a CTB without intra TBs ends the run: the next intra CTB of the row starts a run of its own
      }
      cx++;
      rr = rr_next; nb = nb_next; pre = pre_next;
    }
    if (yielded) continue;
    // ---- run finished ----
    unsigned old = 0;
    if (lane == 0) old = atomicAdd(&sched->done, 1u);
    old = __shfl_sync(B265_FULL_MASK, old, 0);
    if (old == n_runs - 1) {
      // every run is finished: at most total_warps warps sleep or are about to, on consecutive indices >= tail
      const unsigned tail = ld_relaxed_u32(&sched->tail);
      for (unsigned i = lane; i < total_warps; i += 32) queue_put(queue, qmask, tail + i, B265_QUEUE_EXIT);
      return;
    }
  }
}

static int tile_bytes_for(int log2_ctb) { return (ctb_edges_bytes(1 << log2_ctb) + 15) & ~15; }

int intra_max_resident_ctas(int sm_count, int log2_ctb) {
  int per_sm = 0;
  const size_t dyn = (size_t)INTRA_WARPS * tile_bytes_for(log2_ctb);
  // the opt-in limit is a property of the function, not of a launch: always allow the largest CTB size
  const size_t dyn_max = (size_t)INTRA_WARPS * tile_bytes_for(6);
  cudaFuncSetAttribute(k_intra_wavefront<false, INTRA_SEG_SIZES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn_max);
  cudaFuncSetAttribute(k_intra_wavefront<true, INTRA_SEG_SIZES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn_max);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_intra_wavefront<false, INTRA_SEG_SIZES>, INTRA_WARPS * 32, dyn);
  if (per_sm < 1) per_sm = 1;
  return per_sm * sm_count;
}

// queue entries beyond the rows themselves: one exit marker per warp that can be resident
int intra_queue_slack(int sm_count) { return intra_max_resident_ctas(sm_count, 4) * INTRA_WARPS + 64; }

void launch_intra(cudaStream_t st, const FrameDev* frames, int n_frames, int max_slots, int max_rows, int n_runs, IntraSched* sched,
                  unsigned long long* queue, unsigned queue_entries, int sm_count, int log2_ctb, int busy_peers) {
  if (n_runs <= 0) return;
  static int max_ctas[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (!max_ctas[log2_ctb & 7]) max_ctas[log2_ctb & 7] = intra_max_resident_ctas(sm_count, log2_ctb);
  const unsigned qmask = queue_entries - 1;
  {
    int gx = (max_slots + 255) / 256;
    if (gx > 4 * sm_count) gx = 4 * sm_count;
    k_intra_sched_prep<<<dim3(gx, n_frames), 256, 0, st>>>(frames, sched, queue, qmask, max_rows);
  }
  // Sleeping warps cost next to nothing, but a grid that fills every CTA slot keeps a second batch (another context of the
  // same process) off the SMs until it drains.  Batches that run side by side fill each other's ramps and tails and let
  // the residual / deblocking / SAO stages of one use the issue slots the latency-bound wavefront of the other leaves idle
  // (4K, batches of 32: 3.83 ms per batch alone, 2.78 ms with four contexts at one CTA per SM each), so a launch takes
  // INTRA_MIN_CTAS / (1 + busy peer contexts) CTAs per SM.  B265_INTRA_CTAS_PER_SM overrides.
  static int cap_per_sm = -1, sleep_ns = -1, yield_polls = -1;
  if (cap_per_sm < 0) {
    const char* e = getenv("B265_INTRA_CTAS_PER_SM");
    cap_per_sm = e ? atoi(e) : 0;
    const char* s = getenv("B265_INTRA_SLEEP_NS");
    sleep_ns = s ? atoi(s) : 100;
    // measured on B200 (4K, batches of 1..32): giving blocked rows back to the queue costs more (a resumed row pays a full
    // border fetch and its warp's instruction stream restarts) than it gains, so the default is never; the mechanism stays
    // for workloads with many more rows than warps
    const char* y = getenv("B265_INTRA_YIELD_POLLS");
    yield_polls = y ? atoi(y) : 0x7fffffff;
    if (yield_polls < 1) yield_polls = 1;
  }
  // one warp per run is all a launch can use
  int ctas = (n_runs + INTRA_WARPS - 1) / INTRA_WARPS;
  if (ctas > max_ctas[log2_ctb & 7]) ctas = max_ctas[log2_ctb & 7];
  const int share = cap_per_sm > 0 ? cap_per_sm : std::max(1, INTRA_MIN_CTAS / (1 + std::max(0, busy_peers)));
  if (ctas > share * sm_count) ctas = share * sm_count;
  const unsigned total_warps = (unsigned)ctas * INTRA_WARPS;
  const int tb = tile_bytes_for(log2_ctb);
  static const char* trace_path = getenv("B265_INTRA_TRACE");
  if (trace_path) {  // developer aid: synchronous, one file per process (last launch wins); one record per (plane, CTB) slot
    unsigned long long* d = nullptr;
    size_t n_slots = 0;
    {
      std::vector<FrameDev> h(n_frames);
      cudaStreamSynchronize(st);
      cudaMemcpy(h.data(), frames, sizeof(FrameDev) * n_frames, cudaMemcpyDeviceToHost);
      for (const FrameDev& f : h) n_slots = std::max(n_slots, (size_t)f.slot_base + (size_t)3 * f.ctbs_w * f.ctbs_h);
    }
    cudaMalloc((void**)&d, n_slots * 96);
    cudaMemsetAsync(d, 0, n_slots * 96, st);
    k_intra_wavefront<true, INTRA_SEG_SIZES><<<ctas, INTRA_WARPS * 32, (size_t)INTRA_WARPS * tb, st>>>(frames, sched, queue, qmask, total_warps,
                                                                                     (unsigned)sleep_ns, (unsigned)yield_polls, tb, d);
    cudaStreamSynchronize(st);
    unsigned long long* h = (unsigned long long*)malloc(n_slots * 96);
    cudaMemcpy(h, d, n_slots * 96, cudaMemcpyDeviceToHost);
    if (FILE* fp = fopen(trace_path, "wb")) { fwrite(h, 96, n_slots, fp); fclose(fp); }
    free(h);
    cudaFree(d);
    return;
  }
  k_intra_wavefront<false, INTRA_SEG_SIZES><<<ctas, INTRA_WARPS * 32, (size_t)INTRA_WARPS * tb, st>>>(
      frames, sched, queue, qmask, total_warps, (unsigned)sleep_ns, (unsigned)yield_polls, tb, nullptr);
}
