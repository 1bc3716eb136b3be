// k_deblock.cu — in-loop deblocking filter: boundary strength (derived inside the edge kernels), luma and chroma edge filters.
//
// Restates libde265/deblock.cc: derive_boundaryStrength (:241-375), edge_filtering_luma_internal
// (:405-699), edge_filtering_chroma_internal (:730-871), tables :389-399.  The edge flags themselves
// (derive_edgeFlags, :130-237) are produced by the host parser while it walks the CB/TU/PU tree and
// arrive in FrameDev::unit_info.
//
// All vertical edges of the picture are filtered before any horizontal edge (deblock.cc:1031-1050):
// here that is two launches.  Edges lie on an 8-sample grid and touch at most 4 samples each side
// (reads) / 3 (writes), so every edge segment of one direction is independent: one thread per
// 4-sample segment, lanes running along the row for coalesced access.
// HBM traffic per picture: each pass reads+writes the plane once (2P per pass, 4P total) + maps.
#include "transform.cuh"
#include "kernels.h"

__constant__ uint8_t c_beta[52] = {0,  0,  0,  0,  0,  0,  0,  0,  0,  0,  0,  0,  0,  0,  0,  0,  6,  7,
                                   8,  9,  10, 11, 12, 13, 14, 15, 16, 17, 18, 20, 22, 24, 26, 28, 30, 32,
                                   34, 36, 38, 40, 42, 44, 46, 48, 50, 52, 54, 56, 58, 60, 62, 64};
__constant__ uint8_t c_tc[54] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0,  0,  0,  0,  0,
                                 1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 3,  3,  3,  3,  4,
                                 4, 4, 5, 5, 6, 6, 7, 8, 9, 10, 11, 13, 14, 16, 18, 20, 22, 24};
__constant__ uint8_t c_qpc[13] = {29, 30, 31, 32, 33, 33, 34, 34, 35, 35, 36, 36, 37};

// ---- PU list -> per-unit motion map (image.h pb_info equivalent, device resident) ----
// entry: x = predFlags | slot0 << 8 | slot1 << 16, y = mv0 (x | y << 16), z = mv1
// eight lanes per piece (a piece is at most 16 x 8 samples = 4 x 2 units, backend.h pu_expand)
__global__ void k_motion_raster(const FrameDev* __restrict__ frames, const uint32_t* __restrict__ items, int n) {
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 3;
  if (w >= n) return;
  const uint32_t o = items[w];
  const FrameDev& f = frames[o >> 24];
  if (!f.motion) return;
  const b265_pu pu = f.pus[o & 0xffffff];
  const int uw = pu.w >> 2, uh = pu.h >> 2, pitch = (f.width + 3) >> 2;
  uint4 e;
  e.x = (pu.flags & 3) | ((uint32_t)pu.ref_slot[0] << 8) | ((uint32_t)pu.ref_slot[1] << 16);
  e.y = (uint16_t)pu.mv[0][0] | ((uint32_t)(uint16_t)pu.mv[0][1] << 16);
  e.z = (uint16_t)pu.mv[1][0] | ((uint32_t)(uint16_t)pu.mv[1][1] << 16);
  e.w = 0;
  for (int i = threadIdx.x & 7; i < uw * uh; i += 8)
    reinterpret_cast<uint4*>(f.motion)[((pu.y >> 2) + i / uw) * pitch + (pu.x >> 2) + i % uw] = e;
}

__device__ __forceinline__ int mvx(uint32_t v) { return (int16_t)(v & 0xffff); }
__device__ __forceinline__ int mvy(uint32_t v) { return (int16_t)(v >> 16); }
__device__ __forceinline__ bool far4(uint32_t a, uint32_t b) {
  return abs(mvx(a) - mvx(b)) >= 4 || abs(mvy(a) - mvy(b)) >= 4;
}

// bS of the edge between unit P (index up) and unit Q (index uq)  (deblock.cc:283-357)
__device__ __forceinline__ int edge_bs(const FrameDev& f, int up, int uq, bool tu_edge) {
  const uint8_t ip = f.unit_info[up], iq = f.unit_info[uq];
  if ((ip | iq) & B265_UNIT_INTRA) return 2;
  if (tu_edge && ((ip | iq) & B265_UNIT_CBF)) return 1;
  if (!f.motion) return 0;
  const uint4 P = reinterpret_cast<const uint4*>(f.motion)[up], Q = reinterpret_cast<const uint4*>(f.motion)[uq];
  const int rp0 = (P.x & 1) ? (int)((P.x >> 8) & 0xff) : -1, rp1 = (P.x & 2) ? (int)((P.x >> 16) & 0xff) : -1;
  const int rq0 = (Q.x & 1) ? (int)((Q.x >> 8) & 0xff) : -1, rq1 = (Q.x & 2) ? (int)((Q.x >> 16) & 0xff) : -1;
  if (!((rp0 == rq0 && rp1 == rq1) || (rp0 == rq1 && rp1 == rq0))) return 1;
  const uint32_t p0 = (P.x & 1) ? P.y : 0, p1 = (P.x & 2) ? P.z : 0, q0 = (Q.x & 1) ? Q.y : 0, q1 = (Q.x & 2) ? Q.z : 0;
  if (rp0 != rp1) {
    if (rp0 == rq0) return (far4(p0, q0) || far4(p1, q1)) ? 1 : 0;
    return (far4(p0, q1) || far4(p1, q0)) ? 1 : 0;
  }
  return ((far4(p0, q0) || far4(p1, q1)) && (far4(p0, q1) || far4(p1, q0))) ? 1 : 0;
}

__device__ __forceinline__ int qp_at(const FrameDev& f, int x, int y) {
  return f.qp_y[(y >> f.log2_min_cb) * (f.width >> f.log2_min_cb) + (x >> f.log2_min_cb)];
}
__device__ __forceinline__ const b265_slice& slice_at(const FrameDev& f, int x, int y) {
  const int ctb = (y >> f.log2_ctb) * f.ctbs_w + (x >> f.log2_ctb);
  return f.slices[f.ctbs ? f.ctbs[ctb].slice : 0];
}

// one 4-sample luma edge segment with boundary strength bs > 0; p[k][i] / q[k][i]: line k, distance i from the edge
template <bool VERT>
__device__ __forceinline__ void deblock_luma_segment(const FrameDev& f, int ux, int uy, int bs) {
  const int uw = f.width >> 2;
  const int uq = uy * uw + ux;
  const uint8_t info = f.unit_info[uq];
  const int x = ux << 2, y = uy << 2, pitch = f.pitch[0];
  uint8_t* ptr = f.plane[0] + (size_t)y * pitch + x;
  int p[4][4], q[4][4];
  if (VERT) {
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const uint32_t a = *reinterpret_cast<const uint32_t*>(ptr + k * pitch - 4);
      const uint32_t b = *reinterpret_cast<const uint32_t*>(ptr + k * pitch);
#pragma unroll
      for (int j = 0; j < 4; j++) { p[k][j] = (a >> (8 * (3 - j))) & 0xff; q[k][j] = (b >> (8 * j)) & 0xff; }
    }
  } else {
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const uint32_t a = *reinterpret_cast<const uint32_t*>(ptr - (j + 1) * pitch);
      const uint32_t b = *reinterpret_cast<const uint32_t*>(ptr + j * pitch);
#pragma unroll
      for (int k = 0; k < 4; k++) { p[k][j] = (a >> (8 * k)) & 0xff; q[k][j] = (b >> (8 * k)) & 0xff; }
    }
  }
  const int xp = VERT ? x - 1 : x, yp = VERT ? y : y - 1;
  const int qpl = (qp_at(f, x, y) + qp_at(f, xp, yp) + 1) >> 1;
  const b265_slice& sl = slice_at(f, x, y);
  const int beta = c_beta[clip3(0, 51, qpl + sl.beta_offset)];
  const int tc = c_tc[clip3(0, 53, qpl + 2 * (bs - 1) + sl.tc_offset)];
  const int dp0 = abs(p[0][2] - 2 * p[0][1] + p[0][0]), dp3 = abs(p[3][2] - 2 * p[3][1] + p[3][0]);
  const int dq0 = abs(q[0][2] - 2 * q[0][1] + q[0][0]), dq3 = abs(q[3][2] - 2 * q[3][1] + q[3][0]);
  const int dpq0 = dp0 + dq0, dpq3 = dp3 + dq3, dp = dp0 + dp3, dq = dq0 + dq3;
  if (dpq0 + dpq3 >= beta) return;
  const bool s0 = 2 * dpq0 < (beta >> 2) && abs(p[0][3] - p[0][0]) + abs(q[0][0] - q[0][3]) < (beta >> 3) &&
                  abs(p[0][0] - q[0][0]) < ((5 * tc + 1) >> 1);
  const bool s3 = 2 * dpq3 < (beta >> 2) && abs(p[3][3] - p[3][0]) + abs(q[3][0] - q[3][3]) < (beta >> 3) &&
                  abs(p[3][0] - q[3][0]) < ((5 * tc + 1) >> 1);
  const bool strong = s0 && s3;
  const bool dEp = dp < ((beta + (beta >> 1)) >> 3), dEq = dq < ((beta + (beta >> 1)) >> 3);
  const int up = VERT ? uq - 1 : uq - uw;
  const bool filterP = !(f.unit_info[up] & B265_UNIT_NOFILT), filterQ = !(info & B265_UNIT_NOFILT);
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const int p0 = p[k][0], p1 = p[k][1], p2 = p[k][2], p3 = p[k][3];
    const int q0 = q[k][0], q1 = q[k][1], q2 = q[k][2], q3 = q[k][3];
    if (strong) {
      if (filterP) {
        p[k][0] = clip3(p0 - 2 * tc, p0 + 2 * tc, (p2 + 2 * p1 + 2 * p0 + 2 * q0 + q1 + 4) >> 3);
        p[k][1] = clip3(p1 - 2 * tc, p1 + 2 * tc, (p2 + p1 + p0 + q0 + 2) >> 2);
        p[k][2] = clip3(p2 - 2 * tc, p2 + 2 * tc, (2 * p3 + 3 * p2 + p1 + p0 + q0 + 4) >> 3);
      }
      if (filterQ) {
        q[k][0] = clip3(q0 - 2 * tc, q0 + 2 * tc, (p1 + 2 * p0 + 2 * q0 + 2 * q1 + q2 + 4) >> 3);
        q[k][1] = clip3(q1 - 2 * tc, q1 + 2 * tc, (p0 + q0 + q1 + q2 + 2) >> 2);
        q[k][2] = clip3(q2 - 2 * tc, q2 + 2 * tc, (p0 + q0 + q1 + 3 * q2 + 2 * q3 + 4) >> 3);
      }
    } else {
      int delta = (9 * (q0 - p0) - 3 * (q1 - p1) + 8) >> 4;
      if (abs(delta) < tc * 10) {
        delta = clip3(-tc, tc, delta);
        if (filterP) p[k][0] = clip8(p0 + delta);
        if (filterQ) q[k][0] = clip8(q0 - delta);
        if (dEp && filterP) p[k][1] = clip8(p1 + clip3(-(tc >> 1), tc >> 1, (((p2 + p0 + 1) >> 1) - p1 + delta) >> 1));
        if (dEq && filterQ) q[k][1] = clip8(q1 + clip3(-(tc >> 1), tc >> 1, (((q2 + q0 + 1) >> 1) - q1 - delta) >> 1));
      }
    }
  }
  if (VERT) {
#pragma unroll
    for (int k = 0; k < 4; k++) {
      *reinterpret_cast<uint32_t*>(ptr + k * pitch - 4) = pack4(p[k][3], p[k][2], p[k][1], p[k][0]);
      *reinterpret_cast<uint32_t*>(ptr + k * pitch) = pack4(q[k][0], q[k][1], q[k][2], q[k][3]);
    }
  } else {
#pragma unroll
    for (int j = 0; j < 3; j++) {
      *reinterpret_cast<uint32_t*>(ptr - (j + 1) * pitch) = pack4(p[0][j], p[1][j], p[2][j], p[3][j]);
      *reinterpret_cast<uint32_t*>(ptr + j * pitch) = pack4(q[0][j], q[1][j], q[2][j], q[3][j]);
    }
  }
}

// Candidate segments are all positions of the 8-sample edge grid, but only those on a TU / PU edge with bS > 0 are
// filtered (one in four for 32x32 TUs), and a warp pays for the filter as soon as ONE lane needs it.  So a warp first
// tests DBK_ROUNDS x 32 candidates (unit flag + boundary strength, derived here from the unit flags: no bS map round
// trip through memory), compacts the hits into a shared-memory queue with ballot / popc, then filters them with
// all lanes busy.
#ifndef DBK_WARPS
#define DBK_WARPS 4
#endif
#ifndef DBK_ROUNDS
#define DBK_ROUNDS 4
#endif
#ifndef DBK_MIN_CTAS
#define DBK_MIN_CTAS 10   // 48 registers: 40 warps per SM (ncu: long-scoreboard bound at 28); measured 8 / 9 / 10: deblock stage of 128 4K pictures 1.42 / 1.37 / 1.35 ms, 1.48 at 72 registers
#endif
template <bool VERT>
__global__ void __launch_bounds__(DBK_WARPS * 32, DBK_MIN_CTAS) k_deblock_luma(const FrameDev* __restrict__ frames) {
  const FrameDev& f = frames[blockIdx.y];
  if (!(f.flags & B265_FRAME_DEBLOCK) || !f.unit_info) return;
  __shared__ uint32_t queue[DBK_WARPS][DBK_ROUNDS * 32];
  const int uw = f.width >> 2, uh = f.height >> 2;
  const int ew = VERT ? (uw + 1) >> 1 : uw, eh = VERT ? uh : (uh + 1) >> 1;
  const int warp = threadIdx.x >> 5, lane = lane_id();
  const int base = (blockIdx.x * DBK_WARPS + warp) * (DBK_ROUNDS * 32);
  if (base >= ew * eh) return;
  uint32_t* q = queue[warp];
  int n = 0;
#pragma unroll
  for (int r = 0; r < DBK_ROUNDS; r++) {
    const int i = base + r * 32 + lane;
    int bs = 0, ux = 0, uy = 0;
    if (i < ew * eh) {
      ux = VERT ? 2 * (i % ew) : i % ew;
      uy = VERT ? i / ew : 2 * (i / ew);
      if ((VERT ? ux : uy) != 0) {
        const int uq = uy * uw + ux;
        const uint8_t info = f.unit_info[uq];
        if (info & (VERT ? B265_UNIT_EDGE_V : B265_UNIT_EDGE_H))
          bs = edge_bs(f, VERT ? uq - 1 : uq - uw, uq, info & (VERT ? B265_UNIT_TU_EDGE_V : B265_UNIT_TU_EDGE_H));
      }
    }
    const unsigned hit = __ballot_sync(B265_FULL_MASK, bs != 0);
    if (bs) q[n + __popc(hit & ((1u << lane) - 1))] = (uint32_t)ux | ((uint32_t)uy << 14) | ((uint32_t)bs << 28);
    n += __popc(hit);
  }
  __syncwarp();
  for (int k = lane; k < n; k += 32) {
    const uint32_t e = q[k];
    deblock_luma_segment<VERT>(f, (int)(e & 0x3fff), (int)((e >> 14) & 0x3fff), (int)(e >> 28));
  }
}

// chroma (4:2:0): edges on an 8-chroma-sample grid, only bS == 2, one sample each side; the bS of a
// 4-row chroma segment is the one of the FIRST luma unit of the corresponding 8 luma rows
// (deblock.cc:741-763)
template <bool VERT>
__global__ void k_deblock_chroma(const FrameDev* __restrict__ frames) {
  const FrameDev& f = frames[blockIdx.y];
  if (!(f.flags & B265_FRAME_DEBLOCK) || !f.unit_info || !f.cw) return;
  const int uw = f.width >> 2, uh = f.height >> 2;
  // luma-unit steps: along the edge normal 4 units, along the edge 2 units
  const int ew = VERT ? (uw + 3) >> 2 : (uw + 1) >> 1, eh = VERT ? (uh + 1) >> 1 : (uh + 3) >> 2;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= ew * eh) return;
  const int ux = (VERT ? 4 : 2) * (i % ew), uy = (VERT ? 2 : 4) * (i / ew);
  // chroma is filtered where bS == 2 only: a flagged edge with an intra unit on either side (deblock.cc:283-290)
  if ((VERT ? ux : uy) == 0) return;
  const int uq = uy * uw + ux, up = VERT ? uq - 1 : uq - uw;
  const uint8_t iq = f.unit_info[uq], ip = f.unit_info[up];
  if (!(iq & (VERT ? B265_UNIT_EDGE_V : B265_UNIT_EDGE_H)) || !((ip | iq) & B265_UNIT_INTRA)) return;
  const int bs = 2;
  const int xl = ux << 2, yl = uy << 2;          // luma position
  const int xc = xl >> 1, yc = yl >> 1;          // chroma position
  const int qpi = (qp_at(f, xl, yl) + qp_at(f, VERT ? xl - 1 : xl, VERT ? yl : yl - 1) + 1) >> 1;
  const int tc_offset = slice_at(f, xl, yl).tc_offset;
  const bool filterP = !(ip & B265_UNIT_NOFILT), filterQ = !(iq & B265_UNIT_NOFILT);
  // rows/columns of the segment that lie inside the plane (picture sizes are multiples of 8 luma)
  for (int c = 1; c <= 2; c++) {
    const int qi = qpi + (c == 1 ? f.cb_qp_offset : f.cr_qp_offset);
    const int qpc = qi < 30 ? qi : (qi >= 43 ? qi - 6 : c_qpc[qi - 30]);
    const int tc = c_tc[clip3(0, 53, qpc + 2 * (bs - 1) + tc_offset)];
    const int pitch = f.pitch[c];
    uint8_t* ptr = f.plane[c] + (size_t)yc * pitch + xc;
    // word accesses (xc is a multiple of 8 for vertical, of 4 for horizontal edges): the kernel waits on its loads
    if (VERT) {
      uint32_t wp[4], wq[4];
#pragma unroll
      for (int k = 0; k < 4; k++) {
        wp[k] = *reinterpret_cast<const uint32_t*>(ptr + k * pitch - 4);
        wq[k] = *reinterpret_cast<const uint32_t*>(ptr + k * pitch);
      }
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const int p1 = (wp[k] >> 16) & 0xff, p0 = wp[k] >> 24, q0 = wq[k] & 0xff, q1 = (wq[k] >> 8) & 0xff;
        const int delta = clip3(-tc, tc, ((((q0 - p0) << 2) + p1 - q1 + 4) >> 3));
        if (filterP) ptr[k * pitch - 1] = (uint8_t)clip8(p0 + delta);
        if (filterQ) ptr[k * pitch] = (uint8_t)clip8(q0 - delta);
      }
    } else {
      const uint32_t P1 = *reinterpret_cast<const uint32_t*>(ptr - 2 * pitch), P0 = *reinterpret_cast<const uint32_t*>(ptr - pitch);
      const uint32_t Q0 = *reinterpret_cast<const uint32_t*>(ptr), Q1 = *reinterpret_cast<const uint32_t*>(ptr + pitch);
      int np[4], nq[4];
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const int p1 = (P1 >> (8 * k)) & 0xff, p0 = (P0 >> (8 * k)) & 0xff, q0 = (Q0 >> (8 * k)) & 0xff, q1 = (Q1 >> (8 * k)) & 0xff;
        const int delta = clip3(-tc, tc, ((((q0 - p0) << 2) + p1 - q1 + 4) >> 3));
        np[k] = clip8(p0 + delta);
        nq[k] = clip8(q0 - delta);
      }
      if (filterP) *reinterpret_cast<uint32_t*>(ptr - pitch) = pack4(np[0], np[1], np[2], np[3]);
      if (filterQ) *reinterpret_cast<uint32_t*>(ptr) = pack4(nq[0], nq[1], nq[2], nq[3]);
    }
  }
}

void launch_deblock(cudaStream_t st, const FrameDev* d_frames, int n_frames, int max_w, int max_h,
                    const uint32_t* mc_items, int n_mc, bool chroma) {
  const int uw = max_w >> 2, uh = max_h >> 2;
  if (n_mc > 0) k_motion_raster<<<(n_mc * 8 + 255) / 256, 256, 0, st>>>(d_frames, mc_items, n_mc);
  const int per_cta = DBK_WARPS * DBK_ROUNDS * 32;   // candidate segments per block
  dim3 gv((((uw + 1) >> 1) * uh + per_cta - 1) / per_cta, n_frames), gh((uw * ((uh + 1) >> 1) + per_cta - 1) / per_cta, n_frames);
  dim3 cv((((uw + 3) >> 2) * ((uh + 1) >> 1) + 127) / 128, n_frames), chh((((uw + 1) >> 1) * ((uh + 3) >> 2) + 127) / 128, n_frames);
  k_deblock_luma<true><<<gv, DBK_WARPS * 32, 0, st>>>(d_frames);
  if (chroma) k_deblock_chroma<true><<<cv, 128, 0, st>>>(d_frames);
  k_deblock_luma<false><<<gh, DBK_WARPS * 32, 0, st>>>(d_frames);
  if (chroma) k_deblock_chroma<false><<<chh, 128, 0, st>>>(d_frames);
}
