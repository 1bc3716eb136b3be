// k_mc.cu — motion compensation: luma 8-tap qpel + chroma 4-tap epel interpolation fused with the
// four prediction writers (unweighted, average, weighted uni, weighted bi).
//
// Restates generate_inter_prediction_samples (libde265/motion.cc:279-661), mc_luma/mc_chroma
// (:49-273, picture-edge clamp = coordinate clamp) and put_qpel/put_epel_hv/put_*_pred fallbacks
// (libde265/fallback-motion.cc:33-158, 298-407, 482-625).
//
// One formula covers every fractional phase at bit depth 8: tmp = sum_k tx[k]*px (no shift),
// out14 = (sum_k ty[k]*tmp) >> 6, with the phase-0 filter written as {..,64,..}.  This equals the
// reference's case split exactly (a multiple of 64 shifted right by 6 is exact), including its
// "no shift after the first stage when xFrac==0" rule.
//
// Work mapping: one warp per 16x16 luma tile (+ the two 8x8 chroma tiles) of a PU. The reference
// tile with its 3/4-sample halo is staged in shared memory, filtered horizontally into an int16
// tile, then each lane produces 8 horizontally adjacent luma samples (one 8-byte store) and 4
// chroma samples.
#include "transform.cuh"
#include "kernels.h"

__constant__ int8_t c_qpel[4][8] = {{0, 0, 0, 64, 0, 0, 0, 0},
                                    {-1, 4, -10, 58, 17, -5, 1, 0},
                                    {-1, 4, -11, 40, 40, -11, 4, -1},
                                    {0, 1, -5, 17, 58, -10, 4, -1}};
__constant__ int8_t c_epel[8][4] = {{0, 64, 0, 0},   {-2, 58, 10, -2}, {-4, 54, 16, -2}, {-6, 46, 28, -4},
                                    {-4, 36, 36, -4}, {-4, 28, 46, -6}, {-2, 16, 54, -4}, {-2, 10, 58, -2}};

struct __align__(16) McSmem {
  __align__(16) int16_t tmp[23][16];  // after the horizontal pass; chroma: [2][11][8] (read as uint4/uint2)
  uint8_t ref[23][24];                // luma reference tile (16+7)^2; reused for chroma: two 11x12 tiles
};

// luma: 14-bit prediction of 8 samples (row = lane/2, x = 8*(lane&1)..+7) of the 16x16 tile at (tx,ty)
__device__ __forceinline__ void luma_tile(McSmem& s, const uint8_t* __restrict__ ref, int pitch, int W, int H,
                                          int x0, int y0, int fx, int fy, int (&out)[8]) {
  const int lane = lane_id();
  // stage (x0-3 .. x0+19) x (y0-3 .. y0+19), coordinates clamped to the picture
  if (lane < 23) {
    const int cx = clip3(0, W - 1, x0 - 3 + lane);
#pragma unroll 1
    for (int r = 0; r < 23; r++) {
      const int cy = clip3(0, H - 1, y0 - 3 + r);
      s.ref[r][lane] = __ldg(ref + (size_t)cy * pitch + cx);
    }
  }
  __syncwarp();
  int tx[8], ty[8];
#pragma unroll
  for (int k = 0; k < 8; k++) { tx[k] = c_qpel[fx][k]; ty[k] = c_qpel[fy][k]; }
  // horizontal: 23 rows x 16 columns
#pragma unroll 1
  for (int i = lane; i < 23 * 16; i += 32) {
    const int r = i >> 4, x = i & 15;
    int acc = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) acc += tx[k] * s.ref[r][x + k];
    s.tmp[r][x] = (int16_t)acc;
  }
  __syncwarp();
  const int row = lane >> 1, xb = (lane & 1) * 8;
#pragma unroll
  for (int j = 0; j < 8; j++) out[j] = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) {
    const uint4 w = *reinterpret_cast<const uint4*>(&s.tmp[row + k][xb]);
    const int v[8] = {(int16_t)(w.x & 0xffff), (int16_t)(w.x >> 16), (int16_t)(w.y & 0xffff), (int16_t)(w.y >> 16),
                      (int16_t)(w.z & 0xffff), (int16_t)(w.z >> 16), (int16_t)(w.w & 0xffff), (int16_t)(w.w >> 16)};
#pragma unroll
    for (int j = 0; j < 8; j++) out[j] += ty[k] * v[j];
  }
#pragma unroll
  for (int j = 0; j < 8; j++) out[j] >>= 6;
  __syncwarp();
}

// chroma: both planes at once. lane -> plane = lane>>4, row = (lane&15)>>1, x = 4*(lane&1)..+3 of the 8x8 tile
__device__ __forceinline__ void chroma_tile(McSmem& s, const uint8_t* __restrict__ cb, const uint8_t* __restrict__ cr,
                                            int pitch, int W, int H, int x0, int y0, int fx, int fy, int (&out)[4]) {
  const int lane = lane_id();
  uint8_t (*ref)[11][12] = reinterpret_cast<uint8_t (*)[11][12]>(&s.ref[0][0]);
  int16_t (*tmp)[11][8] = reinterpret_cast<int16_t (*)[11][8]>(&s.tmp[0][0]);
  {
    const int p = lane >> 4, c = lane & 15;
    if (c < 11) {
      const uint8_t* src = p ? cr : cb;
      const int cx = clip3(0, W - 1, x0 - 1 + c);
#pragma unroll 1
      for (int r = 0; r < 11; r++) {
        const int cy = clip3(0, H - 1, y0 - 1 + r);
        ref[p][r][c] = __ldg(src + (size_t)cy * pitch + cx);
      }
    }
  }
  __syncwarp();
  int tx[4], ty[4];
#pragma unroll
  for (int k = 0; k < 4; k++) { tx[k] = c_epel[fx][k]; ty[k] = c_epel[fy][k]; }
#pragma unroll 1
  for (int i = lane; i < 2 * 11 * 8; i += 32) {
    const int p = i / 88, rem = i - p * 88, r = rem >> 3, x = rem & 7;
    int acc = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) acc += tx[k] * ref[p][r][x + k];
    tmp[p][r][x] = (int16_t)acc;
  }
  __syncwarp();
  const int p = lane >> 4, row = (lane & 15) >> 1, xb = (lane & 1) * 4;
#pragma unroll
  for (int j = 0; j < 4; j++) out[j] = 0;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const uint2 w = *reinterpret_cast<const uint2*>(&tmp[p][row + k][xb]);
    const int v[4] = {(int16_t)(w.x & 0xffff), (int16_t)(w.x >> 16), (int16_t)(w.y & 0xffff), (int16_t)(w.y >> 16)};
#pragma unroll
    for (int j = 0; j < 4; j++) out[j] += ty[k] * v[j];
  }
#pragma unroll
  for (int j = 0; j < 4; j++) out[j] >>= 6;
  __syncwarp();
}

// the four writers of fallback-motion.cc:33-158 on one 14-bit sample (pair)
struct WParams {
  int mode;  // 0 unweighted uni, 1 average, 2 weighted uni, 3 weighted bi
  int w0, w1, o0, o1, log2wd;
};
__device__ __forceinline__ int write_sample(const WParams& p, int a, int b) {
  switch (p.mode) {
    case 0: return clip8((a + 32) >> 6);
    case 1: return clip8((a + b + 64) >> 7);
    case 2: return clip8(((a * p.w0 + (1 << (p.log2wd - 1))) >> p.log2wd) + p.o0);
    default: return clip8((a * p.w0 + b * p.w1 + ((p.o0 + p.o1 + 1) << p.log2wd)) >> (p.log2wd + 1));
  }
}

__global__ void __launch_bounds__(B265_WARPS_PER_CTA * 32)
k_mc(const FrameDev* __restrict__ frames, const uint32_t* __restrict__ items, int n_items) {
  __shared__ McSmem smem[B265_WARPS_PER_CTA];
  McSmem& s = smem[threadIdx.x >> 5];
  const int lane = lane_id();
  const int warps_total = gridDim.x * B265_WARPS_PER_CTA;
  for (int it = blockIdx.x * B265_WARPS_PER_CTA + (threadIdx.x >> 5); it < n_items; it += warps_total) {
    const uint32_t o = items[it];
    const FrameDev& f = frames[o >> 24];
    const b265_pu pu = f.pus[o & 0xffffff];
    const b265_slice& sl = f.slices[pu.slice];
    int use0 = pu.flags & B265_PU_L0, use1 = (pu.flags & B265_PU_L1) ? 1 : 0;
    // identical-motion bi-prediction collapses to uni-prediction (motion.cc:326-335)
    if (use0 && use1 && !sl.pps_weighted_pred && pu.mv[0][0] == pu.mv[1][0] && pu.mv[0][1] == pu.mv[1][1] &&
        pu.ref_slot[0] == pu.ref_slot[1])
      use1 = 0;
    if (!use0 && !use1) continue;
    const int la = use0 ? 0 : 1;   // first (or only) list
    const bool bi = use0 && use1;
    WParams wl, wc[2];
    if (!sl.weighted) {
      wl.mode = bi ? 1 : 0;
      wl.w0 = wl.w1 = wl.o0 = wl.o1 = 0; wl.log2wd = 6;
      wc[0] = wc[1] = wl;
    } else {
      const int ra = pu.ref_idx[la] & 15, rb = pu.ref_idx[1] & 15;
      wl.mode = bi ? 3 : 2;
      wl.log2wd = sl.luma_log2_denom + 6;
      wl.w0 = sl.weight[la][ra][0]; wl.o0 = sl.offset[la][ra][0];
      wl.w1 = sl.weight[1][rb][0];  wl.o1 = sl.offset[1][rb][0];
      for (int c = 0; c < 2; c++) {
        wc[c].mode = wl.mode;
        wc[c].log2wd = sl.chroma_log2_denom + 6;
        wc[c].w0 = sl.weight[la][ra][1 + c]; wc[c].o0 = sl.offset[la][ra][1 + c];
        wc[c].w1 = sl.weight[1][rb][1 + c];  wc[c].o1 = sl.offset[1][rb][1 + c];
      }
    }
    const int W = f.width, H = f.height;
    for (int ty = 0; ty < pu.h; ty += 16)
      for (int tx = 0; tx < pu.w; tx += 16) {
        const int px = pu.x + tx, py = pu.y + ty;
        const int tw = min(16, pu.w - tx), th = min(16, pu.h - ty);
        // ---- luma ----
        int a[8], b[8];
        {
          const int mvx = pu.mv[la][0], mvy = pu.mv[la][1];
          luma_tile(s, f.ref_plane[pu.ref_slot[la]][0], f.pitch[0], W, H, px + (mvx >> 2), py + (mvy >> 2), mvx & 3,
                    mvy & 3, a);
        }
        if (bi) {
          const int mvx = pu.mv[1][0], mvy = pu.mv[1][1];
          luma_tile(s, f.ref_plane[pu.ref_slot[1]][0], f.pitch[0], W, H, px + (mvx >> 2), py + (mvy >> 2), mvx & 3,
                    mvy & 3, b);
        } else {
#pragma unroll
          for (int j = 0; j < 8; j++) b[j] = 0;
        }
        {
          const int row = lane >> 1, xb = (lane & 1) * 8;
          if (row < th && xb < tw) {
            int v[8];
#pragma unroll
            for (int j = 0; j < 8; j++) v[j] = write_sample(wl, a[j], b[j]);
            uint8_t* d = f.plane[0] + (size_t)(py + row) * f.pitch[0] + px + xb;
            // PB x positions are multiples of 4 only (AMP), so 8-byte alignment is checked at run time
            const uint32_t lo = pack4(v[0], v[1], v[2], v[3]), hi = pack4(v[4], v[5], v[6], v[7]);
            if (xb + 8 <= tw) {
              if ((reinterpret_cast<uintptr_t>(d) & 7) == 0) *reinterpret_cast<uint2*>(d) = make_uint2(lo, hi);
              else { reinterpret_cast<uint32_t*>(d)[0] = lo; reinterpret_cast<uint32_t*>(d)[1] = hi; }
            } else *reinterpret_cast<uint32_t*>(d) = lo;
          }
        }
        // ---- chroma (4:2:0) ----
        if (f.cw > 0) {
          int ca[4], cb4[4];
          const int cpx = px >> 1, cpy = py >> 1;
          {
            const int mvx = pu.mv[la][0], mvy = pu.mv[la][1];
            chroma_tile(s, f.ref_plane[pu.ref_slot[la]][1], f.ref_plane[pu.ref_slot[la]][2], f.pitch[1], f.cw, f.ch,
                        cpx + (mvx >> 3), cpy + (mvy >> 3), mvx & 7, mvy & 7, ca);
          }
          if (bi) {
            const int mvx = pu.mv[1][0], mvy = pu.mv[1][1];
            chroma_tile(s, f.ref_plane[pu.ref_slot[1]][1], f.ref_plane[pu.ref_slot[1]][2], f.pitch[1], f.cw, f.ch,
                        cpx + (mvx >> 3), cpy + (mvy >> 3), mvx & 7, mvy & 7, cb4);
          } else {
#pragma unroll
            for (int j = 0; j < 4; j++) cb4[j] = 0;
          }
          const int p = lane >> 4, row = (lane & 15) >> 1, xb = (lane & 1) * 4;
          const int cw = tw >> 1, chh = th >> 1;
          if (row < chh && xb < cw) {
            int v[4];
#pragma unroll
            for (int j = 0; j < 4; j++) v[j] = write_sample(wc[p], ca[j], cb4[j]);
            uint8_t* d = f.plane[1 + p] + (size_t)(cpy + row) * f.pitch[1 + p] + cpx + xb;
            const uint16_t lo = (uint16_t)(v[0] | (v[1] << 8)), hi = (uint16_t)(v[2] | (v[3] << 8));
            if (xb + 4 <= cw) {
              if ((reinterpret_cast<uintptr_t>(d) & 3) == 0) *reinterpret_cast<uint32_t*>(d) = lo | ((uint32_t)hi << 16);
              else { reinterpret_cast<uint16_t*>(d)[0] = lo; reinterpret_cast<uint16_t*>(d)[1] = hi; }
            } else *reinterpret_cast<uint16_t*>(d) = lo;
          }
        }
      }
  }
}

void launch_mc(cudaStream_t st, const FrameDev* frames, const uint32_t* items, int n, int sm_count) {
  if (n <= 0) return;
  int ctas = (n + B265_WARPS_PER_CTA - 1) / B265_WARPS_PER_CTA;
  int grid = min(ctas, sm_count * 8);
  k_mc<<<grid, B265_WARPS_PER_CTA * 32, 0, st>>>(frames, items, n);
}
