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
// Work mapping: one warp per PU, walking its 16x16 luma tiles (+ the two 8x8 chroma tiles).  Lane r holds
// reference ROW r of the tile (with the 3/4-sample halo) in registers, fetched with independent 32-bit loads
// (one memory round trip per tile); the horizontal filter runs on the packed bytes with dp4a (u8 x s8), the
// int16 rows go through shared memory once, and the vertical filter gives every lane 8 horizontally adjacent
// luma samples (one 8-byte store) / 4 chroma samples.
#include "transform.cuh"
#include "kernels.h"

// filter taps packed four to a word (byte k = tap k), phase 0 written as {..,64,..}
__constant__ int8_t c_qpel[4][8] = {{0, 0, 0, 64, 0, 0, 0, 0},
                                    {-1, 4, -10, 58, 17, -5, 1, 0},
                                    {-1, 4, -11, 40, 40, -11, 4, -1},
                                    {0, 1, -5, 17, 58, -10, 4, -1}};
__constant__ int8_t c_epel[8][4] = {{0, 64, 0, 0},   {-2, 58, 10, -2}, {-4, 54, 16, -2}, {-6, 46, 28, -4},
                                    {-4, 36, 36, -4}, {-4, 28, 46, -6}, {-2, 16, 54, -4}, {-2, 10, 58, -2}};

struct __align__(16) McSmem {
  __align__(16) int16_t tmp[23][16];  // after the horizontal pass; chroma: [2][11][8] (read as uint4/uint2)
};

// sum of four products, unsigned bytes of a times signed bytes of b
__device__ __forceinline__ int dp4a_us(uint32_t a, int b, int c) {
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

// NW words of reference row `y` starting at the 4-byte aligned column xa (all clamped to the picture)
template <int NW>
__device__ __forceinline__ void load_row(const uint8_t* __restrict__ ref, int pitch, int W, int H, int xa, int y, uint32_t (&w)[NW]) {
  const uint8_t* row = ref + (size_t)clip3(0, H - 1, y) * pitch;
  if (xa >= 0 && xa + 4 * NW <= W && (reinterpret_cast<uintptr_t>(row) & 3) == 0) {
#pragma unroll
    for (int i = 0; i < NW; i++) w[i] = __ldg(reinterpret_cast<const uint32_t*>(row + xa) + i);
  } else {   // the row crosses the left or right picture edge: per-sample coordinate clamp (mc_luma, motion.cc:146-149)
#pragma unroll
    for (int i = 0; i < NW; i++) {
      uint32_t v = 0;
#pragma unroll
      for (int k = 0; k < 4; k++) v |= (uint32_t)__ldg(row + clip3(0, W - 1, xa + 4 * i + k)) << (8 * k);
      w[i] = v;
    }
  }
}

// luma, step 1: lane r < th+7 fetches reference row r of the tile at (x0,y0) (7 words from the aligned column)
__device__ __forceinline__ void luma_fetch(const uint8_t* __restrict__ ref, int pitch, int W, int H, int x0, int y0, int th,
                                           uint32_t (&w)[7]) {
  const int lane = lane_id();
  const int xs = x0 - 3;
  if (lane < th + 7) load_row<7>(ref, pitch, W, H, xs - (xs & 3), y0 - 3 + lane, w);
}

// luma, step 2: 14-bit prediction of 8 samples (row = lane/2, x = 8*(lane&1)..+7) of the tw x th (<= 16x16) tile
__device__ __forceinline__ void luma_filter(McSmem& s, const uint32_t (&w)[7], int x0, int fx, int fy, int tw, int th, int (&out)[8]) {
  const int lane = lane_id();
  const int off = (x0 - 3) & 3;
  if (lane < th + 7) {
    uint32_t a[6];
#pragma unroll
    for (int i = 0; i < 6; i++) a[i] = __funnelshift_r(w[i], w[i + 1], 8 * off);
    const int tlo = reinterpret_cast<const int*>(c_qpel[fx])[0], thi = reinterpret_cast<const int*>(c_qpel[fx])[1];
    uint32_t packed[8];
#pragma unroll
    for (int x = 0; x < 16; x += 2) {
      int v[2];
#pragma unroll
      for (int j = 0; j < 2; j++) {
        const int q = (x + j) >> 2, m = (x + j) & 3;
        const uint32_t lo = m ? __funnelshift_r(a[q], a[q + 1], 8 * m) : a[q];
        const uint32_t hi = m ? __funnelshift_r(a[q + 1], a[q + 2 < 6 ? q + 2 : 5], 8 * m) : a[q + 1];
        v[j] = dp4a_us(hi, thi, dp4a_us(lo, tlo, 0));
      }
      packed[x >> 1] = __byte_perm((uint32_t)v[0], (uint32_t)v[1], 0x5410);
    }
    uint4* dst = reinterpret_cast<uint4*>(&s.tmp[lane][0]);
    dst[0] = make_uint4(packed[0], packed[1], packed[2], packed[3]);
    if (tw > 8) dst[1] = make_uint4(packed[4], packed[5], packed[6], packed[7]);
  }
  __syncwarp();
  int ty[8];
#pragma unroll
  for (int k = 0; k < 8; k++) ty[k] = c_qpel[fy][k];
  const int row = lane >> 1, xb = (lane & 1) * 8;
#pragma unroll
  for (int j = 0; j < 8; j++) out[j] = 0;
  if (row < th && xb < tw) {
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const uint4 w4 = *reinterpret_cast<const uint4*>(&s.tmp[row + k][xb]);
      const int v[8] = {(int16_t)(w4.x & 0xffff), (int)w4.x >> 16, (int16_t)(w4.y & 0xffff), (int)w4.y >> 16,
                        (int16_t)(w4.z & 0xffff), (int)w4.z >> 16, (int16_t)(w4.w & 0xffff), (int)w4.w >> 16};
#pragma unroll
      for (int j = 0; j < 8; j++) out[j] += ty[k] * v[j];
    }
#pragma unroll
    for (int j = 0; j < 8; j++) out[j] >>= 6;
  }
  __syncwarp();
}

// chroma, both planes at once (tile cw x chh <= 8x8): lane -> row lane&15 of plane lane>>4.  Step 1: fetch.
__device__ __forceinline__ void chroma_fetch(const uint8_t* __restrict__ cb, const uint8_t* __restrict__ cr, int pitch, int W, int H,
                                             int x0, int y0, int chh, uint32_t (&w)[4]) {
  const int lane = lane_id();
  const int xs = x0 - 1;
  if ((lane & 15) < chh + 3) load_row<4>(lane >> 4 ? cr : cb, pitch, W, H, xs - (xs & 3), y0 - 1 + (lane & 15), w);
}

// step 2. Result: lane -> plane = lane>>4, row = (lane&15)>>1, x = 4*(lane&1)..+3
__device__ __forceinline__ void chroma_filter(McSmem& s, const uint32_t (&w)[4], int x0, int fx, int fy, int cw, int chh, int (&out)[4]) {
  const int lane = lane_id();
  int16_t (*tmp)[11][8] = reinterpret_cast<int16_t (*)[11][8]>(&s.tmp[0][0]);
  const int p = lane >> 4, r = lane & 15;
  const int off = (x0 - 1) & 3;
  if (r < chh + 3) {
    uint32_t a[3];
#pragma unroll
    for (int i = 0; i < 3; i++) a[i] = __funnelshift_r(w[i], w[i + 1], 8 * off);
    const int t = reinterpret_cast<const int*>(c_epel[fx])[0];
    uint32_t packed[4];
#pragma unroll
    for (int x = 0; x < 8; x += 2) {
      int v[2];
#pragma unroll
      for (int j = 0; j < 2; j++) {
        const int q = (x + j) >> 2, m = (x + j) & 3;
        v[j] = dp4a_us(m ? __funnelshift_r(a[q], a[q + 1], 8 * m) : a[q], t, 0);
      }
      packed[x >> 1] = __byte_perm((uint32_t)v[0], (uint32_t)v[1], 0x5410);
    }
    *reinterpret_cast<uint4*>(&tmp[p][r][0]) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
  }
  __syncwarp();
  int ty[4];
#pragma unroll
  for (int k = 0; k < 4; k++) ty[k] = c_epel[fy][k];
  const int row = (lane & 15) >> 1, xb = (lane & 1) * 4;
#pragma unroll
  for (int j = 0; j < 4; j++) out[j] = 0;
  if (row < chh && xb < cw) {
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const uint2 w2 = *reinterpret_cast<const uint2*>(&tmp[p][row + k][xb]);
      const int v[4] = {(int16_t)(w2.x & 0xffff), (int)w2.x >> 16, (int16_t)(w2.y & 0xffff), (int)w2.y >> 16};
#pragma unroll
      for (int j = 0; j < 4; j++) out[j] += ty[k] * v[j];
    }
#pragma unroll
    for (int j = 0; j < 4; j++) out[j] >>= 6;
  }
  __syncwarp();
}

// the four writers of fallback-motion.cc:33-158 on one 14-bit sample (pair)
struct WParams {   // out = clip8(((a*w0 + b*w1 + rnd) >> sh) + post) covers the four writers
  int w0, w1, rnd, sh, post;
};
__device__ __forceinline__ WParams make_writer(bool weighted, bool bi, int w0, int w1, int o0, int o1, int log2wd) {
  WParams p;
  if (!weighted) { p.w0 = 1; p.w1 = bi ? 1 : 0; p.rnd = bi ? 64 : 32; p.sh = bi ? 7 : 6; p.post = 0; }
  else if (!bi) { p.w0 = w0; p.w1 = 0; p.rnd = 1 << (log2wd - 1); p.sh = log2wd; p.post = o0; }
  else { p.w0 = w0; p.w1 = w1; p.rnd = (o0 + o1 + 1) << log2wd; p.sh = log2wd + 1; p.post = 0; }
  return p;
}
__device__ __forceinline__ int write_sample(const WParams& p, int a, int b) {
  return clip8(((a * p.w0 + b * p.w1 + p.rnd) >> p.sh) + p.post);
}

// PU records are fetched MC_CHUNK at a time, one per lane (items are consecutive, so are the records), and then
// broadcast one by one: the dependent item -> record -> slice round trips are paid once per chunk, not per PU.
#define MC_CHUNK 8
#ifndef MC_MIN_CTAS
#define MC_MIN_CTAS 2
#endif

__global__ void __launch_bounds__(B265_WARPS_PER_CTA * 32, MC_MIN_CTAS)
k_mc(const FrameDev* __restrict__ frames, const uint32_t* __restrict__ items, int n_items) {
  __shared__ McSmem smem[B265_WARPS_PER_CTA];
  McSmem& s = smem[threadIdx.x >> 5];
  const int lane = lane_id();
  const int warps_total = gridDim.x * B265_WARPS_PER_CTA;
  for (int base = (blockIdx.x * B265_WARPS_PER_CTA + (threadIdx.x >> 5)) * MC_CHUNK; base < n_items; base += warps_total * MC_CHUNK) {
    const int cnt = min(MC_CHUNK, n_items - base);
    uint32_t my_o = 0;
    uint2 rec[3] = {make_uint2(0, 0), make_uint2(0, 0), make_uint2(0, 0)};
    if (lane < cnt) {
      my_o = __ldg(items + base + lane);
      const uint2* r = reinterpret_cast<const uint2*>(frames[my_o >> 24].pus + (my_o & 0xffffff));   // 24-byte records, 8-byte aligned
      rec[0] = __ldg(r); rec[1] = __ldg(r + 1); rec[2] = __ldg(r + 2);
    }
    for (int j = 0; j < cnt; j++) {
      const uint32_t o = __shfl_sync(B265_FULL_MASK, my_o, j);
      const uint32_t r0 = __shfl_sync(B265_FULL_MASK, rec[0].x, j), r1 = __shfl_sync(B265_FULL_MASK, rec[0].y, j);
      const uint32_t r2 = __shfl_sync(B265_FULL_MASK, rec[1].x, j), r3 = __shfl_sync(B265_FULL_MASK, rec[1].y, j);
      const uint32_t r4 = __shfl_sync(B265_FULL_MASK, rec[2].x, j);
      const FrameDev& f = frames[o >> 24];
      const int pu_x = (int)(r0 & 0xffff), pu_y = (int)(r0 >> 16);
      const int pu_w = (int)(r1 & 0xff), pu_h = (int)((r1 >> 8) & 0xff), pu_flags = (int)((r1 >> 16) & 0xff);
      const b265_slice& sl = f.slices[r1 >> 24];
      const int mv0x = (int16_t)(r2 & 0xffff), mv0y = (int)r2 >> 16, mv1x = (int16_t)(r3 & 0xffff), mv1y = (int)r3 >> 16;
      const int ridx0 = (int8_t)(r4 & 0xff), ridx1 = (int8_t)((r4 >> 8) & 0xff);
      const int slot0 = (int)((r4 >> 16) & 0xff), slot1 = (int)(r4 >> 24);
      int use0 = pu_flags & B265_PU_L0, use1 = (pu_flags & B265_PU_L1) ? 1 : 0;
      const int weighted = sl.weighted;
      // identical-motion bi-prediction collapses to uni-prediction (motion.cc:326-335)
      if (use0 && use1 && mv0x == mv1x && mv0y == mv1y && slot0 == slot1 && !sl.pps_weighted_pred) use1 = 0;
      if (!use0 && !use1) continue;
      const bool first_l1 = !use0;   // first (or only) list is L1
      const bool bi = use0 && use1;
      const int mvax = first_l1 ? mv1x : mv0x, mvay = first_l1 ? mv1y : mv0y, slota = first_l1 ? slot1 : slot0;
      WParams wl, wc;   // luma; this lane's chroma plane (lane >> 4)
      if (!weighted) {
        wl = make_writer(false, bi, 0, 0, 0, 0, 6);
        wc = wl;
      } else {
        const int la = first_l1 ? 1 : 0, ra = (first_l1 ? ridx1 : ridx0) & 15, rb = ridx1 & 15;
        const int c = lane >= 16 ? 2 : 1;
        wl = make_writer(true, bi, sl.weight[la][ra][0], sl.weight[1][rb][0], sl.offset[la][ra][0], sl.offset[1][rb][0],
                         sl.luma_log2_denom + 6);
        wc = make_writer(true, bi, sl.weight[la][ra][c], sl.weight[1][rb][c], sl.offset[la][ra][c], sl.offset[1][rb][c],
                         sl.chroma_log2_denom + 6);
      }
      const int W = f.width, H = f.height;
      const uint8_t* const ra0 = f.ref_plane[slota][0];
      const uint8_t* const ra1 = f.ref_plane[slota][1];
      const uint8_t* const ra2 = f.ref_plane[slota][2];
      const uint8_t* const rb0 = f.ref_plane[slot1][0];
      const uint8_t* const rb1 = f.ref_plane[slot1][1];
      const uint8_t* const rb2 = f.ref_plane[slot1][2];
      const int pitch0 = f.pitch[0], pitch1 = f.pitch[1];
      uint8_t* const out0 = f.plane[0];
      uint8_t* const outc = lane >= 16 ? f.plane[2] : f.plane[1];
      const int fcw = f.cw, fch = f.ch;
      for (int ty = 0; ty < pu_h; ty += 16)
        for (int tx = 0; tx < pu_w; tx += 16) {
          const int px = pu_x + tx, py = pu_y + ty;
          const int tw = min(16, pu_w - tx), th = min(16, pu_h - ty);
          // every reference row this tile needs is requested before any of them is used: one memory round trip
          const int lax = px + (mvax >> 2), lay = py + (mvay >> 2), lbx = px + (mv1x >> 2), lby = py + (mv1y >> 2);
          const int cpx = px >> 1, cpy = py >> 1;
          const int cax = cpx + (mvax >> 3), cay = cpy + (mvay >> 3), cbx = cpx + (mv1x >> 3), cby = cpy + (mv1y >> 3);
          uint32_t wa[7], wb[7], wca[4], wcb[4];
          luma_fetch(ra0, pitch0, W, H, lax, lay, th, wa);
          if (bi) luma_fetch(rb0, pitch0, W, H, lbx, lby, th, wb);
          if (fcw > 0) {
            chroma_fetch(ra1, ra2, pitch1, fcw, fch, cax, cay, th >> 1, wca);
            if (bi) chroma_fetch(rb1, rb2, pitch1, fcw, fch, cbx, cby, th >> 1, wcb);
          }
          // ---- luma ----
          int a[8], b[8];
          luma_filter(s, wa, lax, mvax & 3, mvay & 3, tw, th, a);
          if (bi) {
            luma_filter(s, wb, lbx, mv1x & 3, mv1y & 3, tw, th, b);
          } else {
#pragma unroll
            for (int k = 0; k < 8; k++) b[k] = 0;
          }
          {
            const int row = lane >> 1, xb = (lane & 1) * 8;
            if (row < th && xb < tw) {
              int v[8];
#pragma unroll
              for (int k = 0; k < 8; k++) v[k] = write_sample(wl, a[k], b[k]);
              uint8_t* d = out0 + (size_t)(py + row) * pitch0 + px + xb;
              // PB x positions are multiples of 4 only (AMP), so 8-byte alignment is checked at run time
              const uint32_t lo = pack4(v[0], v[1], v[2], v[3]), hi = pack4(v[4], v[5], v[6], v[7]);
              if (xb + 8 <= tw) {
                if ((reinterpret_cast<uintptr_t>(d) & 7) == 0) *reinterpret_cast<uint2*>(d) = make_uint2(lo, hi);
                else { reinterpret_cast<uint32_t*>(d)[0] = lo; reinterpret_cast<uint32_t*>(d)[1] = hi; }
              } else *reinterpret_cast<uint32_t*>(d) = lo;
            }
          }
          // ---- chroma (4:2:0) ----
          if (fcw > 0) {
            int ca[4], cb4[4];
            chroma_filter(s, wca, cax, mvax & 7, mvay & 7, tw >> 1, th >> 1, ca);
            if (bi) {
              chroma_filter(s, wcb, cbx, mv1x & 7, mv1y & 7, tw >> 1, th >> 1, cb4);
            } else {
#pragma unroll
              for (int k = 0; k < 4; k++) cb4[k] = 0;
            }
            const int row = (lane & 15) >> 1, xb = (lane & 1) * 4;
            const int cw = tw >> 1, chh = th >> 1;
            if (row < chh && xb < cw) {
              int v[4];
#pragma unroll
              for (int k = 0; k < 4; k++) v[k] = write_sample(wc, ca[k], cb4[k]);
              uint8_t* d = outc + (size_t)(cpy + row) * pitch1 + cpx + xb;
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
}

void launch_mc(cudaStream_t st, const FrameDev* frames, const uint32_t* items, int n, int sm_count) {
  if (n <= 0) return;
  const int chunks = (n + MC_CHUNK - 1) / MC_CHUNK;
  int ctas = (chunks + B265_WARPS_PER_CTA - 1) / B265_WARPS_PER_CTA;
  int grid = min(ctas, sm_count * MC_MIN_CTAS * 2);
  k_mc<<<grid, B265_WARPS_PER_CTA * 32, 0, st>>>(frames, items, n);
}
