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
// Work mapping: one HALF-WARP per piece of at most 16 x 8 luma samples (+ its two 8 x 4 chroma tiles; prediction blocks
// are cut into such pieces when the lists are staged, backend.h pu_expand), two pieces per warp pass.  Lane r of a half
// holds reference ROW r of its piece (8 + 7 rows, with the 3/4-sample halo) in registers, fetched with independent
// 32-bit loads (one memory round trip per pass); the horizontal filter runs on the packed bytes with dp4a (u8 x s8), the
// int16 rows go through shared memory once, and the vertical filter gives every lane 8 horizontally adjacent luma samples
// (one 8-byte store) / 4 chroma samples.  A 16 x 16 block costs what it cost as one 23-row tile (the idle lanes of that
// mapping now carry the second piece), everything up to 16 x 8 costs half.
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
  __align__(16) int16_t tmp[2][18 * 16];  // after the horizontal pass, per half (tmp_row); chroma: [4][7][8] (piece, plane)
};

// sum of four products, unsigned bytes of a times signed bytes of b
__device__ __forceinline__ int dp4a_us(uint32_t a, int b, int c) {
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

// NW words of reference row `y` starting at the 4-byte aligned column xa (all clamped to the picture).  Lanes hold different
// ROWS, so every load instruction of the warp touches ~30 cache lines: the kernel is bound by L1 tag look-ups, not by bytes.
// The row piece is therefore fetched as NV 16-byte vectors from the enclosing 16-byte aligned column and the words wanted
// (offset 0..3 words) are picked with two levels of selects: 3 loads instead of 7 for luma, 2 instead of 4 for chroma.
template <int NW, int NV>
__device__ __forceinline__ void load_row(const uint8_t* __restrict__ ref, int pitch, int W, int H, int xa, int y, uint32_t (&w)[NW]) {
  static_assert(NW + 3 <= 4 * NV, "vectors must cover every word offset");
  const uint8_t* row = ref + (size_t)clip3(0, H - 1, y) * pitch;
  const int xv = xa & ~15;
  if (xa >= 0 && xa + 4 * NW <= W && xv + 16 * NV <= pitch && (reinterpret_cast<uintptr_t>(row) & 15) == 0) {
    uint32_t u[4 * NV];
#pragma unroll
    for (int i = 0; i < NV; i++) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(row + xv) + i);
      u[4 * i] = v.x; u[4 * i + 1] = v.y; u[4 * i + 2] = v.z; u[4 * i + 3] = v.w;
    }
    const bool o1 = (xa >> 2) & 1, o2 = (xa >> 3) & 1;
    uint32_t t[NW + 2];
#pragma unroll
    for (int i = 0; i < NW + 2; i++) t[i] = o1 ? u[i + 1] : u[i];
#pragma unroll
    for (int i = 0; i < NW; i++) w[i] = o2 ? t[i + 2] : t[i];
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

// luma, step 1: lane l < th+7 of the half fetches reference row l of the piece at (x0,y0) (7 words from the aligned column)
// Without a vertical phase only the th rows of the piece itself are read (rows 3 .. th+2 of the window): the vertical tap
// vector is then {0,0,0,64,..}, whatever is left in the other rows of the shared buffer is multiplied by zero.
__device__ __forceinline__ bool row_wanted(int l, int n, int halo_before, int halo, int frac) {
  return frac ? l < n + halo : (l >= halo_before && l < n + halo_before);
}
__device__ __forceinline__ void luma_fetch(bool on, int l, const uint8_t* __restrict__ ref, int pitch, int W, int H, int x0, int y0,
                                           int th, int fy, uint32_t (&w)[7]) {
  const int xs = x0 - 3;
  if (on && row_wanted(l, th, 3, 7, fy)) load_row<7, 3>(ref, pitch, W, H, xs - (xs & 3), y0 - 3 + l, w);
}

// two products of int16 halves of a with the two low signed bytes of b
__device__ __forceinline__ int dp2a_ss(uint32_t a, int b, int c) {
  int d;
  asm("dp2a.lo.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

// row r of a half's int16 buffer: 32 bytes of padding after every four rows keep the column reads of the vertical pass
// (rows 4g+k of both row groups and both halves in one instruction) on distinct banks
__device__ __forceinline__ int tmp_row(int r) { return (r + (r >> 2)) * 16; }

// luma, step 2: 14-bit prediction of the tw x th (<= 16x8) piece; lane l of the half gets columns 2*(l&7), +1 of rows
// 4*(l>>3) .. +3: out[2*j + c].  Every lane of the warp calls this (the barriers are warp-wide); `on` says whether its half
// has a piece for this list.
//   horizontal: lane l holds row l; output x = 4q+m is the dot product of bytes x..x+7, taken as three dp4a of the words
//   q, q+1, q+2 with the tap vector shifted by m bytes (no data shifts); the 16 int16 results go to shared memory
//   vertical: a lane reads its two columns of 11 rows (one 32-bit load per row), interleaves vertically adjacent int16
//   (PRMT) and accumulates tap pairs with dp2a
__device__ __forceinline__ void luma_filter(int16_t* tmp, bool on, int l, const uint32_t (&w)[7], int x0, int fx, int fy, int tw,
                                            int th, int (&out)[8]) {
  const int off = (x0 - 3) & 3;
  if (on && row_wanted(l, th, 3, 7, fy)) {
    uint32_t a[6];
#pragma unroll
    for (int i = 0; i < 6; i++) a[i] = __funnelshift_r(w[i], w[i + 1], 8 * off);
    const uint32_t tlo = reinterpret_cast<const uint32_t*>(c_qpel[fx])[0], thi = reinterpret_cast<const uint32_t*>(c_qpel[fx])[1];
    int t0[4], t1[4], t2[4];   // the 8 taps at byte offset m of a 12-byte window
    t0[0] = (int)tlo; t1[0] = (int)thi; t2[0] = 0;
#pragma unroll
    for (int m = 1; m < 4; m++) {
      t0[m] = (int)(tlo << (8 * m));
      t1[m] = (int)__funnelshift_l(tlo, thi, 8 * m);
      t2[m] = (int)(thi >> (32 - 8 * m));
    }
    uint32_t packed[8];
#pragma unroll
    for (int x = 0; x < 16; x += 2) {
      int v[2];
#pragma unroll
      for (int j = 0; j < 2; j++) {
        const int q = (x + j) >> 2, m = (x + j) & 3;
        v[j] = dp4a_us(a[q + 1], t1[m], dp4a_us(a[q], t0[m], 0));
        if (m) v[j] = dp4a_us(a[q + 2], t2[m], v[j]);
      }
      packed[x >> 1] = __byte_perm((uint32_t)v[0], (uint32_t)v[1], 0x5410);
    }
    uint4* dst = reinterpret_cast<uint4*>(tmp + tmp_row(l));
    dst[0] = make_uint4(packed[0], packed[1], packed[2], packed[3]);
    if (tw > 8) dst[1] = make_uint4(packed[4], packed[5], packed[6], packed[7]);
  }
  __syncwarp();
  const int r0 = (l >> 3) * 4, xb = (l & 7) * 2;
#pragma unroll
  for (int j = 0; j < 8; j++) out[j] = 0;
  if (on && r0 < th && xb < tw) {
    int tv[4];   // vertical taps in pairs (2t, 2t+1) as the two low bytes
#pragma unroll
    for (int t = 0; t < 4; t++) tv[t] = (int)(uint8_t)c_qpel[fy][2 * t] | ((int)(uint8_t)c_qpel[fy][2 * t + 1] << 8);
    uint32_t c[11];
#pragma unroll
    for (int k = 0; k < 11; k++) c[k] = *reinterpret_cast<const uint32_t*>(tmp + tmp_row(r0 + k) + xb);
#pragma unroll
    for (int k = 0; k < 10; k++) {   // pair k = rows (r0+k, r0+k+1) feeds output row j = k - 2t with tap pair t
      const uint32_t p0 = __byte_perm(c[k], c[k + 1], 0x5410), p1 = __byte_perm(c[k], c[k + 1], 0x7632);
#pragma unroll
      for (int t = 0; t < 4; t++) {
        const int j = k - 2 * t;
        if (j >= 0 && j < 4) {
          out[2 * j] = dp2a_ss(p0, tv[t], out[2 * j]);
          out[2 * j + 1] = dp2a_ss(p1, tv[t], out[2 * j + 1]);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 8; j++) out[j] >>= 6;
  }
  __syncwarp();
}

// chroma: a group of 8 lanes per (piece, plane), tile cw x chh <= 8x4.  Step 1: lane l8 < chh+3 fetches row l8.
__device__ __forceinline__ void chroma_fetch(bool on, int l8, const uint8_t* __restrict__ ref, int pitch, int W, int H, int x0, int y0,
                                             int chh, int fy, uint32_t (&w)[4]) {
  const int xs = x0 - 1;
  if (on && row_wanted(l8, chh, 1, 3, fy)) load_row<4, 2>(ref, pitch, W, H, xs - (xs & 3), y0 - 1 + l8, w);
}

// step 2. Result: row = l8>>1, x = 4*(l8&1)..+3 of the group's tile
__device__ __forceinline__ void chroma_filter(int16_t (*tmp)[8], bool on, int l8, const uint32_t (&w)[4], int x0, int fx, int fy, int cw,
                                              int chh, int (&out)[4]) {
  const int off = (x0 - 1) & 3;
  if (on && row_wanted(l8, chh, 1, 3, fy)) {
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
    *reinterpret_cast<uint4*>(&tmp[l8][0]) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
  }
  __syncwarp();
  const int row = l8 >> 1, xb = (l8 & 1) * 4;
#pragma unroll
  for (int j = 0; j < 4; j++) out[j] = 0;
  if (on && row < chh && xb < cw) {
    int ty[4];
#pragma unroll
    for (int k = 0; k < 4; k++) ty[k] = c_epel[fy][k];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const uint2 w2 = *reinterpret_cast<const uint2*>(&tmp[row + k][xb]);
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

// Piece records are fetched MC_CHUNK at a time, one per lane (items are consecutive, so are the records), and handed to
// the halves pass by pass: the dependent item -> record -> slice round trips are paid once per chunk, not per piece.
#define MC_CHUNK 16
#ifndef MC_WARPS
#define MC_WARPS 8
#endif
#ifndef MC_MIN_CTAS
#define MC_MIN_CTAS 2
#endif

__global__ void __launch_bounds__(MC_WARPS * 32, MC_MIN_CTAS)
k_mc(const FrameDev* __restrict__ frames, const uint32_t* __restrict__ items, int n_items) {
  __shared__ McSmem smem[MC_WARPS];
  McSmem& s = smem[threadIdx.x >> 5];
  const int lane = lane_id();
  const int half = lane >> 4, l = lane & 15;          // luma: half-warp per piece, lane l of the half
  const int plane = (lane >> 3) & 1, l8 = lane & 7;   // chroma: 8 lanes per (piece, plane)
  int16_t* const ltmp = s.tmp[half];
  int16_t (*const ctmp)[8] = reinterpret_cast<int16_t (*)[8]>(&s.tmp[0][0]) + (lane >> 3) * 7;
  const int warps_total = gridDim.x * MC_WARPS;
  for (int base = (blockIdx.x * MC_WARPS + (threadIdx.x >> 5)) * MC_CHUNK; base < n_items; base += warps_total * MC_CHUNK) {
    const int cnt = min(MC_CHUNK, n_items - base);
    uint32_t my_o = 0;
    uint2 rec[3] = {make_uint2(0, 0), make_uint2(0, 0), make_uint2(0, 0)};
    if (lane < cnt) {
      my_o = __ldg(items + base + lane);
      const uint2* r = reinterpret_cast<const uint2*>(frames[my_o >> 24].pus + (my_o & 0xffffff));   // 24-byte records, 8-byte aligned
      rec[0] = __ldg(r); rec[1] = __ldg(r + 1);
      rec[2].x = __ldg(reinterpret_cast<const uint32_t*>(r + 2));   // the last word is padding: no dead destination register
    }
    for (int j0 = 0; j0 < cnt; j0 += 2) {
      const bool have = j0 + half < cnt;
      const int src = have ? j0 + half : j0;
      const uint32_t o = __shfl_sync(B265_FULL_MASK, my_o, src);
      const uint32_t r0 = __shfl_sync(B265_FULL_MASK, rec[0].x, src), r1 = __shfl_sync(B265_FULL_MASK, rec[0].y, src);
      const uint32_t r2 = __shfl_sync(B265_FULL_MASK, rec[1].x, src), r3 = __shfl_sync(B265_FULL_MASK, rec[1].y, src);
      const uint32_t r4 = __shfl_sync(B265_FULL_MASK, rec[2].x, src);
      const FrameDev& f = frames[o >> 24];
      const int px = (int)(r0 & 0xffff), py = (int)(r0 >> 16);
      const int tw = (int)(r1 & 0xff), th = (int)((r1 >> 8) & 0xff), pu_flags = (int)((r1 >> 16) & 0xff);   // <= 16 x 8
      const b265_slice& sl = f.slices[r1 >> 24];
      const int mv0x = (int16_t)(r2 & 0xffff), mv0y = (int)r2 >> 16, mv1x = (int16_t)(r3 & 0xffff), mv1y = (int)r3 >> 16;
      const int ridx0 = (int8_t)(r4 & 0xff), ridx1 = (int8_t)((r4 >> 8) & 0xff);
      const int slot0 = (int)((r4 >> 16) & 0xff), slot1 = (int)(r4 >> 24);
      int use0 = pu_flags & B265_PU_L0, use1 = (pu_flags & B265_PU_L1) ? 1 : 0;
      const int weighted = sl.weighted;
      // identical-motion bi-prediction collapses to uni-prediction (motion.cc:326-335)
      if (use0 && use1 && mv0x == mv1x && mv0y == mv1y && slot0 == slot1 && !sl.pps_weighted_pred) use1 = 0;
      const bool on = have && (use0 || use1);
      const bool first_l1 = !use0;   // first (or only) list is L1
      const bool bi = on && use0 && use1;
      const int mvax = first_l1 ? mv1x : mv0x, mvay = first_l1 ? mv1y : mv0y, slota = first_l1 ? slot1 : slot0;
      WParams wl, wc;   // luma; this lane's chroma plane
      if (!weighted) {
        wl = make_writer(false, bi, 0, 0, 0, 0, 6);
        wc = wl;
      } else {
        const int la = first_l1 ? 1 : 0, ra = (first_l1 ? ridx1 : ridx0) & 15, rb = ridx1 & 15;
        const int c = 1 + plane;
        wl = make_writer(true, bi, sl.weight[la][ra][0], sl.weight[1][rb][0], sl.offset[la][ra][0], sl.offset[1][rb][0],
                         sl.luma_log2_denom + 6);
        wc = make_writer(true, bi, sl.weight[la][ra][c], sl.weight[1][rb][c], sl.offset[la][ra][c], sl.offset[1][rb][c],
                         sl.chroma_log2_denom + 6);
      }
      const int W = f.width, H = f.height;
      const int pitch0 = f.pitch[0], pitch1 = f.pitch[1];
      const int fcw = f.cw, fch = f.ch;
      const bool con = on && fcw > 0, cbi = bi && fcw > 0;
      // every reference row this pass needs is requested before any of them is used: one memory round trip
      const int lax = px + (mvax >> 2), lay = py + (mvay >> 2), lbx = px + (mv1x >> 2), lby = py + (mv1y >> 2);
      const int cpx = px >> 1, cpy = py >> 1;
      const int cax = cpx + (mvax >> 3), cay = cpy + (mvay >> 3), cbx = cpx + (mv1x >> 3), cby = cpy + (mv1y >> 3);
      uint32_t wa[7], wb[7], wca[4], wcb[4];
      luma_fetch(on, l, f.ref_plane[slota][0], pitch0, W, H, lax, lay, th, mvay & 3, wa);
      luma_fetch(bi, l, f.ref_plane[slot1][0], pitch0, W, H, lbx, lby, th, mv1y & 3, wb);
      chroma_fetch(con, l8, f.ref_plane[slota][1 + plane], pitch1, fcw, fch, cax, cay, th >> 1, mvay & 7, wca);
      chroma_fetch(cbi, l8, f.ref_plane[slot1][1 + plane], pitch1, fcw, fch, cbx, cby, th >> 1, mv1y & 7, wcb);
      const bool any_bi = __any_sync(B265_FULL_MASK, bi), any_c = __any_sync(B265_FULL_MASK, con);
      // ---- luma ----
      int a[8], b[8];
      luma_filter(ltmp, on, l, wa, lax, mvax & 3, mvay & 3, tw, th, a);
      if (any_bi) {
        luma_filter(ltmp, bi, l, wb, lbx, mv1x & 3, mv1y & 3, tw, th, b);
      } else {
#pragma unroll
        for (int k = 0; k < 8; k++) b[k] = 0;
      }
      {
        const int r0 = (l >> 3) * 4, xb = (l & 7) * 2;
        if (on && xb < tw) {
          uint8_t* d = f.plane[0] + (size_t)(py + r0) * pitch0 + px + xb;   // PB x positions are multiples of 4: 2-byte aligned
#pragma unroll
          for (int j = 0; j < 4; j++)
            if (r0 + j < th)
              *reinterpret_cast<uint16_t*>(d + (size_t)j * pitch0) =
                  (uint16_t)(write_sample(wl, a[2 * j], b[2 * j]) | (write_sample(wl, a[2 * j + 1], b[2 * j + 1]) << 8));
        }
      }
      // ---- chroma (4:2:0) ----
      if (any_c) {
        int ca[4], cb4[4];
        chroma_filter(ctmp, con, l8, wca, cax, mvax & 7, mvay & 7, tw >> 1, th >> 1, ca);
        if (any_bi) {
          chroma_filter(ctmp, cbi, l8, wcb, cbx, mv1x & 7, mv1y & 7, tw >> 1, th >> 1, cb4);
        } else {
#pragma unroll
          for (int k = 0; k < 4; k++) cb4[k] = 0;
        }
        const int row = l8 >> 1, xb = (l8 & 1) * 4;
        const int cw = tw >> 1, chh = th >> 1;
        if (con && row < chh && xb < cw) {
          int v[4];
#pragma unroll
          for (int k = 0; k < 4; k++) v[k] = write_sample(wc, ca[k], cb4[k]);
          uint8_t* d = f.plane[1 + plane] + (size_t)(cpy + row) * pitch1 + cpx + xb;
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
  const int chunks = (n + MC_CHUNK - 1) / MC_CHUNK;
  int ctas = (chunks + MC_WARPS - 1) / MC_WARPS;
  int grid = min(ctas, sm_count * MC_MIN_CTAS * 2);   // two waves of resident CTAs, grid-stride over the chunks
  k_mc<<<grid, MC_WARPS * 32, 0, st>>>(frames, items, n);
}
