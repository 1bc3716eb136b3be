/*
 * oracle/mc.c — CPU restatement of inter prediction.  TEST INFRASTRUCTURE ONLY (see oracle.h).
 * Follows libde265/motion.cc:49-273 (mc_luma / mc_chroma: integer/fraction split, picture-edge
 * clamp), libde265/fallback-motion.cc:298-407 and :482-625 (separable 4/8-tap filters with the
 * reference's case split and shifts), :33-158 (the four writers) and motion.cc:279-661 (dispatch).
 */
#include <string.h>
#include "oracle.h"

static int clip3(int lo, int hi, int v) { return v < lo ? lo : (v > hi ? hi : v); }
/* coordinate clamp of mc_luma/mc_chroma (motion.cc:146-149); n<0 = caller guarantees the halo */
static int cc(int n, int v) { return n < 0 ? v : clip3(0, n - 1, v); }
static uint8_t clip8(int v) { return (uint8_t)clip3(0, 255, v); }

static int qpel_filter(const int* p, int frac) { /* p[0] is the sample at offset -3 */
  switch (frac) {
    case 1: return -p[0] + 4 * p[1] - 10 * p[2] + 58 * p[3] + 17 * p[4] - 5 * p[5] + p[6];
    case 2: return -p[0] + 4 * p[1] - 11 * p[2] + 40 * p[3] + 40 * p[4] - 11 * p[5] + 4 * p[6] - p[7];
    case 3: return p[1] - 5 * p[2] + 17 * p[3] + 58 * p[4] - 10 * p[5] + 4 * p[6] - p[7];
    default: return p[3];
  }
}
static int epel_filter(const int* p, int frac) { /* p[0] is the sample at offset -1 */
  static const int t[8][4] = {{0, 1, 0, 0},     {-2, 58, 10, -2}, {-4, 54, 16, -2}, {-6, 46, 28, -4},
                              {-4, 36, 36, -4}, {-4, 28, 46, -6}, {-2, 16, 54, -4}, {-2, 10, 58, -2}};
  return t[frac][0] * p[0] + t[frac][1] * p[1] + t[frac][2] * p[2] + t[frac][3] * p[3];
}

/* 14-bit prediction of a w x h block whose integer position is (xi,yi) in a W x H plane */
static void interp(int luma, const uint8_t* ref, int pitch, int W, int H, int xi, int yi, int fx, int fy,
                   int w, int h, int16_t* out /* stride w */) {
  const int before = luma ? 3 : 1, taps = luma ? 8 : 4;
  if (fx == 0 && fy == 0) {
    for (int y = 0; y < h; y++)
      for (int x = 0; x < w; x++)
        out[y * w + x] = (int16_t)(ref[cc(H, yi + y) * (ptrdiff_t)pitch + cc(W, xi + x)] << 6);
    return;
  }
  /* first stage: rows -before .. h+taps-before-1, int16 like the reference's mcbuffer */
  static _Thread_local int16_t tmp[(64 + 8) * 64];
  int rows = h + taps - 1;
  for (int r = 0; r < rows; r++)
    for (int x = 0; x < w; x++) {
      int p[8];
      int yy = cc(H, yi + r - before);
      for (int k = 0; k < taps; k++) p[k] = ref[yy * (ptrdiff_t)pitch + cc(W, xi + x + k - before)];
      tmp[r * w + x] = (int16_t)(luma ? qpel_filter(p, fx) : epel_filter(p, fx)); /* shift1 = 0 at 8 bit */
    }
  int vshift = fx == 0 ? 0 : 6;
  for (int y = 0; y < h; y++)
    for (int x = 0; x < w; x++) {
      int p[8];
      for (int k = 0; k < taps; k++) p[k] = tmp[(y + k) * w + x];
      int v = fy == 0 ? p[before] : ((luma ? qpel_filter(p, fy) : epel_filter(p, fy)) >> vshift);
      out[y * w + x] = (int16_t)v;
    }
}

void orc_interp_block(int luma, int w, int h, const uint8_t* src, ptrdiff_t stride, int fx, int fy, int16_t* out) {
  /* src points at the block's integer position inside a buffer with the halo present */
  interp(luma, src, (int)stride, -1, -1, 0, 0, fx, fy, w, h, out);
}

static void write_block(uint8_t* dst, int pitch, int w, int h, const int16_t* a, const int16_t* b, int mode, int w0,
                        int o0, int w1, int o1, int log2wd) {
  for (int y = 0; y < h; y++)
    for (int x = 0; x < w; x++) {
      int va = a[y * w + x], vb = b ? b[y * w + x] : 0, v;
      switch (mode) {
        case 0: v = (va + 32) >> 6; break;
        case 1: v = (va + vb + 64) >> 7; break;
        case 2: v = ((va * w0 + (1 << (log2wd - 1))) >> log2wd) + o0; break;
        default: v = (va * w0 + vb * w1 + ((o0 + o1 + 1) << log2wd)) >> (log2wd + 1); break;
      }
      dst[y * (ptrdiff_t)pitch + x] = clip8(v);
    }
}

void orc_mc_batch(const b265_pu* pus, int n_pus, const b265_slice* slices, const uint8_t* const* ref_planes,
                  int n_refs, uint8_t* const* dst_planes, int width, int height, int pitch, int pitch_c) {
  static _Thread_local int16_t pa[64 * 64], pb[64 * 64];
  (void)n_refs;
  for (int i = 0; i < n_pus; i++) {
    const b265_pu* pu = &pus[i];
    const b265_slice* sl = &slices[pu->slice];
    int use0 = (pu->flags & B265_PU_L0) != 0, use1 = (pu->flags & B265_PU_L1) != 0;
    if (use0 && use1 && !sl->pps_weighted_pred && pu->mv[0][0] == pu->mv[1][0] && pu->mv[0][1] == pu->mv[1][1] &&
        pu->ref_slot[0] == pu->ref_slot[1])
      use1 = 0; /* motion.cc:326-335 */
    if (!use0 && !use1) continue;
    int la = use0 ? 0 : 1, bi = use0 && use1;
    for (int c = 0; c < 3; c++) {
      if (c && !dst_planes[c]) continue;
      int sh = c ? 1 : 0, w = pu->w >> sh, h = pu->h >> sh;
      int W = c ? (width + 1) / 2 : width, H = c ? (height + 1) / 2 : height, pt = c ? pitch_c : pitch;
      int fmask = c ? 7 : 3, ishift = c ? 3 : 2;
      int16_t* bufs[2] = {pa, pb};
      int nl = 0;
      for (int l = la; l < 2 && nl < (bi ? 2 : 1); l++, nl++) {
        int mx = pu->mv[l][0], my = pu->mv[l][1];
        interp(c == 0, ref_planes[3 * pu->ref_slot[l] + c], pt, W, H, (pu->x >> sh) + (mx >> ishift),
               (pu->y >> sh) + (my >> ishift), mx & fmask, my & fmask, w, h, bufs[nl]);
      }
      uint8_t* d = dst_planes[c] + (pu->y >> sh) * (ptrdiff_t)pt + (pu->x >> sh);
      if (!sl->weighted) {
        write_block(d, pt, w, h, pa, bi ? pb : NULL, bi ? 1 : 0, 0, 0, 0, 0, 6);
      } else {
        int ra = pu->ref_idx[la] & 15, rb = pu->ref_idx[1] & 15;
        int log2wd = (c ? sl->chroma_log2_denom : sl->luma_log2_denom) + 6;
        write_block(d, pt, w, h, pa, bi ? pb : NULL, bi ? 3 : 2, sl->weight[la][ra][c], sl->offset[la][ra][c],
                    sl->weight[1][rb][c], sl->offset[1][rb][c], log2wd);
      }
    }
  }
}
