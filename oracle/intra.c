/*
 * oracle/intra.c — CPU restatement of intra prediction.  TEST INFRASTRUCTURE ONLY (see oracle.h).
 * Follows libde265/intrapred.cc: intra_border_computer::preproc/fill_from_image (:437-688),
 * reference_sample_substitution (:395-431), intra_prediction_sample_filtering (:816-889),
 * intra_prediction_angular (:903-1005), _planar (:1009-1032), _DC (:1036-1069),
 * decode_intra_prediction_internal (:1074-1111).
 */
#include <stdlib.h>
#include <string.h>
#include "oracle.h"

static int clip3(int lo, int hi, int v) { return v < lo ? lo : (v > hi ? hi : v); }
static int imin(int a, int b) { return a < b ? a : b; }

static const int kAngle[35] = {0,  0,  32, 26, 21, 17, 13, 9,  5,  2,  0,  -2, -5, -9, -13, -17, -21, -26,
                               -32, -26, -21, -17, -13, -9, -5, -2, 0,  2,  5,  9,  13, 17,  21,  26,  32};
static const int kInvAngle[15] = {-4096, -1638, -910, -630, -482, -390, -315, -256,
                                  -315,  -390,  -482, -630, -910, -1638, -4096};

/* prediction from a complete border; p = centre of the (4nT+1)-sample array (modified in place by
 * the smoothing filter exactly like the reference's border_pixels) */
void orc_intra_from_border(int log2, int cidx, int mode, const uint8_t* border_centre, int strong,
                           int no_edge_filter, uint8_t* dst, int stride) {
  int nT = 1 << log2;
  uint8_t mem[4 * 32 + 1], fmem[4 * 32 + 1];
  uint8_t* p = mem + 2 * 32;
  memcpy(p - 2 * nT, border_centre - 2 * nT, 4 * nT + 1);
  /* 8.4.4.2.3 — luma only in 4:2:0 (intrapred.cc:1085-1089) */
  if (cidx == 0 && mode != 1 && nT != 4) {
    int d = imin(abs(mode - 26), abs(mode - 10));
    int filt = nT == 8 ? d > 7 : nT == 16 ? d > 1 : d > 0;
    if (filt) {
      uint8_t* pF = fmem + 2 * 32;
      int bi = strong && nT == 32 && abs(p[0] + p[64] - 2 * p[32]) < 8 && abs(p[0] + p[-64] - 2 * p[-32]) < 8;
      pF[-2 * nT] = p[-2 * nT];
      pF[2 * nT] = p[2 * nT];
      if (bi) {
        pF[0] = p[0];
        for (int i = 1; i <= 63; i++) {
          pF[-i] = (uint8_t)(p[0] + ((i * (p[-64] - p[0]) + 32) >> 6));
          pF[i] = (uint8_t)(p[0] + ((i * (p[64] - p[0]) + 32) >> 6));
        }
      } else {
        for (int i = -(2 * nT - 1); i <= 2 * nT - 1; i++) pF[i] = (uint8_t)((p[i + 1] + 2 * p[i] + p[i - 1] + 2) >> 2);
      }
      memcpy(p - 2 * nT, pF - 2 * nT, 4 * nT + 1);
    }
  }
  if (mode == 0) {
    for (int y = 0; y < nT; y++)
      for (int x = 0; x < nT; x++)
        dst[x + y * stride] = (uint8_t)(((nT - 1 - x) * p[-1 - y] + (x + 1) * p[1 + nT] + (nT - 1 - y) * p[1 + x] +
                                         (y + 1) * p[-1 - nT] + nT) >> (log2 + 1));
  } else if (mode == 1) {
    int dc = 0;
    for (int i = 0; i < nT; i++) dc += p[i + 1] + p[-i - 1];
    dc = (dc + nT) >> (log2 + 1);
    for (int y = 0; y < nT; y++)
      for (int x = 0; x < nT; x++) dst[x + y * stride] = (uint8_t)dc;
    if (cidx == 0 && nT < 32) {
      dst[0] = (uint8_t)((p[-1] + 2 * dc + p[1] + 2) >> 2);
      for (int x = 1; x < nT; x++) dst[x] = (uint8_t)((p[x + 1] + 3 * dc + 2) >> 2);
      for (int y = 1; y < nT; y++) dst[y * stride] = (uint8_t)((p[-y - 1] + 3 * dc + 2) >> 2);
    }
  } else {
    int rmem[4 * 32 + 1];
    int* ref = rmem + 2 * 32;
    int ang = kAngle[mode];
    if (mode >= 18) {
      for (int x = 0; x <= nT; x++) ref[x] = p[x];
      if (ang < 0) {
        int inv = kInvAngle[mode - 11];
        if (((nT * ang) >> 5) < -1)
          for (int x = (nT * ang) >> 5; x <= -1; x++) ref[x] = p[0 - ((x * inv + 128) >> 8)];
      } else
        for (int x = nT + 1; x <= 2 * nT; x++) ref[x] = p[x];
      for (int y = 0; y < nT; y++)
        for (int x = 0; x < nT; x++) {
          int idx = ((y + 1) * ang) >> 5, fact = ((y + 1) * ang) & 31;
          dst[x + y * stride] = (uint8_t)(fact ? ((32 - fact) * ref[x + idx + 1] + fact * ref[x + idx + 2] + 16) >> 5
                                               : ref[x + idx + 1]);
        }
      if (mode == 26 && cidx == 0 && nT < 32 && !no_edge_filter)
        for (int y = 0; y < nT; y++) dst[y * stride] = (uint8_t)clip3(0, 255, p[1] + ((p[-1 - y] - p[0]) >> 1));
    } else {
      for (int x = 0; x <= nT; x++) ref[x] = p[-x];
      if (ang < 0) {
        int inv = kInvAngle[mode - 11];
        if (((nT * ang) >> 5) < -1)
          for (int x = (nT * ang) >> 5; x <= -1; x++) ref[x] = p[(x * inv + 128) >> 8];
      } else
        for (int x = nT + 1; x <= 2 * nT; x++) ref[x] = p[-x];
      for (int y = 0; y < nT; y++)
        for (int x = 0; x < nT; x++) {
          int idx = ((x + 1) * ang) >> 5, fact = ((x + 1) * ang) & 31;
          dst[x + y * stride] = (uint8_t)(fact ? ((32 - fact) * ref[y + idx + 1] + fact * ref[y + idx + 2] + 16) >> 5
                                               : ref[y + idx + 1]);
        }
      if (mode == 10 && cidx == 0 && nT < 32 && !no_edge_filter)
        for (int x = 0; x < nT; x++) dst[x] = (uint8_t)clip3(0, 255, p[-1] + ((p[1 + x] - p[0]) >> 1));
    }
  }
}

/* ---- availability (intrapred.cc:437-688) ---- */
static unsigned interleave(unsigned x, unsigned y) {
  unsigned r = 0;
  for (int b = 0; b < 8; b++) r |= ((x >> b) & 1u) << (2 * b) | ((y >> b) & 1u) << (2 * b + 1);
  return r;
}
static int ctb_of(const b265_frame_job* J, int x, int y) {
  int w = (J->width + (1 << J->log2_ctb) - 1) >> J->log2_ctb;
  return (y >> J->log2_ctb) * w + (x >> J->log2_ctb);
}
static unsigned min_tb_addr_zs(const b265_frame_job* J, int x, int y) {
  int ctb = ctb_of(J, x, y), d = J->log2_ctb - J->log2_min_tb, m = (1 << J->log2_ctb) - 1;
  unsigned ts = J->ctbs ? (unsigned)J->ctbs[ctb].addr_ts : (unsigned)ctb;
  return (ts << (2 * d)) + interleave((x & m) >> J->log2_min_tb, (y & m) >> J->log2_min_tb);
}
static int same_slice_tile(const b265_frame_job* J, int xa, int ya, int xb, int yb) {
  if (!J->ctbs) return 1;
  const b265_ctb* a = &J->ctbs[ctb_of(J, xa, ya)];
  const b265_ctb* b = &J->ctbs[ctb_of(J, xb, yb)];
  return a->slice == b->slice && a->tile == b->tile;
}
static int is_intra(const b265_frame_job* J, int x, int y) {
  return J->unit_info ? (J->unit_info[(y >> 2) * (J->width >> 2) + (x >> 2)] & B265_UNIT_INTRA) != 0 : 1;
}

void orc_intra_tu(const b265_frame_job* J, const b265_tu* tu, uint8_t* const* planes, const int* pitch) {
  int cidx = tu->cidx_log2 & 3, log2 = (tu->cidx_log2 >> 2) & 7, nT = 1 << log2;
  int sh = (cidx && J->chroma_format_idc == 1) ? 1 : 0;
  int xB = tu->x, yB = tu->y, xL = xB << sh, yL = yB << sh;
  const uint8_t* img = planes[cidx];
  int stride = pitch[cidx];
  int constrained = (J->flags & B265_FRAME_CONSTRAINED_INTRA) != 0;
  int availL = xL > 0, availT = yL > 0, availTL = xL > 0 && yL > 0;
  int availTR = yL > 0 && (xL + (nT << sh) < J->width);
  if (availL && !same_slice_tile(J, xL, yL, xL - 1, yL)) availL = 0;
  if (availT && !same_slice_tile(J, xL, yL, xL, yL - 1)) availT = 0;
  if (availTL && !same_slice_tile(J, xL, yL, xL - 1, yL - 1)) availTL = 0;
  if (availTR && !same_slice_tile(J, xL, yL, xL + (nT << sh), yL - 1)) availTR = 0;
  int nBottom = J->height - yL;
  nBottom = (nBottom + (1 << sh) - 1) >> sh;
  if (nBottom > 2 * nT) nBottom = 2 * nT;
  int nRight = J->width - xL;
  nRight = (nRight + (1 << sh) - 1) >> sh;
  if (nRight > 2 * nT) nRight = 2 * nT;
  uint8_t bmem[4 * 32 + 1], amem[4 * 32 + 1];
  uint8_t* b = bmem + 2 * 32;
  uint8_t* av = amem + 2 * 32;
  memset(av - 2 * nT, 0, 4 * nT + 1);
  memset(b - 2 * nT, 0, 4 * nT + 1);
  unsigned cur = min_tb_addr_zs(J, xL, yL);
  int nAvail = 0, first = 0;
  for (int y = nBottom - 1; y >= 0; y -= 4)
    if (availL) {
      int ok = min_tb_addr_zs(J, (xB - 1) << sh, (yB + y) << sh) <= cur;
      if (constrained && !is_intra(J, (xB - 1) << sh, (yB + y) << sh)) ok = 0;
      if (ok) {
        if (!nAvail) first = img[xB - 1 + (yB + y) * stride];
        for (int i = 0; i < 4; i++) { av[-y + i - 1] = 1; b[-y + i - 1] = img[xB - 1 + (yB + y - i) * stride]; }
        nAvail += 4;
      }
    }
  if (availTL) {
    int ok = min_tb_addr_zs(J, (xB - 1) << sh, (yB - 1) << sh) <= cur;
    if (constrained && !is_intra(J, (xB - 1) << sh, (yB - 1) << sh)) ok = 0;
    if (ok) {
      if (!nAvail) first = img[xB - 1 + (yB - 1) * stride];
      b[0] = img[xB - 1 + (yB - 1) * stride];
      av[0] = 1;
      nAvail++;
    }
  }
  for (int x = 0; x < nRight; x += 4) {
    int border_ok = x < nT ? availT : availTR;
    if (border_ok) {
      int ok = min_tb_addr_zs(J, (xB + x) << sh, (yB - 1) << sh) <= cur;
      if (constrained && !is_intra(J, (xB + x) << sh, (yB - 1) << sh)) ok = 0;
      if (ok) {
        if (!nAvail) first = img[xB + x + (yB - 1) * stride];
        for (int i = 0; i < 4; i++) { b[x + i + 1] = img[xB + x + i + (yB - 1) * stride]; av[x + i + 1] = 1; }
        nAvail += 4;
      }
    }
  }
  if (nAvail != 4 * nT + 1) {
    if (nAvail == 0) memset(b - 2 * nT, 128, 4 * nT + 1);
    else {
      if (!av[-2 * nT]) b[-2 * nT] = (uint8_t)first;
      for (int i = -2 * nT + 1; i <= 2 * nT; i++)
        if (!av[i]) b[i] = b[i - 1];
    }
  }
  orc_intra_from_border(log2, cidx, tu->intra_mode, b, (J->flags & B265_FRAME_STRONG_INTRA_SMOOTHING) != 0,
                        (tu->flags & B265_TU_NO_EDGE_FILT) != 0, planes[cidx] + xB + yB * stride, stride);
}

/* the six call sites in decoding order: per TU predict (S1) then residual (S2); PUs first (S3) */
void orc_reconstruct_frame(const b265_frame_job* J, uint8_t* const* planes, const int* pitch,
                           const uint8_t* const* ref_planes) {
  if (J->n_pus > 0)
    orc_mc_batch(J->pus, J->n_pus, J->slices, ref_planes, J->n_refs, planes, J->width, J->height, pitch[0], pitch[1]);
  /* S4: PCM blocks (read_pcm_samples, slice.cc:4356-4406; samples are already at the bit depth) */
  for (int i = 0; i < J->n_pcms; i++) {
    const b265_pcm* p = &J->pcms[i];
    const uint8_t* s = J->pcm_samples + p->sample_off;
    int n = 1 << p->log2_size;
    for (int y = 0; y < n; y++)
      for (int x = 0; x < n; x++) planes[0][(size_t)(p->y + y) * pitch[0] + p->x + x] = *s++;
    if (!J->chroma_format_idc) continue;
    for (int c = 1; c < 3; c++)
      for (int y = 0; y < n / 2; y++)
        for (int x = 0; x < n / 2; x++) planes[c][(size_t)(p->y / 2 + y) * pitch[c] + p->x / 2 + x] = *s++;
  }
  const b265_scaling* sl = (J->flags & B265_FRAME_SCALING_LIST) ? J->scaling : NULL;
  for (int i = 0; i < J->n_tus; i++) {
    const b265_tu* tu = &J->tus[i];
    int cidx = tu->cidx_log2 & 3;
    if (tu->intra_mode != B265_INTRA_NONE) orc_intra_tu(J, tu, planes, pitch);
    orc_residual_tu(tu, J->coeffs, sl, planes[cidx], pitch[cidx]);
  }
}
