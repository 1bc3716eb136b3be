/*
 * oracle/filters.c — CPU restatement of the in-loop filters.  TEST INFRASTRUCTURE ONLY (see oracle.h).
 * Deblocking: libde265/deblock.cc derive_boundaryStrength (:241-375), edge_filtering_luma_internal
 * (:405-699), edge_filtering_chroma_internal (:730-871), driver apply_deblocking_filter (:1020-1058).
 * SAO: libde265/sao.cc apply_sao_internal (:29-254), apply_sample_adaptive_offset_sequential (:318-370).
 * Edge flags arrive precomputed in job->unit_info (the parser's derive_edgeFlags equivalent).
 */
#include <stdlib.h>
#include <string.h>
#include "oracle.h"

static int clip3(int lo, int hi, int v) { return v < lo ? lo : (v > hi ? hi : v); }
static uint8_t clip8(int v) { return (uint8_t)clip3(0, 255, v); }

static const uint8_t kBeta[52] = {0,  0,  0,  0,  0,  0,  0,  0,  0,  0,  0,  0,  0,  0,  0,  0,  6,  7,
                                  8,  9,  10, 11, 12, 13, 14, 15, 16, 17, 18, 20, 22, 24, 26, 28, 30, 32,
                                  34, 36, 38, 40, 42, 44, 46, 48, 50, 52, 54, 56, 58, 60, 62, 64};
static const uint8_t kTc[54] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0,  0,  0,  0,  0,
                                1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 3,  3,  3,  3,  4,
                                4, 4, 5, 5, 6, 6, 7, 8, 9, 10, 11, 13, 14, 16, 18, 20, 22, 24};
static const uint8_t kQpc[13] = {29, 30, 31, 32, 33, 33, 34, 34, 35, 35, 36, 36, 37};

typedef struct { int flags, slot[2]; int mv[2][2]; } Motion;

static Motion* raster_motion(const b265_frame_job* J) {
  int uw = J->width >> 2, uh = J->height >> 2;
  Motion* m = (Motion*)calloc((size_t)uw * uh, sizeof(Motion));
  for (int i = 0; i < J->n_pus; i++) {
    const b265_pu* pu = &J->pus[i];
    for (int y = pu->y >> 2; y < (pu->y + pu->h) >> 2; y++)
      for (int x = pu->x >> 2; x < (pu->x + pu->w) >> 2; x++) {
        Motion* e = &m[y * uw + x];
        e->flags = pu->flags & 3;
        for (int l = 0; l < 2; l++) { e->slot[l] = pu->ref_slot[l]; e->mv[l][0] = pu->mv[l][0]; e->mv[l][1] = pu->mv[l][1]; }
      }
  }
  return m;
}
static int far4(const int* a, const int* b) { return abs(a[0] - b[0]) >= 4 || abs(a[1] - b[1]) >= 4; }

static int edge_bs(const b265_frame_job* J, const Motion* mo, int up, int uq, int tu_edge) {
  uint8_t ip = J->unit_info[up], iq = J->unit_info[uq];
  if ((ip | iq) & B265_UNIT_INTRA) return 2;
  if (tu_edge && ((ip | iq) & B265_UNIT_CBF)) return 1;
  if (!mo) return 0;
  const Motion *P = &mo[up], *Q = &mo[uq];
  int rp0 = (P->flags & 1) ? P->slot[0] : -1, rp1 = (P->flags & 2) ? P->slot[1] : -1;
  int rq0 = (Q->flags & 1) ? Q->slot[0] : -1, rq1 = (Q->flags & 2) ? Q->slot[1] : -1;
  if (!((rp0 == rq0 && rp1 == rq1) || (rp0 == rq1 && rp1 == rq0))) return 1;
  int z[2] = {0, 0};
  const int* p0 = (P->flags & 1) ? P->mv[0] : z; const int* p1 = (P->flags & 2) ? P->mv[1] : z;
  const int* q0 = (Q->flags & 1) ? Q->mv[0] : z; const int* q1 = (Q->flags & 2) ? Q->mv[1] : z;
  if (rp0 != rp1) {
    if (rp0 == rq0) return far4(p0, q0) || far4(p1, q1);
    return far4(p0, q1) || far4(p1, q0);
  }
  return (far4(p0, q0) || far4(p1, q1)) && (far4(p0, q1) || far4(p1, q0));
}

void orc_boundary_strength(const b265_frame_job* J, uint8_t* bs) {
  int uw = J->width >> 2, uh = J->height >> 2;
  memset(bs, 0, (size_t)uw * uh);
  if (!J->unit_info) return;
  Motion* mo = J->n_pus > 0 ? raster_motion(J) : NULL;
  for (int uy = 0; uy < uh; uy++)
    for (int ux = 0; ux < uw; ux++) {
      int i = uy * uw + ux, v = 0;
      uint8_t info = J->unit_info[i];
      if (!(ux & 1) && ux > 0 && (info & B265_UNIT_EDGE_V)) v |= edge_bs(J, mo, i - 1, i, info & B265_UNIT_TU_EDGE_V);
      if (!(uy & 1) && uy > 0 && (info & B265_UNIT_EDGE_H)) v |= edge_bs(J, mo, i - uw, i, info & B265_UNIT_TU_EDGE_H) << 2;
      bs[i] = (uint8_t)v;
    }
  free(mo);
}

static int qp_at(const b265_frame_job* J, int x, int y) {
  return J->qp_y[(y >> J->log2_min_cb) * (J->width >> J->log2_min_cb) + (x >> J->log2_min_cb)];
}
static const b265_slice* slice_at(const b265_frame_job* J, int x, int y) {
  int w = (J->width + (1 << J->log2_ctb) - 1) >> J->log2_ctb;
  int ctb = (y >> J->log2_ctb) * w + (x >> J->log2_ctb);
  return &J->slices[J->ctbs ? J->ctbs[ctb].slice : 0];
}
static int nofilt(const b265_frame_job* J, int x, int y) {
  return (J->unit_info[(y >> 2) * (J->width >> 2) + (x >> 2)] & B265_UNIT_NOFILT) != 0;
}

static void luma_pass(const b265_frame_job* J, const uint8_t* bsmap, uint8_t* img, int stride, int vertical) {
  int uw = J->width >> 2, uh = J->height >> 2;
  for (int y = 0; y < uh; y += vertical ? 1 : 2)
    for (int x = 0; x < uw; x += vertical ? 2 : 1) {
      int xDi = x << 2, yDi = y << 2;
      int bS = (bsmap[y * uw + x] >> (vertical ? 0 : 2)) & 3;
      if (!bS) continue;
      uint8_t* ptr = img + yDi * stride + xDi;
      int q[4][4], p[4][4];
      for (int k = 0; k < 4; k++)
        for (int i = 0; i < 4; i++) {
          if (vertical) { q[k][i] = ptr[i + k * stride]; p[k][i] = ptr[-i - 1 + k * stride]; }
          else { q[k][i] = ptr[k + i * stride]; p[k][i] = ptr[k - (i + 1) * stride]; }
        }
      int xp = vertical ? xDi - 1 : xDi, yp = vertical ? yDi : yDi - 1;
      int qP_L = (qp_at(J, xDi, yDi) + qp_at(J, xp, yp) + 1) >> 1;
      const b265_slice* sl = slice_at(J, xDi, yDi);
      int beta = kBeta[clip3(0, 51, qP_L + sl->beta_offset)];
      int tc = kTc[clip3(0, 53, qP_L + 2 * (bS - 1) + sl->tc_offset)];
      int dp0 = abs(p[0][2] - 2 * p[0][1] + p[0][0]), dp3 = abs(p[3][2] - 2 * p[3][1] + p[3][0]);
      int dq0 = abs(q[0][2] - 2 * q[0][1] + q[0][0]), dq3 = abs(q[3][2] - 2 * q[3][1] + q[3][0]);
      int dpq0 = dp0 + dq0, dpq3 = dp3 + dq3, dp = dp0 + dp3, dq = dq0 + dq3, d = dpq0 + dpq3;
      if (d >= beta) continue;
      int dSam0 = 2 * dpq0 < (beta >> 2) && abs(p[0][3] - p[0][0]) + abs(q[0][0] - q[0][3]) < (beta >> 3) &&
                  abs(p[0][0] - q[0][0]) < ((5 * tc + 1) >> 1);
      int dSam3 = 2 * dpq3 < (beta >> 2) && abs(p[3][3] - p[3][0]) + abs(q[3][0] - q[3][3]) < (beta >> 3) &&
                  abs(p[3][0] - q[3][0]) < ((5 * tc + 1) >> 1);
      int dE = (dSam0 && dSam3) ? 2 : 1;
      int dEp = dp < ((beta + (beta >> 1)) >> 3), dEq = dq < ((beta + (beta >> 1)) >> 3);
      int filterP = !nofilt(J, xp, yp), filterQ = !nofilt(J, xDi, yDi);
      for (int k = 0; k < 4; k++) {
        int p0 = p[k][0], p1 = p[k][1], p2 = p[k][2], p3 = p[k][3], q0 = q[k][0], q1 = q[k][1], q2 = q[k][2], q3 = q[k][3];
#define PP(i) (vertical ? ptr[-(i) - 1 + k * stride] : ptr[k - ((i) + 1) * stride])
#define QQ(i) (vertical ? ptr[(i) + k * stride] : ptr[k + (i) * stride])
        if (dE == 2) {
          int pn[3], qn[3];
          pn[0] = clip3(p0 - 2 * tc, p0 + 2 * tc, (p2 + 2 * p1 + 2 * p0 + 2 * q0 + q1 + 4) >> 3);
          pn[1] = clip3(p1 - 2 * tc, p1 + 2 * tc, (p2 + p1 + p0 + q0 + 2) >> 2);
          pn[2] = clip3(p2 - 2 * tc, p2 + 2 * tc, (2 * p3 + 3 * p2 + p1 + p0 + q0 + 4) >> 3);
          qn[0] = clip3(q0 - 2 * tc, q0 + 2 * tc, (p1 + 2 * p0 + 2 * q0 + 2 * q1 + q2 + 4) >> 3);
          qn[1] = clip3(q1 - 2 * tc, q1 + 2 * tc, (p0 + q0 + q1 + q2 + 2) >> 2);
          qn[2] = clip3(q2 - 2 * tc, q2 + 2 * tc, (p0 + q0 + q1 + 3 * q2 + 2 * q3 + 4) >> 3);
          for (int i = 0; i < 3; i++) {
            if (filterP) { if (vertical) ptr[-i - 1 + k * stride] = (uint8_t)pn[i]; else ptr[k - (i + 1) * stride] = (uint8_t)pn[i]; }
            if (filterQ) { if (vertical) ptr[i + k * stride] = (uint8_t)qn[i]; else ptr[k + i * stride] = (uint8_t)qn[i]; }
          }
        } else {
          int delta = (9 * (q0 - p0) - 3 * (q1 - p1) + 8) >> 4;
          if (abs(delta) < tc * 10) {
            delta = clip3(-tc, tc, delta);
            uint8_t np0 = clip8(p0 + delta), nq0 = clip8(q0 - delta);
            if (filterP) { if (vertical) ptr[-1 + k * stride] = np0; else ptr[k - stride] = np0; }
            if (filterQ) { if (vertical) ptr[k * stride] = nq0; else ptr[k] = nq0; }
            if (dEp && filterP) {
              uint8_t v = clip8(p1 + clip3(-(tc >> 1), tc >> 1, (((p2 + p0 + 1) >> 1) - p1 + delta) >> 1));
              if (vertical) ptr[-2 + k * stride] = v; else ptr[k - 2 * stride] = v;
            }
            if (dEq && filterQ) {
              uint8_t v = clip8(q1 + clip3(-(tc >> 1), tc >> 1, (((q2 + q0 + 1) >> 1) - q1 - delta) >> 1));
              if (vertical) ptr[1 + k * stride] = v; else ptr[k + stride] = v;
            }
          }
        }
#undef PP
#undef QQ
      }
    }
}

static void chroma_pass(const b265_frame_job* J, const uint8_t* bsmap, uint8_t* const* planes, const int* pitch,
                        int vertical) {
  int uw = J->width >> 2, uh = J->height >> 2;
  int xIncr = (vertical ? 2 : 1) * 2, yIncr = (vertical ? 1 : 2) * 2;
  for (int y = 0; y < uh; y += yIncr)
    for (int x = 0; x < uw; x += xIncr) {
      int xDi = x << 1, yDi = y << 1; /* chroma position */
      int bS = (bsmap[y * uw + x] >> (vertical ? 0 : 2)) & 3;
      if (bS <= 1) continue;
      for (int c = 1; c <= 2; c++) {
        int stride = pitch[c];
        uint8_t* ptr = planes[c] + yDi * stride + xDi;
        int xl = 2 * xDi, yl = 2 * yDi, xp = vertical ? xl - 1 : xl, yp = vertical ? yl : yl - 1;
        int qP_i = ((qp_at(J, xl, yl) + qp_at(J, xp, yp) + 1) >> 1) + (c == 1 ? J->cb_qp_offset : J->cr_qp_offset);
        int QP_C = qP_i < 30 ? qP_i : (qP_i >= 43 ? qP_i - 6 : kQpc[qP_i - 30]);
        int tc = kTc[clip3(0, 53, QP_C + 2 * (bS - 1) + slice_at(J, xl, yl)->tc_offset)];
        int filterP = !nofilt(J, xp, yp), filterQ = !nofilt(J, xl, yl);
        for (int k = 0; k < 4; k++) {
          uint8_t* p0 = vertical ? ptr - 1 + k * stride : ptr + k - stride;
          uint8_t* p1 = vertical ? ptr - 2 + k * stride : ptr + k - 2 * stride;
          uint8_t* q0 = vertical ? ptr + k * stride : ptr + k;
          uint8_t* q1 = vertical ? ptr + 1 + k * stride : ptr + k + stride;
          int delta = clip3(-tc, tc, ((((*q0 - *p0) << 2) + *p1 - *q1 + 4) >> 3));
          int np = *p0 + delta, nq = *q0 - delta;
          if (filterP) *p0 = clip8(np);
          if (filterQ) *q0 = clip8(nq);
        }
      }
    }
}

void orc_deblock(const b265_frame_job* J, uint8_t* const* planes, const int* pitch) {
  if (!(J->flags & B265_FRAME_DEBLOCK) || !J->unit_info) return;
  int uw = J->width >> 2, uh = J->height >> 2;
  uint8_t* bs = (uint8_t*)malloc((size_t)uw * uh);
  orc_boundary_strength(J, bs);
  luma_pass(J, bs, planes[0], pitch[0], 1);
  if (J->chroma_format_idc) chroma_pass(J, bs, planes, pitch, 1);
  luma_pass(J, bs, planes[0], pitch[0], 0);
  if (J->chroma_format_idc) chroma_pass(J, bs, planes, pitch, 0);
  free(bs);
}
