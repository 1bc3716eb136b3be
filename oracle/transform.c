/*
 * oracle/transform.c — CPU restatement of the reference's dequantisation and inverse transforms.
 * TEST INFRASTRUCTURE ONLY (see oracle.h).  Direct matrix form, exactly the arithmetic of
 *   libde265/fallback-dct.cc:552-692 (transform_idct_add), :270-337 (transform_4x4_luma_add_8),
 *   :80-90 (transform_skip_residual), :160-224 (bypass / rdpcm), libde265/transform.cc:353-625.
 */
#include <string.h>
#include "oracle.h"

static int clip3(int lo, int hi, int v) { return v < lo ? lo : (v > hi ? hi : v); }
static uint8_t clip8(int v) { return (uint8_t)clip3(0, 255, v); }

/* first 16 columns of the odd rows are enough to define the matrix; we store the magnitudes
 * cos-table (angle m*pi/64 -> integer) and fold by symmetry. */
static const int8_t kMag[33] = {64, 90, 90, 90, 89, 88, 87, 85, 83, 82, 80, 78, 75, 73, 70, 67, 64,
                                61, 57, 54, 50, 46, 43, 38, 36, 31, 25, 22, 18, 13, 9,  4,  0};
static int8_t g_mat[32][32];
static int g_mat_ready = 0;
static void build_matrix(void) {
  if (g_mat_ready) return;
  for (int k = 0; k < 32; k++)
    for (int n = 0; n < 32; n++) {
      int m = (k * (2 * n + 1)) % 128;
      if (m > 64) m = 128 - m;
      g_mat[k][n] = (int8_t)((m > 32) ? -kMag[64 - m] : kMag[m]);
    }
  g_mat_ready = 1;
}
/* exported for the matrix self-check in tests */
const int8_t* orc_dct_matrix(void) { build_matrix(); return &g_mat[0][0]; }

static const int kLevelScale[6] = {40, 45, 51, 57, 64, 72};

void orc_dequant_flat(int16_t* out, const int16_t* in, int log2, int qp) {
  int shift = 8 + log2 - 5 - 4;
  int offset = 1 << (shift - 1);
  int fact = kLevelScale[qp % 6] << (qp / 6);
  for (int i = 0; i < (1 << (2 * log2)); i++)
    out[i] = (int16_t)clip3(-32768, 32767, (in[i] * fact + offset) >> shift);
}

void orc_idct_add(uint8_t* dst, ptrdiff_t stride, int nT, const int16_t* coeffs) {
  build_matrix();
  int16_t g[32 * 32];
  int fact = 32 / nT;
  for (int c = 0; c < nT; c++)
    for (int i = 0; i < nT; i++) {
      int sum = 0;
      for (int j = 0; j < nT; j++) sum += g_mat[fact * j][i] * coeffs[c + j * nT];
      g[c + i * nT] = (int16_t)clip3(-32768, 32767, (sum + 64) >> 7);
    }
  for (int y = 0; y < nT; y++)
    for (int i = 0; i < nT; i++) {
      int sum = 0;
      for (int j = 0; j < nT; j++) sum += g_mat[fact * j][i] * g[y * nT + j];
      int out = (sum + (1 << 11)) >> 12; /* not clipped to 16 bit (fallback-dct.cc:681-682) */
      dst[y * stride + i] = clip8(dst[y * stride + i] + out);
    }
}

void orc_idst4_add(uint8_t* dst, ptrdiff_t stride, const int16_t* coeffs) {
  static const int8_t M[4][4] = {{29, 55, 74, 84}, {74, 74, 0, -74}, {84, -29, -74, 55}, {55, -84, 74, -29}};
  int16_t g[4][4];
  for (int c = 0; c < 4; c++)
    for (int i = 0; i < 4; i++) {
      int sum = 0;
      for (int j = 0; j < 4; j++) sum += M[j][i] * coeffs[c + j * 4];
      g[i][c] = (int16_t)clip3(-32768, 32767, (sum + 64) >> 7);
    }
  for (int y = 0; y < 4; y++)
    for (int i = 0; i < 4; i++) {
      int sum = 0;
      for (int j = 0; j < 4; j++) sum += M[j][i] * g[y][j];
      int out = clip3(-32768, 32767, (sum + (1 << 11)) >> 12);
      dst[y * stride + i] = clip8(dst[y * stride + i] + out);
    }
}

static void add_residual(uint8_t* dst, ptrdiff_t stride, int nT, const int32_t* r) {
  for (int y = 0; y < nT; y++)
    for (int x = 0; x < nT; x++) dst[y * stride + x] = clip8(dst[y * stride + x] + r[y * nT + x]);
}

void orc_skip_add(uint8_t* dst, ptrdiff_t stride, int nT, const int16_t* coeffs, int rdpcm) {
  int log2 = 0;
  while ((1 << log2) < nT) log2++;
  int tsShift = 5 + log2, bdShift = 12, rnd = 1 << (bdShift - 1);
  int32_t r[32 * 32];
  if (rdpcm == 0) {
    for (int i = 0; i < nT * nT; i++) r[i] = ((coeffs[i] << tsShift) + rnd) >> bdShift;
  } else if (rdpcm == 1) {
    for (int y = 0; y < nT; y++) {
      int sum = 0;
      for (int x = 0; x < nT; x++) { sum += ((coeffs[x + y * nT] << tsShift) + rnd) >> bdShift; r[y * nT + x] = sum; }
    }
  } else {
    for (int x = 0; x < nT; x++) {
      int sum = 0;
      for (int y = 0; y < nT; y++) { sum += ((coeffs[x + y * nT] << tsShift) + rnd) >> bdShift; r[y * nT + x] = sum; }
    }
  }
  add_residual(dst, stride, nT, r);
}

void orc_bypass_add(uint8_t* dst, ptrdiff_t stride, int nT, const int16_t* coeffs, int rdpcm) {
  int32_t r[32 * 32];
  if (rdpcm == 0) {
    for (int i = 0; i < nT * nT; i++) r[i] = coeffs[i];
  } else if (rdpcm == 1) {
    for (int y = 0; y < nT; y++) { int s = 0; for (int x = 0; x < nT; x++) { s += coeffs[x + y * nT]; r[y * nT + x] = s; } }
  } else {
    for (int x = 0; x < nT; x++) { int s = 0; for (int y = 0; y < nT; y++) { s += coeffs[x + y * nT]; r[y * nT + x] = s; } }
  }
  add_residual(dst, stride, nT, r);
}

void orc_residual_batch(int kind, int log2, int rdpcm, int n_blocks, const int16_t* coeffs,
                        const uint8_t* qp, uint8_t* dst) {
  int nT = 1 << log2, n2 = nT * nT;
  int16_t tmp[32 * 32];
  for (int i = 0; i < n_blocks; i++) {
    const int16_t* c = coeffs + (size_t)i * n2;
    if (qp) { orc_dequant_flat(tmp, c, log2, qp[i]); c = tmp; }
    uint8_t* d = dst + (size_t)i * n2;
    switch (kind) {
      case B265_TR_DCT: orc_idct_add(d, nT, nT, c); break;
      case B265_TR_DST: orc_idst4_add(d, nT, c); break;
      case B265_TR_SKIP: orc_skip_add(d, nT, nT, c, rdpcm); break;
      default: orc_bypass_add(d, nT, nT, c, rdpcm); break;
    }
  }
}

/* scale_coefficients_internal (transform.cc:353-625) for one work-list TU */
void orc_residual_tu(const b265_tu* tu, const b265_coeff* coeffs, const b265_scaling* sl,
                     uint8_t* plane, int pitch) {
  if (tu->ncoeff == 0) return;
  int log2 = (tu->cidx_log2 >> 2) & 7, nT = 1 << log2, cidx = tu->cidx_log2 & 3;
  int16_t buf[32 * 32];
  memset(buf, 0, sizeof(int16_t) * nT * nT);
  const b265_coeff* c = coeffs + tu->coeff_off;
  int qp = tu->qp;
  for (int i = 0; i < tu->ncoeff; i++) {
    int v;
    if (tu->flags & B265_TU_BYPASS) v = c[i].level;
    else if (sl) {
      int mid = cidx + ((tu->flags & B265_TU_INTRA_CU) ? 0 : (nT < 32 ? 3 : 1));
      const uint8_t* m = log2 == 2 ? sl->s4[mid] : log2 == 3 ? sl->s8[mid] : log2 == 4 ? sl->s16[mid] : sl->s32[mid];
      int shift = 8 + log2 - 5;
      int64_t fact = (int64_t)(m[c[i].pos] * kLevelScale[qp % 6]) << (qp / 6);
      int64_t t = ((int64_t)c[i].level * fact + ((int64_t)1 << (shift - 1))) >> shift;
      v = (int)(t < -32768 ? -32768 : (t > 32767 ? 32767 : t));
    } else {
      int shift = 8 + log2 - 9;
      int fact = kLevelScale[qp % 6] << (qp / 6);
      v = clip3(-32768, 32767, (c[i].level * fact + (1 << (shift - 1))) >> shift);
    }
    buf[c[i].pos] = (int16_t)v;
  }
  int skipby = tu->flags & (B265_TU_SKIP | B265_TU_BYPASS);
  if (skipby && (tu->flags & B265_TU_ROTATE)) { /* rotate_coefficients_fallback, fallback-dct.cc:251-257 */
    for (int i = 0; i < nT * nT / 2; i++) { int16_t t = buf[i]; buf[i] = buf[nT * nT - 1 - i]; buf[nT * nT - 1 - i] = t; }
  }
  uint8_t* d = plane + (size_t)tu->y * pitch + tu->x;
  int rd = (tu->flags & B265_TU_RDPCM_H) ? 1 : (tu->flags & B265_TU_RDPCM_V) ? 2 : 0;
  if (tu->flags & B265_TU_BYPASS) orc_bypass_add(d, pitch, nT, buf, rd);
  else if (tu->flags & B265_TU_SKIP) orc_skip_add(d, pitch, nT, buf, rd);
  else if (tu->flags & B265_TU_DST) orc_idst4_add(d, pitch, buf);
  else orc_idct_add(d, pitch, nT, buf);
}
