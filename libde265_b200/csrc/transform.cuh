// transform.cuh — dequantisation + inverse DST/DCT/skip/bypass as warp-cooperative device code.
//
// Reference behaviour restated (not copied): libde265/transform.cc:353-625 (scale_coefficients_internal),
// libde265/fallback-dct.cc:270-337 (DST 4x4), :552-692 (IDCT, direct matrix form), :80-90 / :160-224
// (skip, bypass, rdpcm).  The reference multiplies by the full 32x32 matrix; here every 1-D transform
// is an even/odd partial butterfly whose integer sums are identical (no intermediate rounding, no
// overflow: |sum| <= 32*90*32768 < 2^31), so results are bit-exact.
//
// Mapping: a transform block of size N=2^LOG2 is processed by a GROUP of N lanes (32/N blocks per
// warp). Pass 1 (columns): lane c owns column c.  Pass 2 (rows): lane y owns row y and ends up with
// the N residuals of its row in registers, ready to be added to the prediction row and stored as one
// contiguous N-byte segment.  The block lives in shared memory with row stride N+2 int16 which makes
// both the column reads and the row reads bank-conflict free for all groups of a warp.
#pragma once
#include "common.cuh"

__host__ __device__ constexpr int dct_T(int m) {
  // HEVC core transform magnitudes at angle m*pi/64 (table of mat_dct, fallback-dct.cc:513-546)
  constexpr int t[33] = {64, 90, 90, 90, 89, 88, 87, 85, 83, 82, 80, 78, 75, 73, 70, 67, 64,
                         61, 57, 54, 50, 46, 43, 38, 36, 31, 25, 22, 18, 13, 9,  4,  0};
  return t[m];
}
// element [k32][n] of the 32-point matrix (row k32 = basis index, n = sample index)
__host__ __device__ constexpr int dct_coef(int k32, int n) {
  int m = (k32 * (2 * n + 1)) % 128;
  if (m > 64) m = 128 - m;
  return (m > 32) ? -dct_T(64 - m) : dct_T(m);
}

// Inverse N-point transform of x[0..L) (x[k>=L] == 0): out[n] = sum_k mat[k*(32/N)][n] * x[k]
template <int N, int L>
struct Idct {
  static __device__ __forceinline__ void run(const int (&x)[N], int (&out)[N]) {
    constexpr int H = N / 2;
    constexpr int F = 32 / N;
    int e_in[H], E[H];
#pragma unroll
    for (int k = 0; k < H; k++) e_in[k] = x[2 * k];
    Idct<H, (L + 1) / 2>::run(e_in, E);
#pragma unroll
    for (int n = 0; n < H; n++) {
      int o = 0;
#pragma unroll
      for (int k = 1; k < L; k += 2) o += dct_coef(k * F, n) * x[k];
      out[n] = E[n] + o;
      out[N - 1 - n] = E[n] - o;
    }
  }
};
template <int L>
struct Idct<2, L> {
  static __device__ __forceinline__ void run(const int (&x)[2], int (&out)[2]) {
    int a = 64 * x[0];
    int b = (L > 1) ? 64 * x[1] : 0;
    out[0] = a + b;
    out[1] = a - b;
  }
};

template <int LOG2>
struct TrBlk {
  static constexpr int N = 1 << LOG2;
  static constexpr int S = N + 2;       // int16 row stride in shared memory
  static constexpr int ELEMS = N * S;   // int16 elements per block
};
// shared bytes one warp needs for its 32/N blocks (largest: one 32x32 block)
#define B265_TR_SMEM_PER_WARP (32 * 34 * 2)

__device__ __forceinline__ int dequant_flat(int level, int qp, int log2) {
  // transform.cc:443-475 with m_x_y folded into the shift (bit depth 8)
  const int ls[6] = {40, 45, 51, 57, 64, 72};
  int shift = 8 + log2 - 9;
  int fact = ls[qp % 6] << (qp / 6);
  return clip16((level * fact + (1 << (shift - 1))) >> shift);
}
__device__ __forceinline__ int dequant_list(int level, int qp, int log2, int m) {
  // transform.cc:477-510 (64-bit because m can be 255)
  const int ls[6] = {40, 45, 51, 57, 64, 72};
  int shift = 8 + log2 - 5;
  long long fact = (long long)(m * ls[qp % 6]) << (qp / 6);
  long long v = ((long long)level * fact + (1ll << (shift - 1))) >> shift;
  return (int)min(max(v, -32768ll), 32767ll);
}

// ---- pass 1 / pass 2 of the 2-D inverse DCT for the lane `sub` of a group ----
template <int LOG2, int L>
__device__ __forceinline__ void idct_cols(int16_t* blk, int sub) {
  constexpr int N = TrBlk<LOG2>::N, S = TrBlk<LOG2>::S;
  int x[N], o[N];
#pragma unroll
  for (int k = 0; k < N; k++) x[k] = (k < L) ? blk[k * S + sub] : 0;
  Idct<N, L>::run(x, o);
#pragma unroll
  for (int n = 0; n < N; n++) blk[n * S + sub] = (int16_t)clip16((o[n] + 64) >> 7);
}
template <int LOG2, int L>
__device__ __forceinline__ void idct_rows(const int16_t* blk, int sub, int (&r)[1 << LOG2]) {
  constexpr int N = TrBlk<LOG2>::N, S = TrBlk<LOG2>::S;
  int x[N];
  const uint32_t* row = reinterpret_cast<const uint32_t*>(blk + sub * S);
#pragma unroll
  for (int k = 0; k < N; k += 2) {
    if (k < L) {
      uint32_t w = row[k >> 1];
      x[k] = (int16_t)(w & 0xffff);
      x[k + 1] = (int16_t)(w >> 16);
    } else {
      x[k] = 0;
      x[k + 1] = 0;
    }
  }
  Idct<N, L>::run(x, r);
#pragma unroll
  for (int n = 0; n < N; n++) r[n] = (r[n] + 2048) >> 12;  // 20-bitDepth = 12, not clipped (fallback-dct.cc:681)
}

template <int LOG2>
__device__ __forceinline__ void idct_cols_dispatch(int16_t* blk, int sub, int nrows) {
  constexpr int N = 1 << LOG2;
  if (N >= 32 && nrows > 16) idct_cols<LOG2, (N >= 32 ? 32 : N)>(blk, sub);
  else if (N >= 16 && nrows > 8) idct_cols<LOG2, (N >= 16 ? 16 : N)>(blk, sub);
  else if (N >= 8 && nrows > 4) idct_cols<LOG2, (N >= 8 ? 8 : N)>(blk, sub);
  else idct_cols<LOG2, 4>(blk, sub);
}
template <int LOG2>
__device__ __forceinline__ void idct_rows_dispatch(const int16_t* blk, int sub, int ncols,
                                                   int (&r)[1 << LOG2]) {
  constexpr int N = 1 << LOG2;
  if (N >= 32 && ncols > 16) idct_rows<LOG2, (N >= 32 ? 32 : N)>(blk, sub, r);
  else if (N >= 16 && ncols > 8) idct_rows<LOG2, (N >= 16 ? 16 : N)>(blk, sub, r);
  else if (N >= 8 && ncols > 4) idct_rows<LOG2, (N >= 8 ? 8 : N)>(blk, sub, r);
  else idct_rows<LOG2, 4>(blk, sub, r);
}

// 4x4 inverse DST-VII, both passes clipped to 16 bit (fallback-dct.cc:297,328)
__device__ __forceinline__ void idst4_cols(int16_t* blk, int sub) {
  constexpr int S = TrBlk<2>::S;
  int c0 = blk[0 * S + sub], c1 = blk[1 * S + sub], c2 = blk[2 * S + sub], c3 = blk[3 * S + sub];
  blk[0 * S + sub] = (int16_t)clip16((29 * c0 + 74 * c1 + 84 * c2 + 55 * c3 + 64) >> 7);
  blk[1 * S + sub] = (int16_t)clip16((55 * c0 + 74 * c1 - 29 * c2 - 84 * c3 + 64) >> 7);
  blk[2 * S + sub] = (int16_t)clip16((74 * c0 - 74 * c2 + 74 * c3 + 64) >> 7);
  blk[3 * S + sub] = (int16_t)clip16((84 * c0 - 74 * c1 + 55 * c2 - 29 * c3 + 64) >> 7);
}
__device__ __forceinline__ void idst4_rows(const int16_t* blk, int sub, int (&r)[4]) {
  constexpr int S = TrBlk<2>::S;
  int c0 = blk[sub * S + 0], c1 = blk[sub * S + 1], c2 = blk[sub * S + 2], c3 = blk[sub * S + 3];
  r[0] = clip16((29 * c0 + 74 * c1 + 84 * c2 + 55 * c3 + 2048) >> 12);
  r[1] = clip16((55 * c0 + 74 * c1 - 29 * c2 - 84 * c3 + 2048) >> 12);
  r[2] = clip16((74 * c0 - 74 * c2 + 74 * c3 + 2048) >> 12);
  r[3] = clip16((84 * c0 - 74 * c1 + 55 * c2 - 29 * c3 + 2048) >> 12);
}

// Residual of the block held (dequantised, or raw for bypass) in `blk`, for lane `sub` of the group:
// fills r[0..N) with the residuals of ROW `sub`.  `flags` are B265_TU_*; nrows/ncols bound the
// non-zero region (nrows = last non-zero row + 1).  All lanes of the warp must call this together
// (it contains warp-level barriers); lanes of idle groups pass active=false.
// RDPCM_V is handled by the caller (column-wise accumulation), see residual_apply().
template <int LOG2>
__device__ __forceinline__ void residual_rows(int16_t* blk, int sub, int flags, int nrows, int ncols,
                                              bool active, int (&r)[1 << LOG2]) {
  constexpr int N = TrBlk<LOG2>::N, S = TrBlk<LOG2>::S;
  const bool tr = !(flags & (B265_TU_SKIP | B265_TU_BYPASS));
  // every lane of the warp executes the same three barriers (groups may take different branches)
  __syncwarp();
  if (active && tr) {
    if (LOG2 == 2 && (flags & B265_TU_DST)) idst4_cols(blk, sub);
    else if (sub < ncols) idct_cols_dispatch<LOG2>(blk, sub, nrows);
  }
  __syncwarp();
  if (active) {
    if (!tr) {
      int acc = 0;
#pragma unroll
      for (int x = 0; x < N; x++) {
        int c = blk[sub * S + x];
        int v = (flags & B265_TU_BYPASS) ? c : (((c << (5 + LOG2)) + 2048) >> 12);
        if (flags & B265_TU_RDPCM_H) { acc += v; v = acc; }
        r[x] = v;
      }
    } else if (LOG2 == 2 && (flags & B265_TU_DST)) {
      int r4[4];
      idst4_rows(blk, sub, r4);
#pragma unroll
      for (int x = 0; x < N; x++) r[x] = r4[x & 3];
    } else {
      idct_rows_dispatch<LOG2>(blk, sub, ncols, r);
    }
  }
  __syncwarp();
}

// pack/unpack helpers for one row of N pixels held as ints
template <int N>
__device__ __forceinline__ void load_row_u8(const uint8_t* p, int (&v)[N]) {
  if (N == 4) {
    uint32_t w = *reinterpret_cast<const uint32_t*>(p);
#pragma unroll
    for (int i = 0; i < 4; i++) v[i] = (w >> (8 * i)) & 0xff;
  } else if (N == 8) {
    uint2 w = *reinterpret_cast<const uint2*>(p);
#pragma unroll
    for (int i = 0; i < 4; i++) { v[i] = (w.x >> (8 * i)) & 0xff; v[4 + i] = (w.y >> (8 * i)) & 0xff; }
  } else {
#pragma unroll
    for (int q = 0; q < N / 16; q++) {
      uint4 w = reinterpret_cast<const uint4*>(p)[q];
#pragma unroll
      for (int i = 0; i < 4; i++) {
        v[16 * q + i] = (w.x >> (8 * i)) & 0xff;
        v[16 * q + 4 + i] = (w.y >> (8 * i)) & 0xff;
        v[16 * q + 8 + i] = (w.z >> (8 * i)) & 0xff;
        v[16 * q + 12 + i] = (w.w >> (8 * i)) & 0xff;
      }
    }
  }
}
__device__ __forceinline__ uint32_t pack4(int a, int b, int c, int d) {
  return (uint32_t)a | ((uint32_t)b << 8) | ((uint32_t)c << 16) | ((uint32_t)d << 24);
}
template <int N>
__device__ __forceinline__ void store_row_u8(uint8_t* p, const int (&v)[N]) {
  if (N == 4) {
    *reinterpret_cast<uint32_t*>(p) = pack4(v[0], v[1], v[2], v[3]);
  } else if (N == 8) {
    *reinterpret_cast<uint2*>(p) = make_uint2(pack4(v[0], v[1], v[2], v[3]), pack4(v[4], v[5], v[6], v[7]));
  } else {
#pragma unroll
    for (int q = 0; q < N / 16; q++) {
      const int* w = v + 16 * q;
      reinterpret_cast<uint4*>(p)[q] =
          make_uint4(pack4(w[0], w[1], w[2], w[3]), pack4(w[4], w[5], w[6], w[7]),
                     pack4(w[8], w[9], w[10], w[11]), pack4(w[12], w[13], w[14], w[15]));
    }
  }
}
