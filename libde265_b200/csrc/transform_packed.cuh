// transform_packed.cuh — inverse DCT / DST on PACKED int16 pairs with dp2a (two int16 x int8 products per
// instruction), the arithmetic of the dense residual kernels (k_residual_dense.cu).
//
// Reference behaviour restated (not copied): libde265/fallback-dct.cc:552-692 (IDCT as a direct matrix
// product, first pass over the columns with clip16((s+64)>>7), second pass over the rows with (s+2048)>>12)
// and :270-337 (DST 4x4).  Every 1-D transform here is the even/odd partial butterfly; all sums are exact
// in int32 (|s| <= 32*90*32768 < 2^31), so the results are bit-identical.
//
// Pair order.  An N-point transform consumes its N inputs as N/2 words {lo = x[ka], hi = x[kb]}:
//   words [0, N/4)      the odd inputs (1,3) (5,7) (9,11) ...      -> odd half of the butterfly
//   words [N/4, N/2)    the pair list of the N/2-point transform on the even inputs (recursively),
//   last word           (x[0], x[N/2])                             -> the 2-point base
// pair_first(N, p) is x-index ka of word p; pair_slot(N, k) is the int16 slot (2*word + half) of input k.
#pragma once
#include "transform.cuh"

__host__ __device__ constexpr int pair_first(int n, int p, int f = 1) {
  return n == 2 ? 0 : (p < n / 4 ? (4 * p + 1) * f : pair_first(n / 2, p - n / 4, 2 * f));
}
__host__ __device__ constexpr int pair_second(int n, int p, int f = 1) {
  return n == 2 ? f : (p < n / 4 ? (4 * p + 3) * f : pair_second(n / 2, p - n / 4, 2 * f));
}
// int16 slot of input k in the pair list (runtime, evaluated once per lane)
__device__ __forceinline__ int pair_slot(int n, int k) {
  int off = 0;
  while (n > 2) {
    if (k & 1) return 2 * (off + (k >> 2)) + ((k >> 1) & 1);
    off += n >> 2;
    n >>= 1;
    k >>= 1;
  }
  return 2 * off + k;
}
// rows (first | second << 8) of pair word p (runtime form of pair_first / pair_second)
__device__ __forceinline__ int pair_rows(int n, int p) {
  int f = 1;
  while (n > 2) {
    if (p < (n >> 2)) return ((4 * p + 1) * f) | (((4 * p + 3) * f) << 8);
    p -= n >> 2;
    n >>= 1;
    f <<= 1;
  }
  return f << 8;
}

__host__ __device__ constexpr uint32_t coef2(int a, int b) { return (uint32_t)(a & 0xff) | ((uint32_t)(b & 0xff) << 8); }

// Coefficient pairs live in constant memory and reach dp2a through uniform registers.  One 32-bit word holds TWO
// pairs, consumed by dp2a.lo (bytes 0,1) and dp2a.hi (bytes 2,3), so the 32-point butterfly needs one uniform
// load per two dp2a.  (As immediates every distinct pair cost a UMOV = a second issue slot per dp2a.)
//   bf[n][p/2]  = pair(n, p even) | pair(n, p odd) << 16      butterfly form  (o = sum over the odd pairs)
//   pn[n][p]    = pair(n, p) | -pair(n, p) << 16              chain form for <= 2 pairs (a = E + o, b = E - o)
// with pair(n, p) of the N-point level = {mat[(4p+1)*32/N][n], mat[(4p+3)*32/N][n]}.
struct CoefTab {
  uint32_t bf32[16 * 4], bf16[8 * 2], bf8[4 * 1];
  uint32_t pn32[16 * 2], pn16[8 * 2], pn8[4 * 2], pn4[2 * 1];
  uint32_t base;
};
__host__ __device__ constexpr uint32_t level_pair(int n_pt, int n, int p, int sign) {
  return coef2(sign * dct_coef((4 * p + 1) * (32 / n_pt), n), sign * dct_coef((4 * p + 3) * (32 / n_pt), n));
}
__host__ __device__ constexpr void fill_bf(uint32_t* t, int n_pt) {
  for (int n = 0; n < n_pt / 2; n++)
    for (int q = 0; q < n_pt / 8; q++)
      t[n * (n_pt / 8) + q] = level_pair(n_pt, n, 2 * q, 1) | (level_pair(n_pt, n, 2 * q + 1, 1) << 16);
}
__host__ __device__ constexpr void fill_pn(uint32_t* t, int n_pt) {
  const int w = n_pt >= 8 ? 2 : 1;
  for (int n = 0; n < n_pt / 2; n++)
    for (int p = 0; p < w; p++) t[n * w + p] = level_pair(n_pt, n, p, 1) | (level_pair(n_pt, n, p, -1) << 16);
}
__host__ __device__ constexpr CoefTab make_coef_tab() {
  CoefTab t = {};
  fill_bf(t.bf32, 32); fill_bf(t.bf16, 16); fill_bf(t.bf8, 8);
  fill_pn(t.pn32, 32); fill_pn(t.pn16, 16); fill_pn(t.pn8, 8); fill_pn(t.pn4, 4);
  t.base = coef2(64, 64) | (coef2(64, -64) << 16);
  return t;
}
__constant__ CoefTab c_coef = make_coef_tab();
template <int N> __device__ __forceinline__ uint32_t coef_bf(int n, int q) {
  return N == 32 ? c_coef.bf32[n * 4 + q] : N == 16 ? c_coef.bf16[n * 2 + q] : c_coef.bf8[n];
}
template <int N> __device__ __forceinline__ uint32_t coef_pn(int n, int p) {
  return N == 32 ? c_coef.pn32[n * 2 + p] : N == 16 ? c_coef.pn16[n * 2 + p] : N == 8 ? c_coef.pn8[n * 2 + p] : c_coef.pn4[n];
}

__device__ __forceinline__ int dp2a(uint32_t pair, uint32_t coefs, int acc) {
  return __dp2a_lo((int)pair, (int)coefs, acc);
}
__device__ __forceinline__ int dp2a_hi(uint32_t pair, uint32_t coefs, int acc) {
  return __dp2a_hi((int)pair, (int)coefs, acc);
}
// {hi = sat16(b), lo = sat16(a)}
__device__ __forceinline__ uint32_t pack_sat16(int a, int b) {
  uint32_t d;
  asm("cvt.pack.sat.s16.s32 %0, %1, %2;" : "=r"(d) : "r"(b), "r"(a));
  return d;
}
// bytes {sat8(a), sat8(b), sat8(c), sat8(d)} from low to high
__device__ __forceinline__ uint32_t pack_sat8(int a, int b, int c, int d) {
  uint32_t t, r;
  asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(t) : "r"(d), "r"(c), "r"(0));
  asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(b), "r"(a), "r"(t));
  return r;
}

// Inverse N-point DCT of the pair list P (inputs k >= L are zero and their pairs are not read):
// out[n] = bias + sum_k mat[k*32/N*F...][n] * x[k].   F = stride of this level's inputs in the top-level input
// numbering (only used to decide which pairs lie below L).
template <int N, int L, int F = 1>
struct IdctP {
  static constexpr int Q = N / 4, H = N / 2;
  static constexpr int USED = (L <= F) ? 0 : ((L / F + 2) / 4 < Q ? (L / F + 2) / 4 : Q);   // odd pairs with (4p+1)*F < L
  static __device__ __forceinline__ void run(const uint32_t* P, int* out, int bias) {
    int E[H];
    IdctP<H, L, 2 * F>::run(P + Q, E, bias);
#pragma unroll
    for (int n = 0; n < H; n++) {
      if (USED == 0) {
        out[n] = E[n];
        out[N - 1 - n] = E[n];
      } else if (USED <= 2) {
        // few pairs: accumulate straight onto E with +/- coefficients (no separate add)
        int a = E[n], b = E[n];
#pragma unroll
        for (int p = 0; p < USED; p++) {
          a = dp2a(P[p], coef_pn<N>(n, p), a);
          b = dp2a_hi(P[p], coef_pn<N>(n, p), b);
        }
        out[n] = a;
        out[N - 1 - n] = b;
      } else {
        int o = 0;
#pragma unroll
        for (int p = 0; p < USED; p++) {
          o = (p & 1) ? dp2a_hi(P[p], coef_bf<N>(n, p >> 1), o) : dp2a(P[p], coef_bf<N>(n, p >> 1), o);
        }
        out[n] = E[n] + o;
        out[N - 1 - n] = E[n] - o;
      }
    }
  }
};
template <int L, int F>
struct IdctP<2, L, F> {
  static __device__ __forceinline__ void run(const uint32_t* P, int* out, int bias) {
    out[0] = dp2a(P[0], c_coef.base, bias);
    out[1] = dp2a_hi(P[0], c_coef.base, bias);
  }
};

// 4-point inverse DST-VII on the same pair order as the 4-point DCT: P[0] = (x1, x3), P[1] = (x0, x2)
__device__ __forceinline__ void idst4_packed(const uint32_t* P, int* out, int bias) {
  out[0] = dp2a(P[0], coef2(74, 55), dp2a(P[1], coef2(29, 84), bias));
  out[1] = dp2a(P[0], coef2(74, -84), dp2a(P[1], coef2(55, -29), bias));
  out[2] = dp2a(P[0], coef2(0, 74), dp2a(P[1], coef2(74, -74), bias));
  out[3] = dp2a(P[0], coef2(-74, -29), dp2a(P[1], coef2(84, 55), bias));
}

// Flat dequantisation (transform.cc:443-475, bit depth 8) as level*f + rnd >> sh with per-block constants:
//   (level*ls*2^k + 2^(s-1)) >> s  ==  level*(ls << (k-s))                      for k >= s
//                                  ==  (level*ls + 2^(s-1-k)) >> (s-k)          for k <  s      (s = log2-1)
// The level is taken out of its packed word by the multiply itself: dp2a(word, {ls, 0}) = lo*ls, dp2a(word, {0, ls})
// = hi*ls (ls <= 72 fits the signed byte), so no unpack instruction runs on the ALU pipe.
struct DequantK {
  uint32_t sel_lo, sel_hi;   // dp2a selectors {ls, 0} / {0, ls}
  int m, rnd, sh;            // (v*m + rnd) >> sh
};
__device__ __forceinline__ DequantK dequant_consts(int qp, int log2) {
  DequantK d;
  if (qp < 0) {  // coefficients are already dequantised
    d.sel_lo = 1; d.sel_hi = 1 << 8; d.m = 1; d.rnd = 0; d.sh = 0;
    return d;
  }
  const int k = (qp * 43) >> 8, rem = qp - 6 * k, s = log2 - 1;
  const uint32_t ls = (uint32_t)((0x484039332D28ull >> (8 * rem)) & 0xff);   // {40,45,51,57,64,72}
  d.sel_lo = ls;
  d.sel_hi = ls << 8;
  d.m = 1 << max(k - s, 0);
  d.sh = max(s - k, 0);
  d.rnd = d.sh ? 1 << (d.sh - 1) : 0;
  return d;
}
__device__ __forceinline__ int dq_lo(uint32_t w, const DequantK& d) { return (dp2a(w, d.sel_lo, 0) * d.m + d.rnd) >> d.sh; }
__device__ __forceinline__ int dq_hi(uint32_t w, const DequantK& d) { return (dp2a(w, d.sel_hi, 0) * d.m + d.rnd) >> d.sh; }
// byte n of w, zero-extended, without touching the ALU pipe
template <int n> __device__ __forceinline__ int byte_of(uint32_t w) { return (int)__dp4a(w, 1u << (8 * n), 0u); }
template <int s> __device__ __forceinline__ int sra(int x) { return x >> s; }   // (mulhi by 2^(32-s) measured: no gain)
