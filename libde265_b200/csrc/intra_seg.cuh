// intra_seg.cuh — intra prediction of one transform block by a warp, SIZE-INDEPENDENT code.
//
// Same arithmetic as intra.cuh (libde265/intrapred.cc:395-431 substitution, :816-889 filtering, :903-1069 angular /
// planar / DC), organised differently: the block size N is a run-time value and the unit of work of a lane is one
// SEGMENT of 8 samples of a line (a row for planar, DC and modes 18-34, a column for modes 2-17):
//     N = 4: 4 items (half a segment each), N = 8: 8, N = 16: 32, N = 32: 128 = four passes of the warp.
// Why: the wavefront (k_intra.cu) is bound by the latency of single warps and by instruction fetch - 16 warps of an SM
// sit at 16 different places of the code.  One body for all sizes instead of four unrolled template instantiations cuts
// the kernel's code by more than half, needs a third of the registers per item (4 packed pairs instead of up to 16),
// and a 16x16 block keeps all 32 lanes busy with half the chain length per lane.
#pragma once
#include "intra.cuh"

// border buffer of a block of size N (same layout as IntraBuf<N>): [3 pad][ext N][T 2N+1][3 pad][ext N][L 2N+1]
__device__ __forceinline__ int ibuf_top(int N) { return N + 3; }
__device__ __forceinline__ int ibuf_left(int N) { return 4 * N + 7; }
__device__ __forceinline__ int ibuf_stride(int N) { return 6 * N + 8; }

// gather_border<LOG2> of intra.cuh with N at run time.  WIDE: N == 32 (33 scan units need a 64-bit mask).
template <bool WIDE>
__device__ __forceinline__ void gather_border_rt(const CtbEdges& E, int N, int xr, int yr, uint64_t M64, int lane, uint8_t* buf) {
  typedef typename std::conditional<WIDE, uint64_t, uint32_t>::type mask_t;
  const int H = N >> 1;
  const mask_t M = (mask_t)M64;
  const uint8_t* colp = E.cols + (xr >> 2) * E.S + yr;                   // column xr-1, from the TB's first row down
  const uint8_t* rowp = E.rows + (yr >> 2) * E.rstride + 16 + xr;        // row yr-1, from the TB's first column on
  const bool left = lane < H;
  const int wi = left ? lane : lane - H;                           // word of the line: samples 1+4w .. 4+4w
  const int s = lane < N ? (left ? H - 1 - lane : lane + 1) : H;   // scan unit of this lane
  // Nearly always the available units are ONE run [ra, rb] of the scan (missing: a piece of the lower-left end and / or of
  // the upper-right end): the source unit is then the scan unit clamped to the run - two warp-uniform bit scans instead of a
  // masked scan per lane on the chain.
  const int ra = WIDE ? __ffsll((long long)M) - 1 : __ffs((int)M) - 1;
  const int rb = WIDE ? 63 - __clzll((long long)M) : 31 - __clz((int)M);
  uint32_t v, c;
  if (M != 0 && M == (mask_t)((((mask_t)2 << rb) - 1) & ~(((mask_t)1 << ra) - 1))) {
    auto fetch_run = [&](int u) -> uint32_t {
      const int t = min(max(u, ra), rb);
      const uint8_t* a = t < H ? colp + 4 * (H - 1 - t) : rowp + 4 * (t - H - 1);
      const uint32_t w = *reinterpret_cast<const uint32_t*>(a);
      // below the run: first sample of unit ra; above it: last sample of unit rb; scan order runs UP the left column (last
      // sample = byte 0) and RIGHT along the top row (last = byte 3); the corner unit is the last byte of its word
      const bool byte0 = (t < H) ? (u > rb) : (u < ra && t != H);
      const uint32_t sel = (u == t && t != H) ? 0x3210u : (byte0 ? 0x0000u : 0x3333u);
      return __byte_perm(w, 0, sel);
    };
    v = fetch_run(s);
    c = fetch_run(H);
  } else {
    v = gather_fetch_general((uint64_t)M, s, H, colp, rowp);
    c = gather_fetch_general((uint64_t)M, H, H, colp, rowp);
  }
  if (lane < N) *reinterpret_cast<uint32_t*>(buf + (left ? ibuf_left(N) : ibuf_top(N)) + 1 + 4 * wi) = v;
  if (lane == 0) {
    buf[ibuf_top(N)] = (uint8_t)c;
    buf[ibuf_left(N)] = (uint8_t)c;
  }
  __syncwarp();
}

// filter_border<LOG2> of intra.cuh with N at run time
__device__ __forceinline__ uint8_t* filter_border_rt(uint8_t* src, uint8_t* dst, int N, int lane, int mode, int cidx, bool strong_enabled) {
  if (cidx != 0 || mode == 1 || N == 4) return src;  // warp-uniform
  const int H = N >> 1;
  const int d = min(abs(mode - 26), abs(mode - 10));
  const bool filt = (N == 8) ? d > 7 : (N == 16) ? d > 1 : d > 0;
  if (!filt) return src;
  const uint8_t* T = src + ibuf_top(N);
  const uint8_t* L = src + ibuf_left(N);
  bool bi = false;
  if (N == 32 && strong_enabled)
    bi = abs(T[0] + T[2 * N] - 2 * T[N]) < 8 && abs(L[0] + L[2 * N] - 2 * L[N]) < 8;  // 1 << (bitDepth-5)
  if (lane < N) {
    const bool left = lane < H;
    const int w = left ? lane : lane - H;
    const uint8_t* X = left ? L : T;
    uint32_t out;
    if (bi) {
      const int p0 = X[0], pe = X[2 * N];
      out = 0;
#pragma unroll
      for (int i = 0; i < 4; i++) {
        const int k = 1 + 4 * w + i;
        out |= (uint32_t)(p0 + ((k * (pe - p0) + 32) >> 6)) << (8 * i);  // k = 64 gives pe
      }
    } else {
      const uint32_t* xw = reinterpret_cast<const uint32_t*>(X + 1 + 4 * w);
      const uint32_t prev = xw[-1], cur = xw[0], next = (w == H - 1) ? 0u : xw[1];
      const uint32_t a = __funnelshift_l(prev, cur, 8);   // samples k-1
      const uint32_t c = __funnelshift_r(cur, next, 8);   // samples k+1
      out = avg4(cur, hadd4(a, c));
      if (w == H - 1) out = (out & 0x00ffffffu) | (cur & 0xff000000u);  // p[+-2N] stays
    }
    *reinterpret_cast<uint32_t*>(dst + (left ? ibuf_left(N) : ibuf_top(N)) + 1 + 4 * w) = out;
  }
  if (lane == (N < 32 ? N : 0)) {
    const int c = bi ? T[0] : (L[1] + 2 * T[0] + T[1] + 2) >> 2;
    dst[ibuf_top(N)] = (uint8_t)c;
    dst[ibuf_left(N)] = (uint8_t)c;
  }
  __syncwarp();
  return dst;
}

// Warp-uniform state of one block's prediction
struct IntraPred {
  int N, log2, mode;
  bool dc_filter;            // cIdx == 0 && N < 32: DC smooths the first row and column (intrapred.cc:1049-1066, not switchable)
  bool angular_filter;       // ... && boundary filters enabled: modes 10 / 26 smooth the first column / row (:985-1003)
  bool by_column;            // modes 2..17: a line is a COLUMN (the main reference is the left line)
  uint8_t* T;                // T[k] = p[k][-1], k = 0 (corner) .. 2N
  uint8_t* L;                // L[k] = p[-1][k]
  uint8_t* Mp;               // main line of an angular mode (T or L), extended to negative indices when the angle is negative
  const uint8_t* Op;         // the other line
  int ang, dc;
};

// Called by all lanes of the warp (it contains warp-wide shuffles / barriers).
__device__ __forceinline__ void intra_pred_setup(IntraPred& P, uint8_t* pb, int N, int log2, int lane, int mode, int cidx, bool edge_filters) {
  P.N = N; P.log2 = log2; P.mode = mode;
  P.T = pb + ibuf_top(N);
  P.L = pb + ibuf_left(N);
  P.dc_filter = cidx == 0 && N < 32;
  P.angular_filter = P.dc_filter && edge_filters;
  P.by_column = mode >= 2 && mode < 18;
  P.Mp = P.by_column ? P.L : P.T;
  P.Op = P.by_column ? P.T : P.L;
  P.ang = 0; P.dc = 0;
  if (mode >= 2) {
    P.ang = c_intra_angle[mode];
    if (P.ang < 0) {  // extend the main line to negative indices by projecting the other line (intrapred.cc:925-944)
      const int inv = c_intra_inv_angle[mode - 11];
      const int k = -1 - lane;
      if (lane < N && k > ((N * P.ang) >> 5)) P.Mp[k] = P.Op[(k * inv + 128) >> 8];  // lowest index read is ((N*ang)>>5) + 1
      __syncwarp();
    }
  } else if (mode == 1) {
    int dc = lane < N ? P.T[1 + lane] + P.L[1 + lane] : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dc += __shfl_xor_sync(B265_FULL_MASK, dc, o);
    P.dc = (dc + N) >> (log2 + 1);
  }
}

// Samples 8s .. 8s+7 of line l (row y = l, or column x = l when P.by_column) as 4 words of two 16-bit values.
// For N == 4 only the first two words mean anything.
__device__ __forceinline__ void intra_pred_seg(const IntraPred& P, int l, int s, uint32_t (&h)[4]) {
  const int N = P.N;
  if (P.mode >= 2) {  // angular (:903-1005)
    const int t = (l + 1) * P.ang, idx = t >> 5, f = t & 31;
    const uint8_t* a = P.Mp + idx + 1 + 8 * s;   // samples a[0..8] are needed
    const unsigned m = (unsigned)(reinterpret_cast<uintptr_t>(a) & 3u);
    const uint32_t* aw = reinterpret_cast<const uint32_t*>(a - m);
    const uint32_t w0 = aw[0], w1 = aw[1], w2 = aw[2];
    uint32_t lo[3];
    lo[0] = __funnelshift_r(w0, w1, 8 * m);
    lo[1] = __funnelshift_r(w1, w2, 8 * m);
    lo[2] = w2 >> (8 * m);
    const uint32_t W0 = (uint32_t)(32 - f) | ((uint32_t)f << 8);
#pragma unroll
    for (int j = 0; j < 2; j++) {
      const uint32_t b34 = __funnelshift_r(lo[j], lo[j + 1], 24);
      const uint32_t v0 = __dp4a(lo[j], W0, 16u), v1 = __dp4a(lo[j], W0 << 8, 16u);
      const uint32_t v2 = __dp4a(lo[j], W0 << 16, 16u), v3 = __dp4a(b34, W0, 16u);
      h[2 * j] = (__byte_perm(v0, v1, 0x5410) >> 5) & 0x00ff00ffu;
      h[2 * j + 1] = (__byte_perm(v2, v3, 0x5410) >> 5) & 0x00ff00ffu;
    }
    if ((P.mode == 26 || P.mode == 10) && P.angular_filter && s == 0)
      h[0] = (h[0] & 0xffff0000u) | (uint32_t)clip8(P.Mp[1] + ((P.Op[1 + l] - P.Mp[0]) >> 1));
    return;
  }
  if (P.mode == 0) {  // planar (:1009-1032), line = row y
    const int y = l;
    const uint32_t abd = (uint32_t)P.L[1 + y] | ((uint32_t)P.T[N + 1] << 8) | ((uint32_t)P.L[N + 1] << 24);
    const uint32_t Wy = ((uint32_t)(N - 1 - y) << 16) | ((uint32_t)(y + 1) << 24);
    const uint32_t* tw = reinterpret_cast<const uint32_t*>(P.T + 1 + 8 * s);
#pragma unroll
    for (int j = 0; j < 2; j++) {
      const uint32_t t4 = tw[j];
      uint32_t v[4];
#pragma unroll
      for (int i = 0; i < 4; i++) {
        const int x = 8 * s + 4 * j + i;
        const uint32_t op = __byte_perm(abd, t4, 0x3010u | ((4u + i) << 8));   // [L[1+y], T[N+1], T[1+x], L[N+1]]
        v[i] = __dp4a(op, Wy | (uint32_t)((N - 1 - x) & 0xff) | ((uint32_t)((x + 1) & 0xff) << 8), (uint32_t)N) >> (P.log2 + 1);
      }
      h[2 * j] = v[0] | (v[1] << 16);
      h[2 * j + 1] = v[2] | (v[3] << 16);
    }
    return;
  }
  // DC (:1036-1069), line = row y
  const int dc = P.dc;
#pragma unroll
  for (int i = 0; i < 4; i++) h[i] = (uint32_t)dc * 0x00010001u;
  if (P.dc_filter) {
    if (l == 0) {
      const uint8_t* tp = P.T + 1 + 8 * s;
#pragma unroll
      for (int i = 0; i < 4; i++)
        h[i] = (uint32_t)((tp[2 * i] + 3 * dc + 2) >> 2) | ((uint32_t)((tp[2 * i + 1] + 3 * dc + 2) >> 2) << 16);
      if (s == 0) h[0] = (h[0] & 0xffff0000u) | (uint32_t)((P.L[1] + 2 * dc + P.T[1] + 2) >> 2);
    } else if (s == 0) {
      h[0] = (h[0] & 0xffff0000u) | (uint32_t)((P.L[1 + l] + 3 * dc + 2) >> 2);
    }
  }
}

// Clip1(pred + residual) on the 16-bit pairs, then pack the 8 samples to two words of bytes.  r = residual of the segment as
// int16 pairs (values pre-clamped to +-1024 by the residual kernel, so the 16-bit add cannot wrap).
__device__ __forceinline__ void finish_seg(uint32_t (&h)[4], const uint32_t (&r)[4], bool coded, uint32_t (&px)[2]) {
  if (coded) {
#pragma unroll
    for (int i = 0; i < 4; i++) h[i] = __viaddmin_s16x2_relu(h[i], r[i], 0x00ff00ffu);
  }
  px[0] = __byte_perm(h[0], h[1], 0x6420);
  px[1] = __byte_perm(h[2], h[3], 0x6420);
}
