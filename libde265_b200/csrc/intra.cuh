// intra.cuh — intra prediction as warp-cooperative device code (lanes 0..N-1 of a warp = one N x N TB).
//
// Restates libde265/intrapred.cc: intra_border_computer (:344-688: availability by picture edge,
// CTB slice/tile ownership, z-scan order and constrained_intra_pred; reference-sample substitution
// :395-431), intra_prediction_sample_filtering (:816-889), angular (:903-1005), planar (:1009-1032)
// and DC (:1036-1069).
//
// The code is written for the LATENCY of one warp walking the TBs of a CTB one after the other
// (k_intra.cu): few instructions per TB, no global memory access on the way.
//  * availability of the neighbouring 4-sample units is a 64-bit mask per TB, computed beforehand by a
//    parallel kernel (intra_avail_mask, k_intra_prep);
//  * neighbouring samples come from the CTB's "edge store" in shared memory: only the bottom row and
//    the right column of a finished TB can ever be read by later TBs, so the store keeps the rows
//    y = -1, 3, 7, ... and the columns x = -1, 3, 7, ... of the CTB;
//  * the border lives in two byte lines T[k] = p[k] (corner, top, top-right) and L[k] = p[-k] (corner,
//    left, bottom-left) whose samples 1.. are word aligned: gather, [1 2 1] filter and the angular
//    interpolation work on 4 samples per instruction (funnel shifts, byte-wise averages, dp4a);
//  * modes 2..17 are predicted with the lane owning a COLUMN (the main reference is the left line), all
//    other modes with the lane owning a ROW.
#pragma once
#include <type_traits>
#include "common.cuh"

#define B265_BORDER_BYTES 512  // per warp: 2 buffers x (6N+8) bytes, N <= 32

// border buffer: [3 pad][ext N][T 2N+1][3 pad][ext N][L 2N+1]; T[1] and L[1] are 4-byte aligned
template <int N> struct IntraBuf {
  static constexpr int TOP = N + 3;
  static constexpr int LEFT = 4 * N + 7;
  static constexpr int STRIDE = 6 * N + 8;
};

// Picture / CTB constants of a ticket, held in registers (the FrameDev record lives in global memory;
// re-reading its fields per TB would cost an L2 round trip each).
struct IntraGeom {
  int sh;                    // 0 luma, 1 chroma (4:2:0)
  int width, height;         // luma picture size
  int log2_ctb, log2_min_tb;
  int cx, cy;                // CTB position in CTB units
  int ctbs_w;                // CTBs per row (resid_block)
  unsigned nb;               // neighbouring CTBs usable for prediction (same slice, same tile, earlier in decoding
                             // order; intrapred.cc:486-508): bit0 left, bit1 upper-left, bit2 upper, bit3 upper-right
  const uint8_t* unit_info;  // non-NULL only with constrained_intra_pred
  int unit_pitch;
  uint8_t* plane;            // destination plane of this component
  const int16_t* resid;
  int pitch;
  bool strong;               // strong_intra_smoothing_enabled
};

// z-scan order of the min-TB containing luma sample (x,y) inside its CTB (pps MinTbAddrZS minus the CTB term)
__device__ __forceinline__ unsigned zs_in_ctb(const IntraGeom& g, int x, int y) {
  const int m = (1 << g.log2_ctb) - 1;
  return morton4((x & m) >> g.log2_min_tb, (y & m) >> g.log2_min_tb);
}

// availability of the neighbouring 4-sample unit whose defining luma sample is (nx,ny); cur_zs = zs_in_ctb of
// the current block.  Pure ALU unless constrained_intra_pred is on.
__device__ __forceinline__ bool unit_available(const IntraGeom& g, unsigned cur_zs, int nx, int ny) {
  if (nx < 0 || ny < 0 || nx >= g.width || ny >= g.height) return false;
  const int dx = (nx >> g.log2_ctb) - g.cx, dy = (ny >> g.log2_ctb) - g.cy;
  if (dx == 0 && dy == 0) {
    if (zs_in_ctb(g, nx, ny) > cur_zs) return false;   // not decoded yet
  } else {
    // CTBs below or to the right come later in decoding order (or lie in another tile)
    int bit;
    if (dy == 0 && dx == -1) bit = 0;
    else if (dy == -1 && dx >= -1 && dx <= 1) bit = 2 + dx;
    else return false;
    if (!((g.nb >> bit) & 1u)) return false;
  }
  if (g.unit_info) {
    if (!(__ldg(&g.unit_info[(ny >> 2) * g.unit_pitch + (nx >> 2)]) & B265_UNIT_INTRA)) return false;
  }
  return true;
}

// Availability mask of the TB at component position (xB,yB), size N = 1 << log2, in scan order
// (intrapred.cc:395-431 walks bottom-left -> corner -> top-right): bits 0..N/2-1 = left units bottom-up,
// bit N/2 = corner, bits N/2+1..N = top units left to right.
__device__ __forceinline__ uint64_t intra_avail_mask(const IntraGeom& g, int xB, int yB, int log2) {
  const int N = 1 << log2, H = N >> 1, sh = g.sh;
  const unsigned cur_zs = zs_in_ctb(g, xB << sh, yB << sh);
  if (!g.unit_info) {
    // Without constrained_intra_pred five tests decide everything: the units next to the left column, next to the top row,
    // below-left and above-right each lie in ONE aligned N x N block (all in one CTB, all before or all after the current
    // block in z-scan), so each group shares its answer; only the picture's bottom / right edge can cut the two outer groups.
    const int Q = N >> 2;   // units per group
    uint64_t M = 0;
    if (unit_available(g, cur_zs, (xB - 1) << sh, (yB + N) << sh)) {            // below-left: units s = Q-1 (nearest) .. 0
      // the reference tests a left unit at its LAST row: units that lie completely above the picture's bottom edge
      const int rows_left = (((g.height + (1 << sh) - 1) >> sh) - (yB + N)) >> 2;
      const int n = min(Q, max(rows_left, 0));
      M |= (((uint64_t)1 << n) - 1) << (Q - n);
    }
    if (unit_available(g, cur_zs, (xB - 1) << sh, yB << sh)) M |= (((uint64_t)1 << Q) - 1) << Q;          // left: s = Q .. H-1
    if (unit_available(g, cur_zs, (xB - 1) << sh, (yB - 1) << sh)) M |= (uint64_t)1 << H;                  // corner
    if (unit_available(g, cur_zs, xB << sh, (yB - 1) << sh)) M |= (((uint64_t)1 << Q) - 1) << (H + 1);    // top
    if (unit_available(g, cur_zs, (xB + N) << sh, (yB - 1) << sh)) {                                       // above-right
      // ... and a top unit at its FIRST column
      const int cols_left = (g.width - ((xB + N) << sh) + (4 << sh) - 1) >> (2 + sh);
      const int n = min(Q, max(cols_left, 0));
      M |= (((uint64_t)1 << n) - 1) << (H + 1 + Q);
    }
    return M;
  }
  uint64_t M = 0;
  for (int s = 0; s < H; s++)
    if (unit_available(g, cur_zs, (xB - 1) << sh, (yB + 2 * N - 1 - 4 * s) << sh)) M |= 1ull << s;
  if (unit_available(g, cur_zs, (xB - 1) << sh, (yB - 1) << sh)) M |= 1ull << H;
  for (int w = 0; w < H; w++)
    if (unit_available(g, cur_zs, (xB + 4 * w) << sh, (yB - 1) << sh)) M |= 1ull << (H + 1 + w);
  return M;
}

// Edge store of the CTB plane a warp is walking (shared memory).
//   rows[q][16 + c] = sample (c, 4q-1) relative to the CTB origin, c = -1..2S-1   (q = 0: row above the CTB)
//   cols[q][r]      = sample (4q-1, r),                            r = 0..S-1     (q = 0: column left of the CTB)
struct CtbEdges {
  uint8_t* rows;
  uint8_t* cols;
  int S, rstride;   // rstride = 2S + 16
  int x0, y0;       // component position of the CTB origin
};
__host__ __device__ constexpr int ctb_edges_bytes(int S) { return (S / 4 + 1) * (2 * S + 16) + (S / 4 + 1) * S; }

// General reference-sample substitution for one scan unit (any availability pattern; constrained_intra_pred can punch holes
// into the run): kept out of line, the wavefront's hot path is the single-run form in gather_border.
static __device__ __noinline__ uint32_t gather_fetch_general(uint64_t M, int s, int H, const uint8_t* colp, const uint8_t* rowp) {
  const bool av = (M >> s) & 1;
  const uint64_t lower = M & (((uint64_t)1 << s) - 1);
  int t = av ? s : (lower ? 63 - __clzll((long long)lower) : __ffsll((long long)M) - 1);
  const int pick = av ? 0 : (lower ? 1 : 2);   // 0 = the unit itself, 1 = its last sample in scan order, 2 = its first sample
  if (M == 0) t = H;
  const uint8_t* a = t < H ? colp + 4 * (H - 1 - t) : rowp + 4 * (t - H - 1);
  const uint32_t w = *reinterpret_cast<const uint32_t*>(a);
  const bool byte0 = (t < H) ? (pick == 1) : (pick == 2 && t != H);
  const uint32_t sel = (pick == 0 && t != H) ? 0x3210u : (byte0 ? 0x0000u : 0x3333u);
  return M ? __byte_perm(w, 0, sel) : 0x80808080u;
}

// Border of the TB at (xr,yr) relative to the CTB origin -> buf (IntraBuf layout), with the reference
// sample substitution of intrapred.cc:395-431: a missing unit copies the nearest available sample before it
// in scan order, leading missing units copy the first available sample, nothing available -> 128.
// Branch-free: every lane resolves its scan unit to a source unit t (itself, the nearest available one
// before it, or the first available one), loads that unit's word from the edge store and picks the bytes
// with one PRMT.  Called by all lanes of the warp.
template <int LOG2>
__device__ __forceinline__ void gather_border(const CtbEdges& E, int xr, int yr, uint64_t M64, int lane, uint8_t* buf) {
  constexpr int N = 1 << LOG2, H = N / 2;
  typedef typename std::conditional<(N == 32), uint64_t, uint32_t>::type mask_t;   // N + 1 scan units
  const mask_t M = (mask_t)M64;
  const uint8_t* colp = E.cols + (xr >> 2) * E.S + yr;                   // column xr-1, from the TB's first row down
  const uint8_t* rowp = E.rows + (yr >> 2) * E.rstride + 16 + xr;        // row yr-1, from the TB's first column on
  const bool left = lane < H;
  const int wi = left ? lane : lane - H;                           // word of the line: samples 1+4w .. 4+4w
  const int s = lane < N ? (left ? H - 1 - lane : lane + 1) : H;   // scan unit of this lane
  // Nearly always the available units are ONE run [ra, rb] of the scan (missing: a piece of the lower-left end and / or of
  // the upper-right end): the source unit is then the scan unit clamped to the run - two warp-uniform bit scans instead of a
  // masked scan per lane on the chain.
  const int ra = (N == 32) ? __ffsll((long long)M) - 1 : __ffs((int)M) - 1;
  const int rb = (N == 32) ? 63 - __clzll((long long)M) : 31 - __clz((int)M);
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
  if (lane < N) *reinterpret_cast<uint32_t*>(buf + (left ? IntraBuf<N>::LEFT : IntraBuf<N>::TOP) + 1 + 4 * wi) = v;
  if (lane == 0) {
    buf[IntraBuf<N>::TOP] = (uint8_t)c;
    buf[IntraBuf<N>::LEFT] = (uint8_t)c;
  }
  __syncwarp();
}

// byte-wise (a + b) >> 1 and (a + b + 1) >> 1 on four packed samples
__device__ __forceinline__ uint32_t hadd4(uint32_t a, uint32_t b) { return (a & b) + (((a ^ b) & 0xfefefefeu) >> 1); }
__device__ __forceinline__ uint32_t avg4(uint32_t a, uint32_t b) { return (a | b) - (((a ^ b) & 0xfefefefeu) >> 1); }

// ---- intra_prediction_sample_filtering (intrapred.cc:816-889) ----
// Writes the filtered border to dst (same layout) when the mode/size rule asks for it and returns the
// buffer to predict from.  (a + 2b + c + 2) >> 2 == avg4(b, hadd4(a, c)) exactly.  All lanes call it.
template <int LOG2>
__device__ __forceinline__ uint8_t* filter_border(uint8_t* src, uint8_t* dst, int lane, int mode, int cidx, bool strong_enabled) {
  constexpr int N = 1 << LOG2, H = N / 2;
  if (cidx != 0 || mode == 1 || N == 4) return src;  // warp-uniform
  const int d = min(abs(mode - 26), abs(mode - 10));
  const bool filt = (N == 8) ? d > 7 : (N == 16) ? d > 1 : d > 0;
  if (!filt) return src;
  const uint8_t* T = src + IntraBuf<N>::TOP;
  const uint8_t* L = src + IntraBuf<N>::LEFT;
  bool bi = false;
  if (N == 32 && strong_enabled)
    bi = abs(T[0] + T[2 * N] - 2 * T[N]) < 8 && abs(L[0] + L[2 * N] - 2 * L[N]) < 8;  // 1 << (bitDepth-5)
  if (lane < N) {
    const bool left = lane < H;
    const int w = left ? lane : lane - H;
    const uint8_t* X = left ? L : T;
    uint32_t out;
    if (N == 32 && bi) {
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
    *reinterpret_cast<uint32_t*>(dst + (left ? IntraBuf<N>::LEFT : IntraBuf<N>::TOP) + 1 + 4 * w) = out;
  }
  if (lane == (N < 32 ? N : 0)) {
    const int c = bi ? T[0] : (L[1] + 2 * T[0] + T[1] + 2) >> 2;
    dst[IntraBuf<N>::TOP] = (uint8_t)c;
    dst[IntraBuf<N>::LEFT] = (uint8_t)c;
  }
  __syncwarp();
  return dst;
}

__constant__ int8_t c_intra_angle[35] = {0,  0,  32, 26, 21, 17, 13, 9,  5,  2,  0,  -2, -5, -9, -13, -17, -21, -26,
                                         -32, -26, -21, -17, -13, -9, -5, -2, 0,  2,  5,  9,  13, 17,  21,  26,  32};
__constant__ int16_t c_intra_inv_angle[15] = {-4096, -1638, -910, -630, -482, -390, -315, -256,
                                              -315,  -390,  -482, -630, -910, -1638, -4096};

// Prediction of line `lane` of the TB from the border buffer pb: a ROW for planar, DC and modes 18..34,
// a COLUMN for modes 2..17 (returns true).  h[i] holds samples 2i (low half) and 2i+1 (high half) of
// the line as 16-bit values.  Contains warp-wide shuffles / barriers: all lanes of the warp call it.
template <int LOG2>
__device__ __forceinline__ bool predict_line(uint8_t* pb, int lane, int mode, int cidx, bool edge_filters,
                                             uint32_t (&h)[(1 << LOG2) / 2]) {
  constexpr int N = 1 << LOG2;
  uint8_t* T = pb + IntraBuf<N>::TOP;
  uint8_t* L = pb + IntraBuf<N>::LEFT;
  const bool active = lane < N;
  const int l = active ? lane : 0;
  if (mode >= 2) {  // angular (:903-1005); the main line M is T (vertical modes) or L (horizontal modes)
    const bool horiz = mode < 18;
    uint8_t* Mp = horiz ? L : T;
    const uint8_t* Op = horiz ? T : L;
    const int ang = c_intra_angle[mode];
    if (ang < 0) {  // extend the main line to negative indices by projecting the other line
      const int inv = c_intra_inv_angle[mode - 11];
      const int k = -1 - l;
      if (k > ((N * ang) >> 5)) Mp[k] = Op[(k * inv + 128) >> 8];  // lowest index read below is ((N*ang)>>5) + 1
      __syncwarp();
    }
    const int t = (l + 1) * ang, idx = t >> 5, f = t & 31;
    const uint8_t* a = Mp + idx + 1;   // samples a[0..N] are needed
    const unsigned m = (unsigned)(reinterpret_cast<uintptr_t>(a) & 3u);
    const uint32_t* aw = reinterpret_cast<const uint32_t*>(a - m);
    uint32_t w[N / 4 + 1], lo[N / 4 + 1];
#pragma unroll
    for (int j = 0; j <= N / 4; j++) w[j] = aw[j];
#pragma unroll
    for (int j = 0; j < N / 4; j++) lo[j] = __funnelshift_r(w[j], w[j + 1], 8 * m);
    lo[N / 4] = w[N / 4] >> (8 * m);
    const uint32_t W0 = (uint32_t)(32 - f) | ((uint32_t)f << 8);
#pragma unroll
    for (int j = 0; j < N / 4; j++) {
      const uint32_t b34 = __funnelshift_r(lo[j], lo[j + 1], 24);
      const uint32_t v0 = __dp4a(lo[j], W0, 16u), v1 = __dp4a(lo[j], W0 << 8, 16u);
      const uint32_t v2 = __dp4a(lo[j], W0 << 16, 16u), v3 = __dp4a(b34, W0, 16u);
      h[2 * j] = (__byte_perm(v0, v1, 0x5410) >> 5) & 0x00ff00ffu;
      h[2 * j + 1] = (__byte_perm(v2, v3, 0x5410) >> 5) & 0x00ff00ffu;
    }
    if ((mode == 26 || mode == 10) && cidx == 0 && N < 32 && edge_filters)
      h[0] = (h[0] & 0xffff0000u) | (uint32_t)clip8(Mp[1] + ((Op[1 + l] - Mp[0]) >> 1));
    return horiz;
  }
  if (mode == 0) {  // planar (:1009-1032), lane = row y
    const int y = l;
    const uint32_t abd = (uint32_t)L[1 + y] | ((uint32_t)T[N + 1] << 8) | ((uint32_t)L[N + 1] << 24);
    const uint32_t Wy = ((uint32_t)(N - 1 - y) << 16) | ((uint32_t)(y + 1) << 24);
    const uint32_t* tw = reinterpret_cast<const uint32_t*>(T + 1);
#pragma unroll
    for (int j = 0; j < N / 4; j++) {
      const uint32_t t4 = tw[j];
      uint32_t v[4];
#pragma unroll
      for (int i = 0; i < 4; i++) {
        const int x = 4 * j + i;
        const uint32_t op = __byte_perm(abd, t4, 0x3010u | ((4u + i) << 8));   // [L[1+y], T[N+1], T[1+x], L[N+1]]
        v[i] = __dp4a(op, Wy | (uint32_t)(N - 1 - x) | ((uint32_t)(x + 1) << 8), (uint32_t)N) >> (LOG2 + 1);
      }
      h[2 * j] = v[0] | (v[1] << 16);
      h[2 * j + 1] = v[2] | (v[3] << 16);
    }
    return false;
  }
  // DC (:1036-1069), lane = row y
  int dc = active ? T[1 + l] + L[1 + l] : 0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) dc += __shfl_xor_sync(B265_FULL_MASK, dc, o);
  dc = (dc + N) >> (LOG2 + 1);
#pragma unroll
  for (int i = 0; i < N / 2; i++) h[i] = (uint32_t)dc * 0x00010001u;
  if (cidx == 0 && N < 32) {
    if (l == 0) {
#pragma unroll
      for (int i = 0; i < N / 2; i++)
        h[i] = (uint32_t)((T[1 + 2 * i] + 3 * dc + 2) >> 2) | ((uint32_t)((T[2 + 2 * i] + 3 * dc + 2) >> 2) << 16);
      h[0] = (h[0] & 0xffff0000u) | (uint32_t)((L[1] + 2 * dc + T[1] + 2) >> 2);
    } else {
      h[0] = (h[0] & 0xffff0000u) | (uint32_t)((L[1 + l] + 3 * dc + 2) >> 2);
    }
  }
  return false;
}

// Clip1(pred + residual) on the 16-bit pairs, then pack the line to bytes.  r = residual line as int16 pairs
// (values pre-clamped to +-1024 by the residual kernel, so the 16-bit add cannot wrap).
template <int N>
__device__ __forceinline__ void finish_line(uint32_t (&h)[N / 2], const uint32_t (&r)[N / 2], bool coded, uint32_t (&px)[N / 4]) {
  if (coded) {
#pragma unroll
    for (int i = 0; i < N / 2; i++) h[i] = __viaddmin_s16x2_relu(h[i], r[i], 0x00ff00ffu);
  }
#pragma unroll
  for (int j = 0; j < N / 4; j++) px[j] = __byte_perm(h[2 * j], h[2 * j + 1], 0x6420);
}
