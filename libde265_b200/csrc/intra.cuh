// intra.cuh — intra prediction as group-cooperative device code (a group = N lanes = one N x N TB).
//
// Restates libde265/intrapred.cc: intra_border_computer (:344-688: availability by picture edge,
// CTB slice/tile ownership, z-scan order and constrained_intra_pred; reference-sample substitution
// :395-431), intra_prediction_sample_filtering (:816-889), angular (:903-1005), planar (:1009-1032)
// and DC (:1036-1069).  Border convention as in the reference: p[0] = top-left corner,
// p[1..2N] = top row then top-right, p[-1..-2N] = left column then bottom-left.
//
// Neighbouring samples are produced either by the same warp (earlier TBs of the same CTB) or by
// other warps of the SAME launch (CTB wavefront, k_intra.cu): the warp waits for the neighbouring
// CTBs' completion flags once per CTB and then reads all border pixels through L2 (ld.cg).
#pragma once
#include "common.cuh"

#define B265_BORDER_BYTES 512  // per warp: 32/N groups x 2 buffers x (4N+8) bytes

template <int N> struct IntraBuf { static constexpr int STRIDE = 4 * N + 8; };

__device__ __forceinline__ unsigned morton4(unsigned x, unsigned y) {
  // interleave the low 4 bits of x (even positions) and y (odd positions)
  x = (x | (x << 2)) & 0x33; x = (x | (x << 1)) & 0x55;
  y = (y | (y << 2)) & 0x33; y = (y | (y << 1)) & 0x55;
  return x | (y << 1);
}

struct IntraGeom {
  int sh;            // 0 luma, 1 chroma (4:2:0)
  int xL, yL;        // luma position of the block
  int cur_ctb, cur_slice, cur_tile;
  unsigned cur_zs;   // z-scan order address of the block at min-TB granularity
};

// z-scan address of the min-TB containing luma sample (x,y)  (pps MinTbAddrZS)
__device__ __forceinline__ unsigned zs_addr(const FrameDev& f, int x, int y, int ctb) {
  const int d = f.log2_ctb - f.log2_min_tb;
  const int m = (1 << f.log2_ctb) - 1;
  const unsigned ts = f.ctbs ? (unsigned)f.ctbs[ctb].addr_ts : (unsigned)ctb;
  return (ts << (2 * d)) + morton4((x & m) >> f.log2_min_tb, (y & m) >> f.log2_min_tb);
}

// availability of the neighbouring 4-sample unit whose defining luma sample is (nx,ny)
__device__ __forceinline__ bool unit_available(const FrameDev& f, const IntraGeom& g, int nx, int ny) {
  if (nx < 0 || ny < 0 || nx >= f.width || ny >= f.height) return false;
  const int ctb = (ny >> f.log2_ctb) * f.ctbs_w + (nx >> f.log2_ctb);
  if (ctb != g.cur_ctb && f.ctbs) {
    if (f.ctbs[ctb].slice != g.cur_slice || f.ctbs[ctb].tile != g.cur_tile) return false;
  }
  if (zs_addr(f, nx, ny, ctb) > g.cur_zs) return false;
  if ((f.flags & B265_FRAME_CONSTRAINED_INTRA) && f.unit_info) {
    if (!(f.unit_info[(ny >> 2) * (f.width >> 2) + (nx >> 2)] & B265_UNIT_INTRA)) return false;
  }
  return true;
}

// Gather p[-2N..2N] of the TB at component position (xB,yB) of plane c into buf[i + 2N].
// Returns nothing; all lanes of the warp call it (ballots). `active` false for idle groups.
template <int LOG2>
__device__ __forceinline__ void gather_border(const FrameDev& f, int c, int xB, int yB, int sub, int grp,
                                              bool active, uint8_t* buf) {
  constexpr int N = 1 << LOG2, NU = N / 4;  // NU units per N samples
  const int sh = (c && f.cw) ? 1 : 0;
  IntraGeom g;
  g.sh = sh; g.xL = xB << sh; g.yL = yB << sh;
  g.cur_ctb = (g.yL >> f.log2_ctb) * f.ctbs_w + (g.xL >> f.log2_ctb);
  g.cur_slice = f.ctbs ? f.ctbs[active ? g.cur_ctb : 0].slice : 0;
  g.cur_tile = f.ctbs ? f.ctbs[active ? g.cur_ctb : 0].tile : 0;
  g.cur_zs = active ? zs_addr(f, g.xL, g.yL, g.cur_ctb) : 0;
  const uint8_t* plane = f.plane[c];
  const int pitch = f.pitch[c];

  // lane `sub` owns scan unit: sub < 2NU -> left column unit (bottom-up), else top row unit
  const bool is_left = sub < 2 * NU;
  int ux, uy;  // component position of the unit's defining sample (reference: bottom row / left col)
  if (is_left) { ux = xB - 1; uy = yB + 2 * N - 1 - 4 * sub; }
  else         { ux = xB + 4 * (sub - 2 * NU); uy = yB - 1; }
  bool av = active && unit_available(f, g, ux << sh, uy << sh);
  // (the reference tests slice/tile ownership once per neighbouring CTB, intrapred.cc:486-508; a
  //  left/top/top-right run of N samples never straddles two CTBs, so the per-unit test is the same)
  bool avc = false;
  if (active && sub == 0) avc = unit_available(f, g, (xB - 1) << sh, (yB - 1) << sh);

  uint32_t px = 0;
  if (av) {
    if (is_left) {
#pragma unroll
      for (int i = 0; i < 4; i++)  // border index -2N+4*sub+i  <->  row uy-(3-i)... ascending index = going up
        px |= (uint32_t)ld_cg_u8(plane + (size_t)(uy - i) * pitch + ux) << (8 * i);
    } else {
      const uint8_t* s = plane + (size_t)uy * pitch + ux;
      px = ((reinterpret_cast<uintptr_t>(s) & 3) == 0) ? ld_cg_u32(s)
           : ((uint32_t)ld_cg_u8(s) | ((uint32_t)ld_cg_u8(s + 1) << 8) | ((uint32_t)ld_cg_u8(s + 2) << 16) |
              ((uint32_t)ld_cg_u8(s + 3) << 24));
    }
  }
  uint32_t pc = 0;
  if (avc) {
    pc = ld_cg_u8(plane + (size_t)(yB - 1) * pitch + xB - 1);
  }
  // scan-order availability mask of this group: bits 0..2NU-1 left, bit 2NU corner, then top
  const uint32_t ball = __ballot_sync(B265_FULL_MASK, av);
  const uint32_t ballc = __ballot_sync(B265_FULL_MASK, avc);
  const uint32_t gm = (N == 32) ? ball : ((ball >> (grp * N)) & ((1u << N) - 1u));
  const uint64_t corner = (ballc >> (grp * N)) & 1u;
  const uint64_t leftm = gm & ((1u << (2 * NU)) - 1u);
  const uint64_t topm = (uint64_t)(gm >> (2 * NU));
  const uint64_t M = leftm | (corner << (2 * NU)) | (topm << (2 * NU + 1));

  // unit -> border index of its first (lowest-index) sample
  auto first_index = [&](int s) -> int {  // scan unit s
    if (s < 2 * NU) return -2 * N + 4 * s;
    if (s == 2 * NU) return 0;
    return 1 + 4 * (s - 2 * NU - 1);
  };
  auto last_index = [&](int s) -> int { return (s == 2 * NU) ? 0 : first_index(s) + 3; };

  // left units store bottom-up: border index -2N+4*sub+i holds row uy-i?  index grows towards the corner
  if (active) {
    if (av) {
      const int base = is_left ? (-2 * N + 4 * sub) : (1 + 4 * (sub - 2 * NU));
      uint8_t* d = buf + 2 * N + base;
#pragma unroll
      for (int i = 0; i < 4; i++) d[i] = (px >> (8 * i)) & 0xff;
    }
    if (sub == 0 && avc) buf[2 * N] = (uint8_t)pc;
  }
  __syncwarp();
  {
    // reference_sample_substitution (intrapred.cc:395-431): missing samples copy the nearest
    // available sample below/left in scan order, leading ones copy the first available sample
    const bool incomplete = active && M != ((1ull << (4 * NU + 1)) - 1ull);
    auto subst = [&](int s) -> int {
      if (M == 0) return 128;
      const uint64_t lower = M & ((1ull << s) - 1ull);
      if (lower) return buf[2 * N + last_index(63 - __clzll((long long)lower))];
      return buf[2 * N + first_index(__ffsll((long long)M) - 1)];
    };
    const int s_me = is_left ? sub : sub + 1;
    int vme = 0, vc = 0;
    const bool need_me = incomplete && !av, need_c = incomplete && (sub == 0) && !corner;
    if (need_me) vme = subst(s_me);
    if (need_c) vc = subst(2 * NU);
    __syncwarp();
    if (need_me) {
      uint8_t* d = buf + 2 * N + first_index(s_me);
#pragma unroll
      for (int i = 0; i < 4; i++) d[i] = (uint8_t)vme;
    }
    if (need_c) buf[2 * N] = (uint8_t)vc;
  }
  __syncwarp();
}

// ---- intra_prediction_sample_filtering (intrapred.cc:816-889) ----
// p = src + 2N (border centre).  Writes the filtered border to dst (same layout) when the mode/size
// rule asks for it and returns the buffer to predict from.  Called by all lanes of the warp.
template <int LOG2>
__device__ __forceinline__ const uint8_t* filter_border(const uint8_t* src, uint8_t* dst, int sub, int mode,
                                                        int cidx, bool strong_enabled, bool active) {
  constexpr int N = 1 << LOG2;
  bool filt = false;
  if (active && cidx == 0 && mode != 1 && N != 4) {
    const int d = min(abs(mode - 26), abs(mode - 10));
    filt = (N == 8) ? d > 7 : (N == 16) ? d > 1 : d > 0;
  }
  const uint8_t* p = src + 2 * N;
  if (filt) {
    bool bi = false;
    if (N == 32 && strong_enabled)
      bi = abs(p[0] + p[64] - 2 * p[32]) < 8 && abs(p[0] + p[-64] - 2 * p[-32]) < 8;  // 1 << (bitDepth-5)
    uint8_t* q = dst + 2 * N;
    for (int j = sub; j < 4 * N + 1; j += N) {
      const int i = j - 2 * N;
      int v;
      if (i == -2 * N || i == 2 * N) v = p[i];
      else if (bi) v = (i == 0) ? p[0] : (i < 0 ? p[0] + ((-i * (p[-64] - p[0]) + 32) >> 6)
                                                : p[0] + ((i * (p[64] - p[0]) + 32) >> 6));
      else v = (p[i + 1] + 2 * p[i] + p[i - 1] + 2) >> 2;
      q[i] = (uint8_t)v;
    }
  }
  __syncwarp();
  return filt ? dst : src;
}

__constant__ int8_t c_intra_angle[35] = {0,  0,  32, 26, 21, 17, 13, 9,  5,  2,  0,  -2, -5, -9, -13, -17, -21, -26,
                                         -32, -26, -21, -17, -13, -9, -5, -2, 0,  2,  5,  9,  13, 17,  21,  26,  32};
__constant__ int16_t c_intra_inv_angle[15] = {-4096, -1638, -910, -630, -482, -390, -315, -256,
                                              -315,  -390,  -482, -630, -910, -1638, -4096};

// Prediction of row `sub` (N samples) from the border buffer b (b[i+2N] = p[i]).
// Contains group-wide shuffles (DC sum): all lanes of the warp call it.
template <int LOG2>
__device__ __forceinline__ void predict_row(const uint8_t* b, int sub, int mode, int cidx, bool edge_filters,
                                            int (&pred)[1 << LOG2]) {
  constexpr int N = 1 << LOG2;
  const uint8_t* p = b + 2 * N;
  const int y = sub;
  // DC value is needed by the whole group; compute it unconditionally (cheap) to keep shuffles uniform
  int dc = p[1 + sub] + p[-1 - sub];
#pragma unroll
  for (int o = N / 2; o > 0; o >>= 1) dc += __shfl_xor_sync(B265_FULL_MASK, dc, o);
  dc = (dc + N) >> (LOG2 + 1);
  if (mode == 0) {  // planar (intrapred.cc:1009-1032)
#pragma unroll
    for (int x = 0; x < N; x++)
      pred[x] = ((N - 1 - x) * p[-1 - y] + (x + 1) * p[N + 1] + (N - 1 - y) * p[1 + x] + (y + 1) * p[-N - 1] + N) >>
                (LOG2 + 1);
  } else if (mode == 1) {  // DC (:1036-1069)
#pragma unroll
    for (int x = 0; x < N; x++) pred[x] = dc;
    if (cidx == 0 && N < 32) {
      if (y == 0) {
#pragma unroll
        for (int x = 1; x < N; x++) pred[x] = (p[x + 1] + 3 * dc + 2) >> 2;
        pred[0] = (p[-1] + 2 * dc + p[1] + 2) >> 2;
      } else {
        pred[0] = (p[-y - 1] + 3 * dc + 2) >> 2;
      }
    }
  } else {  // angular (:903-1005)
    const int ang = c_intra_angle[mode];
    const int inv = (mode >= 11 && mode <= 25) ? c_intra_inv_angle[mode - 11] : 0;
    if (mode >= 18) {
      const int idx = ((y + 1) * ang) >> 5, fact = ((y + 1) * ang) & 31;
      auto ref = [&](int k) -> int { return k >= 0 ? p[k] : p[-((k * inv + 128) >> 8)]; };
#pragma unroll
      for (int x = 0; x < N; x++)
        pred[x] = fact ? ((32 - fact) * ref(x + idx + 1) + fact * ref(x + idx + 2) + 16) >> 5 : ref(x + idx + 1);
      if (mode == 26 && cidx == 0 && N < 32 && edge_filters) pred[0] = clip8(p[1] + ((p[-1 - y] - p[0]) >> 1));
    } else {
      auto ref = [&](int k) -> int { return k >= 0 ? p[-k] : p[(k * inv + 128) >> 8]; };
#pragma unroll
      for (int x = 0; x < N; x++) {
        const int idx = ((x + 1) * ang) >> 5, fact = ((x + 1) * ang) & 31;
        pred[x] = fact ? ((32 - fact) * ref(y + idx + 1) + fact * ref(y + idx + 2) + 16) >> 5 : ref(y + idx + 1);
      }
      if (mode == 10 && cidx == 0 && N < 32 && edge_filters && y == 0) {
#pragma unroll
        for (int x = 0; x < N; x++) pred[x] = clip8(p[-1] + ((p[1 + x] - p[0]) >> 1));
      }
    }
  }
}
