// k_sao.cu — sample adaptive offset: reads the deblocked picture (FrameDev::plane), writes the output /
// DPB surface (FrameDev::out).  Restates apply_sao_internal (libde265/sao.cc:29-254) and its driver
// (:318-370): edge offset with picture/slice/tile boundary rules, band offset, pcm/bypass skip.
// Quirk kept on purpose: the CTB's own slice address is looked up with COMPONENT coordinates through
// a luma accessor (sao.cc:50-56), which only matters for chroma in multi-slice pictures.
// One thread per aligned word of 4 horizontally adjacent samples x 8 rows, all four samples filtered at once (half2
// compares, PRMT as a byte look-up table, 16-bit SIMD add+clip); neighbours come from up to six more word loads
// that hit L1.  Words on the rim of a CTB whose neighbouring CTB belongs to another slice
// or tile take the per-sample path with the reference's boundary rules.  Traffic 2P per picture (+ 18 B per CTB).
#include <cuda_fp16.h>
#include "transform.cuh"
#include "kernels.h"

__device__ __forceinline__ int sgn(int v) { return (v > 0) - (v < 0); }

// the reference's per-sample edge-offset path (sao.cc:104-180) for one word; returns the filtered word
__device__ __noinline__ uint32_t sao_edge_slow(const FrameDev& f, const b265_sao& s, const b265_slice& sl, int c, int x0, int y, int nvalid,
                                               uint32_t cen, uint32_t nofilt) {
  const int sh = c ? 1 : 0;
  const int W = c ? f.cw : f.width, H = c ? f.ch : f.height;
  const int pitch = f.pitch[c];
  const uint8_t* in = f.plane[c];
  const int lc = f.log2_ctb - sh;
  const int xCtb = x0 >> lc, yCtb = y >> lc;
  const int ctb = yCtb * f.ctbs_w + xCtb;
  const int qx = (xCtb << lc) >> f.log2_ctb, qy = (yCtb << lc) >> f.log2_ctb;
  const int ctb_slice_addr = f.slices[f.ctbs ? f.ctbs[qy * f.ctbs_w + qx].slice : 0].slice_addr_rs;
  const int tile = f.ctbs ? f.ctbs[ctb].tile : 0;
  int dx0 = 0, dy0 = 0, dx1 = 0, dy1 = 0;
  switch (s.band_eo[c]) {
    case 0: dx0 = -1; dx1 = 1; break;
    case 1: dy0 = -1; dy1 = 1; break;
    case 2: dx0 = -1; dy0 = -1; dx1 = 1; dy1 = 1; break;
    default: dx0 = 1; dy0 = -1; dx1 = -1; dy1 = 1; break;
  }
  uint32_t res = cen;
  for (int k = 0; k < nvalid; k++) {
    if ((nofilt >> (8 * k)) & 0xff) continue;
    const int x = x0 + k;
    const int v = (cen >> (8 * k)) & 0xff;
    bool keep = false;
    int nb[2];
    const int cx = x - (xCtb << lc), cy = y - (yCtb << lc);
    const int cw_ = min(1 << lc, W - (xCtb << lc)), ch_ = min(1 << lc, H - (yCtb << lc));
    const bool on_rim = cx == 0 || cy == 0 || cx == cw_ - 1 || cy == ch_ - 1;
#pragma unroll
    for (int t = 0; t < 2; t++) {
      const int xs = x + (t ? dx1 : dx0), ys = y + (t ? dy1 : dy0);
      if (xs < 0 || ys < 0 || xs >= W || ys >= H) { keep = true; nb[t] = 0; continue; }
      nb[t] = in[(size_t)ys * pitch + xs];
      const int nctb = (ys >> lc) * f.ctbs_w + (xs >> lc);
      if (on_rim && f.ctbs) {  // the reference runs these tests for rim samples only (sao.cc:119)
        const b265_slice& ns = f.slices[f.ctbs[nctb].slice];
        if (ns.slice_addr_rs < ctb_slice_addr && !sl.lf_across_slices) keep = true;
        if (ns.slice_addr_rs > ctb_slice_addr && !ns.lf_across_slices) keep = true;
        if (!(f.flags & B265_FRAME_LF_ACROSS_TILES) && f.ctbs[nctb].tile != tile) keep = true;
      }
    }
    if (keep) continue;
    const int e = sgn(v - nb[0]) + sgn(v - nb[1]);
    const int off = (e == 0) ? 0 : s.offset[c][e < 0 ? e + 2 : e + 1];
    res = (res & ~(0xffu << (8 * k))) | ((uint32_t)clip8(v + off) << (8 * k));
  }
  return res;
}

#define SAO_ROWS 8   // rows per thread (divides every CTB size of every component)

// ---- the per-word arithmetic.  sm_100 has no byte-SIMD ALU (vcmp4 / vadd4 / vaddus4 are emulated with 5-8
// instructions each), so the edge category is computed on half2 lanes — 1024+b is the half 0x6400|b, one PRMT per two
// samples — where [c > a] = sat(c - a) and [c >= a] = sat(c + 1 - a) are single HADD2.SAT instructions on the FMA
// pipe, and the final clip8(c + offset) is one VIADDMNMX.S16x2.RELU per two samples. ----
__device__ __forceinline__ uint32_t h2u(__half2 h) { return *reinterpret_cast<uint32_t*>(&h); }
__device__ __forceinline__ __half2 u2h(uint32_t u) { return *reinterpret_cast<__half2*>(&u); }
// samples 0,1 / 2,3 of a word as half2 (1024 + sample)
__device__ __forceinline__ __half2 half_lo(uint32_t w) { return u2h(__byte_perm(w, 0x64646464u, 0x4140)); }
__device__ __forceinline__ __half2 half_hi(uint32_t w) { return u2h(__byte_perm(w, 0x64646464u, 0x4342)); }
// prmt with the full selector (__byte_perm masks the nibbles to 3 bits, which drops the sign-replication mode)
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t d;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
  return d;
}
// clip8(sample + signed offset byte) for the four samples of a word
__device__ __forceinline__ uint32_t add_offsets(uint32_t v, uint32_t off) {
  // offsets sign-extended to 16 bit: selector nibble 8|k = byte k's sign replicated
  const uint32_t r01 = __viaddmin_s16x2_relu(__byte_perm(v, 0, 0x4140), prmt(off, 0, 0x9180), 0x00ff00ffu);
  const uint32_t r23 = __viaddmin_s16x2_relu(__byte_perm(v, 0, 0x4342), prmt(off, 0, 0xb3a2), 0x00ff00ffu);
  return __byte_perm(r01, r23, 0x6420);
}
// edge offset of one word: c = centre samples, a / b = the two neighbours of every sample (sao.cc:168-175);
// tab_lo/tab_hi = offsets indexed by edgeIdx + 2
__device__ __forceinline__ uint32_t sao_edge_word(uint32_t c, uint32_t a, uint32_t b, uint32_t tab_lo, uint32_t tab_hi) {
  const __half2 one = __floats2half2_rn(1.f, 1.f);
  const __half2 wgt = __floats2half2_rn(1.f, 16.f), bias = __floats2half2_rn(1024.f, 1024.f);
  uint32_t w[2];
#pragma unroll
  for (int h = 0; h < 2; h++) {
    const __half2 ch = h ? half_hi(c) : half_lo(c), ah = h ? half_hi(a) : half_lo(a), bh = h ? half_hi(b) : half_lo(b);
    const __half2 c1 = __hadd2(ch, one);
    // edgeIdx + 2 = [c > a] + [c >= a] + [c > b] + [c >= b]
    const __half2 t = __hadd2(__hsub2_sat(ch, ah), __hsub2_sat(c1, ah));
    const __half2 u = __hadd2(__hsub2_sat(ch, bh), __hsub2_sat(c1, bh));
    w[h] = h2u(__hfma2(__hadd2(t, u), wgt, bias));   // lanes {1024 + idx, 1024 + 16*idx}: the low bytes are idx, 16*idx
  }
  const uint32_t g = __byte_perm(w[0], w[1], 0x6240);   // {idx0, idx2, 16*idx1, 16*idx3}
  const uint32_t sel = g | (g >> 16);                    // PRMT selector nibbles idx0..idx3 (upper bits are ignored)
  return add_offsets(c, __byte_perm(tab_lo, tab_hi, sel));
}
// band offset of one word (sao.cc:182-251): kb = ((v >> 3) - band_position) & 31, offset[kb] for kb < 4
__device__ __forceinline__ uint32_t sao_band_word(uint32_t c, uint32_t negpos4, uint32_t o4) {
  const uint32_t kb = (((c >> 3) & 0x1f1f1f1fu) + negpos4) & 0x1f1f1f1fu;
  const uint32_t ge4 = ((kb + 0x7c7c7c7cu) & 0x80808080u) >> 5;           // 4 where kb >= 4
  const uint32_t idx = (kb & 0x03030303u) | ge4;                          // 0..3 -> offsets, 4..7 -> 0
  const uint32_t n = (idx | (idx >> 4)) & 0x00ff00ffu;                    // nibble pairs in bytes 0 and 2
  return add_offsets(c, __byte_perm(o4, 0u, __byte_perm(n, 0, 0x4420)));
}

struct SaoEdgeArgs {
  const FrameDev* f; const b265_sao* s; const b265_slice* sl;
  const uint8_t* in; uint8_t* out;
  int pitch, c, x0, y0, H, rows, cy0, ch_;
  uint32_t keep_x, tab_lo, tab_hi, nf[4];
  bool has_l, has_r, any_nf, foreign, rim_lr;
};
// Threads of a CTB that has a neighbour CTB in another slice or tile (rare: multi-slice / tile pictures): plain row
// loop, rim rows / rim words sample by sample with the reference's boundary rules, the others like the fast path.
__device__ __noinline__ void sao_edge_rows_foreign(const SaoEdgeArgs& e, int cls) {
  const int pitch = e.pitch, sh = e.c ? 1 : 0;
  const int W = e.c ? e.f->cw : e.f->width;
  for (int r = 0; r < e.rows; r++) {
    const int y = e.y0 + r, cy = e.cy0 + r;
    const uint8_t* p = e.in + (ptrdiff_t)r * pitch;
    const uint32_t cc = *reinterpret_cast<const uint32_t*>(p);
    uint32_t nofilt = e.any_nf ? (sh ? e.nf[r >> 1] : e.nf[r >> 2]) : 0u;
    uint32_t res;
    if (e.rim_lr || cy == 0 || cy == e.ch_ - 1) {
      res = sao_edge_slow(*e.f, *e.s, *e.sl, e.c, e.x0, y, 4, cc, nofilt);
      nofilt = 0;
    } else {
      // interior of the CTB: all neighbours exist (the word is not on the picture edge either)
      auto word_at = [&](int dx, int dy) {   // the four samples at (x0 + dx .. x0 + dx + 3, y + dy), |dx| <= 1
        const uint8_t* q = p + (ptrdiff_t)dy * pitch;
        const uint32_t m = *reinterpret_cast<const uint32_t*>(q);
        if (dx < 0) return __funnelshift_r(e.x0 > 0 ? *reinterpret_cast<const uint32_t*>(q - 4) : 0u, m, 24);
        if (dx > 0) return __funnelshift_r(m, e.x0 + 4 < W ? *reinterpret_cast<const uint32_t*>(q + 4) : 0u, 8);
        return m;
      };
      uint32_t a, b;
      if (cls == 0) { a = word_at(-1, 0); b = word_at(1, 0); }
      else if (cls == 1) { a = word_at(0, -1); b = word_at(0, 1); }
      else if (cls == 2) { a = word_at(-1, -1); b = word_at(1, 1); }
      else { a = word_at(1, -1); b = word_at(-1, 1); }
      res = sao_edge_word(cc, a, b, e.tab_lo, e.tab_hi);
    }
    res = (res & ~nofilt) | (cc & nofilt);
    *reinterpret_cast<uint32_t*>(e.out + (ptrdiff_t)r * pitch) = res;
  }
}

// The thread's SAO_ROWS rows of one word, edge class as DATA (0: horizontal, 1: vertical, 2: 135 deg, 3: 45 deg): the
// lanes of a warp may belong to CTBs with different classes (chroma CTBs are 16 samples wide), and a class-specialised
// walker would run once per class present in the warp.  Neighbour a of the word's samples is at (dx, dy), neighbour b at
// (-dx, -dy); each neighbour word is funnel-shifted out of the two aligned words that hold it (both hit L1).
__device__ __forceinline__ void sao_edge_rows(const uint8_t* in, uint8_t* out, int pitch, const uint32_t (&v)[SAO_ROWS], int cls,
                                              int rows, int y0, int H, bool has_l, bool has_r, uint32_t keep_x,
                                              uint32_t tab_lo, uint32_t tab_hi, bool any_nf, const uint32_t (&nf)[4], int sh) {
  const int dx = (cls == 1) ? 0 : (cls == 3) ? 1 : -1, dy = (cls == 0) ? 0 : -1;
  // words holding neighbour a: lo at byte offset a_lo, hi at a_hi (relative to the centre word), shifted right by a_sh
  // bits; words outside the picture are never dereferenced (their samples are masked by keep_x)
  const int l4 = has_l ? -4 : 0, r4 = has_r ? 4 : 0;
  const int a_lo = dx < 0 ? l4 : 0, a_hi = dx > 0 ? r4 : 0, a_sh = dx < 0 ? 24 : dx > 0 ? 8 : 0;
  const int b_lo = dx > 0 ? l4 : 0, b_hi = dx < 0 ? r4 : 0, b_sh = dx > 0 ? 24 : dx < 0 ? 8 : 0;
  const bool top = y0 == 0, bot = y0 + rows == H;   // rows is 4 or 8
  const int up = dy * pitch;   // 0 or -pitch
  const uint8_t* prow = in;
  uint8_t* orow = out;
  auto ldw = [](const uint8_t* p) { return __ldg(reinterpret_cast<const uint32_t*>(p)); };
#pragma unroll
  for (int r = 0; r < SAO_ROWS; r++) {
    if (r == 4 && rows == 4) break;
    const bool last = (r == SAO_ROWS - 1) || (r == 3 && rows == 4);
    // rows above / below the picture are not read: the whole row stays unfiltered then
    const bool edge_row = dy != 0 && ((r == 0 && top) || (last && bot));
    const uint8_t* pa = prow + ((r == 0 && top) ? 0 : up);
    const uint8_t* pb = prow - ((last && bot) ? 0 : up);
    const uint32_t cc = v[r];
    const uint32_t a = __funnelshift_r(ldw(pa + a_lo), ldw(pa + a_hi), a_sh);
    const uint32_t b = __funnelshift_r(ldw(pb + b_lo), ldw(pb + b_hi), b_sh);
    uint32_t nofilt = any_nf ? (sh ? nf[r >> 1] : nf[r >> 2]) : 0u;
    uint32_t res = sao_edge_word(cc, a, b, tab_lo, tab_hi);
    nofilt |= keep_x;
    if ((r == 0 || r == 3 || r == SAO_ROWS - 1) && edge_row) nofilt = 0xffffffffu;
    res = (res & ~nofilt) | (cc & nofilt);
    *reinterpret_cast<uint32_t*>(orow) = res;
    prow += pitch;
    orow += pitch;
  }
}

#ifndef SAO_WARPS
#define SAO_WARPS 4   // warps per block: 4 = one row of four 32x32 regions (8 = two rows: measured 5 % slower, blocks retire later)
#endif
#define SAO_WY (SAO_WARPS / 4)
__global__ void __launch_bounds__(SAO_WARPS * 32, 1280 / (SAO_WARPS * 32)) k_sao(const FrameDev* __restrict__ frames) {
  const FrameDev& f = frames[blockIdx.z / 3];
  const uint32_t fflags = f.flags;
  if (!(fflags & B265_FRAME_SAO) || !f.sao_rec) return;
  const int fcw = f.cw;
  const int cz = blockIdx.z % 3;
  if (cz && !fcw) return;
  // A block covers 128 x 64 samples of a plane with 8 warps.  Luma: each warp a 32 x 32 region (8 words x 4 row groups),
  // so the lanes of a warp share one CTB's parameters wherever the CTB is at least 32 samples wide (no divergence on
  // type / class).  Chroma: Cb and Cr share the SAO type and edge class (one syntax element, slice.cc:2735-2760), so a
  // warp takes the SAME 32 x 16 region in both planes (lanes 0-15 Cb, 16-31 Cr); the blocks of grid plane z=1 cover the
  // even 32-row bands, z=2 the odd ones.
  const int wrp = threadIdx.y, lane = threadIdx.x;
  int c, xw, yrow;
  if (cz == 0) {
    c = 0;
    xw = (wrp & 3) * 8 + (lane & 7);
    yrow = (blockIdx.y * (4 * SAO_WY) + (wrp >> 2) * 4 + (lane >> 3)) * SAO_ROWS;
  } else {
    c = 1 + (lane >> 4);
    xw = (wrp & 3) * 8 + (lane & 7);
    yrow = ((blockIdx.y * 2 + (cz - 1)) * (2 * SAO_WY) + (wrp >> 2) * 2 + ((lane >> 3) & 1)) * SAO_ROWS;
  }
  const int sh = c ? 1 : 0;
  const int W = c ? fcw : f.width, H = c ? f.ch : f.height;
  const int x0 = (blockIdx.x * 32 + xw) << 2, y0 = yrow;
  if (x0 >= W || y0 >= H) return;
  const int pitch = f.pitch[c];
  const uint8_t* in = f.plane[c] + (size_t)y0 * pitch + x0;
  uint8_t* out = f.out[c] + (size_t)y0 * pitch + x0;
  // plane widths and heights are multiples of 4 (the coded size is a multiple of the 8x8 minimum coding block), so
  // every word is complete and a thread has 4 or 8 rows
  const int rows = min(SAO_ROWS, H - y0);
  // the centre words do not depend on the CTB's parameters: request them before the parameter record
  uint32_t v[SAO_ROWS];
#pragma unroll
  for (int r = 0; r < SAO_ROWS; r++) v[r] = (r < 4 || rows > 4) ? __ldg(reinterpret_cast<const uint32_t*>(in + (ptrdiff_t)r * pitch)) : 0u;
  const int lc = f.log2_ctb - sh;  // CTB size in this component
  const int xCtb = x0 >> lc, yCtb = y0 >> lc;
  const int ctbs_w = f.ctbs_w;
  const int ctb = yCtb * ctbs_w + xCtb;
  const uint2 rec = __ldg(f.sao_rec + (size_t)c * (ctbs_w * f.ctbs_h) + ctb);
  const int type = rec.x & 3;
  if (type == 0) {
#pragma unroll
    for (int r = 0; r < SAO_ROWS; r++) if (r < 4 || rows > 4) *reinterpret_cast<uint32_t*>(out + (ptrdiff_t)r * pitch) = v[r];
    return;
  }
  const uint32_t o4 = rec.y;
  const int cls = (rec.x >> 2) & 31;   // band position (type 1) or edge class (type 2)
  // pcm (with pcm_loop_filter_disabled) / transquant-bypass samples are not filtered (sao.cc:112-117): the unit flags
  // under this thread's samples, fetched once.  nf[q] covers rows [q*SAO_ROWS/4, (q+1)*SAO_ROWS/4) for chroma (one luma
  // unit row = 2 chroma rows) and rows [4*(q&1), ..) for luma (nf[0], nf[1]).
  uint32_t nf[4] = {0, 0, 0, 0};
  bool any_nf = false;
  if (const uint8_t* unit_info = f.unit_info) {
    const int upitch = f.width >> 2;
    if (!sh) {
      const uint8_t* ui = unit_info + (size_t)(y0 >> 2) * upitch + (x0 >> 2);
      nf[0] = (__ldg(ui) & B265_UNIT_NOFILT) ? 0xffffffffu : 0u;
      if (rows > 4) nf[1] = (__ldg(ui + upitch) & B265_UNIT_NOFILT) ? 0xffffffffu : 0u;
    } else {
#pragma unroll
      for (int q = 0; q < 4; q++) {
        if (2 * q < rows) {
          const uint8_t* ui = unit_info + (size_t)((y0 >> 1) + q) * upitch + (x0 >> 1);
          nf[q] = ((__ldg(ui) & B265_UNIT_NOFILT) ? 0x0000ffffu : 0u) |
                  ((__ldg(ui + 1) & B265_UNIT_NOFILT) ? 0xffff0000u : 0u);
        }
      }
    }
    any_nf = (nf[0] | nf[1] | nf[2] | nf[3]) != 0;
  }
  if (type == 1) {
    const uint32_t negpos4 = 0x01010101u * (uint32_t)((32 - cls) & 31);
#pragma unroll
    for (int r = 0; r < SAO_ROWS; r++) {
      if (r == 4 && rows == 4) break;
      uint32_t res = sao_band_word(v[r], negpos4, o4);
      if (any_nf) {
        const uint32_t m = sh ? nf[r >> 1] : nf[r >> 2];
        res = (res & ~m) | (v[r] & m);
      }
      *reinterpret_cast<uint32_t*>(out + (ptrdiff_t)r * pitch) = res;
    }
    return;
  }
  const int cx = x0 - (xCtb << lc), cy0 = y0 - (yCtb << lc);
  const int cw_ = min(1 << lc, W - (xCtb << lc)), ch_ = min(1 << lc, H - (yCtb << lc));
  // Edge offset next to a CTB of another slice or tile follows the reference's boundary rules sample by sample
  // (record flag, sao_records()); only words on the CTB rim can be affected.
  const bool rim_lr = cx == 0 || cx + 4 >= cw_;
  const bool foreign = (rec.x & B265_SAOREC_FOREIGN) && (rim_lr || cy0 == 0 || cy0 + SAO_ROWS >= ch_);
  const uint32_t tab_lo = (o4 & 0xffffu) | ((o4 & 0x00ff0000u) << 8);   // edgeIdx + 2 -> {o0, o1, 0, o2 | o3}
  const uint32_t tab_hi = o4 >> 24;
  uint32_t keep_x = 0;   // bytes whose horizontal neighbour lies outside the picture stay unfiltered
  if (cls != 1) {
    if (x0 == 0) keep_x |= 0x000000ffu;
    if (x0 + 4 >= W) keep_x |= 0xff000000u;
  }
  if (foreign) {
    const b265_ctb* ctbs = f.ctbs;
    SaoEdgeArgs e;
    e.f = &f; e.s = &f.sao[ctb]; e.sl = &f.slices[ctbs ? ctbs[ctb].slice : 0];
    e.in = in; e.out = out; e.pitch = pitch; e.c = c; e.x0 = x0; e.y0 = y0; e.H = H;
    e.rows = rows; e.has_l = x0 > 0; e.has_r = x0 + 4 < W; e.keep_x = keep_x; e.tab_lo = tab_lo;
    e.tab_hi = tab_hi; e.any_nf = any_nf; e.foreign = foreign; e.rim_lr = rim_lr; e.cy0 = cy0; e.ch_ = ch_;
#pragma unroll
    for (int q = 0; q < 4; q++) e.nf[q] = nf[q];
    sao_edge_rows_foreign(e, cls);
    return;
  }
  sao_edge_rows(in, out, pitch, v, cls, rows, y0, H, x0 > 0, x0 + 4 < W, keep_x, tab_lo, tab_hi, any_nf, nf, sh);
}

void launch_sao(cudaStream_t st, const FrameDev* d_frames, int n_frames, int max_w, int max_h) {
  dim3 g((((max_w + 3) >> 2) + 31) / 32, (max_h + 4 * SAO_WY * SAO_ROWS - 1) / (4 * SAO_WY * SAO_ROWS), 3 * n_frames);
  k_sao<<<g, dim3(32, SAO_WARPS), 0, st>>>(d_frames);
}
