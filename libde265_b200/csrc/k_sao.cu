// k_sao.cu — sample adaptive offset: reads the deblocked picture (FrameDev::plane), writes the output /
// DPB surface (FrameDev::out).  Restates apply_sao_internal (libde265/sao.cc:29-254) and its driver
// (:318-370): edge offset with picture/slice/tile boundary rules, band offset, pcm/bypass skip.
// Quirk kept on purpose: the CTB's own slice address is looked up with COMPONENT coordinates through
// a luma accessor (sao.cc:50-56), which only matters for chroma in multi-slice pictures.
// One thread per aligned word of 4 horizontally adjacent samples, all four filtered at once with the byte-SIMD
// instructions (compare masks, saturating add/sub, PRMT as a byte look-up table); neighbours come from up to
// six more word loads that hit L1.  Words on the rim of a CTB whose neighbouring CTB belongs to another slice
// or tile take the per-sample path with the reference's boundary rules.  Traffic 2P per picture (+ 18 B per CTB).
#include "transform.cuh"
#include "kernels.h"

__device__ __forceinline__ int sgn(int v) { return (v > 0) - (v < 0); }

// bytes of `idx` (0..7 each) select bytes of the 8-entry table {lo, hi}
__device__ __forceinline__ uint32_t lut8(uint32_t lo, uint32_t hi, uint32_t idx) {
  const uint32_t sel = (idx & 0xf) | ((idx >> 4) & 0xf0) | ((idx >> 8) & 0xf00) | ((idx >> 12) & 0xf000);
  return __byte_perm(lo, hi, sel);
}
// clip8(v + off) per byte, off signed
__device__ __forceinline__ uint32_t add_offsets(uint32_t v, uint32_t off) {
  const uint32_t pos = __vmaxs4(off, 0u), neg = __vmaxs4(__vneg4(off), 0u);
  return __vsubus4(__vaddus4(v, pos), neg);
}

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

__global__ void __launch_bounds__(256, 4) k_sao(const FrameDev* __restrict__ frames) {
  const FrameDev& f = frames[blockIdx.z / 3];
  const uint32_t fflags = f.flags;
  const b265_sao* sao = f.sao;
  if (!(fflags & B265_FRAME_SAO) || !sao) return;
  const int c = blockIdx.z % 3;
  const int fcw = f.cw;
  if (c && !fcw) return;
  const int sh = c ? 1 : 0;
  const int W = c ? fcw : f.width, H = c ? f.ch : f.height;
  // a block covers 128 x 64 samples; each warp a 32 x 32 region of it (8 words x 4 row groups), so that the lanes of
  // a warp share one CTB's parameters wherever the CTB is at least 32 samples wide (no divergence on type / class)
  const int xw = (threadIdx.y & 3) * 8 + (threadIdx.x & 7), rg = (threadIdx.y >> 2) * 4 + (threadIdx.x >> 3);
  const int x0 = (blockIdx.x * 32 + xw) << 2, y0 = (blockIdx.y * 8 + rg) * SAO_ROWS;
  if (x0 >= W || y0 >= H) return;
  const int pitch = f.pitch[c];
  const uint8_t* in = f.plane[c] + (size_t)y0 * pitch + x0;
  uint8_t* out = f.out[c] + (size_t)y0 * pitch + x0;
  const int lc = f.log2_ctb - sh;  // CTB size in this component
  const int xCtb = x0 >> lc, yCtb = y0 >> lc;
  const int ctbs_w = f.ctbs_w;
  const int ctb = yCtb * ctbs_w + xCtb;
  const b265_ctb* ctbs = f.ctbs;
  const b265_ctb me = ctbs ? ctbs[ctb] : b265_ctb{0, 0, 0};
  const b265_slice& sl = f.slices[me.slice];
  const b265_sao& s = sao[ctb];
  int type = s.type[c];
  if (!(c ? sl.sao_chroma : sl.sao_luma)) type = 0;
  const int nvalid = min(4, W - x0);
  const int rows = min(SAO_ROWS, H - y0);
  // rows are padded to the pitch (a multiple of 4), so whole words can be read and written
  if (type == 0) {
    uint32_t v[SAO_ROWS];
#pragma unroll
    for (int r = 0; r < SAO_ROWS; r++) if (r < rows) v[r] = __ldg(reinterpret_cast<const uint32_t*>(in + (size_t)r * pitch));
#pragma unroll
    for (int r = 0; r < SAO_ROWS; r++) if (r < rows) *reinterpret_cast<uint32_t*>(out + (size_t)r * pitch) = v[r];
    return;
  }
  const uint8_t* unit_info = f.unit_info;
  const int upitch = f.width >> 2;
  const uint32_t o4 = (uint32_t)(uint8_t)s.offset[c][0] | ((uint32_t)(uint8_t)s.offset[c][1] << 8) |
                      ((uint32_t)(uint8_t)s.offset[c][2] << 16) | ((uint32_t)(uint8_t)s.offset[c][3] << 24);
  const int cls = s.band_eo[c];   // band position (type 1) or edge class (type 2)
  const int cx = x0 - (xCtb << lc), cy0 = y0 - (yCtb << lc);
  const int cw_ = min(1 << lc, W - (xCtb << lc)), ch_ = min(1 << lc, H - (yCtb << lc));
  // Edge offset next to a CTB of another slice or tile follows the reference's boundary rules sample by sample.
  // `foreign`: some CTB around this one (or, for chroma, the CTB the reference takes this one's slice from,
  // sao.cc:50-56) differs in slice or tile; only words on the CTB rim can be affected.
  bool foreign = false;
  const bool rim_lr = cx == 0 || cx + 4 >= cw_;
  if (type == 2 && ctbs && (rim_lr || cy0 == 0 || cy0 + SAO_ROWS >= ch_)) {
    const int ctbs_h = f.ctbs_h;
    for (int dy = -1; dy <= 1; dy++)
      for (int dx = -1; dx <= 1; dx++) {
        const int nx = xCtb + dx, ny = yCtb + dy;
        if ((dx | dy) == 0 || nx < 0 || ny < 0 || nx >= ctbs_w || ny >= ctbs_h) continue;
        const b265_ctb o = ctbs[ny * ctbs_w + nx];
        foreign |= o.slice != me.slice || o.tile != me.tile;
      }
    if (sh) {
      const int qx = (xCtb << lc) >> f.log2_ctb, qy = (yCtb << lc) >> f.log2_ctb;
      foreign |= ctbs[qy * ctbs_w + qx].slice != me.slice;
    }
  }
  const uint32_t one = 0x01010101u;
  const uint32_t tab_lo = (o4 & 0xffffu) | ((o4 & 0x00ff0000u) << 8);   // edgeIdx + 2 -> {o0, o1, 0, o2 | o3}
  const uint32_t tab_hi = o4 >> 24;
  uint32_t keep_x = 0;   // bytes whose horizontal neighbour lies outside the picture stay unfiltered
  if (type == 2 && cls != 1) {
    if (x0 == 0) keep_x |= 0x000000ffu;
    if (x0 + nvalid >= W) keep_x |= 0xffu << (8 * (nvalid - 1));
  }
  const bool has_l = x0 > 0, has_r = x0 + 4 < W;
  const bool need_lr = type == 2 && cls != 1, need_ud = type == 2 && cls != 0, need_diag = type == 2 && cls >= 2;
  // rolling window over the rows: previous / current / next (left, centre, right words)
  uint32_t pl = 0, pc = 0, pr = 0, cl = 0, cc, cr = 0, nl = 0, nc = 0, nr = 0;
  auto load_row3 = [&](int r, uint32_t& l, uint32_t& m, uint32_t& rt, bool lr) {
    const uint8_t* p = in + (ptrdiff_t)r * pitch;
    m = __ldg(reinterpret_cast<const uint32_t*>(p));
    if (lr) {
      l = has_l ? __ldg(reinterpret_cast<const uint32_t*>(p - 4)) : 0u;
      rt = has_r ? __ldg(reinterpret_cast<const uint32_t*>(p + 4)) : 0u;
    }
  };
  if (need_ud && y0 > 0) load_row3(-1, pl, pc, pr, need_diag);
  load_row3(0, cl, cc, cr, need_lr);
#pragma unroll
  for (int r = 0; r < SAO_ROWS; r++) {
    if (r >= rows) break;
    const int y = y0 + r;
    const bool has_dn = y + 1 < H;
    if (r + 1 < SAO_ROWS || need_ud) {
      if ((r + 1 < rows) || (need_ud && has_dn)) load_row3(r + 1, nl, nc, nr, need_lr);
    }
    uint32_t nofilt = 0;
    if (unit_info) {   // pcm (with pcm_loop_filter_disabled) / transquant-bypass samples are not filtered (sao.cc:112-117)
      const uint8_t* ui = unit_info + ((y << sh) >> 2) * upitch + ((x0 << sh) >> 2);
      if (!sh) nofilt = (ui[0] & B265_UNIT_NOFILT) ? 0xffffffffu : 0u;
      else nofilt = ((ui[0] & B265_UNIT_NOFILT) ? 0x0000ffffu : 0u) | ((nvalid > 2 && (ui[1] & B265_UNIT_NOFILT)) ? 0xffff0000u : 0u);
    }
    uint32_t res;
    if (type == 1) {
      // band offset: kb = ((v >> 3) - band_position) & 31, offset[kb] for kb < 4
      const uint32_t kb = __vsub4((cc >> 3) & 0x1f1f1f1fu, one * cls) & 0x1f1f1f1fu;
      res = add_offsets(cc, lut8(o4, 0u, __vminu4(kb, 0x04040404u)));
    } else {
      const int cy = cy0 + r;
      if (foreign && (rim_lr || cy == 0 || cy == ch_ - 1)) {
        res = sao_edge_slow(f, s, sl, c, x0, y, nvalid, cc, nofilt);
        nofilt = 0;
      } else {
        uint32_t a, b, keep = keep_x;
        if (cls == 0) {
          a = __funnelshift_r(cl, cc, 24);   // sample x-1
          b = __funnelshift_r(cc, cr, 8);    // sample x+1
        } else {
          if (y == 0 || !has_dn) keep = 0xffffffffu;
          if (cls == 1) { a = pc; b = nc; }
          else if (cls == 2) { a = __funnelshift_r(pl, pc, 24); b = __funnelshift_r(nc, nr, 8); }
          else { a = __funnelshift_r(pc, pr, 8); b = __funnelshift_r(nl, nc, 24); }
        }
        // edgeIdx = sign(c-a) + sign(c-b)
        uint32_t idx = 0x02020202u;
        idx = __vadd4(idx, __vcmpgtu4(cc, a) & one);
        idx = __vsub4(idx, __vcmpltu4(cc, a) & one);
        idx = __vadd4(idx, __vcmpgtu4(cc, b) & one);
        idx = __vsub4(idx, __vcmpltu4(cc, b) & one);
        res = add_offsets(cc, lut8(tab_lo, tab_hi, idx));
        nofilt |= keep;
      }
    }
    res = (res & ~nofilt) | (cc & nofilt);
    uint8_t* o = out + (size_t)r * pitch;
    if (nvalid == 4) *reinterpret_cast<uint32_t*>(o) = res;
    else for (int k = 0; k < nvalid; k++) o[k] = (res >> (8 * k)) & 0xff;
    pl = cl; pc = cc; pr = cr;
    cl = nl; cc = nc; cr = nr;
  }
}

void launch_sao(cudaStream_t st, const FrameDev* d_frames, int n_frames, int max_w, int max_h) {
  dim3 g((((max_w + 3) >> 2) + 31) / 32, (max_h + 8 * SAO_ROWS - 1) / (8 * SAO_ROWS), 3 * n_frames);
  k_sao<<<g, dim3(32, 8), 0, st>>>(d_frames);
}
