// slice_syntax.h — CABAC + syntax decoding of slice segment data (H.265 7.3.8, 9.3) on the host,
// emitting per-picture WORK LISTS instead of reconstructing pixels.  The syntax walk is a template over
// the arithmetic-coding ENGINE: the decoder instantiates it with Cabac (cabac.h); the test-stream
// generator (tools/streamgen) instantiates it with an engine that DRAWS every bin at random, encodes it
// and returns it, so that generator and parser share one statement of the syntax while the unmodified
// reference decoder judges the result.
//
// Counterpart of the reference's slice.cc: read_sao :2700-2850, read_coding_quadtree :4590-4700,
// read_coding_unit :4245-4576, read_transform_tree :3810-3990, read_transform_unit :3510-3800,
// residual_coding :2905-3421, decode_substream :4720-4830, and of decode_quantization_parameters
// (transform.cc:31-200).  Where the reference calls into the pixel path, this file appends records:
//   S1 decode_intra_prediction (slice.cc:3453)  -> b265_tu.intra_mode
//   S2 scale_coefficients     (slice.cc:3471)   -> b265_tu {qp, flags, coeff_off, ncoeff} + b265_coeff[]
//   S4 read_pcm_samples       (slice.cc:4356)   -> b265_pcm + sample bytes
//   deblocking edge flags (deblock.cc:130-225), QP_Y / pred-mode / pcm / bypass maps -> unit_info, qp_y
//   sao_info (slice.h:265-273)                  -> b265_sao per CTB
// Inter prediction syntax (prediction_unit, merge/AMVP) is not implemented yet: P/B slices are
// rejected with DE265_ERROR_NOT_IMPLEMENTED_YET (502).
#pragma once
#include <string.h>
#include <algorithm>
#include "cabac.h"
#ifdef B265_PROF
#include <x86intrin.h>
#include <stdio.h>
extern unsigned long long g_prof[16];
struct ProfScope { int i; unsigned long long t; ProfScope(int k) : i(k), t(__rdtsc()) {} ~ProfScope() { g_prof[i] += __rdtsc() - t; } };
#define PROF(k) ProfScope prof_scope_##k(k)
#define PROF_T(v) unsigned long long v = __rdtsc()
#define PROF_ADD(k, a, b) g_prof[k] += (b) - (a)
#else
#define PROF(k)
#define PROF_T(v)
#define PROF_ADD(k, a, b)
#endif
#include "picture.h"

namespace b265h {

enum { E_OK = 0, E_CTB_OUTSIDE = 6, E_PREMATURE_END = 17, E_DECODING = 18, E_NOT_IMPLEMENTED = 502 };

namespace syn {


// ---- scan tables (6.5.3 - 6.5.5) -------------------------------------------------------------------
struct ScanTables {
  // [log2 size 0..3][scanIdx][i] -> x | y<<4 ; sizes 1,2,4,8 (sub-block grids and the 4x4 coefficient scan)
  uint8_t pos[4][3][64];
  // inverse for the 4x4 scan and the sub-block grids: [log2][scanIdx][y*n+x] -> i
  uint8_t inv[4][3][64];
  ScanTables() {
    for (int lg = 0; lg < 4; lg++) {
      const int n = 1 << lg;
      int i = 0, x = 0, y = 0;
      for (;;) {  // up-right diagonal
        while (y >= 0) {
          if (x < n && y < n) pos[lg][0][i++] = (uint8_t)(x | (y << 4));
          y--; x++;
        }
        y = x; x = 0;
        if (i >= n * n) break;
      }
      i = 0;
      for (y = 0; y < n; y++) for (x = 0; x < n; x++) pos[lg][1][i++] = (uint8_t)(x | (y << 4));  // horizontal
      i = 0;
      for (x = 0; x < n; x++) for (y = 0; y < n; y++) pos[lg][2][i++] = (uint8_t)(x | (y << 4));  // vertical
      for (int s = 0; s < 3; s++)
        for (i = 0; i < n * n; i++) inv[lg][s][(pos[lg][s][i] >> 4) * n + (pos[lg][s][i] & 15)] = (uint8_t)i;
    }
  }
};
const ScanTables g_scan;

// sig_coeff_flag context for 4x4 TBs (9.3.4.2.5, ctxIdxMap)
const uint8_t kCtxIdxMap4x4[16] = {0, 1, 4, 5, 2, 3, 4, 5, 6, 6, 8, 8, 7, 7, 8, 8};
// base sigCtx by (prevCsbf, yP*4+xP)
struct SigLut {
  uint8_t v[4][16];
  SigLut() {
    for (int yP = 0; yP < 4; yP++)
      for (int xP = 0; xP < 4; xP++) {
        const int i = yP * 4 + xP;
        v[0][i] = (xP + yP == 0) ? 2 : (xP + yP < 3) ? 1 : 0;
        v[1][i] = (yP == 0) ? 2 : (yP == 1) ? 1 : 0;
        v[2][i] = (xP == 0) ? 2 : (xP == 1) ? 1 : 0;
        v[3][i] = 2;
      }
  }
};
const SigLut g_sig;

// sig_coeff_flag context index (9.3.4.2.5) per SCAN POSITION of a 4x4 sub-block, including the context base:
// [chroma][log2 size - 2][scanIdx][prevCsbf][sub-block index > 0][k]
struct SigCtxTab {
  uint8_t t[2][4][3][4][2][16];
  SigCtxTab() {
    for (int ch = 0; ch < 2; ch++)
      for (int lg = 2; lg <= 5; lg++)
        for (int scan = 0; scan < 3; scan++)
          for (int prev = 0; prev < 4; prev++)
            for (int nz = 0; nz < 2; nz++)
              for (int k = 0; k < 16; k++) {
                const int xP = g_scan.pos[2][scan][k] & 15, yP = g_scan.pos[2][scan][k] >> 4;
                int sc;
                if (lg == 2) sc = kCtxIdxMap4x4[(yP << 2) + xP];
                else if (!nz && k == 0) sc = 0;
                else {
                  sc = g_sig.v[prev][(yP << 2) + xP];
                  if (!ch) {
                    if (nz) sc += 3;
                    sc += lg == 3 ? (scan == 0 ? 9 : 15) : 21;
                  } else {
                    sc += lg == 3 ? 9 : 12;
                  }
                }
                t[ch][lg - 2][scan][prev][nz][k] = (uint8_t)((ch ? CTX_SIG_COEFF + 27 : CTX_SIG_COEFF) + sc);
              }
  }
};
const SigCtxTab g_sigctx;

const uint8_t kQpC[58] = {0,  1,  2,  3,  4,  5,  6,  7,  8,  9,  10, 11, 12, 13, 14, 15, 16, 17, 18, 19,
                          20, 21, 22, 23, 24, 25, 26, 27, 28, 29, 29, 30, 31, 32, 33, 33, 34, 34, 35, 35,
                          36, 36, 37, 37, 38, 39, 40, 41, 42, 43, 44, 45, 46, 47, 48, 49, 50, 51};

inline int clip3(int lo, int hi, int v) { return v < lo ? lo : (v > hi ? hi : v); }

template <class Eng>
struct SliceDec {
  Picture& pic;
  const Sps& sps;
  const Pps& pps;
  const SliceHeader& sh;
  const SliceUnit& unit;
  Eng cabac;
  Ctx ctx[CTX_COUNT];
  Ctx wpp_ctx[CTX_COUNT];
  bool wpp_ctx_valid = false;

  int ctb_rs = 0, ctb_ts = 0;
  int slice_index = 0;
  // quantisation state (transform.cc:31-200)
  int is_cu_qp_delta_coded = 0, cu_qp_delta = 0;
  int is_cu_chroma_qp_offset_coded = 0, cu_qp_offset_cb = 0, cu_qp_offset_cr = 0;
  int cur_qg_x = -1, cur_qg_y = -1, last_qpy_prev_qg = 0, cur_qpy = 0;
  int qp_y_prime = 0, qp_cb_prime = 0, qp_cr_prime = 0;
  // current CU
  int cu_tq_bypass = 0;
  int cu_pred_mode = MODE_INTRA;
  int intra_chroma_mode = 0;   // derived chroma mode of the CU (4:2:0)
  int stat_rice[4] = {0, 0, 0, 0};

  SliceDec(Picture& p, const SliceUnit& u)
      : pic(p), sps(*u.pps->sps), pps(*u.pps), sh(u.hdr), unit(u) {}

  // ---- availability (6.4.1) at CTB granularity + picture bounds; (xN,yN) must precede the current
  // position in decoding order when it lies in the same slice and tile, which holds for the left and
  // upper neighbours this parser asks about.
  inline bool available(int xN, int yN) const {
    if (xN < 0 || yN < 0 || xN >= sps.width || yN >= sps.height) return false;
    const int ctbN = (yN >> sps.log2_ctb) * sps.ctbs_w + (xN >> sps.log2_ctb);
    if (ctbN == ctb_rs) return true;
    if (pic.ctb_slice_addr[ctbN] != sh.slice_addr_rs) return false;
    if (pps.tile_id_rs[ctbN] != pps.tile_id_rs[ctb_rs]) return false;
    return true;
  }
  inline int cb_idx(int x, int y) const { return (y >> sps.log2_min_cb) * pic.cb_w + (x >> sps.log2_min_cb); }

  int decode_substreams();
  int decode_ctu();
  void read_sao(int rx, int ry);
  int coding_quadtree(int x0, int y0, int log2_cb, int depth);
  int coding_unit(int x0, int y0, int log2_cb);
  int transform_tree(int x0, int y0, int xBase, int yBase, int xCU, int yCU, int log2_ts, int depth, int blk_idx,
                     int max_depth, int intra_split, int parent_cbf_cb, int parent_cbf_cr, int log2_cb);
  int transform_unit(int x0, int y0, int xBase, int yBase, int xCU, int yCU, int log2_ts, int depth, int blk_idx,
                     int cbf_luma, int cbf_cb, int cbf_cr, int log2_cb);
  int residual_coding(int x0, int y0, int log2_ts, int cidx, int intra_mode, b265_tu& tu);
  void emit_tb(int x, int y, int log2_ts, int cidx, int intra_mode, bool cbf, int scan_mode);
  void decode_qp(int xCU, int yCU, int log2_cb);
  int read_pcm(int x0, int y0, int log2_cb);
  void mark_tu_edges(int x0, int y0, int log2_ts, int xCU, int yCU, int log2_cb, bool cbf_luma);
};

// ---- SAO syntax (7.3.8.3) --------------------------------------------------------------------------
template <class Eng>
void SliceDec<Eng>::read_sao(int rx, int ry) {
  b265_sao& out = pic.sao[ctb_rs];
  memset(&out, 0, sizeof(out));
  if (!sh.sao_luma && !sh.sao_chroma) return;
  int merge_left = 0, merge_up = 0;
  if (rx > 0) {
    const bool in_slice = ctb_rs > sh.slice_addr_rs;
    const bool in_tile = pps.tile_id_rs[ctb_rs] == pps.tile_id_rs[ctb_rs - 1];
    if (in_slice && in_tile) merge_left = cabac.bin(ctx[CTX_SAO_MERGE]);
  }
  if (ry > 0 && !merge_left) {
    const bool in_slice = (ctb_rs - sps.ctbs_w) >= sh.slice_addr_rs;
    const bool in_tile = pps.tile_id_rs[ctb_rs] == pps.tile_id_rs[ctb_rs - sps.ctbs_w];
    if (in_slice && in_tile) merge_up = cabac.bin(ctx[CTX_SAO_MERGE]);
  }
  if (merge_left) { out = pic.sao[ctb_rs - 1]; return; }
  if (merge_up) { out = pic.sao[ctb_rs - sps.ctbs_w]; return; }
  const int n_comp = sps.chroma_array_type != 0 ? 3 : 1;
  for (int c = 0; c < n_comp; c++) {
    if ((c == 0 && !sh.sao_luma) || (c > 0 && !sh.sao_chroma)) continue;
    if (c == 0 || c == 1) {
      int t = 0;
      if (cabac.bin(ctx[CTX_SAO_TYPE])) t = cabac.bypass() ? 2 : 1;
      out.type[c] = (uint8_t)t;
    } else {
      out.type[2] = out.type[1];
    }
    if (out.type[c] == 0) continue;
    const int bd = c == 0 ? sps.bit_depth_luma : sps.bit_depth_chroma;
    const int cmax = (1 << (std::min(bd, 10) - 5)) - 1;
    const int scale = c == 0 ? pps.log2_sao_offset_scale_luma : pps.log2_sao_offset_scale_chroma;
    int absv[4];
    for (int i = 0; i < 4; i++) {
      int v = 0;
      while (v < cmax && cabac.bypass()) v++;
      absv[i] = v;
    }
    if (out.type[c] == 1) {
      for (int i = 0; i < 4; i++)
        if (absv[i] && cabac.bypass()) absv[i] = -absv[i];
      out.band_eo[c] = (uint8_t)cabac.bypass_bits(5);
    } else {
      if (c == 0 || c == 1) out.band_eo[c] = (uint8_t)cabac.bypass_bits(2);
      else out.band_eo[2] = out.band_eo[1];
      absv[2] = -absv[2];
      absv[3] = -absv[3];
    }
    for (int i = 0; i < 4; i++) out.offset[c][i] = (int8_t)(absv[i] * (1 << scale));
    pic.any_sao = true;
  }
}

// ---- coding quadtree (7.3.8.4) ---------------------------------------------------------------------
template <class Eng>
int SliceDec<Eng>::coding_quadtree(int x0, int y0, int log2_cb, int depth) {
  const int size = 1 << log2_cb;
  int split;
  if (x0 + size <= sps.width && y0 + size <= sps.height && log2_cb > sps.log2_min_cb) {
    int inc = 0;
    if (available(x0 - 1, y0) && pic.ct_depth[cb_idx(x0 - 1, y0)] > depth) inc++;
    if (available(x0, y0 - 1) && pic.ct_depth[cb_idx(x0, y0 - 1)] > depth) inc++;
    split = cabac.bin(ctx[CTX_SPLIT_CU + inc]);
  } else {
    split = log2_cb > sps.log2_min_cb;
  }
  if (pps.cu_qp_delta_enabled && log2_cb >= pps.log2_min_cu_qp_delta) {
    is_cu_qp_delta_coded = 0;
    cu_qp_delta = 0;
  }
  if (sh.cu_chroma_qp_offset_enabled && log2_cb >= pps.log2_min_cu_chroma_qp_offset) is_cu_chroma_qp_offset_coded = 0;
  if (split) {
    const int h = size >> 1, x1 = x0 + h, y1 = y0 + h;
    int r = coding_quadtree(x0, y0, log2_cb - 1, depth + 1);
    if (r) return r;
    if (x1 < sps.width) { r = coding_quadtree(x1, y0, log2_cb - 1, depth + 1); if (r) return r; }
    if (y1 < sps.height) { r = coding_quadtree(x0, y1, log2_cb - 1, depth + 1); if (r) return r; }
    if (x1 < sps.width && y1 < sps.height) { r = coding_quadtree(x1, y1, log2_cb - 1, depth + 1); if (r) return r; }
    return E_OK;
  }
  // leaf: record the coding-tree depth for later context selection
  const int n_cb = 1 << (log2_cb - sps.log2_min_cb);
  const int ci = cb_idx(x0, y0);
  for (int j = 0; j < n_cb; j++) memset(&pic.ct_depth[ci + j * pic.cb_w], depth, n_cb);
  return coding_unit(x0, y0, log2_cb);
}

// ---- QP derivation (8.6.1; transform.cc:31-200) ------------------------------------------------------
template <class Eng>
void SliceDec<Eng>::decode_qp(int xCU, int yCU, int log2_cb) {
  PROF(3);
  const int qg_mask = (1 << pps.log2_min_cu_qp_delta) - 1;
  const int xQG = xCU & ~qg_mask, yQG = yCU & ~qg_mask;
  if (xQG != cur_qg_x || yQG != cur_qg_y) {
    last_qpy_prev_qg = cur_qpy;
    cur_qg_x = xQG;
    cur_qg_y = yQG;
  }
  const int ctb_mask = (1 << sps.log2_ctb) - 1;
  const bool first_in_ctb_row = xQG == 0 && (yQG & ctb_mask) == 0;
  const int slice_x = (sh.slice_addr_rs % sps.ctbs_w) << sps.log2_ctb, slice_y = (sh.slice_addr_rs / sps.ctbs_w) << sps.log2_ctb;
  const bool first_qg_in_slice = slice_x == xQG && slice_y == yQG;
  bool first_qg_in_tile = false;
  if (pps.tiles_enabled && (xQG & ctb_mask) == 0 && (yQG & ctb_mask) == 0) {
    const int cx = xQG >> sps.log2_ctb, cy = yQG >> sps.log2_ctb;
    bool col = false, row = false;
    for (int v : pps.col_bd) col |= (v == cx);
    for (int v : pps.row_bd) row |= (v == cy);
    first_qg_in_tile = col && row;
  }
  int pred;
  if (first_qg_in_slice || first_qg_in_tile || (first_in_ctb_row && pps.entropy_coding_sync)) pred = sh.slice_qp;
  else pred = last_qpy_prev_qg;
  int qa = pred, qb = pred;
  if (available(xQG - 1, yQG)) {
    const int ctbA = (yQG >> sps.log2_ctb) * sps.ctbs_w + ((xQG - 1) >> sps.log2_ctb);
    if (ctbA == ctb_rs) qa = pic.qp_y[cb_idx(xQG - 1, yQG)];
  }
  if (available(xQG, yQG - 1)) {
    const int ctbB = ((yQG - 1) >> sps.log2_ctb) * sps.ctbs_w + (xQG >> sps.log2_ctb);
    if (ctbB == ctb_rs) qb = pic.qp_y[cb_idx(xQG, yQG - 1)];
  }
  pred = (qa + qb + 1) >> 1;
  const int qpy = ((pred + cu_qp_delta + 52 + 2 * sps.qp_bd_offset_y) % (52 + sps.qp_bd_offset_y)) - sps.qp_bd_offset_y;
  qp_y_prime = qpy + sps.qp_bd_offset_y;
  const int qi_cb = clip3(-sps.qp_bd_offset_c, 57, qpy + pps.cb_qp_offset + sh.cb_qp_offset + cu_qp_offset_cb);
  const int qi_cr = clip3(-sps.qp_bd_offset_c, 57, qpy + pps.cr_qp_offset + sh.cr_qp_offset + cu_qp_offset_cr);
  int qc_cb = qi_cb, qc_cr = qi_cr;
  if (sps.chroma_array_type == 1) {
    qc_cb = qi_cb < 0 ? qi_cb : kQpC[qi_cb];
    qc_cr = qi_cr < 0 ? qi_cr : kQpC[qi_cr];
  }
  qp_cb_prime = qc_cb + sps.qp_bd_offset_c;
  qp_cr_prime = qc_cr + sps.qp_bd_offset_c;
  const int n_cb = 1 << (std::max(log2_cb, 3) - sps.log2_min_cb);
  const int ci = cb_idx(xCU, yCU);
  for (int j = 0; j < n_cb && (yCU >> sps.log2_min_cb) + j < pic.cb_h; j++)
    for (int i = 0; i < n_cb && (xCU >> sps.log2_min_cb) + i < pic.cb_w; i++) pic.qp_y[ci + j * pic.cb_w + i] = (int8_t)qpy;
  cur_qpy = qpy;
}

// ---- deblocking edge flags for one transform block (deblock.cc:31-61 + :130-225) ----------------------
// Called for every leaf TU. Edges on the CU boundary use the CU's left/top filter decision; edges
// inside the CU are always filtered. Nothing is marked when the slice has deblocking disabled.
template <class Eng>
void SliceDec<Eng>::mark_tu_edges(int x0, int y0, int log2_ts, int xCU, int yCU, int log2_cb, bool cbf_luma) {
  PROF(4);
  const int n_u = 1 << (log2_ts - 2);
  const int ux = x0 >> 2, uy = y0 >> 2;
  uint8_t* ui = pic.unit_info.data();
  if (cbf_luma)
    for (int j = 0; j < n_u; j++)
      for (int i = 0; i < n_u; i++) ui[(uy + j) * pic.u_w + ux + i] |= B265_UNIT_CBF;
  if (sh.deblocking_disabled) return;
  pic.any_deblock = true;
  bool filter_left = true, filter_top = true;
  const int ctb_mask = (1 << sps.log2_ctb) - 1;
  if (x0 == xCU) {
    if (x0 == 0) filter_left = false;
    else if ((x0 & ctb_mask) == 0) {
      const int ctbN = (y0 >> sps.log2_ctb) * sps.ctbs_w + ((x0 - 1) >> sps.log2_ctb);
      if (!sh.lf_across_slices && pic.ctb_slice_addr[ctbN] >= 0 && pic.ctb_slice_addr[ctbN] != sh.slice_addr_rs) filter_left = false;
      else if (!pps.lf_across_tiles && pps.tile_id_rs[ctbN] != pps.tile_id_rs[ctb_rs]) filter_left = false;
    }
  }
  if (y0 == yCU) {
    if (y0 == 0) filter_top = false;
    else if ((y0 & ctb_mask) == 0) {
      const int ctbN = ((y0 - 1) >> sps.log2_ctb) * sps.ctbs_w + (x0 >> sps.log2_ctb);
      if (!sh.lf_across_slices && pic.ctb_slice_addr[ctbN] >= 0 && pic.ctb_slice_addr[ctbN] != sh.slice_addr_rs) filter_top = false;
      else if (!pps.lf_across_tiles && pps.tile_id_rs[ctbN] != pps.tile_id_rs[ctb_rs]) filter_top = false;
    }
  }
  (void)log2_cb;
  if (filter_left)
    for (int j = 0; j < n_u; j++) ui[(uy + j) * pic.u_w + ux] |= B265_UNIT_EDGE_V | B265_UNIT_TU_EDGE_V;
  if (filter_top)
    for (int i = 0; i < n_u; i++) ui[uy * pic.u_w + ux + i] |= B265_UNIT_EDGE_H | B265_UNIT_TU_EDGE_H;
}

// ---- PCM samples (7.3.8.7; slice.cc:4143-4213) ---------------------------------------------------------
template <class Eng>
int SliceDec<Eng>::read_pcm(int x0, int y0, int log2_cb) {
  const int n = 1 << log2_cb;
  b265_pcm rec = {};
  rec.x = (uint16_t)x0; rec.y = (uint16_t)y0; rec.log2_size = (uint8_t)log2_cb;
  rec.sample_off = (uint32_t)pic.pcm_samples.size();
  const int n_chroma = sps.chroma_array_type != 0 ? (n / 2) * (n / 2) * 2 : 0;
  // raw samples at the byte position the arithmetic decoder stopped at, then the engine restarts (9.3.2.5)
  if (!cabac.pcm_samples(n * n, sps.pcm_bits_luma, sps.bit_depth_luma - sps.pcm_bits_luma, n_chroma, sps.pcm_bits_chroma,
                         sps.bit_depth_chroma - sps.pcm_bits_chroma, pic.pcm_samples))
    return E_PREMATURE_END;
  pic.pcms.push_back(rec);
  return E_OK;
}

// ---- coding unit (7.3.8.5) — intra only ---------------------------------------------------------------
template <class Eng>
int SliceDec<Eng>::coding_unit(int x0, int y0, int log2_cb) {
  PROF(2);
  const int n_cbs = 1 << log2_cb;
  decode_qp(x0, y0, log2_cb);
  cu_tq_bypass = 0;
  if (pps.transquant_bypass_enabled) cu_tq_bypass = cabac.bin(ctx[CTX_CU_TQ_BYPASS]);
  cu_pred_mode = MODE_INTRA;
  int nxn = 0;
  if (log2_cb == sps.log2_min_cb) nxn = !cabac.bin(ctx[CTX_PART_MODE]);
  if (nxn && log2_cb == 3 && sps.log2_min_tb > 2) return E_DECODING;
  // metadata of the CB
  {
    const int n_cb = 1 << (log2_cb - sps.log2_min_cb), ci = cb_idx(x0, y0);
    for (int j = 0; j < n_cb; j++) {
      memset(&pic.pred_mode[ci + j * pic.cb_w], MODE_INTRA, n_cb);
      memset(&pic.cu_flags[ci + j * pic.cb_w], cu_tq_bypass ? 2 : 0, n_cb);
    }
  }
  int pcm = 0;
  if (!nxn && sps.pcm_enabled && log2_cb >= sps.log2_min_pcm && log2_cb <= sps.log2_max_pcm) {
    cabac.hint(HINT_PCM_FLAG, 0);
    pcm = cabac.terminate();
  }
  const int n_u = n_cbs >> 2, ux = x0 >> 2, uy = y0 >> 2;
  uint8_t unit_bits = B265_UNIT_INTRA;
  if (cu_tq_bypass || (pcm && sps.pcm_lf_disabled)) unit_bits |= B265_UNIT_NOFILT;
  if (pcm) unit_bits |= B265_UNIT_DONE;
  for (int j = 0; j < n_u; j++) memset(&pic.unit_info[(uy + j) * pic.u_w + ux], unit_bits, n_u);
  if (pcm) {
    const int n_cb = 1 << (log2_cb - sps.log2_min_cb), ci = cb_idx(x0, y0);
    for (int j = 0; j < n_cb; j++)
      for (int i = 0; i < n_cb; i++) pic.cu_flags[ci + j * pic.cb_w + i] |= 1;
    for (int j = 0; j < n_u; j++) memset(&pic.ipm[(uy + j) * pic.u_w + ux], 1, n_u);
    int r = read_pcm(x0, y0, log2_cb);
    if (r) return r;
    mark_tu_edges(x0, y0, log2_cb, x0, y0, log2_cb, false);
    return E_OK;
  }
  // intra prediction modes (7.3.8.5, 8.4.2)
  const int n_pu = nxn ? 4 : 1, pb = nxn ? n_cbs >> 1 : n_cbs;
  int prev_flag[4], mpm_idx[4] = {0, 0, 0, 0}, rem[4] = {0, 0, 0, 0};
  for (int i = 0; i < n_pu; i++) prev_flag[i] = cabac.bin(ctx[CTX_PREV_INTRA_LUMA]);
  for (int i = 0; i < n_pu; i++) {
    if (prev_flag[i]) {
      mpm_idx[i] = cabac.bypass() ? (cabac.bypass() ? 2 : 1) : 0;
    } else {
      rem[i] = (int)cabac.bypass_bits(5);
    }
  }
  int luma_mode0 = 1;
  for (int i = 0; i < n_pu; i++) {
    const int xP = x0 + (i & 1) * pb, yP = y0 + (i >> 1) * pb;
    int candA = 1, candB = 1;
    if (available(xP - 1, yP)) {
      const int ci = cb_idx(xP - 1, yP);
      if (pic.pred_mode[ci] == MODE_INTRA && !(pic.cu_flags[ci] & 1)) candA = pic.ipm[(yP >> 2) * pic.u_w + ((xP - 1) >> 2)];
    }
    if (available(xP, yP - 1)) {
      const int ci = cb_idx(xP, yP - 1);
      if (pic.pred_mode[ci] == MODE_INTRA && !(pic.cu_flags[ci] & 1) && (yP - 1) >= ((yP >> sps.log2_ctb) << sps.log2_ctb))
        candB = pic.ipm[((yP - 1) >> 2) * pic.u_w + (xP >> 2)];
    }
    int cand[3];
    if (candA == candB) {
      if (candA < 2) { cand[0] = 0; cand[1] = 1; cand[2] = 26; }
      else { cand[0] = candA; cand[1] = 2 + ((candA - 2 - 1 + 32) % 32); cand[2] = 2 + ((candA - 2 + 1) % 32); }
    } else {
      cand[0] = candA; cand[1] = candB;
      if (candA != 0 && candB != 0) cand[2] = 0;
      else if (candA != 1 && candB != 1) cand[2] = 1;
      else cand[2] = 26;
    }
    int mode;
    if (prev_flag[i]) mode = cand[mpm_idx[i]];
    else {
      if (cand[0] > cand[1]) std::swap(cand[0], cand[1]);
      if (cand[0] > cand[2]) std::swap(cand[0], cand[2]);
      if (cand[1] > cand[2]) std::swap(cand[1], cand[2]);
      mode = rem[i];
      for (int k = 0; k < 3; k++) if (mode >= cand[k]) mode++;
    }
    if (i == 0) luma_mode0 = mode;
    const int pu = pb >> 2;
    for (int j = 0; j < pu; j++) memset(&pic.ipm[((yP >> 2) + j) * pic.u_w + (xP >> 2)], mode, pu);
  }
  intra_chroma_mode = luma_mode0;
  if (sps.chroma_array_type != 0) {
    int v = 4;
    if (cabac.bin(ctx[CTX_INTRA_CHROMA])) v = (int)cabac.bypass_bits(2);
    if (v < 4) {
      static const int tab[4] = {0, 26, 10, 1};
      intra_chroma_mode = tab[v] == luma_mode0 ? 34 : tab[v];
    }
  }
  const int max_depth = sps.max_th_depth_intra + nxn;
  return transform_tree(x0, y0, x0, y0, x0, y0, log2_cb, 0, 0, max_depth, nxn, 1, 1, log2_cb);
}

// ---- transform tree (7.3.8.8) --------------------------------------------------------------------------
template <class Eng>
int SliceDec<Eng>::transform_tree(int x0, int y0, int xBase, int yBase, int xCU, int yCU, int log2_ts, int depth, int blk_idx,
                             int max_depth, int intra_split, int parent_cbf_cb, int parent_cbf_cr, int log2_cb) {
  int split;
  if (log2_ts <= sps.log2_max_tb && log2_ts > sps.log2_min_tb && depth < max_depth && !(intra_split && depth == 0))
    split = cabac.bin(ctx[CTX_SPLIT_TRANSFORM + 5 - log2_ts]);
  else
    split = (log2_ts > sps.log2_max_tb || (intra_split && depth == 0)) ? 1 : 0;
  int cbf_cb = 0, cbf_cr = 0;
  if (log2_ts > 2 && sps.chroma_array_type != 0) {
    if (depth == 0 || parent_cbf_cb) cbf_cb = cabac.bin(ctx[CTX_CBF_CHROMA + depth]);
    if (depth == 0 || parent_cbf_cr) cbf_cr = cabac.bin(ctx[CTX_CBF_CHROMA + depth]);
  } else if (log2_ts == 2 && depth > 0 && sps.chroma_array_type != 0) {
    cbf_cb = parent_cbf_cb;
    cbf_cr = parent_cbf_cr;
  }
  if (split) {
    const int h = 1 << (log2_ts - 1);
    for (int i = 0; i < 4; i++) {
      int r = transform_tree(x0 + (i & 1) * h, y0 + (i >> 1) * h, x0, y0, xCU, yCU, log2_ts - 1, depth + 1, i, max_depth,
                             intra_split, cbf_cb, cbf_cr, log2_cb);
      if (r) return r;
    }
    return E_OK;
  }
  int cbf_luma = 1;
  if (cu_pred_mode == MODE_INTRA || depth != 0 || cbf_cb || cbf_cr) cbf_luma = cabac.bin(ctx[CTX_CBF_LUMA + (depth == 0 ? 1 : 0)]);
  return transform_unit(x0, y0, xBase, yBase, xCU, yCU, log2_ts, depth, blk_idx, cbf_luma, cbf_cb, cbf_cr, log2_cb);
}

// ---- transform unit (7.3.8.10) -------------------------------------------------------------------------
template <class Eng>
int SliceDec<Eng>::transform_unit(int x0, int y0, int xBase, int yBase, int xCU, int yCU, int log2_ts, int depth, int blk_idx,
                             int cbf_luma, int cbf_cb, int cbf_cr, int log2_cb) {
  (void)depth;
  const bool chroma420 = sps.chroma_array_type == 1;
  if (cbf_luma || cbf_cb || cbf_cr) {
    bool redo = false;
    if (pps.cu_qp_delta_enabled && !is_cu_qp_delta_coded) {
      int v = 0;
      if (cabac.bin(ctx[CTX_CU_QP_DELTA])) {
        v = 1;
        while (v < 5 && cabac.bin(ctx[CTX_CU_QP_DELTA + 1])) v++;
        if (v == 5) {  // EG0 suffix
          cabac.hint(HINT_QP_DELTA_SUFFIX, 0);
          int k = 0;
          while (k < 16 && cabac.bypass()) k++;
          v += ((1 << k) - 1) + (int)cabac.bypass_bits(k);
        }
      }
      int sign = 0;
      if (v) sign = cabac.bypass();
      is_cu_qp_delta_coded = 1;
      cu_qp_delta = sign ? -v : v;
      redo = true;
    }
    if (sh.cu_chroma_qp_offset_enabled && (cbf_cb || cbf_cr) && !cu_tq_bypass && !is_cu_chroma_qp_offset_coded) {
      const int flag = cabac.bin(ctx[CTX_CHROMA_QP_OFFSET_FLAG]);
      int idx = 0;
      if (flag && pps.chroma_qp_offset_list_len > 1) idx = cabac.bin(ctx[CTX_CHROMA_QP_OFFSET_IDX]);
      is_cu_chroma_qp_offset_coded = 1;
      cu_qp_offset_cb = flag ? pps.cb_qp_offset_list[idx] : 0;
      cu_qp_offset_cr = flag ? pps.cr_qp_offset_list[idx] : 0;
      redo = true;
    }
    if (redo) decode_qp(xCU, yCU, log2_cb);
  }
  mark_tu_edges(x0, y0, log2_ts, xCU, yCU, log2_cb, cbf_luma != 0);
  // luma
  {
    const int mode = pic.ipm[(y0 >> 2) * pic.u_w + (x0 >> 2)];
    int r = 0;
    b265_tu tu = {};
    tu.x = (uint16_t)x0; tu.y = (uint16_t)y0;
    tu.cidx_log2 = (uint8_t)(0 | (log2_ts << 2));
    tu.intra_mode = (uint8_t)mode;
    tu.qp = (uint8_t)qp_y_prime;
    if (sps.implicit_rdpcm && cu_tq_bypass) tu.flags |= B265_TU_NO_EDGE_FILT;   // intrapred.cc:1102-1104
    tu.coeff_off = (uint32_t)pic.coeffs.size();
    if (cbf_luma) { r = residual_coding(x0, y0, log2_ts, 0, mode, tu); if (r) return r; }
    pic.tus.push_back(tu);
  }
  if (!chroma420) return E_OK;
  int xc, yc, lc;
  if (log2_ts > 2) { xc = x0 >> 1; yc = y0 >> 1; lc = log2_ts - 1; }
  else if (blk_idx == 3) { xc = xBase >> 1; yc = yBase >> 1; lc = 2; }
  else return E_OK;
  for (int c = 1; c <= 2; c++) {
    const int cbf = c == 1 ? cbf_cb : cbf_cr;
    b265_tu tu = {};
    tu.x = (uint16_t)xc; tu.y = (uint16_t)yc;
    tu.cidx_log2 = (uint8_t)(c | (lc << 2));
    tu.intra_mode = (uint8_t)intra_chroma_mode;
    tu.qp = (uint8_t)(c == 1 ? qp_cb_prime : qp_cr_prime);
    tu.coeff_off = (uint32_t)pic.coeffs.size();
    if (cbf) { int r = residual_coding(xc, yc, lc, c, intra_chroma_mode, tu); if (r) return r; }
    pic.tus.push_back(tu);
  }
  return E_OK;
}

// ---- residual coding (7.3.8.11, 9.3.4.2.4-7) -----------------------------------------------------------
// The arithmetic decoder state is copied into a local for the duration of the block: the context
// variables are bytes, so every context update would otherwise force the engine's registers back to
// memory (char stores alias everything).
template <class Eng>
int SliceDec<Eng>::residual_coding(int x0, int y0, int log2_ts, int cidx, int intra_mode, b265_tu& tu) {
  PROF(1);
  (void)x0; (void)y0;
  typename Eng::Local cb(cabac);
  Ctx* const cx = ctx;
  const int n = 1 << log2_ts;
  int tskip = 0;
  if (pps.transform_skip_enabled && !cu_tq_bypass && log2_ts <= pps.log2_max_transform_skip)
    tskip = cb.bin(cx[CTX_TRANSFORM_SKIP + (cidx ? 1 : 0)]);

  // last significant coefficient position
  const int cmax = (log2_ts << 1) - 1;
  int off, shift;
  if (cidx == 0) { off = 3 * (log2_ts - 2) + ((log2_ts - 1) >> 2); shift = (log2_ts + 1) >> 2; }
  else { off = 15; shift = log2_ts - 2; }
  int px = 0, py = 0;
  while (px < cmax && cb.bin(cx[CTX_LAST_X + off + (px >> shift)])) px++;
  while (py < cmax && cb.bin(cx[CTX_LAST_Y + off + (py >> shift)])) py++;
  int last_x = px, last_y = py;
  if (px > 3) { const int nb = (px >> 1) - 1; last_x = (1 << nb) * (2 + (px & 1)) + (int)cb.bypass_bits(nb); }
  if (py > 3) { const int nb = (py >> 1) - 1; last_y = (1 << nb) * (2 + (py & 1)) + (int)cb.bypass_bits(nb); }
  int scan = 0;
  if (cu_pred_mode == MODE_INTRA && (log2_ts == 2 || (log2_ts == 3 && cidx == 0))) {
    if (intra_mode >= 6 && intra_mode <= 14) scan = 2;
    else if (intra_mode >= 22 && intra_mode <= 30) scan = 1;
  }
  if (scan == 2) std::swap(last_x, last_y);
  if (last_x >= n || last_y >= n) { cb.commit(cabac); return E_DECODING; }

  const int lg_sb = log2_ts - 2, sbw = 1 << lg_sb;
  const uint8_t* sb_scan = g_scan.pos[lg_sb][scan];
  const uint8_t* c_scan = g_scan.pos[2][scan];
  const int last_sb = g_scan.inv[lg_sb][scan][(last_y >> 2) * sbw + (last_x >> 2)];
  const int last_pos = g_scan.inv[2][scan][(last_y & 3) * 4 + (last_x & 3)];

  uint8_t csbf[9][9];
  memset(csbf, 0, sizeof(csbf));
  const bool ts_ctx = sps.transform_skip_context && (tskip || cu_tq_bypass);
  const bool implicit_rdpcm_hv = cu_pred_mode == MODE_INTRA && sps.implicit_rdpcm && (tskip || cu_tq_bypass) &&
                                 (intra_mode == 10 || intra_mode == 26);
  const bool sign_hiding_ok = pps.sign_data_hiding && !cu_tq_bypass && !implicit_rdpcm_hv;
  const int ch = cidx ? 1 : 0;
  uint8_t ts_tab[16];
  if (ts_ctx) memset(ts_tab, (cidx == 0 ? CTX_SIG_COEFF + 42 : CTX_SIG_COEFF + 27 + 16), 16);
  const int csb_base = CTX_CODED_SUB_BLOCK + (cidx ? 2 : 0);
  const int gt1_base = CTX_GT1 + (cidx ? 16 : 0), gt2_base = CTX_GT2 + (cidx ? 4 : 0);
  int c1 = 1;
  bool first_sb = true;
  // room for every coefficient of the block; the list grows by what is actually coded
  b265_coeff* const out0 = pic.coeffs.room((size_t)n * n);
  b265_coeff* out = out0;

  for (int i = last_sb; i >= 0; i--) {
    PROF_T(p0);
    const int xS = sb_scan[i] & 15, yS = sb_scan[i] >> 4;
    const int prev_csbf = csbf[yS][xS + 1] | (csbf[yS + 1][xS] << 1);
    int infer_dc = 0, coded = 1;
    if (i < last_sb && i > 0) {
      coded = cb.bin(cx[csb_base + (prev_csbf ? 1 : 0)]);
      infer_dc = 1;
    }
    csbf[yS][xS] = (uint8_t)coded;
    if (!coded) continue;

    uint8_t sig[17];
    int nsig = 0, nstart = 15;
    if (i == last_sb) { sig[nsig++] = (uint8_t)last_pos; nstart = last_pos - 1; }
    const uint8_t* tab = ts_ctx ? ts_tab : g_sigctx.t[ch][lg_sb][scan][prev_csbf][i > 0];
    for (int k = nstart; k >= 1; k--) {   // branch-free append: the flag values are unpredictable
      sig[nsig] = (uint8_t)k;
      nsig += cb.bin(cx[tab[k]]);
    }
    if (nstart >= 0) {
      if (infer_dc && nsig == 0) sig[nsig++] = 0;           // the only coefficient of a coded sub-block
      else if (cb.bin(cx[tab[0]])) sig[nsig++] = 0;
    }
    if (!nsig) continue;
    PROF_T(p1);
    PROF_ADD(6, p0, p1);
    // greater1 / greater2 flags
    int ctx_set = (i == 0 || cidx > 0) ? 0 : 2;
    if (!first_sb && c1 == 0) ctx_set++;
    first_sb = false;
    c1 = 1;
    int absv[16];
    int first_g1 = -1;
    const int ng1 = nsig < 8 ? nsig : 8;
    Ctx* g1 = cx + gt1_base + ctx_set * 4;
    static const uint8_t c1_next[2][4] = {{0, 2, 3, 3}, {0, 0, 0, 0}};   // 9.3.4.2.6
    for (int k = 0; k < ng1; k++) {
      const int g = cb.bin(g1[c1]);
      absv[k] = 1 + g;
      first_g1 = (g & (first_g1 < 0)) ? k : first_g1;
      c1 = c1_next[g][c1];
    }
    for (int k = ng1; k < nsig; k++) absv[k] = 1;
    if (first_g1 >= 0) absv[first_g1] += cb.bin(cx[gt2_base + ctx_set]);
    PROF_T(p2);
    PROF_ADD(7, p1, p2);
    const bool sign_hidden = sign_hiding_ok && (sig[0] - sig[nsig - 1] > 3);
    const int nsign = nsig - (sign_hidden ? 1 : 0);
    uint32_t signs = cb.bypass_bits(nsign) << (16 - nsign);  // bit 15 = first coefficient

    int rice = 0, sum_abs = 0;
    const int sb_pos = (xS << 2) + ((yS << 2) << log2_ts);
    for (int k = 0; k < nsig; k++) {
      const int base = absv[k];
      const int thr = k < 8 ? (k == first_g1 ? 3 : 2) : 1;
      int a = base;
      if (base == thr) {
        cb.hint(HINT_COEFF_REMAINING, rice);
        int prefix = 0;
        while (prefix < 32 && cb.bypass()) prefix++;
        if (prefix >= 32) { cb.commit(cabac); return E_DECODING; }
        int rem;
        if (prefix <= 3) rem = (prefix << rice) + (int)cb.bypass_bits(rice);
        else {
          const int nb = prefix - 3 + rice;
          if (nb > 24) { cb.commit(cabac); return E_DECODING; }
          rem = (((1 << (prefix - 3)) + 3 - 1) << rice) + (int)cb.bypass_bits(nb);
        }
        a = base + rem;
        if (a > 3 * (1 << rice)) rice = rice < 4 ? rice + 1 : 4;
      }
      sum_abs += a;
      int neg;
      if (sign_hidden && k == nsig - 1) neg = sum_abs & 1;
      else { neg = (signs >> 15) & 1; signs <<= 1; }
      int v = neg ? -a : a;
      if (v > 32767) v = 32767;
      if (v < -32768) v = -32768;
      const int p = c_scan[sig[k]];
      out->level = (int16_t)v;
      out->pos = (uint16_t)(sb_pos + (p & 15) + ((p >> 4) << log2_ts));
      out++;
    }
    PROF_T(p3);
    PROF_ADD(8, p2, p3);
  }
  cb.commit(cabac);
  const size_t ncoded = (size_t)(out - out0);
  pic.coeffs.n += ncoded;
  tu.ncoeff = (uint16_t)ncoded;
  uint8_t fl = B265_TU_INTRA_CU;
  if (cidx == 0 && log2_ts == 2) fl |= B265_TU_DST;
  if (tskip) fl |= B265_TU_SKIP;
  if (cu_tq_bypass) fl |= B265_TU_BYPASS;
  if (implicit_rdpcm_hv) fl |= intra_mode == 26 ? B265_TU_RDPCM_V : B265_TU_RDPCM_H;
  if (sps.transform_skip_rotation && log2_ts == 2) fl |= B265_TU_ROTATE;
  tu.flags |= fl;
  return E_OK;
}

// ---- CTU and slice segment data (7.3.8.1, 7.3.8.2; slice.cc:4640-4830) -------------------------------
template <class Eng>
int SliceDec<Eng>::decode_ctu() {
  PROF(0);
  const int rx = ctb_rs % sps.ctbs_w, ry = ctb_rs / sps.ctbs_w;
  pic.ctb_slice_addr[ctb_rs] = sh.slice_addr_rs;
  pic.ctbs[ctb_rs].addr_ts = ctb_ts;
  pic.ctbs[ctb_rs].slice = (uint16_t)slice_index;
  pic.ctbs[ctb_rs].tile = (uint16_t)pps.tile_id_rs[ctb_rs];
  read_sao(rx, ry);
  return coding_quadtree(rx << sps.log2_ctb, ry << sps.log2_ctb, sps.log2_ctb, 0);
}

}  // namespace syn


struct PicDecodeState {
  std::vector<Ctx> wpp;       // per CTB row: contexts after the 2nd CTB of the row
  std::vector<uint8_t> wpp_valid;
  Ctx dep[CTX_COUNT];         // contexts at the end of the previous slice segment
  bool dep_valid = false;
};

template <class Eng>
int decode_slice_unit(Picture& pic, const SliceUnit& unit, PicDecodeState& st, const typename Eng::Setup& setup) {
  syn::SliceDec<Eng> d(pic, unit);
  d.cabac.setup(setup);
  const Sps& sps = d.sps;
  const Pps& pps = d.pps;
  const SliceHeader& sh = d.sh;
  if (sh.slice_type != SLICE_I) return E_NOT_IMPLEMENTED;
  const int n_ctbs = sps.ctbs_w * sps.ctbs_h;
  d.slice_index = sh.slice_index;
  d.ctb_ts = pps.rs_to_ts[sh.segment_address];
  d.ctb_rs = sh.segment_address;
  if (sh.data_offset > unit.rbsp.size()) return E_PREMATURE_END;
  d.cabac.set_data(unit.rbsp.data() + sh.data_offset, unit.rbsp.data() + unit.rbsp.size());
  d.cabac.set_ctx_base(d.ctx);
  d.cabac.start();
  // quantiser state at the start of the segment (decctx.cc:452-490)
  d.cur_qpy = 0;
  if (sh.segment_address > 0) {
    const int prev = pps.ts_to_rs[d.ctb_ts - 1];
    const int x = std::min((((prev % sps.ctbs_w) + 1) << sps.log2_ctb) - 1, sps.width - 1);
    const int y = std::min((((prev / sps.ctbs_w) + 1) << sps.log2_ctb) - 1, sps.height - 1);
    d.cur_qpy = pic.qp_y[d.cb_idx(x, y)];
  }
  const int init_type = sh.slice_type == SLICE_I ? 0 : (sh.slice_type == SLICE_P ? (sh.cabac_init_flag ? 2 : 1) : (sh.cabac_init_flag ? 1 : 2));
  bool tile_start = false;
  {
    const int cx = d.ctb_rs % sps.ctbs_w, cy = d.ctb_rs / sps.ctbs_w;
    bool col = false, row = false;
    for (int v : pps.col_bd) col |= v == cx;
    for (int v : pps.row_bd) row |= v == cy;
    tile_start = col && row;
  }
  if (sh.dependent && !tile_start) {
    if (!st.dep_valid) return E_DECODING;
    memcpy(d.ctx, st.dep, sizeof(d.ctx));
  } else {
    cabac_init_contexts(d.ctx, init_type, sh.slice_qp);
  }
  bool first_ctb = true;
  for (;;) {
    const int cx = d.ctb_rs % sps.ctbs_w, cy = d.ctb_rs / sps.ctbs_w;
    if (pps.entropy_coding_sync && cx == 0 && cy >= 1 && (sh.dependent || !first_ctb)) {
      if (sps.ctbs_w > 1) {
        if (!st.wpp_valid[cy - 1]) return E_DECODING;
        memcpy(d.ctx, &st.wpp[(size_t)(cy - 1) * CTX_COUNT], sizeof(d.ctx));
      } else {
        cabac_init_contexts(d.ctx, init_type, sh.slice_qp);
      }
    }
    first_ctb = false;
    int r = d.decode_ctu();
    if (r) return r;
    if (pps.entropy_coding_sync && cx == 1 && cy < sps.ctbs_h - 1) {
      memcpy(&st.wpp[(size_t)cy * CTX_COUNT], d.ctx, sizeof(d.ctx));
      st.wpp_valid[cy] = 1;
    }
    d.cabac.hint(HINT_END_OF_SEGMENT, d.ctb_ts);
    const int end_of_segment = d.cabac.terminate();
    if (end_of_segment) {
      memcpy(st.dep, d.ctx, sizeof(d.ctx));
      st.dep_valid = true;
      return E_OK;
    }
    d.ctb_ts++;
    if (d.ctb_ts >= n_ctbs) return E_CTB_OUTSIDE;
    d.ctb_rs = pps.ts_to_rs[d.ctb_ts];
    const bool new_tile = pps.tiles_enabled && pps.tile_id_rs[d.ctb_rs] != pps.tile_id_rs[pps.ts_to_rs[d.ctb_ts - 1]];
    const bool new_row = pps.entropy_coding_sync && (d.ctb_rs % sps.ctbs_w) == 0;
    if (new_tile || new_row) {
      d.cabac.hint(HINT_END_OF_SUBSET, 0);
      if (!d.cabac.terminate()) return E_DECODING;  // end_of_subset_one_bit
      d.cabac.start();                              // byte_alignment + re-initialisation
      if (new_tile) cabac_init_contexts(d.ctx, init_type, sh.slice_qp);
    }
  }
}

}  // namespace b265h
