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
//   S3 generate_inter_prediction_samples (slice.cc:4110-4140 -> motion.cc:279) -> b265_pu (motion vectors derived here:
//      merge candidates, AMVP, temporal MVP = the reference's motion.cc:1020-2160)
#pragma once
#include <string.h>
#include <algorithm>
#include <chrono>
#include <thread>
#include "cabac.h"
#ifdef B265_PROF
#include <x86intrin.h>
#include <stdio.h>
extern unsigned long long g_prof[16];
struct ProfScope { int i; unsigned long long t; ProfScope(int k) : i(k), t(__rdtsc()) {} ~ProfScope() { g_prof[i] += __rdtsc() - t; } };
#define PROF(k) ProfScope prof_scope_##k(k)
#define PROF_T(v) unsigned long long v = __rdtsc()
#define PROF_ADD(k, a, b) g_prof[k] += (b) - (a)
#define PROF_CNT(k, v) g_prof[k] += (v)
#else
#define PROF_CNT(k, v)
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
  // ---- inter prediction (7.3.8.6, 7.3.8.9, 8.5.3.2) ----
  int cu_part_mode = 0;       // PartMode of the current CU (0 2Nx2N 1 2NxN 2 Nx2N 3 NxN 4 2NxnU 5 2NxnD 6 nLx2N 7 nRx2N)
  int cu_ct_depth = 0;
  const SliceRefs* refs = nullptr;
  inline MotionInfo& motion_at(int x, int y) { return pic.self->motion[(size_t)(y >> 2) * pic.u_w + (x >> 2)]; }
  bool pb_neighbour(int xCb, int yCb, int nCbS, int xPb, int yPb, int nPbW, int nPbH, int partIdx, int xN, int yN, MotionInfo& out);
  int prediction_unit(int xCb, int yCb, int nCbS, int xPb, int yPb, int nPbW, int nPbH, int partIdx, bool skip, int& merge_out);
  void derive_merge(int xCb, int yCb, int nCbS, int xPb, int yPb, int nPbW, int nPbH, int partIdx, int merge_idx, MotionInfo& out);
  void derive_amvp(int xCb, int yCb, int nCbS, int xPb, int yPb, int nPbW, int nPbH, int partIdx, int X, int ref_idx, int16_t cand[2][2]);
  bool temporal_mv(int xPb, int yPb, int nPbW, int nPbH, int ref_idx, int X, int16_t mv[2]);
  bool col_mv(const RefPic& col, int xCol, int yCol, int ref_idx, int X, int16_t mv[2]);
  void mark_pu_edges(int x0, int y0, int nCbS, int part_mode);
  int read_mvd(int16_t mvd[2]);
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
  {
    const int nx = std::min(n_cb, pic.cb_w - (xCU >> sps.log2_min_cb)), ny = std::min(n_cb, pic.cb_h - (yCU >> sps.log2_min_cb));
    for (int j = 0; j < ny; j++) memset(&pic.qp_y[ci + j * pic.cb_w], qpy, (size_t)std::max(nx, 0));
  }
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
  if (cbf_luma) {   // whole words where the TB is wide enough (x86: unaligned word access is fine; memcpy keeps it legal C++)
    for (int j = 0; j < n_u; j++) {
      uint8_t* row = ui + (uy + j) * pic.u_w + ux;
      if (n_u == 8) {
        uint64_t v; memcpy(&v, row, 8); v |= 0x0101010101010101ull * B265_UNIT_CBF; memcpy(row, &v, 8);
      } else if (n_u == 4) {
        uint32_t v; memcpy(&v, row, 4); v |= 0x01010101u * B265_UNIT_CBF; memcpy(row, &v, 4);
      } else {
        for (int i = 0; i < n_u; i++) row[i] |= B265_UNIT_CBF;
      }
    }
  }
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

// ---- inter prediction: neighbours, merge mode, AMVP, temporal candidates (8.5.3.2.1 - 8.5.3.2.9; the
// reference's motion.cc:1020-2160) ------------------------------------------------------------------------------
inline int mv_scale(int mv, int tb, int td) {   // (8-179)..(8-183)
  td = clip3(-128, 127, td);
  tb = clip3(-128, 127, tb);
  const int tx = (16384 + (abs(td) >> 1)) / td;
  const int dsf = clip3(-4096, 4095, (tb * tx + 32) >> 6);
  const int p = dsf * mv;
  return clip3(-32768, 32767, (p < 0 ? -1 : 1) * ((abs(p) + 127) >> 8));
}
inline bool same_motion(const MotionInfo& a, const MotionInfo& b) {
  if (a.pred_flags != b.pred_flags) return false;
  for (int l = 0; l < 2; l++)
    if (a.pred_flags & (1 << l))
      if (a.mv[l][0] != b.mv[l][0] || a.mv[l][1] != b.mv[l][1] || a.ref_idx[l] != b.ref_idx[l]) return false;
  return true;
}

// 6.4.2 prediction block availability + "is inter coded"; fills `out` when usable
template <class Eng>
bool SliceDec<Eng>::pb_neighbour(int xCb, int yCb, int nCbS, int xPb, int yPb, int nPbW, int nPbH, int partIdx, int xN, int yN,
                                 MotionInfo& out) {
  (void)xPb; (void)yPb;
  const bool same_cb = xCb <= xN && yCb <= yN && xCb + nCbS > xN && yCb + nCbS > yN;
  if (!same_cb) {
    if (!available(xN, yN)) return false;
    if (pic.pred_mode[cb_idx(xN, yN)] == MODE_NONE) return false;   // not decoded yet (z-scan order, 6.4.1)
  } else if ((nPbW << 1) == nCbS && (nPbH << 1) == nCbS && partIdx == 1 && yCb + nPbH <= yN && xCb + nPbW > xN) {
    return false;   // second block of an NxN split asking for the third
  }
  const MotionInfo& m = motion_at(xN, yN);
  if (!m.pred_flags) return false;   // intra
  out = m;
  return true;
}

// collocated motion (8.5.3.2.9) of the block covering (xCol, yCol) in `col`
template <class Eng>
bool SliceDec<Eng>::col_mv(const RefPic& col, int xCol, int yCol, int ref_idx, int X, int16_t mv[2]) {
  xCol = (xCol >> 4) << 4;
  yCol = (yCol >> 4) << 4;
  const MotionInfo& m = col.motion[(size_t)(yCol >> 2) * col.u_w + (xCol >> 2)];
  if (!m.pred_flags) return false;
  const SliceRefs& cr = col.slice_refs[col.ctb_slice[(size_t)(yCol >> col.log2_ctb) * col.ctbs_w + (xCol >> col.log2_ctb)]];
  int l;
  if (!(m.pred_flags & 1)) l = 1;
  else if (!(m.pred_flags & 2)) l = 0;
  else {
    bool no_backward = true;   // DiffPicOrderCnt(aPic, currPic) <= 0 for every picture of both lists
    for (int k = 0; k < 2; k++)
      for (int i = 0; i < refs->n[k]; i++) no_backward &= refs->poc[k][i] <= pic.poc;
    l = no_backward ? X : sh.collocated_from_l0;
  }
  const int ri = m.ref_idx[l];
  if (ri < 0 || ri >= cr.n[l]) return false;
  if ((refs->is_lt[X][ref_idx] != 0) != (cr.is_lt[l][ri] != 0)) return false;
  const int col_diff = col.poc - cr.poc[l][ri], cur_diff = pic.poc - refs->poc[X][ref_idx];
  if (refs->is_lt[X][ref_idx] || col_diff == cur_diff || col_diff == 0) {
    mv[0] = m.mv[l][0]; mv[1] = m.mv[l][1];
  } else {
    mv[0] = (int16_t)mv_scale(m.mv[l][0], cur_diff, col_diff);
    mv[1] = (int16_t)mv_scale(m.mv[l][1], cur_diff, col_diff);
  }
  return true;
}

// temporal luma motion vector prediction (8.5.3.2.8)
template <class Eng>
bool SliceDec<Eng>::temporal_mv(int xPb, int yPb, int nPbW, int nPbH, int ref_idx, int X, int16_t mv[2]) {
  if (!sh.temporal_mvp_enabled) return false;
  const int cl = (sh.slice_type == SLICE_B && !sh.collocated_from_l0) ? 1 : 0;
  if (sh.collocated_ref_idx >= refs->n[cl]) return false;
  const RefPic* col = pic.ref_pics[refs->slot[cl][sh.collocated_ref_idx]].get();
  if (!col || col->failed.load(std::memory_order_relaxed) || col->motion.empty() || col->u_w != pic.u_w || col->u_h != pic.u_h) return false;
  const int xBr = xPb + nPbW, yBr = yPb + nPbH;
  if ((yPb >> sps.log2_ctb) == (yBr >> sps.log2_ctb) && yBr < sps.height && xBr < sps.width)
    if (col_mv(*col, xBr, yBr, ref_idx, X, mv)) return true;
  return col_mv(*col, xPb + (nPbW >> 1), yPb + (nPbH >> 1), ref_idx, X, mv);
}

// merge mode (8.5.3.2.2 - 8.5.3.2.5)
template <class Eng>
void SliceDec<Eng>::derive_merge(int xCb, int yCb, int nCbS, int xPb, int yPb, int nPbW, int nPbH, int partIdx, int merge_idx,
                                 MotionInfo& out) {
  const int orig_w = nPbW, orig_h = nPbH;
  const int lvl = pps.log2_parallel_merge_level;
  int part_mode = cu_part_mode;
  if (lvl > 2 && nCbS == 8) { xPb = xCb; yPb = yCb; nPbW = nPbH = nCbS; partIdx = 0; part_mode = 0; }
  MotionInfo list[6];
  int n = 0;
  const int max_cand = sh.max_num_merge_cand;
  auto same_region = [&](int xN, int yN) { return (xPb >> lvl) == (xN >> lvl) && (yPb >> lvl) == (yN >> lvl); };
  MotionInfo a1, b1, b0, a0, b2;
  bool av_a1 = false, av_b1 = false, av_b0 = false, av_a0 = false, av_b2 = false;
  {
    const int xN = xPb - 1, yN = yPb + nPbH - 1;
    if (!same_region(xN, yN) && !((part_mode == 2 || part_mode == 6 || part_mode == 7) && partIdx == 1))
      av_a1 = pb_neighbour(xCb, yCb, nCbS, xPb, yPb, nPbW, nPbH, partIdx, xN, yN, a1);
    if (av_a1) list[n++] = a1;
  }
  // pruning compares with the neighbour's AVAILABILITY (av_*), not with whether it entered the list (8.5.3.2.3)
  {
    const int xN = xPb + nPbW - 1, yN = yPb - 1;
    if (!same_region(xN, yN) && !((part_mode == 1 || part_mode == 4 || part_mode == 5) && partIdx == 1))
      av_b1 = pb_neighbour(xCb, yCb, nCbS, xPb, yPb, nPbW, nPbH, partIdx, xN, yN, b1);
    if (av_b1 && !(av_a1 && same_motion(a1, b1))) list[n++] = b1;
  }
  {
    const int xN = xPb + nPbW, yN = yPb - 1;
    if (!same_region(xN, yN)) av_b0 = pb_neighbour(xCb, yCb, nCbS, xPb, yPb, nPbW, nPbH, partIdx, xN, yN, b0);
    if (av_b0 && !(av_b1 && same_motion(b1, b0))) list[n++] = b0;
  }
  {
    const int xN = xPb - 1, yN = yPb + nPbH;
    if (!same_region(xN, yN)) av_a0 = pb_neighbour(xCb, yCb, nCbS, xPb, yPb, nPbW, nPbH, partIdx, xN, yN, a0);
    if (av_a0 && !(av_a1 && same_motion(a1, a0))) list[n++] = a0;
  }
  if (n < 4) {
    const int xN = xPb - 1, yN = yPb - 1;
    if (!same_region(xN, yN)) av_b2 = pb_neighbour(xCb, yCb, nCbS, xPb, yPb, nPbW, nPbH, partIdx, xN, yN, b2);
    if (av_b2 && !(av_a1 && same_motion(a1, b2)) && !(av_b1 && same_motion(b1, b2))) list[n++] = b2;
  }
  if (n > max_cand) n = max_cand;
  if (n < max_cand) {   // temporal candidate, refIdx 0 in both lists
    MotionInfo t = {{{0, 0}, {0, 0}}, {-1, -1}, 0, 0};
    if (refs->n[0] > 0 && temporal_mv(xPb, yPb, nPbW, nPbH, 0, 0, t.mv[0])) { t.pred_flags |= 1; t.ref_idx[0] = 0; }
    if (sh.slice_type == SLICE_B && refs->n[1] > 0 && temporal_mv(xPb, yPb, nPbW, nPbH, 0, 1, t.mv[1])) { t.pred_flags |= 2; t.ref_idx[1] = 0; }
    if (t.pred_flags) list[n++] = t;
  }
  if (sh.slice_type == SLICE_B && n > 1 && n < max_cand) {   // combined bi-predictive candidates (8.5.3.2.4)
    static const uint8_t l0c[12] = {0, 1, 0, 2, 1, 2, 0, 3, 1, 3, 2, 3}, l1c[12] = {1, 0, 2, 0, 2, 1, 3, 0, 3, 1, 3, 2};
    const int orig = n;
    for (int comb = 0; comb < orig * (orig - 1) && n < max_cand; comb++) {
      const MotionInfo& c0 = list[l0c[comb]];
      const MotionInfo& c1 = list[l1c[comb]];
      if ((c0.pred_flags & 1) && (c1.pred_flags & 2) &&
          (refs->poc[0][c0.ref_idx[0]] != refs->poc[1][c1.ref_idx[1]] || c0.mv[0][0] != c1.mv[1][0] || c0.mv[0][1] != c1.mv[1][1])) {
        MotionInfo c = {{{c0.mv[0][0], c0.mv[0][1]}, {c1.mv[1][0], c1.mv[1][1]}}, {c0.ref_idx[0], c1.ref_idx[1]}, 3, 0};
        list[n++] = c;
      }
    }
  }
  {   // zero candidates (8.5.3.2.5)
    const int num_ref = sh.slice_type == SLICE_P ? refs->n[0] : std::min(refs->n[0], refs->n[1]);
    int zero = 0;
    while (n < max_cand) {
      const int ri = zero < num_ref ? zero : 0;
      MotionInfo z = {{{0, 0}, {0, 0}}, {(int8_t)ri, (int8_t)(sh.slice_type == SLICE_B ? ri : -1)}, (uint8_t)(sh.slice_type == SLICE_B ? 3 : 1), 0};
      list[n++] = z;
      zero++;
    }
  }
  out = list[merge_idx < n ? merge_idx : n - 1];
  if (out.pred_flags == 3 && orig_w + orig_h == 12) { out.pred_flags = 1; out.ref_idx[1] = -1; }   // 8x4 / 4x8: uni-prediction only
}

// AMVP candidate list (8.5.3.2.6, 8.5.3.2.7) for list X and reference index ref_idx
template <class Eng>
void SliceDec<Eng>::derive_amvp(int xCb, int yCb, int nCbS, int xPb, int yPb, int nPbW, int nPbH, int partIdx, int X, int ref_idx,
                                int16_t cand[2][2]) {
  const int Y = 1 - X;
  const int target_poc = refs->poc[X][ref_idx];
  const bool target_lt = refs->is_lt[X][ref_idx] != 0;
  MotionInfo nb[5];
  bool av[5];
  const int pos[5][2] = {{xPb - 1, yPb + nPbH}, {xPb - 1, yPb + nPbH - 1},                       // A0, A1
                         {xPb + nPbW, yPb - 1}, {xPb + nPbW - 1, yPb - 1}, {xPb - 1, yPb - 1}};  // B0, B1, B2
  for (int k = 0; k < 5; k++) av[k] = pb_neighbour(xCb, yCb, nCbS, xPb, yPb, nPbW, nPbH, partIdx, pos[k][0], pos[k][1], nb[k]);
  // candidate from a neighbour without scaling: same reference picture in either list
  auto exact = [&](const MotionInfo& m, int16_t mv[2]) {
    for (int t = 0; t < 2; t++) {
      const int l = t ? Y : X;
      if ((m.pred_flags & (1 << l)) && refs->poc[l][m.ref_idx[l]] == target_poc) {
        mv[0] = m.mv[l][0]; mv[1] = m.mv[l][1];
        return true;
      }
    }
    return false;
  };
  // candidate with scaling: any reference of the same kind (long / short term)
  auto scaled = [&](const MotionInfo& m, int16_t mv[2]) {
    for (int t = 0; t < 2; t++) {
      const int l = t ? Y : X;
      if ((m.pred_flags & (1 << l)) && (refs->is_lt[l][m.ref_idx[l]] != 0) == target_lt) {
        mv[0] = m.mv[l][0]; mv[1] = m.mv[l][1];
        const int ref_poc = refs->poc[l][m.ref_idx[l]];
        if (!target_lt && !refs->is_lt[l][m.ref_idx[l]] && ref_poc != target_poc) {
          const int td = pic.poc - ref_poc, tb = pic.poc - target_poc;
          if (td != 0) {
            mv[0] = (int16_t)mv_scale(mv[0], tb, td);
            mv[1] = (int16_t)mv_scale(mv[1], tb, td);
          }
        }
        return true;
      }
    }
    return false;
  };
  int16_t mva[2] = {0, 0}, mvb[2] = {0, 0};
  bool has_a = false, has_b = false;
  const bool is_scaled = av[0] || av[1];
  for (int k = 0; k < 2 && !has_a; k++) if (av[k]) has_a = exact(nb[k], mva);
  for (int k = 0; k < 2 && !has_a; k++) if (av[k]) has_a = scaled(nb[k], mva);
  for (int k = 2; k < 5 && !has_b; k++) if (av[k]) has_b = exact(nb[k], mvb);
  if (!is_scaled && has_b) { mva[0] = mvb[0]; mva[1] = mvb[1]; has_a = true; }
  if (!is_scaled) {
    has_b = false;
    for (int k = 2; k < 5 && !has_b; k++) if (av[k]) has_b = scaled(nb[k], mvb);
  }
  int n = 0;
  if (has_a) { cand[n][0] = mva[0]; cand[n][1] = mva[1]; n++; }
  if (has_b && !(has_a && mva[0] == mvb[0] && mva[1] == mvb[1])) { cand[n][0] = mvb[0]; cand[n][1] = mvb[1]; n++; }
  if (n < 2) {
    int16_t mvc[2];
    if (temporal_mv(xPb, yPb, nPbW, nPbH, ref_idx, X, mvc)) { cand[n][0] = mvc[0]; cand[n][1] = mvc[1]; n++; }
  }
  while (n < 2) { cand[n][0] = cand[n][1] = 0; n++; }
}

// mvd_coding (7.3.8.9)
template <class Eng>
int SliceDec<Eng>::read_mvd(int16_t mvd[2]) {
  int gt0[2], gt1[2] = {0, 0};
  gt0[0] = cabac.bin(ctx[CTX_ABS_MVD_GT0]);
  gt0[1] = cabac.bin(ctx[CTX_ABS_MVD_GT0]);
  if (gt0[0]) gt1[0] = cabac.bin(ctx[CTX_ABS_MVD_GT1]);
  if (gt0[1]) gt1[1] = cabac.bin(ctx[CTX_ABS_MVD_GT1]);
  for (int c = 0; c < 2; c++) {
    int v = 0;
    if (gt0[c]) {
      v = 1;
      if (gt1[c]) {   // abs_mvd_minus2: first-order Exp-Golomb, bypass
        cabac.hint(HINT_MVD_SUFFIX, 0);
        int k = 1, base = 0;
        while (cabac.bypass()) {
          base += 1 << k;
          if (++k > 16) return E_DECODING;
        }
        v = 2 + base + (int)cabac.bypass_bits(k);
      }
      if (cabac.bypass()) v = -v;
    }
    if (v < -32768 || v > 32767) return E_DECODING;
    mvd[c] = (int16_t)v;
  }
  return E_OK;
}

// prediction_unit (7.3.8.6) + motion vector derivation (8.5.3.2.1) + the PU work-list record (S3)
template <class Eng>
int SliceDec<Eng>::prediction_unit(int xCb, int yCb, int nCbS, int xPb, int yPb, int nPbW, int nPbH, int partIdx, bool skip,
                                   int& merge_out) {
  int merge = 1, merge_idx = 0;
  if (!skip) merge = cabac.bin(ctx[CTX_MERGE_FLAG]);
  merge_out = merge;
  MotionInfo mi = {{{0, 0}, {0, 0}}, {-1, -1}, 0, 0};
  if (merge) {
    if (sh.max_num_merge_cand > 1) {
      if (cabac.bin(ctx[CTX_MERGE_IDX])) {
        merge_idx = 1;
        while (merge_idx < sh.max_num_merge_cand - 1 && cabac.bypass()) merge_idx++;
      }
    }
    derive_merge(xCb, yCb, nCbS, xPb, yPb, nPbW, nPbH, partIdx, merge_idx, mi);
  } else {
    int idc = 0;   // 0 L0, 1 L1, 2 BI
    if (sh.slice_type == SLICE_B) {
      if (nPbW + nPbH != 12 && cabac.bin(ctx[CTX_INTER_PRED_IDC + cu_ct_depth])) idc = 2;
      else idc = cabac.bin(ctx[CTX_INTER_PRED_IDC + 4]);
    }
    int16_t mvd[2][2] = {{0, 0}, {0, 0}};
    int mvp_flag[2] = {0, 0};
    for (int X = 0; X < 2; X++) {
      if (idc == 1 - X) continue;   // list not used
      int ri = 0;
      const int nref = sh.num_ref_idx[X];
      if (nref > 1) {   // ref_idx_lX: truncated Rice, two context coded bins
        while (ri < nref - 1) {
          const int b = ri < 2 ? cabac.bin(ctx[CTX_REF_IDX + ri]) : cabac.bypass();
          if (!b) break;
          ri++;
        }
      }
      if (ri >= refs->n[X]) return E_DECODING;
      mi.ref_idx[X] = (int8_t)ri;
      mi.pred_flags |= (uint8_t)(1 << X);
      if (!(X == 1 && sh.mvd_l1_zero && idc == 2)) {
        const int r = read_mvd(mvd[X]);
        if (r) return r;
      }
      mvp_flag[X] = cabac.bin(ctx[CTX_MVP_FLAG]);
    }
    for (int X = 0; X < 2; X++) {
      if (!(mi.pred_flags & (1 << X))) continue;
      int16_t cand[2][2];
      derive_amvp(xCb, yCb, nCbS, xPb, yPb, nPbW, nPbH, partIdx, X, mi.ref_idx[X], cand);
      for (int c = 0; c < 2; c++) mi.mv[X][c] = (int16_t)(uint16_t)((int)cand[mvp_flag[X]][c] + mvd[X][c]);   // modulo 2^16 (8-272..8-275)
    }
  }
#ifdef B265_PU_TRACE
  fprintf(stderr, "PU %d %d %dx%d part %d merge %d idx %d pf %d ref %d %d mv %d %d %d %d\n", xPb, yPb, nPbW, nPbH, partIdx, merge, merge_idx,
          mi.pred_flags, mi.ref_idx[0], mi.ref_idx[1], mi.mv[0][0], mi.mv[0][1], mi.mv[1][0], mi.mv[1][1]);
#endif
  // motion field
  for (int y = 0; y < nPbH; y += 4) {
    MotionInfo* row = &motion_at(xPb, yPb + y);
    for (int x = 0; x < (nPbW >> 2); x++) row[x] = mi;
  }
  // work-list record
  b265_pu pu;
  memset(&pu, 0, sizeof(pu));
  pu.x = (uint16_t)xPb; pu.y = (uint16_t)yPb; pu.w = (uint8_t)nPbW; pu.h = (uint8_t)nPbH;
  pu.flags = mi.pred_flags;
  pu.slice = (uint8_t)sh.slice_index;
  for (int l = 0; l < 2; l++) {
    if (!(mi.pred_flags & (1 << l))) continue;
    pu.mv[l][0] = mi.mv[l][0]; pu.mv[l][1] = mi.mv[l][1];
    pu.ref_idx[l] = mi.ref_idx[l];
    pu.ref_slot[l] = refs->slot[l][mi.ref_idx[l]];
  }
  pic.pus.push_back(pu);
  return E_OK;
}

// prediction-block edges inside a CU (deblock.cc:185-225): filtered like transform edges, strength from the motion
template <class Eng>
void SliceDec<Eng>::mark_pu_edges(int x0, int y0, int nCbS, int part_mode) {
  if (sh.deblocking_disabled || part_mode == 0) return;
  uint8_t* ui = pic.unit_info.data();
  const int n_u = nCbS >> 2, ux = x0 >> 2, uy = y0 >> 2;
  const int half = n_u >> 1, quarter = n_u >> 2;
  int v_at = -1, h_at = -1;
  switch (part_mode) {
    case 1: h_at = half; break;
    case 2: v_at = half; break;
    case 3: h_at = half; v_at = half; break;
    case 4: h_at = quarter; break;
    case 5: h_at = half + quarter; break;
    case 6: v_at = quarter; break;
    default: v_at = half + quarter; break;
  }
  if (h_at > 0) for (int i = 0; i < n_u; i++) ui[(uy + h_at) * pic.u_w + ux + i] |= B265_UNIT_EDGE_H;
  if (v_at > 0) for (int j = 0; j < n_u; j++) ui[(uy + j) * pic.u_w + ux + v_at] |= B265_UNIT_EDGE_V;
}

// ---- coding unit (7.3.8.5) ---------------------------------------------------------------------------------
template <class Eng>
int SliceDec<Eng>::coding_unit(int x0, int y0, int log2_cb) {
  PROF(2);
  const int n_cbs = 1 << log2_cb;
  decode_qp(x0, y0, log2_cb);
  cu_tq_bypass = 0;
  if (pps.transquant_bypass_enabled) cu_tq_bypass = cabac.bin(ctx[CTX_CU_TQ_BYPASS]);
  const int n_cb = 1 << (log2_cb - sps.log2_min_cb), ci0 = cb_idx(x0, y0);
  const int n_u = n_cbs >> 2, ux = x0 >> 2, uy = y0 >> 2;
  cu_ct_depth = pic.ct_depth[ci0];
  int skip = 0;
  if (sh.slice_type != SLICE_I) {
    int inc = 0;
    if (available(x0 - 1, y0) && (pic.cu_flags[cb_idx(x0 - 1, y0)] & 4)) inc++;
    if (available(x0, y0 - 1) && (pic.cu_flags[cb_idx(x0, y0 - 1)] & 4)) inc++;
    skip = cabac.bin(ctx[CTX_CU_SKIP + inc]);
  }
  int intra = 1, part_mode = 0;
  if (!skip) {
    if (sh.slice_type != SLICE_I) intra = cabac.bin(ctx[CTX_PRED_MODE]);
    if (intra) {
      if (log2_cb == sps.log2_min_cb) part_mode = cabac.bin(ctx[CTX_PART_MODE]) ? 0 : 3;
    } else if (!cabac.bin(ctx[CTX_PART_MODE])) {   // part_mode binarisation of an inter CU (9.3.3.7)
      if (log2_cb == sps.log2_min_cb) {
        if (cabac.bin(ctx[CTX_PART_MODE + 1])) part_mode = 1;
        else if (log2_cb == 3) part_mode = 2;
        else part_mode = cabac.bin(ctx[CTX_PART_MODE + 2]) ? 2 : 3;
      } else if (!sps.amp_enabled) {
        part_mode = cabac.bin(ctx[CTX_PART_MODE + 1]) ? 1 : 2;
      } else {
        const int horizontal = cabac.bin(ctx[CTX_PART_MODE + 1]);
        if (cabac.bin(ctx[CTX_PART_MODE + 3])) part_mode = horizontal ? 1 : 2;
        else part_mode = (horizontal ? 4 : 6) + cabac.bypass();
      }
    }
  } else {
    intra = 0;
  }
  cu_pred_mode = intra ? MODE_INTRA : (skip ? MODE_SKIP : MODE_INTER);
  cu_part_mode = part_mode;
  const int nxn = intra && part_mode == 3;
  if (nxn && log2_cb == 3 && sps.log2_min_tb > 2) return E_DECODING;
  // metadata of the CB
  for (int j = 0; j < n_cb; j++) {
    memset(&pic.pred_mode[ci0 + j * pic.cb_w], cu_pred_mode, n_cb);
    memset(&pic.cu_flags[ci0 + j * pic.cb_w], (cu_tq_bypass ? 2 : 0) | (skip ? 4 : 0), n_cb);
  }
  if (!intra) {
    // ---- inter CU: prediction units, then the residual quadtree ----
    if (!refs || refs->n[0] == 0) return E_DECODING;
    pic.has_inter = true;
    const uint8_t unit_bits = (uint8_t)(B265_UNIT_DONE | (cu_tq_bypass ? B265_UNIT_NOFILT : 0));
    for (int j = 0; j < n_u; j++) memset(&pic.unit_info[(uy + j) * pic.u_w + ux], unit_bits, n_u);
    int merge0 = 0, r = 0;
    const int h = n_cbs >> 1, q = n_cbs >> 2, f = n_cbs;
    // prediction blocks {x, y, w, h} of the partitioning modes (table 7-10)
    const int parts[8][4][4] = {
        {{0, 0, f, f}},                                             // 2Nx2N
        {{0, 0, f, h}, {0, h, f, h}},                               // 2NxN
        {{0, 0, h, f}, {h, 0, h, f}},                               // Nx2N
        {{0, 0, h, h}, {h, 0, h, h}, {0, h, h, h}, {h, h, h, h}},   // NxN
        {{0, 0, f, q}, {0, q, f, f - q}},                           // 2NxnU
        {{0, 0, f, f - q}, {0, f - q, f, q}},                       // 2NxnD
        {{0, 0, q, f}, {q, 0, f - q, f}},                           // nLx2N
        {{0, 0, f - q, f}, {f - q, 0, q, f}}};                      // nRx2N
    const int n_parts = part_mode == 0 ? 1 : (part_mode == 3 ? 4 : 2);
    for (int i = 0; i < n_parts && !r; i++) {
      int merge = 0;
      r = prediction_unit(x0, y0, n_cbs, x0 + parts[part_mode][i][0], y0 + parts[part_mode][i][1], parts[part_mode][i][2],
                          parts[part_mode][i][3], i, skip != 0, merge);
      if (i == 0) merge0 = merge;
    }
    if (r) return r;
    mark_pu_edges(x0, y0, n_cbs, part_mode);
    int rqt_root_cbf = 0;
    if (!skip) {   // inferred 1 for a merged 2Nx2N CU
      rqt_root_cbf = 1;
      if (!(part_mode == 0 && merge0)) rqt_root_cbf = cabac.bin(ctx[CTX_RQT_ROOT_CBF]);
    }
    if (!rqt_root_cbf) {
      mark_tu_edges(x0, y0, log2_cb, x0, y0, log2_cb, false);   // the CU boundary is still a block edge
      return E_OK;
    }
    const int inter_split = sps.max_th_depth_inter == 0 && part_mode != 0;
    return transform_tree(x0, y0, x0, y0, x0, y0, log2_cb, 0, 0, sps.max_th_depth_inter, inter_split, 1, 1, log2_cb);
  }
  int pcm = 0;
  if (!nxn && sps.pcm_enabled && log2_cb >= sps.log2_min_pcm && log2_cb <= sps.log2_max_pcm) {
    cabac.hint(HINT_PCM_FLAG, 0);
    pcm = cabac.terminate();
  }
  uint8_t unit_bits = B265_UNIT_INTRA;
  if (cu_tq_bypass || (pcm && sps.pcm_lf_disabled)) unit_bits |= B265_UNIT_NOFILT;
  if (pcm) unit_bits |= B265_UNIT_DONE;
  for (int j = 0; j < n_u; j++) memset(&pic.unit_info[(uy + j) * pic.u_w + ux], unit_bits, n_u);
  if (pcm) {
    const int ci = ci0;
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
    const int mode = cu_pred_mode == MODE_INTRA ? pic.ipm[(y0 >> 2) * pic.u_w + (x0 >> 2)] : B265_INTRA_NONE;
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
    tu.intra_mode = (uint8_t)(cu_pred_mode == MODE_INTRA ? intra_chroma_mode : B265_INTRA_NONE);
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
  PROF_CNT(9, 1);
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
    PROF_CNT(10, 1);
    PROF_CNT(11, nsig);
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
        PROF_CNT(12, 1);
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
  const bool intra_cu = cu_pred_mode == MODE_INTRA;
  uint8_t fl = intra_cu ? B265_TU_INTRA_CU : 0;
  if (intra_cu && cidx == 0 && log2_ts == 2) fl |= B265_TU_DST;
  if (tskip) fl |= B265_TU_SKIP;
  if (cu_tq_bypass) fl |= B265_TU_BYPASS;
  if (implicit_rdpcm_hv) fl |= intra_mode == 26 ? B265_TU_RDPCM_V : B265_TU_RDPCM_H;
  // The reference decides the 4x4 rotation from the prediction mode found at the block's COMPONENT coordinates in
  // the luma-resolution map (transform.cc:392-397 with slice.cc:3471): for chroma that is the CU covering
  // (x/2, y/2), already decoded.  Kept as is: the reference is the oracle.
  if (sps.transform_skip_rotation && log2_ts == 2 && pic.pred_mode[cb_idx(x0, y0)] == MODE_INTRA) fl |= B265_TU_ROTATE;
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
  pic.self->ctb_slice[ctb_rs] = (uint16_t)slice_index;
  read_sao(rx, ry);
  return coding_quadtree(rx << sps.log2_ctb, ry << sps.log2_ctb, sps.log2_ctb, 0);
}

}  // namespace syn


inline void wait_a_little() { std::this_thread::sleep_for(std::chrono::microseconds(50)); }

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
  if (sh.slice_type != SLICE_I) {
    if (sps.explicit_rdpcm) return E_NOT_IMPLEMENTED;
    if (sh.slice_index >= (int)pic.slice_refs.size()) return E_DECODING;
    if (pic.self->motion.empty()) pic.self->motion.assign((size_t)pic.u_w * pic.u_h, MotionInfo{{{0, 0}, {0, 0}}, {-1, -1}, 0, 0});
    d.refs = &pic.slice_refs[sh.slice_index];
    if (d.refs->n[0] < sh.num_ref_idx[0] || (sh.slice_type == SLICE_B && d.refs->n[1] < sh.num_ref_idx[1])) return E_DECODING;
    if (sh.temporal_mvp_enabled) {   // the collocated picture's motion field must be complete (it precedes in decoding order)
      const int cl = (sh.slice_type == SLICE_B && !sh.collocated_from_l0) ? 1 : 0;
      if (sh.collocated_ref_idx < d.refs->n[cl]) {
        const RefPic* col = pic.ref_pics[d.refs->slot[cl][sh.collocated_ref_idx]].get();
        while (col && !col->parsed.load(std::memory_order_acquire)) wait_a_little();   // `failed` is published with `parsed`
      }
    }
  }
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
