/*
 * oracle/sao.c — CPU restatement of sample adaptive offset.  TEST INFRASTRUCTURE ONLY (see oracle.h).
 * Follows libde265/sao.cc apply_sao_internal (:29-254) per CTB and component, driven like
 * apply_sample_adaptive_offset_sequential (:318-370): read the deblocked picture, write the output.
 */
#include <string.h>
#include "oracle.h"

static int clip3(int lo, int hi, int v) { return v < lo ? lo : (v > hi ? hi : v); }
static int sign(int v) { return (v > 0) - (v < 0); }

static int ctbs_w(const b265_frame_job* J) { return (J->width + (1 << J->log2_ctb) - 1) >> J->log2_ctb; }
static int ctbs_h(const b265_frame_job* J) { return (J->height + (1 << J->log2_ctb) - 1) >> J->log2_ctb; }
/* slice header of the CTB containing LUMA sample (x,y) (img->get_SliceHeader) */
static const b265_slice* slice_luma(const b265_frame_job* J, int x, int y) {
  int ctb = (y >> J->log2_ctb) * ctbs_w(J) + (x >> J->log2_ctb);
  return &J->slices[J->ctbs ? J->ctbs[ctb].slice : 0];
}
static int tile_of(const b265_frame_job* J, int cx, int cy) { return J->ctbs ? J->ctbs[cy * ctbs_w(J) + cx].tile : 0; }

static void sao_ctb(const b265_frame_job* J, int xCtb, int yCtb, int cIdx, int nSW, int nSH, const uint8_t* in,
                    int in_stride, uint8_t* out, int out_stride) {
  const b265_sao* s = &J->sao[yCtb * ctbs_w(J) + xCtb];
  int type = s->type[cIdx];
  if (type == 0) return;
  int sh = cIdx ? 1 : 0;
  int width = cIdx ? (J->width + 1) / 2 : J->width, height = cIdx ? (J->height + 1) / 2 : J->height;
  int xC = xCtb * nSW, yC = yCtb * nSH;
  /* component coordinates handed to a luma accessor — the reference's quirk (sao.cc:50-56) */
  int ctbSliceAddrRS = slice_luma(J, xC, yC)->slice_addr_rs;
  int ctbshift = J->log2_ctb - sh;
  int ctbW = (xC + nSW > width) ? width - xC : nSW;
  int ctbH = (yC + nSH > height) ? height - yC : nSH;
  int lf_tiles = (J->flags & B265_FRAME_LF_ACROSS_TILES) != 0;
  if (type == 2) {
    int hPos[2], vPos[2];
    switch (s->band_eo[cIdx]) {
      case 0: hPos[0] = -1; hPos[1] = 1; vPos[0] = 0; vPos[1] = 0; break;
      case 1: hPos[0] = 0; hPos[1] = 0; vPos[0] = -1; vPos[1] = 1; break;
      case 2: hPos[0] = -1; hPos[1] = 1; vPos[0] = -1; vPos[1] = 1; break;
      default: hPos[0] = 1; hPos[1] = -1; vPos[0] = -1; vPos[1] = 1; break;
    }
    int off[5] = {s->offset[cIdx][0], s->offset[cIdx][1], 0, s->offset[cIdx][2], s->offset[cIdx][3]};
    for (int j = 0; j < ctbH; j++)
      for (int i = 0; i < ctbW; i++) {
        int xl = (xC + i) << sh, yl = (yC + j) << sh;
        if (J->unit_info && (J->unit_info[(yl >> 2) * (J->width >> 2) + (xl >> 2)] & B265_UNIT_NOFILT)) continue;
        int edgeIdx = -1;
        int test = (i == 0 || j == 0 || i == ctbW - 1 || j == ctbH - 1);
        if (test)
          for (int k = 0; k < 2; k++) {
            int xS = xC + i + hPos[k], yS = yC + j + vPos[k];
            if (xS < 0 || yS < 0 || xS >= width || yS >= height) { edgeIdx = 0; break; }
            const b265_slice* ns = slice_luma(J, xS << sh, yS << sh);
            if (ns->slice_addr_rs < ctbSliceAddrRS && !slice_luma(J, xl, yl)->lf_across_slices) { edgeIdx = 0; break; }
            if (ns->slice_addr_rs > ctbSliceAddrRS && !ns->lf_across_slices) { edgeIdx = 0; break; }
            if (!lf_tiles && tile_of(J, xS >> ctbshift, yS >> ctbshift) != tile_of(J, xC >> ctbshift, yC >> ctbshift)) {
              edgeIdx = 0;
              break;
            }
          }
        if (edgeIdx != 0) {
          const uint8_t* p = in + xC + i + (yC + j) * in_stride;
          edgeIdx = sign(p[0] - p[hPos[0] + vPos[0] * in_stride]) + sign(p[0] - p[hPos[1] + vPos[1] * in_stride]);
          out[xC + i + (yC + j) * out_stride] = (uint8_t)clip3(0, 255, p[0] + off[edgeIdx + 2]);
        }
      }
  } else {
    int bandTable[32];
    memset(bandTable, 0, sizeof(bandTable));
    for (int k = 0; k < 4; k++) bandTable[(k + s->band_eo[cIdx]) & 31] = k + 1;
    for (int j = 0; j < ctbH; j++)
      for (int i = 0; i < ctbW; i++) {
        int xl = (xC + i) << sh, yl = (yC + j) << sh;
        if (J->unit_info && (J->unit_info[(yl >> 2) * (J->width >> 2) + (xl >> 2)] & B265_UNIT_NOFILT)) continue;
        int v = in[xC + i + (yC + j) * in_stride];
        int bandIdx = bandTable[v >> 3];
        if (bandIdx > 0) out[xC + i + (yC + j) * out_stride] = (uint8_t)clip3(0, 255, v + s->offset[cIdx][bandIdx - 1]);
      }
  }
}

void orc_sao(const b265_frame_job* J, const uint8_t* const* in_planes, uint8_t* const* out_planes, const int* pitch) {
  int nc = J->chroma_format_idc ? 3 : 1;
  for (int c = 0; c < nc; c++) {
    int h = c ? (J->height + 1) / 2 : J->height, w = c ? (J->width + 1) / 2 : J->width;
    for (int y = 0; y < h; y++) memcpy(out_planes[c] + (size_t)y * pitch[c], in_planes[c] + (size_t)y * pitch[c], w);
    if (!(J->flags & B265_FRAME_SAO) || !J->sao) continue;
    for (int yCtb = 0; yCtb < ctbs_h(J); yCtb++)
      for (int xCtb = 0; xCtb < ctbs_w(J); xCtb++) {
        const b265_slice* sl = &J->slices[J->ctbs ? J->ctbs[yCtb * ctbs_w(J) + xCtb].slice : 0];
        int n = 1 << J->log2_ctb;
        if (c == 0 && sl->sao_luma) sao_ctb(J, xCtb, yCtb, 0, n, n, in_planes[0], pitch[0], out_planes[0], pitch[0]);
        if (c != 0 && sl->sao_chroma) sao_ctb(J, xCtb, yCtb, c, n / 2, n / 2, in_planes[c], pitch[c], out_planes[c], pitch[c]);
      }
  }
}
