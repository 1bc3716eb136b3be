// k_sao.cu — sample adaptive offset: reads the deblocked picture (FrameDev::plane), writes the output /
// DPB surface (FrameDev::out).  Restates apply_sao_internal (libde265/sao.cc:29-254) and its driver
// (:318-370): edge offset with picture/slice/tile boundary rules, band offset, pcm/bypass skip.
// Quirk kept on purpose: the CTB's own slice address is looked up with COMPONENT coordinates through
// a luma accessor (sao.cc:50-56), which only matters for chroma in multi-slice pictures.
// One thread per 4 horizontally adjacent samples; traffic 2P per picture (+ 18 B per CTB).
#include "transform.cuh"
#include "kernels.h"

__device__ __forceinline__ int sgn(int v) { return (v > 0) - (v < 0); }

__global__ void k_sao(const FrameDev* __restrict__ frames) {
  const FrameDev& f = frames[blockIdx.z];
  if (!(f.flags & B265_FRAME_SAO) || !f.sao) return;
  const int c = blockIdx.y;
  if (c && !f.cw) return;
  const int sh = c ? 1 : 0;
  const int W = c ? f.cw : f.width, H = c ? f.ch : f.height;
  const int qw = (W + 3) >> 2;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= qw * H) return;
  const int x0 = (i % qw) << 2, y = i / qw;
  const int pitch = f.pitch[c];
  const uint8_t* in = f.plane[c];
  uint8_t* out = f.out[c];
  const int lc = f.log2_ctb - sh;  // CTB size in this component
  const int xCtb = x0 >> lc, yCtb = y >> lc;
  const int ctb = yCtb * f.ctbs_w + xCtb;
  const int sidx = f.ctbs ? f.ctbs[ctb].slice : 0;
  const b265_slice& sl = f.slices[sidx];
  const b265_sao s = f.sao[ctb];
  int type = s.type[c];
  if (!(c ? sl.sao_chroma : sl.sao_luma)) type = 0;
  uint32_t cen = 0;
  const int nvalid = min(4, W - x0);
  for (int k = 0; k < nvalid; k++) cen |= (uint32_t)in[(size_t)y * pitch + x0 + k] << (8 * k);
  uint32_t res = cen;
  if (type != 0) {
    // slice address the reference associates with this CTB (component coords through a luma accessor)
    const int qx = (xCtb << lc) >> f.log2_ctb, qy = (yCtb << lc) >> f.log2_ctb;
    const int qctb = qy * f.ctbs_w + qx;
    const int ctb_slice_addr = f.slices[f.ctbs ? f.ctbs[qctb].slice : 0].slice_addr_rs;
    const int tile = f.ctbs ? f.ctbs[ctb].tile : 0;
    int dx0 = 0, dy0 = 0, dx1 = 0, dy1 = 0;
    if (type == 2) {
      switch (s.band_eo[c]) {
        case 0: dx0 = -1; dx1 = 1; break;
        case 1: dy0 = -1; dy1 = 1; break;
        case 2: dx0 = -1; dy0 = -1; dx1 = 1; dy1 = 1; break;
        default: dx0 = 1; dy0 = -1; dx1 = -1; dy1 = 1; break;
      }
    }
    for (int k = 0; k < nvalid; k++) {
      const int x = x0 + k;
      const int v = (cen >> (8 * k)) & 0xff;
      if (f.unit_info && (f.unit_info[((y << sh) >> 2) * (f.width >> 2) + ((x << sh) >> 2)] & B265_UNIT_NOFILT)) continue;
      int off = 0;
      if (type == 2) {
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
        off = (e == 0) ? 0 : s.offset[c][e < 0 ? e + 2 : e + 1];
      } else {
        const int kb = ((v >> 3) - s.band_eo[c]) & 31;
        if (kb >= 4) continue;
        off = s.offset[c][kb];
      }
      res = (res & ~(0xffu << (8 * k))) | ((uint32_t)clip8(v + off) << (8 * k));
    }
  }
  if (nvalid == 4) *reinterpret_cast<uint32_t*>(out + (size_t)y * pitch + x0) = res;
  else for (int k = 0; k < nvalid; k++) out[(size_t)y * pitch + x0 + k] = (res >> (8 * k)) & 0xff;
}

void launch_sao(cudaStream_t st, const FrameDev* d_frames, int n_frames, int max_w, int max_h) {
  dim3 g((((max_w + 3) >> 2) * max_h + 255) / 256, 3, n_frames);
  k_sao<<<g, 256, 0, st>>>(d_frames);
}
