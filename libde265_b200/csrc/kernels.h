// kernels.h — host-callable launchers of the b265 CUDA kernels (internal to libb265.so)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "common.cuh"

void launch_residual_dense(cudaStream_t st, int log2, int kind, int rdpcm, int n_blocks,
                           const int16_t* coeffs, const uint8_t* qp, uint8_t* dst, int sm_count);
// packed dp2a kernels for kind DCT (all sizes) and DST 4x4 (k_residual_dense.cu); false = no fast kernel for this kind
bool launch_residual_dense_dct(cudaStream_t st, int log2, int kind, int n_blocks, const int16_t* coeffs,
                               const uint8_t* qp, uint8_t* dst, int sm_count);
// to_scratch: write the residuals as int16 into FrameDev::resid (intra TBs, added later by the
// wavefront) instead of adding them to the prediction already in the plane (inter TBs)
void launch_residual_sparse(cudaStream_t st, int log2, const FrameDev* frames, const uint32_t* order,
                            int n, int sm_count, bool to_scratch = false);
// plain DCT TUs of 16x16 / 32x32 on the packed dp2a kernels (k_residual_dense.cu); false = size not covered
bool launch_residual_sparse_fast(cudaStream_t st, int log2, const FrameDev* frames, const uint32_t* order, int n,
                                 int sm_count, bool to_scratch);
void launch_mc(cudaStream_t st, const FrameDev* frames, const uint32_t* items, int n, int sm_count);
// PCM sample store (k_blocks.cu): grid (max_pcms, n_frames)
void launch_pcm(cudaStream_t st, const FrameDev* frames, int n_frames, int max_pcms);
// isolated-block kernels behind b265_intra_batch / b265_interp_batch (k_blocks.cu)
void launch_intra_blocks(cudaStream_t st, int log2, int cidx, int n_blocks, const uint8_t* modes, const uint8_t* borders,
                         int strong, int no_edge_filter, uint8_t* out);
void launch_interp_blocks(cudaStream_t st, int luma, int n_blocks, int w, int h, const uint8_t* src, long long src_stride,
                          long long src_block_stride, const uint8_t* frac_xy, int16_t* out);
// availability masks of all intra TBs (k_intra.cu), grid (.., n_frames)
// availability masks of all intra TBs + runs per CTB-row level (k_intra.cu), grid (.., n_frames); `sched` is zero
void launch_intra_prep(cudaStream_t st, const FrameDev* frames, int n_frames, int max_tus, int max_slots, IntraSched* sched);
int intra_max_resident_ctas(int sm_count, int log2_ctb);
// neighbour masks + run queue (grid over the (plane, CTB) slots of every frame), then the wavefront itself; after
// launch_intra_prep on the same frames.  `queue` has `queue_entries` (a power of two >= n_runs + intra_queue_slack())
// 8-byte entries and is zero; busy_peers = other contexts with work on the device (the grid takes its share of the SMs)
void launch_intra(cudaStream_t st, const FrameDev* frames, int n_frames, int max_slots, int max_rows, int n_runs, IntraSched* sched,
                  unsigned long long* queue, unsigned queue_entries, int sm_count, int log2_ctb, int busy_peers);
int intra_queue_slack(int sm_count);
// decoded picture hash of n pictures (k_hash.cu): type 0 MD5, 1 CRC, 2 checksum (sei.h sei_decoded_picture_hash_type);
// out = n * 3 slots of 16 bytes (MD5 digest / CRC or checksum in the first word); returns the number of launches
int launch_hash(cudaStream_t st, int type, const HashJob* jobs, int n, int max_h, uint16_t* row_scratch, uint8_t* out);
struct b265_ctx;
// bS derivation + deblock V/H + SAO for every frame of a batch (k_postfilter.cu); ev (optional) is
// recorded between deblocking and SAO for stage timing.
int launch_postfilters(b265_ctx* ctx, const FrameDev* d_frames, const FrameDev* h_frames, int n_frames,
                       cudaEvent_t ev, const uint32_t* mc_items, int n_mc);
void launch_deblock(cudaStream_t st, const FrameDev* d_frames, int n_frames, int max_w, int max_h,
                    const uint32_t* mc_items, int n_mc, bool chroma);
void launch_sao(cudaStream_t st, const FrameDev* d_frames, int n_frames, int max_w, int max_h);

// Per-plane, per-CTB SAO control records for k_sao (host side, at batch build): resolves the slice's sao_luma / sao_chroma
// switches, packs the offsets, and flags CTBs that touch another slice or tile.  out has 3 * ctbs_w * ctbs_h entries.
inline void sao_records(const b265_frame_job& J, int ctbs_w, int ctbs_h, bool default_on, uint2* out) {
  const int n = ctbs_w * ctbs_h;
  for (int y = 0; y < ctbs_h; y++)
    for (int x = 0; x < ctbs_w; x++) {
      const int i = y * ctbs_w + x;
      const b265_sao& s = J.sao[i];
      const int my_slice = J.ctbs ? J.ctbs[i].slice : 0, my_tile = J.ctbs ? J.ctbs[i].tile : 0;
      const b265_slice* sl = (J.n_slices > 0 && J.slices) ? &J.slices[my_slice] : nullptr;
      bool foreign = false;
      if (J.ctbs)
        for (int dy = -1; dy <= 1; dy++)
          for (int dx = -1; dx <= 1; dx++) {
            const int nx = x + dx, ny = y + dy;
            if ((dx | dy) == 0 || nx < 0 || ny < 0 || nx >= ctbs_w || ny >= ctbs_h) continue;
            const b265_ctb& o = J.ctbs[ny * ctbs_w + nx];
            foreign |= o.slice != my_slice || o.tile != my_tile;
          }
      for (int c = 0; c < 3; c++) {
        int type = s.type[c];
        const bool on = sl ? (c ? sl->sao_chroma : sl->sao_luma) != 0 : default_on;   // no slice list: the caller's default slice
        if (!on) type = 0;
        bool fgn = foreign;
        // chroma: the reference looks this CTB's slice up with component coordinates (sao.cc:50-56)
        if (c && J.ctbs) fgn |= J.ctbs[(y >> 1) * ctbs_w + (x >> 1)].slice != my_slice;
        const uint32_t o4 = (uint32_t)(uint8_t)s.offset[c][0] | ((uint32_t)(uint8_t)s.offset[c][1] << 8) |
                            ((uint32_t)(uint8_t)s.offset[c][2] << 16) | ((uint32_t)(uint8_t)s.offset[c][3] << 24);
        out[c * n + i] = make_uint2((uint32_t)(type & 3) | ((uint32_t)(s.band_eo[c] & 31) << 2) | (fgn ? B265_SAOREC_FOREIGN : 0u), o4);
      }
    }
}
