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
void launch_mc(cudaStream_t st, const FrameDev* frames, const uint32_t* items, int n, int sm_count);
// PCM sample store (k_blocks.cu): grid (max_pcms, n_frames)
void launch_pcm(cudaStream_t st, const FrameDev* frames, int n_frames, int max_pcms);
// isolated-block kernels behind b265_intra_batch / b265_interp_batch (k_blocks.cu)
void launch_intra_blocks(cudaStream_t st, int log2, int cidx, int n_blocks, const uint8_t* modes, const uint8_t* borders,
                         int strong, int no_edge_filter, uint8_t* out);
void launch_interp_blocks(cudaStream_t st, int luma, int n_blocks, int w, int h, const uint8_t* src, long long src_stride,
                          long long src_block_stride, const uint8_t* frac_xy, int16_t* out);
// availability masks of all intra TBs (k_intra.cu), grid (.., n_frames)
void launch_intra_prep(cudaStream_t st, const FrameDev* frames, int n_frames, int max_tus);
int intra_max_resident_ctas(int sm_count, int log2_ctb);
void launch_intra(cudaStream_t st, const FrameDev* frames, const IntraTicket* tickets, int n_tickets,
                  unsigned* ticket, int sm_count, int wave_width, int log2_ctb);
struct b265_ctx;
// bS derivation + deblock V/H + SAO for every frame of a batch (k_postfilter.cu); ev (optional) is
// recorded between deblocking and SAO for stage timing.
int launch_postfilters(b265_ctx* ctx, const FrameDev* d_frames, const FrameDev* h_frames, int n_frames,
                       cudaEvent_t ev, const uint32_t* mc_items, int n_mc);
void launch_deblock(cudaStream_t st, const FrameDev* d_frames, int n_frames, int max_w, int max_h,
                    const uint32_t* mc_items, int n_mc, bool chroma);
void launch_sao(cudaStream_t st, const FrameDev* d_frames, int n_frames, int max_w, int max_h);
