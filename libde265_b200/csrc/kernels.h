// kernels.h — host-callable launchers of the b265 CUDA kernels (internal to libb265.so)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "common.cuh"

void launch_residual_dense(cudaStream_t st, int log2, int kind, int rdpcm, int n_blocks,
                           const int16_t* coeffs, const uint8_t* qp, uint8_t* dst, int sm_count);
void launch_residual_sparse(cudaStream_t st, int log2, const FrameDev* frames, const uint32_t* order,
                            int n, int sm_count);
void launch_mc(cudaStream_t st, const FrameDev* frames, const uint32_t* items, int n, int sm_count);
int intra_max_resident_ctas(int sm_count);
void launch_intra(cudaStream_t st, const FrameDev* frames, const IntraTicket* tickets, int n_tickets,
                  unsigned* ticket, int max_ctas, int wave_width);
struct b265_ctx;
// bS derivation + deblock V/H + SAO for every frame of a batch (k_postfilter.cu); ev (optional) is
// recorded between deblocking and SAO for stage timing.
int launch_postfilters(b265_ctx* ctx, const FrameDev* d_frames, const FrameDev* h_frames, int n_frames,
                       cudaEvent_t ev, const uint32_t* mc_items, int n_mc);
void launch_deblock(cudaStream_t st, const FrameDev* d_frames, int n_frames, int max_w, int max_h,
                    const uint32_t* mc_items, int n_mc, bool chroma);
void launch_sao(cudaStream_t st, const FrameDev* d_frames, int n_frames, int max_w, int max_h);
