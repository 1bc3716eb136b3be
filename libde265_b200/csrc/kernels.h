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
