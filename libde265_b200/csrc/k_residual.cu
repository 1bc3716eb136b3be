// k_residual.cu — residual kernels: dense batch form (acceleration.h transform slots, config 2) and
// sparse work-list form (inter TUs of the frame path, S2 = scale_coefficients, transform.cc:353-625).
#include "residual.cuh"
#include "kernels.h"

// ---------------------------------------------------------------------------------------------
// Dense form: block i has N*N int16 coefficients at coeffs + i*N*N and N*N pixels at dst + i*N*N.
// One warp handles 32/N consecutive blocks per iteration => every global access of the warp is a
// single contiguous span (64*N bytes of coefficients, 32*N bytes of pixels).
// HBM traffic per block: 2*N*N (coeff) + N*N (pred) + N*N (recon) = 4*N*N bytes.
// ---------------------------------------------------------------------------------------------
template <int LOG2>
__global__ void __launch_bounds__(B265_WARPS_PER_CTA * 32)
k_residual_dense(const int16_t* __restrict__ coeffs, const uint8_t* __restrict__ qp,
                 uint8_t* __restrict__ dst, int n_blocks, int kind, int rdpcm) {
  constexpr int N = TrBlk<LOG2>::N, S = TrBlk<LOG2>::S, G = 32 / N;
  __shared__ __align__(16) int16_t smem[B265_WARPS_PER_CTA][B265_TR_SMEM_PER_WARP / 2];
  const int warp = threadIdx.x >> 5, lane = lane_id();
  const int grp = lane / N, sub = lane % N;
  int16_t* blk = smem[warp] + grp * TrBlk<LOG2>::ELEMS;
  int flags = (kind == B265_TR_DST)    ? B265_TU_DST
              : (kind == B265_TR_SKIP)   ? B265_TU_SKIP
              : (kind == B265_TR_BYPASS) ? B265_TU_BYPASS : 0;
  if (rdpcm == 1) flags |= B265_TU_RDPCM_H;
  if (rdpcm == 2) flags |= B265_TU_RDPCM_V;

  const int warps_total = gridDim.x * B265_WARPS_PER_CTA;
  for (long long base = (long long)(blockIdx.x * B265_WARPS_PER_CTA + warp) * G; base < n_blocks;
       base += (long long)warps_total * G) {
    const long long i = base + grp;
    const bool active = i < n_blocks;
    int lastx = 0, anyrow = 0;
    if (active) {
      // row `sub` of block i
      const int16_t* src = coeffs + i * (N * N) + sub * N;
      const int q = qp ? qp[i] : -1;
      uint32_t* drow = reinterpret_cast<uint32_t*>(blk + sub * S);
#pragma unroll
      for (int x = 0; x < N; x += 8) {
        int v[8];
        if (N == 4) {
          uint2 w = *reinterpret_cast<const uint2*>(src);
          v[0] = (int16_t)(w.x & 0xffff); v[1] = (int16_t)(w.x >> 16);
          v[2] = (int16_t)(w.y & 0xffff); v[3] = (int16_t)(w.y >> 16);
          v[4] = v[5] = v[6] = v[7] = 0;
        } else {
          uint4 w = *reinterpret_cast<const uint4*>(src + x);
          v[0] = (int16_t)(w.x & 0xffff); v[1] = (int16_t)(w.x >> 16);
          v[2] = (int16_t)(w.y & 0xffff); v[3] = (int16_t)(w.y >> 16);
          v[4] = (int16_t)(w.z & 0xffff); v[5] = (int16_t)(w.z >> 16);
          v[6] = (int16_t)(w.w & 0xffff); v[7] = (int16_t)(w.w >> 16);
        }
#pragma unroll
        for (int k = 0; k < (N == 4 ? 4 : 8); k++) {
          if (q >= 0) v[k] = dequant_flat(v[k], q, LOG2);
          if (v[k] != 0) lastx = x + k + 1;
        }
#pragma unroll
        for (int k = 0; k < (N == 4 ? 4 : 8); k += 2)
          drow[(x + k) >> 1] = (uint32_t)(v[k] & 0xffff) | ((uint32_t)v[k + 1] << 16);
      }
      anyrow = lastx ? sub + 1 : 0;
    }
    int ncols = group_max<N>(lastx);
    int nrows = group_max<N>(anyrow);
    int pred[N];
    uint8_t* d = dst + (active ? i : 0) * (N * N);
    if (active) load_row_u8<N>(d + sub * N, pred);
    residual_add_store<LOG2>(blk, sub, flags, nrows | (ncols << 8), active, pred, d, N);
  }
}

// ---------------------------------------------------------------------------------------------
// Sparse work-list form. order[] lists TUs (frame<<24 | index) of ONE size; the prediction is read
// from the destination plane (written by the MC kernel of an earlier launch).
// ---------------------------------------------------------------------------------------------
template <int LOG2, bool SCRATCH>
__global__ void __launch_bounds__(B265_WARPS_PER_CTA * 32)
k_residual_sparse(const FrameDev* __restrict__ frames, const uint32_t* __restrict__ order, int n) {
  constexpr int N = TrBlk<LOG2>::N, G = 32 / N;
  __shared__ __align__(16) int16_t smem[B265_WARPS_PER_CTA][B265_TR_SMEM_PER_WARP / 2];
  const int warp = threadIdx.x >> 5, lane = lane_id();
  const int grp = lane / N, sub = lane % N;
  int16_t* blk = smem[warp] + grp * TrBlk<LOG2>::ELEMS;
  const int warps_total = gridDim.x * B265_WARPS_PER_CTA;
  for (int base = (blockIdx.x * B265_WARPS_PER_CTA + warp) * G; base < n; base += warps_total * G) {
    const int i = base + grp;
    const bool active = i < n;
    __syncwarp();
    zero_warp_scratch(smem[warp]);
    __syncwarp();
    b265_tu tu = {};
    const FrameDev* f = frames;
    if (active) {
      uint32_t o = order[i];
      f = frames + (o >> 24);
      tu = f->tus[o & 0xffffff];
    }
    const int cidx = tu.cidx_log2 & 3;
    int bounds = scatter_coeffs<LOG2>(blk, sub, f->coeffs + tu.coeff_off, active ? tu.ncoeff : 0, tu.qp,
                                      tu.flags, cidx, (f->flags & B265_FRAME_SCALING_LIST) ? f->scaling : nullptr);
    if (SCRATCH) {
      int16_t* d = nullptr;
      int pitch = 0;
      if (active) {   // block-linear scratch (common.cuh resid_block): the TB's lines are N apart
        pitch = N;
        d = f->resid[cidx] + resid_block(tu.x, tu.y, f->log2_ctb - ((cidx && f->cw) ? 1 : 0), f->ctbs_w);
      }
      residual_store_scratch<LOG2>(blk, sub, tu.flags, bounds, active, d, pitch, tu.intra_mode >= 2 && tu.intra_mode < 18);
    } else {
      int pred[N];
      uint8_t* d = nullptr;
      int pitch = 0;
      if (active) {
        pitch = f->pitch[cidx];
        d = f->plane[cidx] + (size_t)tu.y * pitch + tu.x;
        load_row_u8<N>(d + sub * pitch, pred);
      }
      residual_add_store<LOG2>(blk, sub, tu.flags, bounds, active, pred, d, pitch);
    }
  }
}

void launch_residual_dense(cudaStream_t st, int log2, int kind, int rdpcm, int n_blocks,
                           const int16_t* coeffs, const uint8_t* qp, uint8_t* dst, int sm_count) {
  if (n_blocks <= 0) return;
  // DCT / DST: the packed dp2a kernels (k_residual_dense.cu); skip / bypass / rdpcm: the generic kernel below
  if (!rdpcm && launch_residual_dense_dct(st, log2, kind, n_blocks, coeffs, qp, dst, sm_count)) return;
  const int G = 32 >> log2;
  long long warps = ((long long)n_blocks + G - 1) / G;
  long long ctas = (warps + B265_WARPS_PER_CTA - 1) / B265_WARPS_PER_CTA;
  int grid = (int)min(ctas, (long long)sm_count * 8);
  dim3 b(B265_WARPS_PER_CTA * 32);
  switch (log2) {
    case 2: k_residual_dense<2><<<grid, b, 0, st>>>(coeffs, qp, dst, n_blocks, kind, rdpcm); break;
    case 3: k_residual_dense<3><<<grid, b, 0, st>>>(coeffs, qp, dst, n_blocks, kind, rdpcm); break;
    case 4: k_residual_dense<4><<<grid, b, 0, st>>>(coeffs, qp, dst, n_blocks, kind, rdpcm); break;
    default: k_residual_dense<5><<<grid, b, 0, st>>>(coeffs, qp, dst, n_blocks, kind, rdpcm); break;
  }
}

void launch_residual_sparse(cudaStream_t st, int log2, const FrameDev* frames, const uint32_t* order,
                            int n, int sm_count, bool to_scratch) {
  if (n <= 0) return;
  const int G = 32 >> log2;
  int warps = (n + G - 1) / G;
  int ctas = (warps + B265_WARPS_PER_CTA - 1) / B265_WARPS_PER_CTA;
  int grid = min(ctas, sm_count * 8);
  dim3 b(B265_WARPS_PER_CTA * 32);
  switch (log2) {
    case 2: to_scratch ? k_residual_sparse<2, true><<<grid, b, 0, st>>>(frames, order, n)
                       : k_residual_sparse<2, false><<<grid, b, 0, st>>>(frames, order, n); break;
    case 3: to_scratch ? k_residual_sparse<3, true><<<grid, b, 0, st>>>(frames, order, n)
                       : k_residual_sparse<3, false><<<grid, b, 0, st>>>(frames, order, n); break;
    case 4: to_scratch ? k_residual_sparse<4, true><<<grid, b, 0, st>>>(frames, order, n)
                       : k_residual_sparse<4, false><<<grid, b, 0, st>>>(frames, order, n); break;
    default: to_scratch ? k_residual_sparse<5, true><<<grid, b, 0, st>>>(frames, order, n)
                        : k_residual_sparse<5, false><<<grid, b, 0, st>>>(frames, order, n); break;
  }
}
