// residual.cuh — loading a transform block into shared memory (dense or sparse work-list form),
// and adding the residual rows to a prediction.  Used by k_residual.cu and k_intra.cu.
#pragma once
#include "transform.cuh"

template <int N>
__device__ __forceinline__ int group_max(int v) {
#pragma unroll
  for (int o = N / 2; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(B265_FULL_MASK, v, o));
  return v;
}

// Zero the warp's transform scratch (all 32 lanes).
__device__ __forceinline__ void zero_warp_scratch(int16_t* wblk) {
  uint4* p = reinterpret_cast<uint4*>(wblk);
  const uint4 z = make_uint4(0, 0, 0, 0);
#pragma unroll
  for (int i = lane_id(); i < B265_TR_SMEM_PER_WARP / 16; i += 32) p[i] = z;
}

// Scatter + dequantise the sparse coefficients of one TU into the group's block.
// Returns (nrows | ncols << 8) bounds of the non-zero region for this group (valid in every lane
// of the group).  Must be called by all lanes of the warp (shuffles); inactive groups pass ncoeff=0.
template <int LOG2>
__device__ __forceinline__ int scatter_coeffs(int16_t* blk, int sub, const b265_coeff* __restrict__ c,
                                              int ncoeff, int qp, int flags, int cidx,
                                              const b265_scaling* __restrict__ sl) {
  constexpr int N = TrBlk<LOG2>::N, S = TrBlk<LOG2>::S;
  int mx = -1, my = -1;
  for (int e = sub; e < ncoeff; e += N) {
    b265_coeff k = c[e];
    int pos = k.pos;
    if ((flags & B265_TU_ROTATE) && (flags & (B265_TU_SKIP | B265_TU_BYPASS))) pos = N * N - 1 - pos;
    int x = pos & (N - 1), y = pos >> LOG2;
    int v;
    if (flags & B265_TU_BYPASS) v = k.level;
    else if (sl) {
      int mid = cidx + ((flags & B265_TU_INTRA_CU) ? 0 : (N < 32 ? 3 : 1));
      const uint8_t* m = (LOG2 == 2) ? sl->s4[mid] : (LOG2 == 3) ? sl->s8[mid]
                       : (LOG2 == 4) ? sl->s16[mid] : sl->s32[mid];
      // the list is indexed with the (un-rotated) coefficient position (transform.cc:497-503)
      v = dequant_list(k.level, qp, LOG2, m[k.pos]);
    } else v = dequant_flat(k.level, qp, LOG2);
    blk[y * S + x] = (int16_t)v;
    mx = max(mx, x);
    my = max(my, y);
  }
  mx = group_max<N>(mx);
  my = group_max<N>(my);
  return (my + 1) | ((mx + 1) << 8);
}

// dst row = Clip1(pred row + residual row); vertical RDPCM accumulates down the columns first.
// pred[] holds row `sub` of the prediction; dst points at the block's top-left sample.
template <int LOG2>
__device__ __forceinline__ void residual_add_store(int16_t* blk, int sub, int flags, int bounds,
                                                   bool active, int (&pred)[1 << LOG2],
                                                   uint8_t* dst, int pitch) {
  constexpr int N = TrBlk<LOG2>::N, S = TrBlk<LOG2>::S;
  const bool vert = active && (flags & B265_TU_RDPCM_V) && (flags & (B265_TU_SKIP | B265_TU_BYPASS));
  if (__any_sync(B265_FULL_MASK, vert)) {
    // rare (range-extension) path: turn the column-wise running sum into plain residuals in place.
    // The sums can exceed 16 bit, so the accumulated values are applied directly column by column.
    if (vert) store_row_u8<N>(dst + sub * pitch, pred);
    __syncwarp();
    if (vert) {
      int acc = 0;
      for (int y = 0; y < N; y++) {
        int c = blk[y * S + sub];
        int v = (flags & B265_TU_BYPASS) ? c : (((c << (5 + LOG2)) + 2048) >> 12);
        acc += v;
        uint8_t* p = dst + y * pitch + sub;
        *p = (uint8_t)clip8((int)*p + acc);
      }
    }
    __syncwarp();
    active = active && !vert;
  }
  int r[N];
  residual_rows<LOG2>(blk, sub, flags, bounds & 0xff, bounds >> 8, active, r);
  if (active) {
#pragma unroll
    for (int x = 0; x < N; x++) pred[x] = clip8(pred[x] + r[x]);
    store_row_u8<N>(dst + sub * pitch, pred);
  }
}

// Scratch form for intra TBs: the residual rows go to an int16 plane (row pitch in samples) and are
// added to the prediction later by the intra wavefront.  Values are clamped to +-1024, which is exact:
// Clip1(pred + r) only depends on r clamped to [-255, 255].  `transposed` stores element (row y, col x) of
// the block at dst[x * pitch + y] (the wavefront predicts modes 2..17 column-wise, intra.cuh).
__device__ __forceinline__ int clip_res(int v) { return min(max(v, -1024), 1023); }

template <int LOG2>
__device__ __forceinline__ void residual_store_scratch(int16_t* blk, int sub, int flags, int bounds, bool active,
                                                       int16_t* dst, int pitch, bool transposed) {
  constexpr int N = TrBlk<LOG2>::N, S = TrBlk<LOG2>::S;
  const bool vert = active && (flags & B265_TU_RDPCM_V) && (flags & (B265_TU_SKIP | B265_TU_BYPASS));
  if (__any_sync(B265_FULL_MASK, vert)) {
    if (vert) {  // lane `sub` accumulates down column `sub`
      int acc = 0;
      for (int y = 0; y < N; y++) {
        int c = blk[y * S + sub];
        int v = (flags & B265_TU_BYPASS) ? c : (((c << (5 + LOG2)) + 2048) >> 12);
        acc += v;
        dst[transposed ? (size_t)sub * pitch + y : (size_t)y * pitch + sub] = (int16_t)clip_res(acc);
      }
    }
    __syncwarp();
    active = active && !vert;
  }
  int r[N];
  residual_rows<LOG2>(blk, sub, flags, bounds & 0xff, bounds >> 8, active, r);
  if (active && transposed) {
#pragma unroll
    for (int x = 0; x < N; x++) dst[(size_t)x * pitch + sub] = (int16_t)clip_res(r[x]);
  } else if (active) {
    uint32_t* d = reinterpret_cast<uint32_t*>(dst + (size_t)sub * pitch);
    if (N == 4) {
      *reinterpret_cast<uint2*>(d) = make_uint2((uint32_t)(clip_res(r[0]) & 0xffff) | ((uint32_t)clip_res(r[1]) << 16),
                                                (uint32_t)(clip_res(r[2]) & 0xffff) | ((uint32_t)clip_res(r[3]) << 16));
    } else {
#pragma unroll
      for (int q = 0; q < N / 8; q++) {
        uint32_t w[4];
#pragma unroll
        for (int k = 0; k < 4; k++)
          w[k] = (uint32_t)(clip_res(r[8 * q + 2 * k]) & 0xffff) | ((uint32_t)clip_res(r[8 * q + 2 * k + 1]) << 16);
        reinterpret_cast<uint4*>(d)[q] = make_uint4(w[0], w[1], w[2], w[3]);
      }
    }
  }
}
