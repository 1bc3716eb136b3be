// k_intra.cu — intra prediction + residual of all intra TBs of a batch of pictures as ONE persistent
// launch: the TU-dependency wavefront.
//
// Host side (frames.cu) sorts the intra TBs by (wave, size) where wave(TB) = 1 + max wave of the TBs
// that produce its reference samples, and cuts the sorted list into groups of <= 32/N TBs of equal
// size and wave.  Warps draw groups in that order from an atomic ticket counter.  Because every
// dependency of a group has a smaller ticket and tickets are only held by running warps, spinning
// on the `done` flags cannot deadlock as long as the grid is resident (it is sized that way).
//
// Per TB (group of N lanes): gather border -> filter -> predict row in registers -> scatter+dequant
// coefficients -> IDCT passes -> add -> ONE store of the finished row -> publish rim flags.
// HBM traffic per TB: N*N bytes written + 4 B per non-zero coefficient read; borders come from L2.
#include "intra.cuh"
#include "residual.cuh"
#include "kernels.h"

template <int LOG2>
__device__ __forceinline__ void intra_group(const FrameDev* __restrict__ frames, const uint32_t* __restrict__ order,
                                            uint32_t first, int count, int16_t* wtr, uint8_t* wborder) {
  constexpr int N = 1 << LOG2, NU = N / 4, STRIDE = IntraBuf<N>::STRIDE;
  const int lane = lane_id(), grp = lane / N, sub = lane % N;
  const bool active = grp < count;
  const FrameDev* f = frames;
  b265_tu tu = {};
  if (active) {
    const uint32_t o = order[first + grp];
    f = frames + (o >> 24);
    tu = f->tus[o & 0xffffff];
  }
  const int cidx = tu.cidx_log2 & 3;
  uint8_t* buf0 = wborder + grp * 2 * STRIDE;
  uint8_t* buf1 = buf0 + STRIDE;
  gather_border<LOG2>(*f, cidx, tu.x, tu.y, sub, grp, active, buf0);
  const uint8_t* pb = filter_border<LOG2>(buf0, buf1, sub, tu.intra_mode, cidx,
                                          (f->flags & B265_FRAME_STRONG_INTRA_SMOOTHING) != 0, active);
  int pred[N];
  predict_row<LOG2>(pb, sub, active ? tu.intra_mode : 1, cidx, !(tu.flags & B265_TU_NO_EDGE_FILT), pred);

  const int pitch = f->pitch[cidx];
  uint8_t* d = active ? f->plane[cidx] + (size_t)tu.y * pitch + tu.x : nullptr;
  const bool has_res = active && tu.ncoeff > 0;
  if (__any_sync(B265_FULL_MASK, has_res)) {
    int16_t* blk = wtr + grp * TrBlk<LOG2>::ELEMS;
    __syncwarp();
    zero_warp_scratch(wtr);
    __syncwarp();
    const int bounds = scatter_coeffs<LOG2>(blk, sub, f->coeffs + tu.coeff_off, has_res ? tu.ncoeff : 0, tu.qp, tu.flags,
                                            cidx, (f->flags & B265_FRAME_SCALING_LIST) ? f->scaling : nullptr);
    if (active && !has_res) store_row_u8<N>(d + sub * pitch, pred);
    residual_add_store<LOG2>(blk, sub, tu.flags, bounds, has_res, pred, d, pitch);
  } else if (active) {
    store_row_u8<N>(d + sub * pitch, pred);
  }
  // publish: rim units of this TB are final (pre-deblock) samples now
  __syncwarp();
  if (active && sub < 2 * NU) {
    __threadfence();
    const int ux = (sub < NU) ? (tu.x >> 2) + sub : ((tu.x + N) >> 2) - 1;
    const int uy = (sub < NU) ? ((tu.y + N) >> 2) - 1 : (tu.y >> 2) + (sub - NU);
    st_release_u8(f->done[cidx] + uy * f->done_pitch[cidx] + ux, 1);
  }
}

struct IntraSmem {
  __align__(16) int16_t tr[B265_WARPS_PER_CTA][B265_TR_SMEM_PER_WARP / 2];
  __align__(16) uint8_t border[B265_WARPS_PER_CTA][B265_BORDER_BYTES];
};

__global__ void __launch_bounds__(B265_WARPS_PER_CTA * 32)
k_intra_wavefront(const FrameDev* __restrict__ frames, const uint32_t* __restrict__ order,
                  const TuGroup* __restrict__ groups, int n_groups, unsigned* __restrict__ ticket) {
  __shared__ IntraSmem sm;
  const int warp = threadIdx.x >> 5;
  for (;;) {
    unsigned t = 0;
    if (lane_id() == 0) t = atomicAdd(ticket, 1u);
    t = __shfl_sync(B265_FULL_MASK, t, 0);
    if (t >= (unsigned)n_groups) break;
    const TuGroup g = groups[t];
    switch (g.log2) {
      case 2: intra_group<2>(frames, order, g.first, g.count, sm.tr[warp], sm.border[warp]); break;
      case 3: intra_group<3>(frames, order, g.first, g.count, sm.tr[warp], sm.border[warp]); break;
      case 4: intra_group<4>(frames, order, g.first, g.count, sm.tr[warp], sm.border[warp]); break;
      default: intra_group<5>(frames, order, g.first, g.count, sm.tr[warp], sm.border[warp]); break;
    }
    __syncwarp();
  }
}

int intra_max_resident_ctas(int sm_count) {
  int per_sm = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_intra_wavefront, B265_WARPS_PER_CTA * 32, 0);
  if (per_sm < 1) per_sm = 1;
  return per_sm * sm_count;
}

void launch_intra(cudaStream_t st, const FrameDev* frames, const uint32_t* order, const TuGroup* groups, int n_groups,
                  unsigned* ticket, int max_ctas) {
  if (n_groups <= 0) return;
  int ctas = (n_groups + B265_WARPS_PER_CTA - 1) / B265_WARPS_PER_CTA;
  if (ctas > max_ctas) ctas = max_ctas;
  cudaMemsetAsync(ticket, 0, sizeof(unsigned), st);
  k_intra_wavefront<<<ctas, B265_WARPS_PER_CTA * 32, 0, st>>>(frames, order, groups, n_groups, ticket);
}
