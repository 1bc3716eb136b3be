// k_intra.cu — intra prediction + residual of all intra TBs of a batch of pictures as ONE persistent
// launch: a CTB-level wavefront.
//
// Intra TBs form dependency chains in z-scan order inside a CTB (every TB reads the reconstructed
// samples left of / above it), so a CTB plane is walked SEQUENTIALLY by one warp — the hand-over
// between consecutive TBs is a warp-level barrier, not a trip through global flags — and
// parallelism comes from (a) the three colour planes, (b) the CTB anti-diagonal
// wave = ctbX + 2*ctbY (a CTB needs its left, upper-left, upper and upper-right neighbours,
// intrapred.cc:597-687), and (c) the pictures of the batch, which are mutually independent.
//
// Host side (frames.cu) emits one ticket per (picture, CTB, plane) that contains intra TBs, sorted by
// wave.  Warps draw tickets from an atomic counter; a warp waits for the (at most four) neighbouring
// CTB flags of its plane with acquire loads, processes the CTB's TBs in decoding order, then publishes
// its own flag with a release store.  Every dependency has a smaller ticket, tickets are only held by
// running warps, so spinning cannot deadlock for any grid size.
//
// Per TB (N lanes of the warp own one row each): gather border -> substitute -> filter -> predict row
// in registers -> scatter+dequant coefficients -> IDCT passes -> add -> ONE store of the finished row.
// HBM traffic per TB: N*N bytes written + 4 B per non-zero coefficient read; borders come from L2.
#include <stdlib.h>
#include "intra.cuh"
#include "residual.cuh"
#include "kernels.h"

#define INTRA_WARPS 4   // one warp per SM sub-partition; several CTAs per SM when the batch is large

template <int LOG2>
__device__ __forceinline__ void intra_tb(const FrameDev* __restrict__ f, const b265_tu tu, int16_t* wtr, uint8_t* wborder) {
  constexpr int N = 1 << LOG2, STRIDE = IntraBuf<N>::STRIDE;
  const int lane = lane_id(), grp = lane / N, sub = lane % N;
  const bool active = grp == 0;
  const int cidx = tu.cidx_log2 & 3;
  uint8_t* buf0 = wborder;
  uint8_t* buf1 = buf0 + STRIDE;
  gather_border<LOG2>(*f, cidx, tu.x, tu.y, sub, grp, active, buf0);
  const uint8_t* pb = filter_border<LOG2>(buf0, buf1, sub, tu.intra_mode, cidx,
                                          (f->flags & B265_FRAME_STRONG_INTRA_SMOOTHING) != 0, active);
  int pred[N];
  predict_row<LOG2>(pb, sub, active ? tu.intra_mode : 1, cidx, !(tu.flags & B265_TU_NO_EDGE_FILT), pred);

  const int pitch = f->pitch[cidx];
  uint8_t* d = f->plane[cidx] + (size_t)tu.y * pitch + tu.x;
  if (tu.ncoeff > 0) {  // warp-uniform
    int16_t* blk = wtr;
    __syncwarp();
    zero_warp_scratch(wtr);
    __syncwarp();
    const int bounds = scatter_coeffs<LOG2>(blk, sub, f->coeffs + tu.coeff_off, active ? tu.ncoeff : 0, tu.qp, tu.flags,
                                            cidx, (f->flags & B265_FRAME_SCALING_LIST) ? f->scaling : nullptr);
    residual_add_store<LOG2>(blk, sub, tu.flags, bounds, active, pred, d, pitch);
  } else if (active) {
    store_row_u8<N>(d + sub * pitch, pred);
  }
  // the next TB of this CTB (same warp) may read these samples through L2
  __threadfence_block();
  __syncwarp();
}

struct IntraSmem {
  __align__(16) int16_t tr[INTRA_WARPS][B265_TR_SMEM_PER_WARP / 2];
  __align__(16) uint8_t border[INTRA_WARPS][B265_BORDER_BYTES];
};

__device__ __forceinline__ void wait_flag(const uint8_t* flag, unsigned sleep_ns) {
  // relaxed polling (no L1 invalidation per iteration); the caller issues one acquire fence afterwards
  while (ld_relaxed_u8(flag) == 0) {
    if (sleep_ns) __nanosleep(sleep_ns);
  }
}

__global__ void __launch_bounds__(INTRA_WARPS * 32, 4)
k_intra_wavefront(const FrameDev* __restrict__ frames, const IntraTicket* __restrict__ tickets, int n_tickets,
                  unsigned* __restrict__ ticket, unsigned sleep_ns) {
  __shared__ IntraSmem sm;
  const int warp = threadIdx.x >> 5, lane = lane_id();
  for (;;) {
    unsigned t = 0;
    if (lane == 0) t = atomicAdd(ticket, 1u);
    t = __shfl_sync(B265_FULL_MASK, t, 0);
    if (t >= (unsigned)n_tickets) break;
    const IntraTicket tk = tickets[t];
    const FrameDev* f = frames + tk.frame;
    const int c = tk.plane;
    const int cx = tk.ctb % f->ctbs_w, cy = tk.ctb / f->ctbs_w;
    // prefetch the first 32 TU records of the CTB while waiting for the neighbours
    const uint4* tus4 = reinterpret_cast<const uint4*>(f->tus);  // 16-byte records, 256-byte aligned list
    uint4 mine = make_uint4(0, 0xff000000u, 0, 0);
    if ((uint32_t)lane < tk.tu_count) mine = tus4[tk.tu_first + lane];
    // neighbours: left, upper-left, upper, upper-right
    if (lane < 4) {
      const int nx = cx + (lane == 3 ? 1 : (lane == 2 ? 0 : -1)), ny = cy - (lane == 0 ? 0 : 1);
      if (nx >= 0 && ny >= 0 && nx < f->ctbs_w) wait_flag(f->done[c] + ny * f->ctbs_w + nx, sleep_ns);
    }
    fence_acq_rel_gpu();
    __syncwarp();
    for (uint32_t base = 0; base < tk.tu_count; base += 32) {
      if (base) {
        mine = make_uint4(0, 0xff000000u, 0, 0);
        if (base + lane < tk.tu_count) mine = tus4[tk.tu_first + base + lane];
      }
      const int n_here = min(32u, tk.tu_count - base);
      // lanes whose TU is an intra TB of this plane
      const bool want = lane < n_here && (mine.y >> 24) != B265_INTRA_NONE && (int)(mine.y & 3) == c;
      unsigned todo = __ballot_sync(B265_FULL_MASK, want);
      while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        b265_tu tu;
        {
          const uint32_t r0 = __shfl_sync(B265_FULL_MASK, mine.x, src), r1 = __shfl_sync(B265_FULL_MASK, mine.y, src);
          tu.x = (uint16_t)(r0 & 0xffff);
          tu.y = (uint16_t)(r0 >> 16);
          tu.cidx_log2 = (uint8_t)(r1 & 0xff);
          tu.flags = (uint8_t)((r1 >> 8) & 0xff);
          tu.qp = (uint8_t)((r1 >> 16) & 0xff);
          tu.intra_mode = (uint8_t)(r1 >> 24);
          tu.coeff_off = __shfl_sync(B265_FULL_MASK, mine.z, src);
          tu.ncoeff = (uint16_t)(__shfl_sync(B265_FULL_MASK, mine.w, src) & 0xffff);
          tu.reserved = 0;
        }
        switch ((tu.cidx_log2 >> 2) & 7) {
          case 2: intra_tb<2>(f, tu, sm.tr[warp], sm.border[warp]); break;
          case 3: intra_tb<3>(f, tu, sm.tr[warp], sm.border[warp]); break;
          case 4: intra_tb<4>(f, tu, sm.tr[warp], sm.border[warp]); break;
          default: intra_tb<5>(f, tu, sm.tr[warp], sm.border[warp]); break;
        }
      }
    }
    // publish: every sample of this CTB plane is final (pre-deblock) now
    __threadfence();
    __syncwarp();
    if (lane == 0) st_release_u8(f->done[c] + tk.ctb, 1);
    __syncwarp();
  }
}

int intra_max_resident_ctas(int sm_count) {
  int per_sm = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_intra_wavefront, INTRA_WARPS * 32, 0);
  if (per_sm < 1) per_sm = 1;
  return per_sm * sm_count;
}

void launch_intra(cudaStream_t st, const FrameDev* frames, const IntraTicket* tickets, int n_tickets, unsigned* ticket,
                  int max_ctas, int wave_width) {
  if (n_tickets <= 0) return;
  // warps beyond ~3 waves' worth would only poll; keep them off the SMs the working warps run on
  int want = 3 * wave_width;
  if (want < 296) want = 296;
  if (want > n_tickets) want = n_tickets;
  int ctas = (want + INTRA_WARPS - 1) / INTRA_WARPS;
  if (ctas > max_ctas) ctas = max_ctas;
  cudaMemsetAsync(ticket, 0, sizeof(unsigned), st);
  static int sleep_ns = -1;
  if (sleep_ns < 0) {
    const char* e = getenv("B265_INTRA_SLEEP_NS");
    sleep_ns = e ? atoi(e) : 100;
  }
  k_intra_wavefront<<<ctas, INTRA_WARPS * 32, 0, st>>>(frames, tickets, n_tickets, ticket, (unsigned)sleep_ns);
}
