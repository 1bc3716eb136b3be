// k_residual_dense.cu — dense batch form of dequant + inverse DCT/DST + add (acceleration.h transform slots
// `transform_add_8[4]` / `transform_4x4_dst_add_8`, acceleration.h:152-153; BASELINE config 2).
// Block i has N*N int16 levels at coeffs + i*N*N and N*N pixels at dst + i*N*N; HBM traffic per block is
// 2*N*N (levels) + N*N (prediction) + N*N (reconstruction) = 4*N*N bytes.
//
// The kernels are issue-bound integer code, so the design minimises instructions per sample:
//  * both 1-D passes run on packed int16 pairs with dp2a (transform_packed.cuh);
//  * the clip to 16 bit after dequantisation / after the first pass is the saturation of the pack instruction;
//  * the final Clip1 + byte packing is cvt.pack.sat.u8;
//  * 4x4: one lane owns a whole block in registers (no shared memory, no shuffles);
//  * 8..32: N lanes per block.  Phase A: each lane loads two ROWS (ra, rb) of a pair word for a chunk of columns,
//    dequantises and writes VERTICAL pairs {x[ra][c], x[rb][c]} to shared memory.  Phase B: lane = column, reads its
//    N/2 pair words, first pass, writes the clipped int16 results to the slot of the row pass' pair order.
//    Phase C: lane = row, reads N/2 pair words with 128-bit loads, second pass, adds the prediction row, stores.
//    Zero-tail bounds (rows for B, columns for C) are reduced over the warp and select a shorter butterfly.
#include "transform_packed.cuh"
#include "kernels.h"

#define DENSE_WARPS 8

// four reconstructed samples: Clip1(pred byte + (r >> 12)), packed
__device__ __forceinline__ uint32_t add_pack4(const int* r, uint32_t pw) {
  return pack_sat8((r[0] >> 12) + byte_of<0>(pw), (r[1] >> 12) + byte_of<1>(pw),
                   (r[2] >> 12) + byte_of<2>(pw), (r[3] >> 12) + byte_of<3>(pw));
}

// ------------------------------------------------------------------------------------------------------
// 4x4: one lane per block
// ------------------------------------------------------------------------------------------------------
template <bool DST>
__global__ void __launch_bounds__(256)
k_dense4(const int16_t* __restrict__ coeffs, const uint8_t* __restrict__ qp, uint8_t* __restrict__ dst, int n_blocks) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_blocks; i += stride) {
    const uint4* src = reinterpret_cast<const uint4*>(coeffs + i * 16);
    const uint4 w0 = src[0], w1 = src[1];   // rows 0,1 / rows 2,3
    const uint4 pr = *reinterpret_cast<const uint4*>(dst + i * 16);
    const DequantK k = dequant_consts(qp ? (int)qp[i] : -1, 2);
    const uint32_t r0[2] = {w0.x, w0.y}, r1[2] = {w0.z, w0.w}, r2[2] = {w1.x, w1.y}, r3[2] = {w1.z, w1.w};
    // vertical pairs per column: P0 = (row1,row3), P1 = (row0,row2)
    uint32_t PA[4], PB[4];
#pragma unroll
    for (int h = 0; h < 2; h++) {
      PB[2 * h] = pack_sat16(dq_lo(r0[h], k), dq_lo(r2[h], k));
      PB[2 * h + 1] = pack_sat16(dq_hi(r0[h], k), dq_hi(r2[h], k));
      PA[2 * h] = pack_sat16(dq_lo(r1[h], k), dq_lo(r3[h], k));
      PA[2 * h + 1] = pack_sat16(dq_hi(r1[h], k), dq_hi(r3[h], k));
    }
    int g[4][4];   // g[c][row]
#pragma unroll
    for (int c = 0; c < 4; c++) {
      const uint32_t P[2] = {PA[c], PB[c]};
      if (DST) idst4_packed(P, g[c], 64);
      else IdctP<4, 4>::run(P, g[c], 64);
    }
    uint32_t o[4];
    const uint32_t pw[4] = {pr.x, pr.y, pr.z, pr.w};
#pragma unroll
    for (int y = 0; y < 4; y++) {
      const uint32_t P[2] = {pack_sat16(sra<7>(g[1][y]), sra<7>(g[3][y])), pack_sat16(sra<7>(g[0][y]), sra<7>(g[2][y]))};
      int r[4];
      if (DST) idst4_packed(P, r, 2048);
      else IdctP<4, 4>::run(P, r, 2048);
      o[y] = add_pack4(r, pw[y]);
    }
    *reinterpret_cast<uint4*>(dst + i * 16) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// ------------------------------------------------------------------------------------------------------
// 8x8: one lane per block as well (64 levels = 32 registers).  Columns are transformed in the pair order of the
// row pass, two at a time, so the clipped first-pass results are packed straight into the row pass' pair words.
// ------------------------------------------------------------------------------------------------------
#ifndef DENSE8_MIN_CTAS
#define DENSE8_MIN_CTAS 7   // 72 registers, 28 warps/SM (the kernel is latency-bound)
#endif
__global__ void __launch_bounds__(128, DENSE8_MIN_CTAS)
k_dense8(const int16_t* __restrict__ coeffs, const uint8_t* __restrict__ qp, uint8_t* __restrict__ dst, int n_blocks) {
  constexpr int KA[4] = {1, 5, 2, 0}, KB[4] = {3, 7, 6, 4};   // pair order of the 8-point transform
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_blocks; i += stride) {
    const uint4* src = reinterpret_cast<const uint4*>(coeffs + i * 64);
    uint32_t x[8][4];   // row r, word w = columns (2w, 2w+1)
#pragma unroll
    for (int r = 0; r < 8; r++) {
      const uint4 w = __ldg(src + r);
      x[r][0] = w.x; x[r][1] = w.y; x[r][2] = w.z; x[r][3] = w.w;
    }
    uint4* dp = reinterpret_cast<uint4*>(dst + i * 64);
    asm volatile("prefetch.global.L2 [%0];" ::"l"(dp));   // the prediction is loaded after the first pass (16 registers less)
    const DequantK k = dequant_consts(qp ? (int)qp[i] : -1, 3);
    uint32_t Q[4][8];
#pragma unroll
    for (int p = 0; p < 4; p++) {
      int o[2][8];
#pragma unroll
      for (int h = 0; h < 2; h++) {
        const int c = h ? KB[p] : KA[p];
        uint32_t P[4];
#pragma unroll
        for (int v = 0; v < 4; v++) {
          const uint32_t wa = x[KA[v]][c >> 1], wb = x[KB[v]][c >> 1];
          P[v] = (c & 1) ? pack_sat16(dq_hi(wa, k), dq_hi(wb, k)) : pack_sat16(dq_lo(wa, k), dq_lo(wb, k));
        }
        IdctP<8, 8>::run(P, o[h], 64);
      }
#pragma unroll
      for (int y = 0; y < 8; y++) Q[p][y] = pack_sat16(sra<7>(o[0][y]), sra<7>(o[1][y]));
    }
    uint32_t pw[16];   // row y = words 2y, 2y+1
#pragma unroll
    for (int q = 0; q < 4; q++) {
      const uint4 w = dp[q];
      pw[4 * q] = w.x; pw[4 * q + 1] = w.y; pw[4 * q + 2] = w.z; pw[4 * q + 3] = w.w;
    }
    uint32_t out[16];
#pragma unroll
    for (int y = 0; y < 8; y++) {
      const uint32_t P[4] = {Q[0][y], Q[1][y], Q[2][y], Q[3][y]};
      int r[8];
      IdctP<8, 8>::run(P, r, 2048);
      out[2 * y] = add_pack4(r, pw[2 * y]);
      out[2 * y + 1] = add_pack4(r + 4, pw[2 * y + 1]);
    }
#pragma unroll
    for (int q = 0; q < 4; q++) dp[q] = make_uint4(out[4 * q], out[4 * q + 1], out[4 * q + 2], out[4 * q + 3]);
  }
}

// ------------------------------------------------------------------------------------------------------
// 16x16, 32x32 (and 8x8 with -DDENSE8_SMEM)
// ------------------------------------------------------------------------------------------------------
template <int LOG2>
struct DenseCfg {
  static constexpr int N = 1 << LOG2;
  static constexpr int G = 32 / N;                       // blocks per warp
  static constexpr int CH = (N == 8) ? 4 : 8;            // columns per phase-A task
  static constexpr int KCH = N / CH;                     // chunks per row
  static constexpr int ROUNDS = (N / 2) * KCH / N;       // phase-A tasks per lane
  static constexpr int GSW = N / 2 + (N >= 16 ? 4 : 0);  // phase-C row stride in words (conflict-free 128-bit reads)
  // words per block: max(pair plane N*N/2, result plane N*GSW) + a bank offset between the blocks of a warp
  static constexpr int BLKW = (N == 8) ? 40 : (N == 16) ? 208 : 640;
  static constexpr int WARPW = G * BLKW;
};

// pair words of column `sub` whose first row lies below L (compile-time selection; the others are zero)
template <int N, int L, int P0 = 0>
struct LoadPairs {
  static __device__ __forceinline__ void run(uint32_t* P, const uint32_t* col) {
    constexpr bool used = pair_first(N, P0) < L;
    P[P0] = used ? col[P0 * N] : 0u;
    LoadPairs<N, L, P0 + 1>::run(P, col);
  }
};
template <int N, int L>
struct LoadPairs<N, L, N / 2> {
  static __device__ __forceinline__ void run(uint32_t*, const uint32_t*) {}
};

template <int LOG2, int L>
__device__ __forceinline__ void dense_cols(uint32_t* blk, int sub, int slot) {
  constexpr int N = 1 << LOG2, GSW = DenseCfg<LOG2>::GSW;
  uint32_t P[N / 2];
  LoadPairs<N, L>::run(P, blk + sub);
  int out[N];
  IdctP<N, L>::run(P, out, 64);
  __syncwarp();   // the result plane overlays the pair plane: every lane has read its words
  int16_t* g = reinterpret_cast<int16_t*>(blk) + slot;
#pragma unroll
  for (int y = 0; y < N; y++) g[y * (2 * GSW)] = (int16_t)pack_sat16(sra<7>(out[y]), 0);
}

template <int LOG2, int L>
__device__ __forceinline__ void dense_rows(const uint32_t* blk, int sub, const uint32_t* pred, uint8_t* drow, bool active) {
  constexpr int N = 1 << LOG2, GSW = DenseCfg<LOG2>::GSW;
  uint32_t P[N / 2];
  const uint4* row = reinterpret_cast<const uint4*>(blk + sub * GSW);
#pragma unroll
  for (int q = 0; q < N / 8; q++) {
    uint4 w = row[q];
    P[4 * q] = w.x; P[4 * q + 1] = w.y; P[4 * q + 2] = w.z; P[4 * q + 3] = w.w;
  }
  int out[N];
  IdctP<N, L>::run(P, out, 2048);
  uint32_t o[N / 4];
#pragma unroll
  for (int q = 0; q < N / 4; q++) {
    const uint32_t pw = pred[q];
    o[q] = add_pack4(out + 4 * q, pw);
  }
  if (active) {
    if (N == 8) *reinterpret_cast<uint2*>(drow) = make_uint2(o[0], o[1]);
    else {
#pragma unroll
      for (int q = 0; q < N / 16; q++) reinterpret_cast<uint4*>(drow)[q] = make_uint4(o[4 * q], o[4 * q + 1], o[4 * q + 2], o[4 * q + 3]);
    }
  }
}

template <int LOG2>
__global__ void __launch_bounds__(DENSE_WARPS * 32)
k_dense(const int16_t* __restrict__ coeffs, const uint8_t* __restrict__ qp, uint8_t* __restrict__ dst, int n_blocks) {
  using Cfg = DenseCfg<LOG2>;
  constexpr int N = Cfg::N, G = Cfg::G, CH = Cfg::CH, KCH = Cfg::KCH, ROUNDS = Cfg::ROUNDS;
  __shared__ __align__(16) uint32_t smem[DENSE_WARPS][Cfg::WARPW];
  const int warp = threadIdx.x >> 5, lane = lane_id();
  const int grp = lane / N, sub = lane % N;
  uint32_t* blk = smem[warp] + grp * Cfg::BLKW;
  const int slot = pair_slot(N, sub);   // where column `sub` goes in the row pass' pair order
  // phase-A tasks of this lane: pair word j (rows ra, rb), column chunk kc
  int t_rows[ROUNDS], t_j[ROUNDS], t_c0[ROUNDS];
#pragma unroll
  for (int r = 0; r < ROUNDS; r++) {
    const int t = sub + N * r;
    t_j[r] = t / KCH;
    t_c0[r] = (t % KCH) * CH;
    t_rows[r] = pair_rows(N, t_j[r]);
  }
  const int warps_total = gridDim.x * DENSE_WARPS;
  for (long long base = (long long)(blockIdx.x * DENSE_WARPS + warp) * G; base < n_blocks; base += (long long)warps_total * G) {
    const long long i = base + grp;
    const bool active = i < n_blocks;
    const long long ib = active ? i : n_blocks - 1;   // idle groups of the last iteration redo the last block (no store)
    // prediction row `sub` (needed last, requested first)
    uint32_t pred[N / 4];
    uint8_t* drow = dst + ib * (N * N) + sub * N;
    if (N == 8) {
      uint2 w = *reinterpret_cast<const uint2*>(drow);
      pred[0] = w.x; pred[1] = w.y;
    } else {
#pragma unroll
      for (int q = 0; q < N / 16; q++) {
        uint4 w = reinterpret_cast<const uint4*>(drow)[q];
        pred[4 * q] = w.x; pred[4 * q + 1] = w.y; pred[4 * q + 2] = w.z; pred[4 * q + 3] = w.w;
      }
    }
    const DequantK k = dequant_consts(qp ? (int)qp[ib] : -1, LOG2);
    // ---- phase A ----
    uint32_t wa[ROUNDS][CH / 2], wb[ROUNDS][CH / 2];
    const int16_t* cblk = coeffs + ib * (N * N);
#pragma unroll
    for (int r = 0; r < ROUNDS; r++) {
      const int ra = t_rows[r] & 0xff, rb = t_rows[r] >> 8;
      if (CH == 4) {
        uint2 a = *reinterpret_cast<const uint2*>(cblk + ra * N + t_c0[r]);
        uint2 b = *reinterpret_cast<const uint2*>(cblk + rb * N + t_c0[r]);
        wa[r][0] = a.x; wa[r][1] = a.y; wb[r][0] = b.x; wb[r][1] = b.y;
      } else {
        uint4 a = *reinterpret_cast<const uint4*>(cblk + ra * N + t_c0[r]);
        uint4 b = *reinterpret_cast<const uint4*>(cblk + rb * N + t_c0[r]);
        wa[r][0] = a.x; wa[r][1] = a.y; wa[r][2] = a.z; wa[r][3] = a.w;
        wb[r][0] = b.x; wb[r][1] = b.y; wb[r][2] = b.z; wb[r][3] = b.w;
      }
    }
    int nrows = 0, ncols = 0;
    __syncwarp();   // previous iteration's phase C has read the result plane
#pragma unroll
    for (int r = 0; r < ROUNDS; r++) {
      const int ra = t_rows[r] & 0xff, rb = t_rows[r] >> 8;
      uint32_t ora = 0, orb = 0;
      uint32_t v[CH];
#pragma unroll
      for (int h = 0; h < CH / 2; h++) {
        const uint32_t a = wa[r][h], b = wb[r][h];
        ora |= a; orb |= b;
        v[2 * h] = pack_sat16(dq_lo(a, k), dq_lo(b, k));
        v[2 * h + 1] = pack_sat16(dq_hi(a, k), dq_hi(b, k));
      }
      {  // column bound in steps of CH/2 columns (the pass variants are selected at 4 / 8 / 16)
        uint32_t hi = 0;
#pragma unroll
        for (int h = CH / 4; h < CH / 2; h++) hi |= wa[r][h] | wb[r][h];
        if (hi) ncols = max(ncols, t_c0[r] + CH);
        else if (ora | orb) ncols = max(ncols, t_c0[r] + CH / 2);
      }
      if (ora) nrows = max(nrows, ra + 1);
      if (orb) nrows = max(nrows, rb + 1);
      uint4* d = reinterpret_cast<uint4*>(blk + t_j[r] * N + t_c0[r]);
      d[0] = make_uint4(v[0], v[1], v[2], v[3]);
      if (CH == 8) d[1] = make_uint4(v[4], v[5], v[6], v[7]);
    }
    nrows = __reduce_max_sync(B265_FULL_MASK, nrows);
    ncols = __reduce_max_sync(B265_FULL_MASK, ncols);
    __syncwarp();
    // ---- phase B: first pass down the columns ----
    if (N >= 32 && nrows > 16) dense_cols<LOG2, (N >= 32 ? 32 : N)>(blk, sub, slot);
    else if (N >= 16 && nrows > 8) dense_cols<LOG2, (N >= 16 ? 16 : N)>(blk, sub, slot);
    else if (nrows > 4) dense_cols<LOG2, 8>(blk, sub, slot);
    else dense_cols<LOG2, 4>(blk, sub, slot);
    __syncwarp();
    // ---- phase C: second pass along the rows, add, store ----
    if (N >= 32 && ncols > 16) dense_rows<LOG2, (N >= 32 ? 32 : N)>(blk, sub, pred, drow, active);
    else if (N >= 16 && ncols > 8) dense_rows<LOG2, (N >= 16 ? 16 : N)>(blk, sub, pred, drow, active);
    else if (ncols > 4) dense_rows<LOG2, 8>(blk, sub, pred, drow, active);
    else dense_rows<LOG2, 4>(blk, sub, pred, drow, active);
  }
}

// returns false when (log2, kind) has no fast kernel (skip / bypass: the generic kernel in k_residual.cu)
bool launch_residual_dense_dct(cudaStream_t st, int log2, int kind, int n_blocks, const int16_t* coeffs,
                               const uint8_t* qp, uint8_t* dst, int sm_count) {
  if (kind != B265_TR_DCT && !(kind == B265_TR_DST && log2 == 2)) return false;
  if (log2 == 2) {
    long long ctas = ((long long)n_blocks + 255) / 256;
    int grid = (int)min(ctas, (long long)sm_count * 16);
    if (kind == B265_TR_DST) k_dense4<true><<<grid, 256, 0, st>>>(coeffs, qp, dst, n_blocks);
    else k_dense4<false><<<grid, 256, 0, st>>>(coeffs, qp, dst, n_blocks);
    return true;
  }
#ifndef DENSE8_SMEM
  if (log2 == 3) {
    long long ctas = ((long long)n_blocks + 127) / 128;
    k_dense8<<<(int)min(ctas, (long long)sm_count * 32), 128, 0, st>>>(coeffs, qp, dst, n_blocks);
    return true;
  }
#endif
  const int G = 32 >> log2;
  long long warps = ((long long)n_blocks + G - 1) / G;
  long long ctas = (warps + DENSE_WARPS - 1) / DENSE_WARPS;
  int grid = (int)min(ctas, (long long)sm_count * 16);
  dim3 b(DENSE_WARPS * 32);
  switch (log2) {
    case 3: k_dense<3><<<grid, b, 0, st>>>(coeffs, qp, dst, n_blocks); break;
    case 4: k_dense<4><<<grid, b, 0, st>>>(coeffs, qp, dst, n_blocks); break;
    default: k_dense<5><<<grid, b, 0, st>>>(coeffs, qp, dst, n_blocks); break;
  }
  return true;
}

// ------------------------------------------------------------------------------------------------------
// Sparse work-list form of the same arithmetic for 16x16 and 32x32 transform blocks (frame path, S2 =
// scale_coefficients, transform.cc:353-625): the non-zero levels {int16 level, uint16 pos} of a TU are dequantised
// and scattered straight into the vertical-pair plane, then the two passes of k_dense run on it.  Only plain DCT
// TUs come here (frames.cu puts transform-skip / bypass TUs and the 4x4 / 8x8 sizes on the generic kernel's list).
// SCRATCH: residual rows go to the picture's int16 plane (intra TBs, added by the wavefront) clamped to +-1024,
// stored transposed for the column-wise predicted modes 2..17; otherwise they are added to the prediction in place.
// ------------------------------------------------------------------------------------------------------
template <int LOG2, int L>
__device__ __forceinline__ void sparse_rows(const uint32_t* blk, int sub, int (&out)[1 << LOG2]) {
  constexpr int N = 1 << LOG2, GSW = DenseCfg<LOG2>::GSW;
  uint32_t P[N / 2];
  const uint4* row = reinterpret_cast<const uint4*>(blk + sub * GSW);
#pragma unroll
  for (int q = 0; q < N / 8; q++) {
    uint4 w = row[q];
    P[4 * q] = w.x; P[4 * q + 1] = w.y; P[4 * q + 2] = w.z; P[4 * q + 3] = w.w;
  }
  IdctP<N, L>::run(P, out, 2048);
}

template <int LOG2, bool SCRATCH>
__global__ void __launch_bounds__(DENSE_WARPS * 32)
k_residual_sparse_fast(const FrameDev* __restrict__ frames, const uint32_t* __restrict__ order, int n) {
  using Cfg = DenseCfg<LOG2>;
  constexpr int N = Cfg::N, G = Cfg::G;
  __shared__ __align__(16) uint32_t smem[DENSE_WARPS][Cfg::WARPW];
  __shared__ uint8_t slot_lut[32];   // row -> int16 slot of the column pass' pair order
  if (threadIdx.x < N) slot_lut[threadIdx.x] = (uint8_t)pair_slot(N, threadIdx.x);
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = lane_id();
  const int grp = lane / N, sub = lane % N;
  uint32_t* blk = smem[warp] + grp * Cfg::BLKW;
  const int slot = pair_slot(N, sub);
  const int stride = gridDim.x * DENSE_WARPS * G;
  // The chain list item -> frame -> TU record -> levels is three dependent memory round trips per TU.  It is walked two
  // iterations ahead: the item of iteration k+2 and the TU record of iteration k+1 are requested while TU k is computed
  // (SPARSE_PREFETCH: also the cache lines of its levels).
  int base = (blockIdx.x * DENSE_WARPS + warp) * G;
  auto item_at = [&](int b, bool& ok) -> uint32_t {
    ok = b < n && b + grp < n;
    return ok ? __ldg(order + b + grp) : 0u;
  };
  auto tu_at = [&](uint32_t o, bool ok, const FrameDev*& f) -> uint4 {
    f = frames + (o >> 24);
    return ok ? __ldg(reinterpret_cast<const uint4*>(f->tus + (o & 0xffffff))) : make_uint4(0, 0, 0, 0);
  };
  bool ok_next, ok_nn;
  uint32_t o_next = item_at(base, ok_next);
  const FrameDev* f_next;
  uint4 tu_next = tu_at(o_next, ok_next, f_next);
  uint32_t o_nn = item_at(base + stride, ok_nn);
  uint32_t kc_pref = 0;
  if (ok_next && sub < (int)(tu_next.w & 0xffff)) kc_pref = __ldg(reinterpret_cast<const uint32_t*>(f_next->coeffs + tu_next.z) + sub);
  for (; base < n; base += stride) {
    const bool active = ok_next;
    const FrameDev* f = f_next;
    b265_tu tu;
    tu.x = (uint16_t)(tu_next.x & 0xffff); tu.y = (uint16_t)(tu_next.x >> 16);
    tu.cidx_log2 = (uint8_t)tu_next.y; tu.flags = (uint8_t)(tu_next.y >> 8); tu.qp = (uint8_t)(tu_next.y >> 16);
    tu.intra_mode = (uint8_t)(tu_next.y >> 24);
    tu.coeff_off = tu_next.z;
    tu.ncoeff = (uint16_t)(tu_next.w & 0xffff);
    tu.reserved = 0;
    ok_next = ok_nn;
    o_next = o_nn;
    o_nn = item_at(base + 2 * stride, ok_nn);
    const int cidx = tu.cidx_log2 & 3;
    __syncwarp();   // the previous iteration's row pass has read the result plane
#pragma unroll
    for (int q = 0; q < N / 8; q++) reinterpret_cast<uint4*>(blk)[sub * (N / 8) + q] = make_uint4(0, 0, 0, 0);
    __syncwarp();
    tu_next = tu_at(o_next, ok_next, f_next);
    // ---- scatter + dequantise ----
    int mx = -1, my = -1;
    {
      const b265_coeff* __restrict__ c = f->coeffs + tu.coeff_off;
      const int ncoeff = active ? tu.ncoeff : 0;
      const b265_scaling* sl = (f->flags & B265_FRAME_SCALING_LIST) ? f->scaling : nullptr;
      const DequantK k = dequant_consts(tu.qp, LOG2);
      const int mid = cidx + ((tu.flags & B265_TU_INTRA_CU) ? 0 : (N < 32 ? 3 : 1));
      int16_t* plane = reinterpret_cast<int16_t*>(blk);
      for (int e = sub; e < ncoeff; e += N) {
        b265_coeff kc;
        if (e == sub) { kc.level = (int16_t)(kc_pref & 0xffff); kc.pos = (uint16_t)(kc_pref >> 16); }   // requested an iteration ago
        else kc = c[e];
        const int x = kc.pos & (N - 1), y = kc.pos >> LOG2;
        int v;
        if (sl) v = dequant_list(kc.level, tu.qp, LOG2, (LOG2 == 4) ? sl->s16[mid][kc.pos] : sl->s32[mid][kc.pos]);
        else v = clip16((kc.level * (int)k.sel_lo * k.m + k.rnd) >> k.sh);
        const int s = slot_lut[y];
        plane[((((s >> 1) * N) + x) << 1) | (s & 1)] = (int16_t)v;
        mx = max(mx, x);
        my = max(my, y);
      }
    }
    const int nrows = __reduce_max_sync(B265_FULL_MASK, my + 1);
    const int ncols = __reduce_max_sync(B265_FULL_MASK, mx + 1);
    __syncwarp();
    // ---- first pass down the columns ----
    if (N >= 32 && nrows > 16) dense_cols<LOG2, (N >= 32 ? 32 : N)>(blk, sub, slot);
    else if (nrows > 8) dense_cols<LOG2, 16>(blk, sub, slot);
    else if (nrows > 4) dense_cols<LOG2, 8>(blk, sub, slot);
    else dense_cols<LOG2, 4>(blk, sub, slot);
    __syncwarp();
    // the next TU's first levels (one per lane; most TUs have no more) travel during the row pass and the stores
    kc_pref = 0;
    if (ok_next && sub < (int)(tu_next.w & 0xffff))
      kc_pref = __ldg(reinterpret_cast<const uint32_t*>(f_next->coeffs + tu_next.z) + sub);
    // ---- second pass along the rows ----
    int out[N];
    if (N >= 32 && ncols > 16) sparse_rows<LOG2, (N >= 32 ? 32 : N)>(blk, sub, out);
    else if (ncols > 8) sparse_rows<LOG2, 16>(blk, sub, out);
    else if (ncols > 4) sparse_rows<LOG2, 8>(blk, sub, out);
    else sparse_rows<LOG2, 4>(blk, sub, out);
    if (!active) continue;
    const int pitch = f->pitch[cidx];
    if (SCRATCH) {
      // block-linear scratch (common.cuh resid_block): line l of the TB at +l*N, the whole TB contiguous
      int16_t* d = f->resid[cidx] + resid_block(tu.x, tu.y, f->log2_ctb - ((cidx && f->cw) ? 1 : 0), f->ctbs_w);
      const bool transposed = tu.intra_mode >= 2 && tu.intra_mode < 18;
      if (transposed) {
#pragma unroll
        for (int x = 0; x < N; x++) d[x * N + sub] = (int16_t)min(max(out[x] >> 12, -1024), 1023);
      } else {
        // Every lane stores another row, so a store instruction costs one L1 wavefront per lane whatever its width (the
        // pre-pass is bound by exactly those, ncu l1tex data-pipe 64-74 %): 32 bytes per instruction (sm_100 STG.256)
        int16_t* dr = d + sub * N;   // 32-byte aligned: a TB starts at a multiple of 256 values, a line is 32 or 64 bytes
#pragma unroll
        for (int q = 0; q < N / 16; q++) {
          uint32_t w[8];
#pragma unroll
          for (int j = 0; j < 8; j++) {
            const uint32_t p2 = pack_sat16(out[16 * q + 2 * j] >> 12, out[16 * q + 2 * j + 1] >> 12);
            w[j] = __vmaxs2(__vmins2(p2, 0x03ff03ffu), 0xfc00fc00u);
          }
          asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(dr + 16 * q), "r"(w[0]), "r"(w[1]), "r"(w[2]),
                       "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7])
                       : "memory");
        }
      }
    } else {
      uint8_t* drow = f->plane[cidx] + (size_t)(tu.y + sub) * pitch + tu.x;
#pragma unroll
      for (int q = 0; q < N / 16; q++) {
        const uint4 pw = reinterpret_cast<const uint4*>(drow)[q];
        reinterpret_cast<uint4*>(drow)[q] = make_uint4(add_pack4(out + 16 * q, pw.x), add_pack4(out + 16 * q + 4, pw.y),
                                                       add_pack4(out + 16 * q + 8, pw.z), add_pack4(out + 16 * q + 12, pw.w));
      }
    }
  }
}

// plain DCT TUs of 16x16 / 32x32; returns false for the other sizes (generic kernel)
bool launch_residual_sparse_fast(cudaStream_t st, int log2, const FrameDev* frames, const uint32_t* order, int n,
                                 int sm_count, bool to_scratch) {
  if (log2 < 4) return false;
  if (n <= 0) return true;
  const int G = 32 >> log2;
  const int warps = (n + G - 1) / G, ctas = (warps + DENSE_WARPS - 1) / DENSE_WARPS;
  const int grid = min(ctas, sm_count * 16);
  dim3 b(DENSE_WARPS * 32);
  if (log2 == 4) {
    if (to_scratch) k_residual_sparse_fast<4, true><<<grid, b, 0, st>>>(frames, order, n);
    else k_residual_sparse_fast<4, false><<<grid, b, 0, st>>>(frames, order, n);
  } else {
    if (to_scratch) k_residual_sparse_fast<5, true><<<grid, b, 0, st>>>(frames, order, n);
    else k_residual_sparse_fast<5, false><<<grid, b, 0, st>>>(frames, order, n);
  }
  return true;
}
