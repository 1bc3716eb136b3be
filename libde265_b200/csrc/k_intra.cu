// k_intra.cu — intra prediction (+ residual add) of all intra TBs of a batch of pictures as ONE
// persistent launch: a CTB-level wavefront.
//
// Intra TBs form dependency chains in z-scan order inside a CTB (every TB reads the reconstructed
// samples left of / above it, including below-left and above-right), so a CTB plane is walked
// SEQUENTIALLY by one warp and parallelism comes from (a) the three colour planes, (b) the CTB
// anti-diagonal wave = ctbX + 2*ctbY (a CTB needs its left, upper-left, upper and upper-right neighbours,
// intrapred.cc:597-687), and (c) the pictures of the batch, which are mutually independent.
// The kernel is therefore bound by the LATENCY of that one warp; everything that does not depend on
// neighbouring samples is kept off its path:
//  * the residuals of the intra TBs (dequant + inverse transform) are computed beforehand by the fully
//    parallel k_residual_sparse<.., true> launches into an int16 residual plane (frames.cu);
//  * the availability mask of every TB is computed beforehand by k_intra_prep (one thread per TB);
//  * the samples a TB needs from earlier TBs of the same CTB come from the CTB's edge store in shared
//    memory (intra.cuh), so global memory is read once per CTB (bordering row/column of the neighbouring
//    CTBs, one L2 round trip after the flag wait) and never per TB.
//
// Host side (frames.cu) emits one ticket per (picture, CTB, plane) that contains intra TBs, sorted by
// wave.  Warps draw tickets from an atomic counter; a warp waits for the (at most four) neighbouring
// CTB flags of its plane, processes the CTB's TBs in decoding order, then publishes its own flag with a
// release store.  Every dependency has a smaller ticket, tickets are only held by running warps, so
// spinning cannot deadlock for any grid size.
//
// HBM traffic per TB: N*N bytes written (+ 2*N*N bytes of residual read when the TB is coded).
#include <stdio.h>
#include <stdlib.h>
#include "transform.cuh"
#include "intra.cuh"
#include "kernels.h"


#define INTRA_WARPS 4   // one warp per SM sub-partition; several CTAs per SM when the batch is large
#ifndef INTRA_MIN_CTAS
#define INTRA_MIN_CTAS 4  // resident CTAs per SM the register budget is sized for
#endif

// ---------------------------------------------------------------------------------------------
// k_intra_prep: availability mask of every intra TB (one thread per TU), FrameDev::tb_mask[tu]
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_intra_prep(const FrameDev* __restrict__ frames) {
  const FrameDev& f = frames[blockIdx.y];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < f.n_tus; i += gridDim.x * blockDim.x) {
    const b265_tu tu = f.tus[i];
    if (tu.intra_mode == B265_INTRA_NONE) continue;
    const int c = tu.cidx_log2 & 3, lg = (tu.cidx_log2 >> 2) & 7;
    IntraGeom g;
    g.sh = (c && f.cw) ? 1 : 0;
    g.width = f.width; g.height = f.height;
    g.log2_ctb = f.log2_ctb; g.log2_min_tb = f.log2_min_tb;
    g.cx = (tu.x << g.sh) >> f.log2_ctb; g.cy = (tu.y << g.sh) >> f.log2_ctb;
    g.unit_info = (f.flags & B265_FRAME_CONSTRAINED_INTRA) ? f.unit_info : nullptr;
    g.unit_pitch = f.width >> 2;
    g.nb = 0;
    const b265_ctb* ctbs = f.ctbs;
    const b265_ctb me = ctbs ? ctbs[g.cy * f.ctbs_w + g.cx] : b265_ctb{0, 0, 0};
#pragma unroll
    for (int k = 0; k < 4; k++) {  // left, upper-left, upper, upper-right
      const int nx = g.cx + (k == 3 ? 1 : (k == 2 ? 0 : -1)), ny = g.cy - (k == 0 ? 0 : 1);
      if (nx < 0 || ny < 0 || nx >= f.ctbs_w) continue;
      bool ok = true;
      if (ctbs) {
        const b265_ctb r = ctbs[ny * f.ctbs_w + nx];
        ok = r.slice == me.slice && r.tile == me.tile && r.addr_ts < me.addr_ts;
      }
      if (ok) g.nb |= 1u << k;
    }
    f.tb_mask[i] = intra_avail_mask(g, tu.x, tu.y, lg);
  }
}

void launch_intra_prep(cudaStream_t st, const FrameDev* frames, int n_frames, int max_tus) {
  if (n_frames <= 0 || max_tus <= 0) return;
  int gx = (max_tus + 255) / 256;
  if (gx > 592) gx = 592;
  k_intra_prep<<<dim3(gx, n_frames), 256, 0, st>>>(frames);
}

// ---------------------------------------------------------------------------------------------
// the wavefront
// ---------------------------------------------------------------------------------------------
// residual line `l` of a coded TB (written by k_residual_sparse<LOG2, true>, transposed for modes 2..17)
template <int N>
__device__ __forceinline__ void load_resid_line(const int16_t* __restrict__ p, uint32_t (&w)[N / 2]) {
  if (N == 4) {
    const uint2 v = __ldg(reinterpret_cast<const uint2*>(p));
    w[0] = v.x; w[1] = v.y;
  } else {
#pragma unroll
    for (int q = 0; q < N / 8; q++) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(p) + q);
      w[4 * q] = v.x; w[4 * q + 1] = v.y; w[4 * q + 2] = v.z; w[4 * q + 3] = v.w;
    }
  }
}

template <int N>
__device__ __forceinline__ void store_words(uint8_t* p, const uint32_t (&px)[N / 4]) {
  if (N == 4) *reinterpret_cast<uint32_t*>(p) = px[0];
  else if (N == 8) *reinterpret_cast<uint2*>(p) = make_uint2(px[0], px[1]);
  else {
#pragma unroll
    for (int q = 0; q < N / 16; q++)
      reinterpret_cast<uint4*>(p)[q] = make_uint4(px[4 * q], px[4 * q + 1], px[4 * q + 2], px[4 * q + 3]);
  }
}
// the edge store is only 4-byte aligned in general
template <int N>
__device__ __forceinline__ void store_words_smem(uint8_t* p, const uint32_t (&px)[N / 4]) {
#pragma unroll
  for (int q = 0; q < N / 4; q++) reinterpret_cast<uint32_t*>(p)[q] = px[q];
}

template <int LOG2, bool TRACE>
__device__ __forceinline__ void intra_tb(const CtbEdges& E, const IntraGeom& g, int cidx, int x, int y, int mode, int flags,
                                         bool coded, uint64_t M, uint8_t* wborder, long long (&acc)[4]) {
  constexpr int N = 1 << LOG2, STRIDE = IntraBuf<N>::STRIDE;
  long long c0 = 0, c1 = 0, c2 = 0, c3 = 0;
  if (TRACE) c0 = clock64();
  const int lane = lane_id();
  const bool active = lane < N;
  const int l = active ? lane : 0;
  // the residual line is independent of the neighbours: issue its load first
  uint32_t r[N / 2];
  if (coded && active) load_resid_line<N>(g.resid + (size_t)(y + l) * g.pitch + x, r);
  const int xr = x - E.x0, yr = y - E.y0;
  uint8_t* buf0 = wborder;
  uint8_t* buf1 = buf0 + STRIDE;
  gather_border<LOG2>(E, xr, yr, M, lane, buf0);
  if (TRACE) c1 = clock64();
  uint8_t* pb = filter_border<LOG2>(buf0, buf1, lane, mode, cidx, g.strong);
  if (TRACE) c2 = clock64();
  uint32_t h[N / 2];
  const bool by_column = predict_line<LOG2>(pb, lane, mode, cidx, !(flags & B265_TU_NO_EDGE_FILT), h);
  if (TRACE) c3 = clock64();
  if (active) {
    uint32_t px[N / 4];
    finish_line<N>(h, r, coded, px);
    uint8_t* d = g.plane + (size_t)y * g.pitch + x;
    uint8_t* erow = E.rows + ((yr + N) >> 2) * E.rstride + 16 + xr;   // row below the TB's last row ... i.e. its bottom row
    uint8_t* ecol = E.cols + ((xr + N) >> 2) * E.S + yr;              // its right column
    // The corner sample of a later TB lies on an edge row AND an edge column but only needs to be this TB's
    // bottom row OR right column: the right column is therefore also entered into the edge rows it crosses.
    if (!by_column) {   // lane = row
      store_words<N>(d + (size_t)l * g.pitch, px);
      ecol[l] = (uint8_t)(px[N / 4 - 1] >> 24);
      if ((l & 3) == 3) E.rows[((yr + l + 1) >> 2) * E.rstride + 16 + xr + N - 1] = (uint8_t)(px[N / 4 - 1] >> 24);
      if (l == N - 1) store_words_smem<N>(erow, px);
    } else {            // lane = column
#pragma unroll
      for (int j = 0; j < N; j++) d[(size_t)j * g.pitch + l] = (uint8_t)(px[j >> 2] >> (8 * (j & 3)));
      erow[l] = (uint8_t)(px[N / 4 - 1] >> 24);
      if (l == N - 1) {
        store_words_smem<N>(ecol, px);
#pragma unroll
        for (int q = 0; q < N / 4; q++) E.rows[((yr + 4 * q + 4) >> 2) * E.rstride + 16 + xr + N - 1] = (uint8_t)(px[q] >> 24);
      }
    }
  }
  __syncwarp();  // the next TB of this CTB reads the edge store
  if (TRACE) {
    acc[0] += c1 - c0; acc[1] += c2 - c1; acc[2] += c3 - c2; acc[3] += clock64() - c3;
  }
}

// ---- border exchange between CTBs: flag-in-data mailboxes ---------------------------------------------
// A CTB's warp posts the CTB's bottom row and right column (the only samples later CTBs can read) as
// 8-byte words {4 samples, tag != 0} written with ONE 64-bit store each: a reader that sees the tag has
// the samples, so neither side needs a fence (the scheme of low-latency collective protocols).  The
// mailboxes are zeroed before the launch.  Layout per CTB: S/4 words bottom row, S/4 words right column.
// Neighbours without a ticket (inter/PCM CTBs, done[] == 1) were finished by earlier launches: their
// samples are read from the plane.
//
// The polling loop is executed by the WHOLE warp with a warp-uniform exit: a per-lane loop leaves the
// warp split into independently scheduled groups, which then walk the TBs separately and pay the slow
// path of every warp collective (measured: 3x).
__device__ __forceinline__ void fetch_borders(const IntraGeom& g, const CtbEdges& E, const uint8_t* done, const uint8_t* mbox,
                                              int ctbs_w, bool interior, unsigned sleep_ns) {
  const int lane = lane_id();
  const int S = E.S, q = S / 4, pitch = g.pitch;
  const int W = (g.width + g.sh) >> g.sh, H = (g.height + g.sh) >> g.sh;
  const uint8_t* org = g.plane + (size_t)E.y0 * pitch + E.x0;
  // slots: [0,q) upper CTB's bottom row, [q,2q) upper-right CTB's bottom row, [2q,3q) left CTB's right column,
  // 3q: corner (last word of the upper-left CTB's bottom row); two slots per lane (3q+1 <= 49)
  const uint8_t* src[2];
  int kind[2];   // 0 nothing, 1 mailbox word, 2 plane (row word), 3 plane (column word, 4 strided bytes), 4 plane corner byte
  int slotv[2];
  const int lq = g.log2_ctb - g.sh - 2;   // q = 1 << lq
  const bool two = 3 * q >= 32;           // more than 32 slots: only for 64-sample CTB planes
#pragma unroll
  for (int k = 0; k < 2; k++) {
    const int slot = lane + 32 * k;
    slotv[k] = slot;
    kind[k] = 0;
    src[k] = nullptr;
    if (k == 1 && !two) break;   // warp-uniform
    if (slot > 3 * q) continue;
    const int piece = slot >> lq, w = slot == 3 * q ? q - 1 : slot & (q - 1);
    // neighbour CTB of the piece
    const int nx = g.cx + (piece == 1 ? 1 : (piece == 0 ? 0 : -1)), ny = g.cy - (piece == 2 ? 0 : 1);
    // not usable for prediction (outside the picture, other slice/tile): never read, never waited for
    if (!((g.nb >> (piece == 0 ? 2 : piece == 1 ? 3 : piece == 2 ? 0 : 1)) & 1)) continue;
    const int nctb = ny * ctbs_w + nx;
    if (__ldg(done + nctb) == 0) {   // the neighbour posts its borders in this launch
      kind[k] = 1;
      src[k] = mbox + (size_t)nctb * (4 * S) + (piece == 2 ? 2 * S : 0) + 8 * w;
    } else if (piece == 2) {
      if (E.y0 + 4 * w < H) { kind[k] = 3; src[k] = org + (size_t)(4 * w) * pitch - 1; }
    } else if (piece == 3) {
      kind[k] = 4; src[k] = org - pitch - 1;
    } else {
      const int x = (piece == 1 ? S : 0) + 4 * w;
      if (E.x0 + x < W) { kind[k] = 2; src[k] = org - pitch + x; }
    }
  }
  uint32_t val[2] = {0, 0};
  bool have[2] = {kind[0] != 1, kind[1] != 1};
  // plane reads (rare: mixed pictures)
#pragma unroll
  for (int k = 0; k < 2; k++) {
    if (k == 1 && !two) break;
    if (kind[k] == 2) val[k] = ld_cg_u32(src[k]);
    else if (kind[k] == 3) {
      for (int i = 0; i < 4; i++) val[k] |= (uint32_t)ld_cg_u8(src[k] + (size_t)i * pitch) << (8 * i);
    } else if (kind[k] == 4) val[k] = (uint32_t)ld_cg_u8(src[k]) << 24;
  }
  // every lane polls its own word(s): restricting the polling to one lane per neighbour, or pausing longer
  // between polls, was measured slower at every batch size (the wavefront is bound by detection latency)
  for (;;) {
#pragma unroll
    for (int k = 0; k < 2; k++)
      if (!have[k] && (k == 0 || two)) {
        const unsigned long long v = ld_relaxed_u64(src[k]);
        if ((uint32_t)(v >> 32) != 0) { val[k] = (uint32_t)v; have[k] = true; }
      }
    if (__all_sync(B265_FULL_MASK, have[0] && have[1])) break;
    if (sleep_ns) __nanosleep(sleep_ns);
  }
  // into the edge store
#pragma unroll
  for (int k = 0; k < 2; k++) {
    if (k == 1 && !two) break;
    if (kind[k] == 0) continue;
    const int slot = slotv[k];
    if (slot < 2 * q) *reinterpret_cast<uint32_t*>(E.rows + 16 + 4 * slot) = val[k];
    else if (slot < 3 * q) {
      const int w = slot - 2 * q;
      *reinterpret_cast<uint32_t*>(E.cols + 4 * w) = val[k];
      E.rows[(w + 1) * E.rstride + 15] = (uint8_t)(val[k] >> 24);   // sample (-1, 4w+3) of the inner edge rows
    } else E.rows[15] = (uint8_t)(val[k] >> 24);
  }
  if (interior) {
    // CTBs that also hold inter/PCM samples: their edge rows 4q-1 (q >= 1) and columns 4q-1 come from the plane
    const int wpr = S / 4;
    for (int i = lane; i < (S / 4) * wpr; i += 32) {
      const int qq = 1 + i / wpr, w = i % wpr, r = 4 * qq - 1;
      if (E.y0 + r < H && E.x0 + 4 * w < W)
        *reinterpret_cast<uint32_t*>(E.rows + qq * E.rstride + 16 + 4 * w) = ld_cg_u32(org + (size_t)r * pitch + 4 * w);
    }
    for (int i = lane; i < (S / 4) * S; i += 32) {
      const int qq = 1 + i / S, r = i % S, cc = 4 * qq - 1;
      if (E.y0 + r < H && E.x0 + cc < W) E.cols[qq * S + r] = ld_cg_u8(org + (size_t)r * pitch + cc);
    }
  }
  __syncwarp();
}

// post this CTB's bottom row and right column (edge store rows[S/4], cols[S/4]) to its mailbox
__device__ __forceinline__ void post_borders(const CtbEdges& E, uint8_t* mbox, int ctb) {
  const int lane = lane_id();
  const int S = E.S, q = S / 4;
  if (lane < 2 * q) {
    const uint8_t* s = lane < q ? E.rows + q * E.rstride + 16 + 4 * lane : E.cols + q * S + 4 * (lane - q);
    const unsigned long long v = (unsigned long long)*reinterpret_cast<const uint32_t*>(s) | (1ull << 32);
    st_relaxed_u64(mbox + (size_t)ctb * (4 * S) + 8 * lane, v);
  }
}

// TRACE (developer aid, B265_INTRA_TRACE=<file>): per ticket 8 x uint64 {t_ticket, t_deps, t_tile, t_tbs, t_pub,
// n_tbs, gather, filter} + 4 more {predict, add+store, globaltimer, sm id}: 12 x uint64 per ticket, SM cycles
template <bool TRACE>
__global__ void __launch_bounds__(INTRA_WARPS * 32, INTRA_MIN_CTAS)
k_intra_wavefront(const FrameDev* __restrict__ frames, const IntraTicket* __restrict__ tickets, int n_tickets,
                  unsigned* __restrict__ ticket, unsigned sleep_ns, int tile_bytes, unsigned long long* trace) {
  extern __shared__ __align__(16) uint8_t dyn_smem[];
  __shared__ __align__(16) uint8_t s_border[INTRA_WARPS][B265_BORDER_BYTES];
  const int warp = threadIdx.x >> 5, lane = lane_id();
  uint8_t* my_edges = dyn_smem + (size_t)warp * tile_bytes;
  // Software pipeline over tickets: the next ticket number is drawn while the current CTB is walked, its
  // record is loaded during the TB loop and its picture constants are loaded before the (slow) release
  // store of the current CTB, so that none of these round trips sits between two dependent CTBs.
  // Holding a reserved ticket cannot deadlock: the smallest unfinished ticket is always being processed
  // (a warp reserves ticket B only while processing a smaller ticket A and then processes B next).
  unsigned t = 0;
  if (lane == 0) t = atomicAdd(ticket, 1u);
  t = __shfl_sync(B265_FULL_MASK, t, 0);
  IntraTicket tk = {0, 0, 0, 0, 0, 0};
  if (t < (unsigned)n_tickets) tk = tickets[t];
  // picture constants of the ticket -> registers (independent loads, one round trip)
  IntraGeom g;
  int ctbs_w = 1;
  const b265_ctb* __restrict__ ctbs = nullptr;
  const uint8_t* done = nullptr;
  uint8_t* mbox = nullptr;
  const uint4* tus4 = nullptr;
  const uint2* masks = nullptr;
  auto load_constants = [&](const IntraTicket& k) {
    const FrameDev* __restrict__ f = frames + k.frame;
    const int c = k.plane;
    const uint32_t fflags = f->flags;
    ctbs_w = f->ctbs_w;
    ctbs = f->ctbs;
    done = f->done[c];
    mbox = f->mbox[c];
    tus4 = reinterpret_cast<const uint4*>(f->tus);  // 16-byte records, 256-byte aligned list
    masks = reinterpret_cast<const uint2*>(f->tb_mask);
    g.sh = (c && f->cw) ? 1 : 0;
    g.width = f->width; g.height = f->height;
    g.log2_ctb = f->log2_ctb; g.log2_min_tb = f->log2_min_tb;
    g.plane = f->plane[c]; g.resid = f->resid[c]; g.pitch = f->pitch[c];
    g.unit_info = (fflags & B265_FRAME_CONSTRAINED_INTRA) ? f->unit_info : nullptr;
    g.unit_pitch = g.width >> 2;
    g.strong = (fflags & B265_FRAME_STRONG_INTRA_SMOOTHING) != 0;
  };
  if (t < (unsigned)n_tickets) load_constants(tk);
  while (t < (unsigned)n_tickets) {
    long long tr0 = 0, tr1 = 0, tr2 = 0, tr3 = 0;
    unsigned long long gt0 = 0;
    int n_tbs = 0;
    long long acc[4] = {0, 0, 0, 0};
    long long lacc = 0;
    if (TRACE) {
      tr0 = clock64();
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt0));
    }
    unsigned t_next = 0;
    if (lane == 0) t_next = atomicAdd(ticket, 1u);   // consumed after fill_edges
    const int c = tk.plane;
    g.cx = tk.ctb % ctbs_w; g.cy = tk.ctb / ctbs_w;
    const int cx = g.cx, cy = g.cy;
    // prefetch the first 32 TU records of the CTB while waiting for the neighbours
    uint4 mine = make_uint4(0, 0xff000000u, 0, 0);
    uint2 mymask = make_uint2(0, 0);
    if ((uint32_t)lane < tk.tu_count) {
      mine = __ldg(tus4 + tk.tu_first + lane);
      mymask = __ldg(masks + tk.tu_first + lane);
    }
    CtbEdges E;
    E.S = (1 << g.log2_ctb) >> g.sh;
    E.rstride = 2 * E.S + 16;
    E.rows = my_edges;
    E.cols = my_edges + (E.S / 4 + 1) * E.rstride;
    E.x0 = cx * E.S;
    E.y0 = cy * E.S;
    // neighbouring CTBs usable for prediction: lane 0 left, 1 upper-left, 2 upper, 3 upper-right; lane 4 = this CTB
    {
      const int nx = cx + (lane == 3 ? 1 : (lane == 2 || lane == 4 ? 0 : -1)), ny = cy - (lane == 0 || lane == 4 ? 0 : 1);
      const bool exists = lane < 5 && nx >= 0 && ny >= 0 && nx < ctbs_w;
      b265_ctb rec = {0, 0, 0};
      if (exists && ctbs) rec = ctbs[ny * ctbs_w + nx];
      const int my_ts = __shfl_sync(B265_FULL_MASK, rec.addr_ts, 4);
      const int my_slice = __shfl_sync(B265_FULL_MASK, (int)rec.slice, 4), my_tile_id = __shfl_sync(B265_FULL_MASK, (int)rec.tile, 4);
      const bool usable = exists && lane < 4 && (!ctbs || (rec.slice == my_slice && rec.tile == my_tile_id && rec.addr_ts < my_ts));
      g.nb = __ballot_sync(B265_FULL_MASK, usable) & 0xfu;
    }
    if (TRACE) tr1 = clock64();
    fetch_borders(g, E, done, mbox, ctbs_w, (tk.flags & B265_TICKET_MIXED) != 0, sleep_ns);
    if (TRACE) tr2 = clock64();
    // the next ticket's record travels while the TBs are processed
    t_next = __shfl_sync(B265_FULL_MASK, t_next, 0);
    IntraTicket tk_next = {0, 0, 0, 0, 0, 0};
    if (t_next < (unsigned)n_tickets) tk_next = tickets[t_next];
    for (uint32_t base = 0; base < tk.tu_count; base += 32) {
      if (base) {
        mine = make_uint4(0, 0xff000000u, 0, 0);
        if (base + lane < tk.tu_count) {
          mine = __ldg(tus4 + tk.tu_first + base + lane);
          mymask = __ldg(masks + tk.tu_first + base + lane);
        }
      }
      const int n_here = min(32u, tk.tu_count - base);
      // lanes whose TU is an intra TB of this plane
      const bool want = lane < n_here && (mine.y >> 24) != B265_INTRA_NONE && (int)(mine.y & 3) == c;
      unsigned todo = __ballot_sync(B265_FULL_MASK, want);
      while (todo) {
        long long lc0 = 0;
        if (TRACE) lc0 = clock64();
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const uint32_t r0 = __shfl_sync(B265_FULL_MASK, mine.x, src), r1 = __shfl_sync(B265_FULL_MASK, mine.y, src);
        const uint32_t r3 = __shfl_sync(B265_FULL_MASK, mine.w, src);
        const int x = (int)(r0 & 0xffff), y = (int)(r0 >> 16);
        const int flags = (int)((r1 >> 8) & 0xff), mode = (int)(r1 >> 24);
        const bool coded = (r3 & 0xffff) != 0;
        const uint64_t M = (uint64_t)__shfl_sync(B265_FULL_MASK, mymask.x, src) |
                           ((uint64_t)__shfl_sync(B265_FULL_MASK, mymask.y, src) << 32);
        if (TRACE) { n_tbs++; lacc += clock64() - lc0; }
        switch ((r1 >> 2) & 7) {
          case 2: intra_tb<2, TRACE>(E, g, c, x, y, mode, flags, coded, M, s_border[warp], acc); break;
          case 3: intra_tb<3, TRACE>(E, g, c, x, y, mode, flags, coded, M, s_border[warp], acc); break;
          case 4: intra_tb<4, TRACE>(E, g, c, x, y, mode, flags, coded, M, s_border[warp], acc); break;
          default: intra_tb<5, TRACE>(E, g, c, x, y, mode, flags, coded, M, s_border[warp], acc); break;
        }
      }
    }
    if (TRACE) tr3 = clock64();
    // next ticket's picture constants: issued before the release store so that both round trips overlap
    uint8_t* my_mbox = mbox;
    const int my_ctb = (int)tk.ctb;
    const unsigned t_cur = t;
    t = t_next;
    tk = tk_next;
    if (t < (unsigned)n_tickets) load_constants(tk);
    // post the borders later CTBs read (the picture itself is complete at the end of the launch)
    post_borders(E, my_mbox, my_ctb);
    if (TRACE && lane == 0) {
      unsigned smid;
      asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
      unsigned long long* o = trace + (size_t)t_cur * 12;
      o[0] = tr0; o[1] = tr1; o[2] = tr2; o[3] = tr3; o[4] = clock64(); o[5] = n_tbs; o[6] = acc[0]; o[7] = acc[1];
      o[8] = acc[2]; o[9] = acc[3]; o[10] = gt0; o[11] = ((unsigned long long)lacc << 16) | smid;
    }
  }
}

static int tile_bytes_for(int log2_ctb) { return (ctb_edges_bytes(1 << log2_ctb) + 15) & ~15; }

int intra_max_resident_ctas(int sm_count, int log2_ctb) {
  int per_sm = 0;
  const size_t dyn = (size_t)INTRA_WARPS * tile_bytes_for(log2_ctb);
  // the opt-in limit is a property of the function, not of a launch: always allow the largest CTB size
  const size_t dyn_max = (size_t)INTRA_WARPS * tile_bytes_for(6);
  cudaFuncSetAttribute(k_intra_wavefront<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn_max);
  cudaFuncSetAttribute(k_intra_wavefront<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn_max);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_intra_wavefront<false>, INTRA_WARPS * 32, dyn);
  if (per_sm < 1) per_sm = 1;
  return per_sm * sm_count;
}

void launch_intra(cudaStream_t st, const FrameDev* frames, const IntraTicket* tickets, int n_tickets, unsigned* ticket,
                  int sm_count, int wave_width, int log2_ctb) {
  if (n_tickets <= 0) return;
  static int max_ctas[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (!max_ctas[log2_ctb & 7]) max_ctas[log2_ctb & 7] = intra_max_resident_ctas(sm_count, log2_ctb);
  // warps beyond ~3 waves' worth would only poll; keep them off the SMs the working warps run on
  int want = 3 * wave_width;
  if (want < 2 * sm_count) want = 2 * sm_count;
  if (want > n_tickets) want = n_tickets;
  int ctas = (want + INTRA_WARPS - 1) / INTRA_WARPS;
  if (ctas > max_ctas[log2_ctb & 7]) ctas = max_ctas[log2_ctb & 7];
  cudaMemsetAsync(ticket, 0, sizeof(unsigned), st);
  static int sleep_ns = -1;
  if (sleep_ns < 0) {
    const char* e = getenv("B265_INTRA_SLEEP_NS");
    sleep_ns = e ? atoi(e) : 100;
  }
  const int tb = tile_bytes_for(log2_ctb);
  static const char* trace_path = getenv("B265_INTRA_TRACE");
  if (trace_path) {  // developer aid: synchronous, one file per process (last launch wins)
    unsigned long long* d = nullptr;
    cudaMalloc((void**)&d, (size_t)n_tickets * 96);
    k_intra_wavefront<true><<<ctas, INTRA_WARPS * 32, (size_t)INTRA_WARPS * tb, st>>>(frames, tickets, n_tickets, ticket,
                                                                                     (unsigned)sleep_ns, tb, d);
    cudaStreamSynchronize(st);
    unsigned long long* h = (unsigned long long*)malloc((size_t)n_tickets * 96);
    cudaMemcpy(h, d, (size_t)n_tickets * 96, cudaMemcpyDeviceToHost);
    if (FILE* fp = fopen(trace_path, "wb")) { fwrite(h, 96, (size_t)n_tickets, fp); fclose(fp); }
    free(h);
    cudaFree(d);
    return;
  }
  k_intra_wavefront<false><<<ctas, INTRA_WARPS * 32, (size_t)INTRA_WARPS * tb, st>>>(frames, tickets, n_tickets, ticket,
                                                                                    (unsigned)sleep_ns, tb, nullptr);
}
