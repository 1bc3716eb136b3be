// k_blocks.cu — small kernels outside the wavefront: PCM sample store for the frame path (S4) and the
// isolated-block forms of intra prediction and interpolation behind b265_intra_batch /
// b265_interp_batch (the per-block slots of acceleration.h, batched).
//
//  * k_pcm: the store half of read_pcm_samples (libde265/slice.cc:4356-4406); the host parser has
//    already shifted the codes to the bit depth (sample = code << (8 - pcm_bits)).
//  * k_intra_blocks: intra_prediction_sample_filtering + planar/DC/angular (intrapred.cc:816-1069)
//    from explicit border arrays, the same device functions the wavefront uses (intra.cuh).
//  * k_interp_blocks: put_hevc_qpel_8 / put_hevc_epel_8 slots (fallback-motion.cc:298-407, 482-625).
#include "intra_seg.cuh"
#include "kernels.h"

// ---------------------------------------------------------------------------------------------
// PCM: one CTA per PCM coding block; samples are stored Y (n*n), Cb, Cr (n/2*n/2 each) at sample_off
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_pcm(const FrameDev* __restrict__ frames) {
  const FrameDev& f = frames[blockIdx.y];
  if ((int)blockIdx.x >= f.n_pcms) return;
  const b265_pcm p = f.pcms[blockIdx.x];
  const int n = 1 << p.log2_size;
  const uint8_t* s = f.pcm_samples + p.sample_off;
  for (int i = threadIdx.x; i < n * n; i += blockDim.x) {
    const int x = i & (n - 1), y = i >> p.log2_size;
    f.plane[0][(size_t)(p.y + y) * f.pitch[0] + p.x + x] = s[i];
  }
  if (!f.cw) return;
  const int nc = n >> 1, lc = p.log2_size - 1;
  s += n * n;
  for (int i = threadIdx.x; i < 2 * nc * nc; i += blockDim.x) {
    const int c = 1 + (i >= nc * nc), j = i - (c - 1) * nc * nc;
    const int x = j & (nc - 1), y = j >> lc;
    f.plane[c][(size_t)((p.y >> 1) + y) * f.pitch[c] + (p.x >> 1) + x] = s[i];
  }
}

void launch_pcm(cudaStream_t st, const FrameDev* frames, int n_frames, int max_pcms) {
  if (n_frames <= 0 || max_pcms <= 0) return;
  k_pcm<<<dim3(max_pcms, n_frames), 128, 0, st>>>(frames);
}

// ---------------------------------------------------------------------------------------------
// intra prediction of isolated blocks: one block per warp, the size-independent code of the wavefront (intra_seg.cuh)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_intra_blocks(int log2, int cidx, int n_blocks, const uint8_t* __restrict__ modes,
                                                      const uint8_t* __restrict__ borders, int strong, int no_edge_filter,
                                                      uint8_t* __restrict__ out) {
  const int N = 1 << log2;
  __shared__ __align__(16) uint8_t sm[4][B265_BORDER_BYTES];
  const int warp = threadIdx.x >> 5, lane = lane_id();
  const int blk = blockIdx.x * 4 + warp;   // one block per warp, as in the wavefront
  if (blk >= n_blocks) return;
  uint8_t* buf0 = sm[warp];
  uint8_t* buf1 = buf0 + ibuf_stride(N);
  const int mode = modes[blk];
  {  // p[-2N..2N] -> the two border lines T[k] = p[k], L[k] = p[-k]
    const uint8_t* p = borders + (size_t)blk * (4 * N + 1) + 2 * N;
    for (int k = lane; k <= 2 * N; k += 32) {
      buf0[ibuf_top(N) + k] = p[k];
      buf0[ibuf_left(N) + k] = p[-k];
    }
  }
  __syncwarp();
  uint8_t* pb = filter_border_rt(buf0, buf1, N, lane, mode, cidx, strong != 0);
  IntraPred P;
  intra_pred_setup(P, pb, N, log2, lane, mode, cidx, !no_edge_filter);
  const int lseg = log2 < 3 ? 0 : log2 - 3, items = N << lseg, nvalid = log2 == 2 ? 4 : 8;
  uint8_t* d = out + (size_t)blk * N * N;
  for (int item = lane; item < items; item += 32) {
    const int l = item >> lseg, sg = item & ((1 << lseg) - 1);
    uint32_t h[4], px[2];
    const uint32_t none[4] = {0, 0, 0, 0};
    intra_pred_seg(P, l, sg, h);
    finish_seg(h, none, false, px);
    for (int j = 0; j < nvalid; j++) {
      const uint8_t v = (uint8_t)(px[j >> 2] >> (8 * (j & 3)));
      if (!P.by_column) d[l * N + 8 * sg + j] = v;
      else d[(8 * sg + j) * N + l] = v;
    }
  }
}

void launch_intra_blocks(cudaStream_t st, int log2, int cidx, int n_blocks, const uint8_t* modes, const uint8_t* borders,
                         int strong, int no_edge_filter, uint8_t* out) {
  if (n_blocks <= 0) return;
  k_intra_blocks<<<(n_blocks + 3) / 4, 128, 0, st>>>(log2, cidx, n_blocks, modes, borders, strong, no_edge_filter, out);
}

// ---------------------------------------------------------------------------------------------
// raw interpolation: one thread per output sample.  Same single formula as k_mc.cu (phase 0 written
// as the tap {..,64,..}): out14 = (sum_j ty[j] * (sum_k tx[k] * px)) >> 6.
// ---------------------------------------------------------------------------------------------
__constant__ int8_t c_blk_qpel[4][8] = {{0, 0, 0, 64, 0, 0, 0, 0},
                                        {-1, 4, -10, 58, 17, -5, 1, 0},
                                        {-1, 4, -11, 40, 40, -11, 4, -1},
                                        {0, 1, -5, 17, 58, -10, 4, -1}};
__constant__ int8_t c_blk_epel[8][4] = {{0, 64, 0, 0},   {-2, 58, 10, -2}, {-4, 54, 16, -2}, {-6, 46, 28, -4},
                                        {-4, 36, 36, -4}, {-4, 28, 46, -6}, {-2, 16, 54, -4}, {-2, 10, 58, -2}};

template <int TAPS>
__global__ void __launch_bounds__(256) k_interp_blocks(int n_blocks, int w, int h, const uint8_t* __restrict__ src,
                                                       long long src_stride, long long src_block_stride,
                                                       const uint8_t* __restrict__ frac_xy, int16_t* __restrict__ out) {
  const long long total = (long long)n_blocks * w * h;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int blk = (int)(i / (w * h)), r = (int)(i - (long long)blk * w * h), y = r / w, x = r - y * w;
    const int fx = frac_xy[2 * blk], fy = frac_xy[2 * blk + 1];
    const uint8_t* p = src + blk * src_block_stride + (y - (TAPS / 2 - 1)) * src_stride + (x - (TAPS / 2 - 1));
    int acc = 0;
#pragma unroll
    for (int j = 0; j < TAPS; j++) {
      const int ty = TAPS == 8 ? c_blk_qpel[fy][j] : c_blk_epel[fy][j];
      if (ty == 0) continue;  // rows outside the filter support are not part of the slot's halo contract
      int hs = 0;
#pragma unroll
      for (int k = 0; k < TAPS; k++) {
        const int tx = TAPS == 8 ? c_blk_qpel[fx][k] : c_blk_epel[fx][k];
        if (tx) hs += tx * p[j * src_stride + k];
      }
      acc += ty * hs;
    }
    out[i] = (int16_t)(acc >> 6);
  }
}

void launch_interp_blocks(cudaStream_t st, int luma, int n_blocks, int w, int h, const uint8_t* src, long long src_stride,
                          long long src_block_stride, const uint8_t* frac_xy, int16_t* out) {
  const long long total = (long long)n_blocks * w * h;
  if (total <= 0) return;
  long long grid = (total + 255) / 256;
  if (grid > 148 * 32) grid = 148 * 32;
  if (luma) k_interp_blocks<8><<<(int)grid, 256, 0, st>>>(n_blocks, w, h, src, src_stride, src_block_stride, frac_xy, out);
  else k_interp_blocks<4><<<(int)grid, 256, 0, st>>>(n_blocks, w, h, src, src_stride, src_block_stride, frac_xy, out);
}
