// k_hash.cu — decoded picture hash (SEI payload 132) of HBM-resident pictures: MD5, CRC and checksum of each
// colour plane as the reference computes them when DE265_DECODER_PARAM_BOOL_SEI_CHECK_HASH is set
// (libde265/sei.cc:160-270 compute_checksum_8bit / compute_CRC_8bit_fast / compute_MD5, verified in
// process_sei_decoded_picture_hash :273-354).  The hashed byte stream of a plane is its rows, width bytes each,
// top to bottom (8-bit samples).
//
//  * checksum: sum of (sample ^ xorMask(x, y)) - embarrassingly parallel, one atomicAdd per warp.
//  * CRC (CCITT polynomial 0x1021, the reference's variant: register 0xFFFF, two zero bytes FIRST, then the data):
//    every row is one thread's serial job from register 0; the register after a whole plane is then the fold
//    r <- M^(8w) r ^ crc0(row) over the rows, M^(8w) being the 16x16 GF(2) matrix of "shift in 8w zero bits"
//    (computed on the device from the single-bit step by repeated squaring).  CRCs are linear, so this is exact.
//  * MD5 is serial by construction (every 64-byte block starts from the previous block's state): one warp per
//    plane, lanes fetch the next 512 bytes of the stream into shared memory while lane 0 runs the 64 steps of
//    each block.  ~8 fps per 4K plane-stream; pictures and planes of a batch run side by side.
#include "kernels.h"

namespace {

__device__ __forceinline__ uint32_t rotl(uint32_t x, int n) { return (x << n) | (x >> (32 - n)); }

__constant__ uint32_t c_md5_k[64] = {
    0xd76aa478, 0xe8c7b756, 0x242070db, 0xc1bdceee, 0xf57c0faf, 0x4787c62a, 0xa8304613, 0xfd469501, 0x698098d8, 0x8b44f7af, 0xffff5bb1,
    0x895cd7be, 0x6b901122, 0xfd987193, 0xa679438e, 0x49b40821, 0xf61e2562, 0xc040b340, 0x265e5a51, 0xe9b6c7aa, 0xd62f105d, 0x02441453,
    0xd8a1e681, 0xe7d3fbc8, 0x21e1cde6, 0xc33707d6, 0xf4d50d87, 0x455a14ed, 0xa9e3e905, 0xfcefa3f8, 0x676f02d9, 0x8d2a4c8a, 0xfffa3942,
    0x8771f681, 0x6d9d6122, 0xfde5380c, 0xa4beea44, 0x4bdecfa9, 0xf6bb4b60, 0xbebfbc70, 0x289b7ec6, 0xeaa127fa, 0xd4ef3085, 0x04881d05,
    0xd9d4d039, 0xe6db99e5, 0x1fa27cf8, 0xc4ac5665, 0xf4292244, 0x432aff97, 0xab9423a7, 0xfc93a039, 0x655b59c3, 0x8f0ccc92, 0xffeff47d,
    0x85845dd1, 0x6fa87e4f, 0xfe2ce6e0, 0xa3014314, 0x4e0811a1, 0xf7537e82, 0xbd3af235, 0x2ad7d2bb, 0xeb86d391};
__constant__ uint8_t c_md5_s[64] = {7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 5, 9,  14, 20, 5, 9,  14, 20, 5, 9,  14, 20, 5, 9,  14, 20,
                                    4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21};

__device__ void md5_block(uint32_t st[4], const uint32_t* m) {
  uint32_t a = st[0], b = st[1], c = st[2], d = st[3];
#pragma unroll 16
  for (int i = 0; i < 64; i++) {
    uint32_t f;
    int g;
    if (i < 16) { f = (b & c) | (~b & d); g = i; }
    else if (i < 32) { f = (d & b) | (~d & c); g = (5 * i + 1) & 15; }
    else if (i < 48) { f = b ^ c ^ d; g = (3 * i + 5) & 15; }
    else { f = c ^ (b | ~d); g = (7 * i) & 15; }
    const uint32_t t = d;
    d = c;
    c = b;
    b = b + rotl(a + f + c_md5_k[i] + m[g], c_md5_s[i]);
    a = t;
  }
  st[0] += a; st[1] += b; st[2] += c; st[3] += d;
}

// byte `o` .. `o + 3` of the plane's row stream (w is a multiple of 4, so a word never straddles rows)
__device__ __forceinline__ uint32_t stream_word(const uint8_t* p, int pitch, int w, long long o) {
  const int y = (int)(o / w), x = (int)(o - (long long)y * w);
  return *reinterpret_cast<const uint32_t*>(p + (size_t)y * pitch + x);
}

// one warp per (plane, picture)
__global__ void __launch_bounds__(32) k_hash_md5(const HashJob* __restrict__ jobs, uint8_t* __restrict__ out) {
  const HashJob j = jobs[blockIdx.y];
  const int c = blockIdx.x;
  if (!j.plane[c]) return;
  const int w = j.w[c], h = j.h[c], pitch = j.pitch[c];
  const uint8_t* p = j.plane[c];
  const long long total = (long long)w * h;
  __shared__ uint32_t buf[2][128];   // two chunks of 8 blocks
  const int lane = threadIdx.x;
  uint32_t st[4] = {0x67452301u, 0xefcdab89u, 0x98badcfeu, 0x10325476u};
  const long long n_full = total / 64;          // whole 64-byte blocks of data
  const long long n_chunks = (n_full + 7) / 8;
  auto load_chunk = [&](long long chunk, int slot) {
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int wi = lane + 32 * k;
      const long long o = chunk * 512 + 4 * wi;
      buf[slot][wi] = (o + 4 <= total) ? stream_word(p, pitch, w, o) : 0u;
    }
  };
  if (n_chunks) load_chunk(0, 0);
  __syncwarp();
  for (long long ch = 0; ch < n_chunks; ch++) {
    const int slot = (int)(ch & 1);
    if (ch + 1 < n_chunks) load_chunk(ch + 1, slot ^ 1);
    if (lane == 0) {
      const int nb = (int)min(8LL, n_full - ch * 8);
      for (int b = 0; b < nb; b++) md5_block(st, &buf[slot][16 * b]);
    }
    __syncwarp();
  }
  if (lane == 0) {   // tail: remaining bytes (a multiple of 4, < 64), 0x80, zero padding, bit length
    uint32_t m[16];
    const int rem = (int)(total - n_full * 64);
    for (int i = 0; i < 16; i++) m[i] = 0;
    for (int i = 0; i < rem / 4; i++) m[i] = stream_word(p, pitch, w, n_full * 64 + 4 * i);
    m[rem / 4] = 0x80u;
    if (rem >= 56) {
      md5_block(st, m);
      for (int i = 0; i < 16; i++) m[i] = 0;
    }
    const unsigned long long bits = (unsigned long long)total * 8;
    m[14] = (uint32_t)bits;
    m[15] = (uint32_t)(bits >> 32);
    md5_block(st, m);
    uint32_t* o = reinterpret_cast<uint32_t*>(out + ((size_t)blockIdx.y * 3 + c) * 16);
    for (int i = 0; i < 4; i++) o[i] = st[i];   // little-endian words = the 16 digest bytes in order
  }
}

// checksum: grid (x, 3, pictures); out word 0 of the plane's 16-byte slot (zeroed before the launch)
__global__ void __launch_bounds__(256) k_hash_checksum(const HashJob* __restrict__ jobs, uint8_t* __restrict__ out) {
  const HashJob j = jobs[blockIdx.z];
  const int c = blockIdx.y;
  if (!j.plane[c]) return;
  const int w = j.w[c], h = j.h[c], pitch = j.pitch[c];
  const int wq = w / 4;
  const long long n = (long long)wq * h;
  uint32_t sum = 0;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int y = (int)(i / wq), x = (int)(i - (long long)y * wq) * 4;
    const uint32_t v = *reinterpret_cast<const uint32_t*>(j.plane[c] + (size_t)y * pitch + x);
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int xx = x + k;
      const uint32_t mask = ((xx & 0xff) ^ (y & 0xff) ^ (xx >> 8) ^ (y >> 8)) & 0xff;
      sum += ((v >> (8 * k)) & 0xff) ^ mask;
    }
  }
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(B265_FULL_MASK, sum, o);
  if ((threadIdx.x & 31) == 0 && sum) atomicAdd(reinterpret_cast<uint32_t*>(out + ((size_t)blockIdx.z * 3 + c) * 16), sum);
}

__device__ __forceinline__ uint32_t crc_byte(uint32_t crc, uint32_t byte) {   // sei.cc:216-225
  const uint32_t s = (byte ^ (crc >> 8)) & 0xff;
  const uint32_t t = s ^ (s >> 4);
  return ((crc << 8) ^ t ^ (t << 5) ^ (t << 12)) & 0xffff;
}

// rows: grid (x, 3, pictures), one thread per row: register after the row's bytes starting from 0 -> row_crc[plane slot][y]
__global__ void __launch_bounds__(128) k_hash_crc_rows(const HashJob* __restrict__ jobs, uint16_t* __restrict__ row_crc, int max_h) {
  const HashJob j = jobs[blockIdx.z];
  const int c = blockIdx.y;
  if (!j.plane[c]) return;
  const int w = j.w[c], h = j.h[c], pitch = j.pitch[c];
  for (int y = blockIdx.x * blockDim.x + threadIdx.x; y < h; y += gridDim.x * blockDim.x) {
    const uint32_t* p = reinterpret_cast<const uint32_t*>(j.plane[c] + (size_t)y * pitch);
    uint32_t crc = 0;
    for (int i = 0; i < w / 4; i++) {
      const uint32_t v = p[i];
      crc = crc_byte(crc, v & 0xff);
      crc = crc_byte(crc, (v >> 8) & 0xff);
      crc = crc_byte(crc, (v >> 16) & 0xff);
      crc = crc_byte(crc, v >> 24);
    }
    row_crc[((size_t)blockIdx.z * 3 + c) * max_h + y] = (uint16_t)crc;
  }
}

// fold: one thread per (plane, picture)
__global__ void k_hash_crc_fold(const HashJob* __restrict__ jobs, const uint16_t* __restrict__ row_crc, int max_h, uint8_t* __restrict__ out) {
  const HashJob j = jobs[blockIdx.y];
  const int c = blockIdx.x;
  if (!j.plane[c] || threadIdx.x) return;
  const int w = j.w[c], h = j.h[c];
  // column k of M = register after shifting in one zero BYTE starting from 1 << k; M^(w) by repeated squaring
  uint16_t m[16], r[16];
  for (int k = 0; k < 16; k++) { m[k] = (uint16_t)crc_byte(1u << k, 0); r[k] = (uint16_t)(1u << k); }
  auto apply = [](const uint16_t* mat, uint32_t v) {
    uint32_t o = 0;
    for (int k = 0; k < 16; k++) if ((v >> k) & 1) o ^= mat[k];
    return o;
  };
  for (int e = w; e; e >>= 1) {
    uint16_t t[16];
    if (e & 1) { for (int k = 0; k < 16; k++) t[k] = (uint16_t)apply(m, r[k]); for (int k = 0; k < 16; k++) r[k] = t[k]; }
    for (int k = 0; k < 16; k++) t[k] = (uint16_t)apply(m, m[k]);
    for (int k = 0; k < 16; k++) m[k] = t[k];
  }
  uint32_t crc = 0xffff;
  crc = crc_byte(crc, 0);
  crc = crc_byte(crc, 0);
  const uint16_t* rows = row_crc + ((size_t)blockIdx.y * 3 + c) * max_h;
  for (int y = 0; y < h; y++) crc = apply(r, crc) ^ rows[y];
  *reinterpret_cast<uint32_t*>(out + ((size_t)blockIdx.y * 3 + c) * 16) = crc;
}

}  // namespace

int launch_hash(cudaStream_t st, int type, const HashJob* jobs, int n, int max_h, uint16_t* row_scratch, uint8_t* out) {
  if (n <= 0) return 0;
  cudaMemsetAsync(out, 0, (size_t)n * 48, st);
  if (type == 0) {
    k_hash_md5<<<dim3(3, n), 32, 0, st>>>(jobs, out);
    return 1;
  }
  if (type == 1) {
    k_hash_crc_rows<<<dim3((max_h + 127) / 128, 3, n), 128, 0, st>>>(jobs, row_scratch, max_h);
    k_hash_crc_fold<<<dim3(3, n), 32, 0, st>>>(jobs, row_scratch, max_h, out);
    return 2;
  }
  k_hash_checksum<<<dim3(64, 3, n), 256, 0, st>>>(jobs, out);
  return 1;
}

// ------------------------------------------------------------------------------------------------------------------
// b265_plane_crc_noalloc: the CRC of one plane for callers that must not allocate (code under CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Same arithmetic as k_hash_crc_rows / k_hash_crc_fold above, but the per-row scratch lives in a file-scope
// device array of fixed size instead of the context's growable buffer, and the job description travels as kernel
// arguments instead of through device memory.  Every GPU that loads the library holds the array until exit (16 KB).
//   * all calls share the array: calls that overlap on the GPU (different streams / contexts / threads) race;
//   * the array keeps the last call's contents: a call writes rows [0, height) before the fold reads them, nothing has
//     to start at zero.
// ------------------------------------------------------------------------------------------------------------------
#include <mutex>
#include <string>
#include "backend.h"

static __device__ __align__(256) uint16_t g_crc_row_scratch[B265_CRC_NOALLOC_MAX_ROWS];

namespace {

__global__ void __launch_bounds__(128) k_crc_rows_static(const uint8_t* __restrict__ plane, int pitch, int w, int h, uint16_t* __restrict__ rows) {
  for (int y = blockIdx.x * blockDim.x + threadIdx.x; y < h; y += gridDim.x * blockDim.x) {
    const uint32_t* p = reinterpret_cast<const uint32_t*>(plane + (size_t)y * pitch);
    uint32_t crc = 0;
    for (int i = 0; i < w / 4; i++) {
      const uint32_t v = p[i];
      crc = crc_byte(crc, v & 0xff);
      crc = crc_byte(crc, (v >> 8) & 0xff);
      crc = crc_byte(crc, (v >> 16) & 0xff);
      crc = crc_byte(crc, v >> 24);
    }
    rows[y] = (uint16_t)crc;
  }
}

__global__ void k_crc_fold_static(const uint16_t* __restrict__ rows, int w, int h, uint32_t* __restrict__ out) {
  if (threadIdx.x || blockIdx.x) return;
  uint16_t m[16], r[16];
  for (int k = 0; k < 16; k++) { m[k] = (uint16_t)crc_byte(1u << k, 0); r[k] = (uint16_t)(1u << k); }
  auto apply = [](const uint16_t* mat, uint32_t v) {
    uint32_t o = 0;
    for (int k = 0; k < 16; k++) if ((v >> k) & 1) o ^= mat[k];
    return o;
  };
  for (int e = w; e; e >>= 1) {
    uint16_t t[16];
    if (e & 1) { for (int k = 0; k < 16; k++) t[k] = (uint16_t)apply(m, r[k]); for (int k = 0; k < 16; k++) r[k] = t[k]; }
    for (int k = 0; k < 16; k++) t[k] = (uint16_t)apply(m, m[k]);
    for (int k = 0; k < 16; k++) m[k] = t[k];
  }
  uint32_t crc = 0xffff;
  crc = crc_byte(crc, 0);
  crc = crc_byte(crc, 0);
  for (int y = 0; y < h; y++) crc = apply(r, crc) ^ rows[y];
  *out = crc;
}

// In host code `g_crc_row_scratch` is a host-side stand-in, not a device address: the address a kernel can use comes from
// cudaGetSymbolAddress, once per device, with that device current.
std::mutex g_scratch_mu;
uint16_t* g_scratch_ptr[64] = {};

}  // namespace

extern "C" int b265_plane_crc_noalloc(b265_ctx* ctx, const uint8_t* dev_plane, int pitch, int width, int height, uint32_t* dev_crc) {
  if (!ctx || !dev_plane || !dev_crc || width <= 0 || height <= 0 || pitch < width) return B265_ERR_INVALID_ARG;
  if ((width & 3) || (pitch & 3) || (reinterpret_cast<uintptr_t>(dev_plane) & 3) || (reinterpret_cast<uintptr_t>(dev_crc) & 3)) {
    ctx->last_error = "b265_plane_crc_noalloc: plane address, pitch and width must be multiples of 4 (rows are read as 32-bit words)";
    return B265_ERR_INVALID_ARG;
  }
  if (height > B265_CRC_NOALLOC_MAX_ROWS) {
    ctx->last_error = "b265_plane_crc_noalloc: height " + std::to_string(height) + " exceeds the static scratch of " +
                      std::to_string(B265_CRC_NOALLOC_MAX_ROWS) + " rows (B265_CRC_NOALLOC_MAX_ROWS); use b265_surface_hash";
    return B265_ERR_UNSUPPORTED;
  }
  if (ctx->device < 0 || ctx->device >= 64) return B265_ERR_INVALID_ARG;
  CK(cudaSetDevice(ctx->device));
  uint16_t* rows = nullptr;
  {
    std::lock_guard<std::mutex> l(g_scratch_mu);
    if (!g_scratch_ptr[ctx->device]) {
      void* p = nullptr;
      CK(cudaGetSymbolAddress(&p, g_crc_row_scratch));
      g_scratch_ptr[ctx->device] = static_cast<uint16_t*>(p);
    }
    rows = g_scratch_ptr[ctx->device];
  }
  k_crc_rows_static<<<(height + 127) / 128, 128, 0, ctx->stream>>>(dev_plane, pitch, width, height, rows);
  k_crc_fold_static<<<1, 32, 0, ctx->stream>>>(rows, width, height, dev_crc);
  ctx->launches += 2;
  CK_LAUNCH("k_crc_rows_static + k_crc_fold_static");
  return B265_OK;
}
