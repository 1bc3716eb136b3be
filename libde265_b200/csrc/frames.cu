// frames.cu — frame-level C ABI: b265_upload_frames / b265_run_batch / b265_submit_frames.
//
// Host work per batch (cheap, linear in the number of TUs): split TUs into the inter-residual lists
// (by size) and the intra wavefront list, derive each intra TB's wave from a per-4x4 wave map, counting
// sort by (wave,size), cut into groups; pack every list of every frame into ONE pinned staging buffer
// and ship it with ONE cudaMemcpyAsync.  Device work: see b265_run_batch.
#include <string.h>
#include <algorithm>
#include <vector>
#include "backend.h"
#include "kernels.h"

struct b265_batch {
  int n_frames = 0;
  uint8_t* dev = nullptr;      // one allocation holding everything below
  size_t dev_bytes = 0;
  uint8_t* host = nullptr;     // pinned staging mirror of the uploaded part
  size_t upload_bytes = 0;
  // device pointers into dev
  FrameDev* d_frames = nullptr;
  uint32_t* d_mc_items = nullptr;     int n_mc = 0;
  uint32_t* d_res_order[4] = {};      int n_res[4] = {0, 0, 0, 0};
  uint32_t* d_intra_order = nullptr;  int n_intra = 0;
  TuGroup* d_groups = nullptr;        int n_groups = 0;
  unsigned* d_ticket = nullptr;
  uint8_t* d_zero_begin = nullptr;    size_t zero_bytes = 0;   // done maps (+ bs) cleared per run
  std::vector<FrameDev> frames;       // host copy (for post-filter launches)
  int max_wave = 0;
};

static inline size_t al(size_t v) { return (v + 255) & ~(size_t)255; }

namespace {
struct Packer {
  size_t off = 0;
  size_t take(size_t bytes) { size_t o = off; off = al(off + bytes); return o; }
};
}  // namespace

// wave of an intra TB = 1 + max wave among the units that hold its reference samples
static inline int tb_wave(std::vector<uint16_t>& map, int mw, int mh, int x, int y, int n) {
  const int ux = x >> 2, uy = y >> 2, nu = n >> 2;
  int w = 0;
  if (ux > 0) {
    const int y1 = std::min(mh, uy + 2 * nu);
    for (int v = std::max(0, uy - 1); v < y1; v++) w = std::max<int>(w, map[v * mw + ux - 1]);
  }
  if (uy > 0) {
    const int x1 = std::min(mw, ux + 2 * nu);
    for (int u = ux; u < x1; u++) w = std::max<int>(w, map[(uy - 1) * mw + u]);
  }
  w += 1;
  if (w > 65535) w = 65535;
  // only the rim is ever looked at by later TBs
  const int xr = std::min(mw, ux + nu) - 1, yb = std::min(mh, uy + nu) - 1;
  for (int u = ux; u <= xr; u++) map[yb * mw + u] = (uint16_t)w;
  for (int v = uy; v <= yb; v++) map[v * mw + xr] = (uint16_t)w;
  return w;
}

extern "C" void b265_free_batch(b265_ctx* ctx, b265_batch* b) {
  (void)ctx;
  if (!b) return;
  if (b->dev) cudaFree(b->dev);
  if (b->host) cudaFreeHost(b->host);
  delete b;
}

static int build_batch(b265_ctx* ctx, const b265_frame_job* jobs, int n_jobs, b265_batch* b) {
  if (n_jobs > 255) return B265_ERR_INVALID_ARG;
  b->n_frames = n_jobs;
  b->frames.assign(n_jobs, FrameDev());
  struct Key { uint32_t key; uint32_t item; };
  std::vector<Key> intra;
  std::vector<uint32_t> res[4], mc;
  std::vector<uint16_t> wmap_l, wmap_c;
  int max_wave = 0;
  // ---- pass 1: classify + waves ----
  for (int j = 0; j < n_jobs; j++) {
    const b265_frame_job& J = jobs[j];
    if (J.bit_depth != 8 || J.chroma_format_idc > 1 || J.n_tus >= (1 << 24) || J.n_pus >= (1 << 24))
      return B265_ERR_UNSUPPORTED;
    const int lw = (J.width + 3) >> 2, lh = (J.height + 3) >> 2;
    const int cw = (((J.width + 1) >> 1) + 3) >> 2, ch = (((J.height + 1) >> 1) + 3) >> 2;
    wmap_l.assign((size_t)lw * lh, 0);
    wmap_c.assign((size_t)cw * ch, 0);
    for (int i = 0; i < J.n_tus; i++) {
      const b265_tu& t = J.tus[i];
      const int c = t.cidx_log2 & 3, lg = (t.cidx_log2 >> 2) & 7;
      if (lg < 2 || lg > 5 || c > 2) return B265_ERR_INVALID_ARG;
      const uint32_t item = ((uint32_t)j << 24) | (uint32_t)i;
      if (t.intra_mode != B265_INTRA_NONE) {
        int w = (c == 0) ? tb_wave(wmap_l, lw, lh, t.x, t.y, 1 << lg)
                : (c == 1) ? tb_wave(wmap_c, cw, ch, t.x, t.y, 1 << lg)
                           : 0;
        if (c == 2) {  // Cr mirrors the Cb TB decoded just before it at the same position
          const int ux = std::min(cw, (t.x >> 2) + (1 << (lg - 2))) - 1, uy = std::min(ch, (t.y >> 2) + (1 << (lg - 2))) - 1;
          w = wmap_c[uy * cw + ux];
          if (w == 0) w = tb_wave(wmap_c, cw, ch, t.x, t.y, 1 << lg);
        }
        max_wave = std::max(max_wave, w);
        intra.push_back({((uint32_t)w << 2) | (uint32_t)(lg - 2), item});
      } else if (t.ncoeff > 0) {
        res[lg - 2].push_back(item);
      }
    }
    for (int i = 0; i < J.n_pus; i++) mc.push_back(((uint32_t)j << 24) | (uint32_t)i);
  }
  b->max_wave = max_wave;
  // ---- counting sort of the intra TBs by (wave,size); stable => decoding order inside a key ----
  std::vector<uint32_t> order(intra.size());
  std::vector<TuGroup> groups;
  {
    std::vector<uint32_t> cnt((size_t)(max_wave + 1) * 4 + 1, 0);
    for (auto& k : intra) cnt[k.key + 1]++;
    for (size_t i = 1; i < cnt.size(); i++) cnt[i] += cnt[i - 1];
    std::vector<uint32_t> pos(cnt.begin(), cnt.end() - 1);
    for (auto& k : intra) order[pos[k.key]++] = k.item;
    for (size_t key = 0; key + 1 < cnt.size(); key++) {
      const uint32_t lg = 2 + (key & 3), cap = 32u >> lg;
      for (uint32_t s = cnt[key]; s < cnt[key + 1]; s += cap)
        groups.push_back({s, (uint16_t)std::min(cap, cnt[key + 1] - s), (uint16_t)lg});
    }
  }
  b->n_intra = (int)order.size();
  b->n_groups = (int)groups.size();
  b->n_mc = (int)mc.size();
  for (int s = 0; s < 4; s++) b->n_res[s] = (int)res[s].size();

  // ---- layout ----
  Packer P;
  const size_t o_frames = P.take(sizeof(FrameDev) * n_jobs);
  const size_t o_mc = P.take(4 * mc.size());
  size_t o_res[4];
  for (int s = 0; s < 4; s++) o_res[s] = P.take(4 * res[s].size());
  const size_t o_order = P.take(4 * order.size());
  const size_t o_groups = P.take(sizeof(TuGroup) * groups.size());
  struct FOff { size_t tus, coeffs, pus, slices, ctbs, sao, unit, qp, scaling, pcms, pcm_samples; };
  std::vector<FOff> fo(n_jobs);
  for (int j = 0; j < n_jobs; j++) {
    const b265_frame_job& J = jobs[j];
    const int nctb = ((J.width + (1 << J.log2_ctb) - 1) >> J.log2_ctb) * ((J.height + (1 << J.log2_ctb) - 1) >> J.log2_ctb);
    fo[j].tus = P.take(sizeof(b265_tu) * (size_t)J.n_tus);
    fo[j].coeffs = P.take(sizeof(b265_coeff) * (size_t)J.n_coeffs);
    fo[j].pus = P.take(sizeof(b265_pu) * (size_t)J.n_pus);
    fo[j].slices = P.take(sizeof(b265_slice) * (size_t)std::max(1, J.n_slices));
    fo[j].ctbs = J.ctbs ? P.take(sizeof(b265_ctb) * (size_t)nctb) : 0;
    fo[j].sao = J.sao ? P.take(sizeof(b265_sao) * (size_t)nctb) : 0;
    fo[j].unit = J.unit_info ? P.take((size_t)(J.width >> 2) * (J.height >> 2)) : 0;
    fo[j].qp = J.qp_y ? P.take((size_t)(J.width >> J.log2_min_cb) * (J.height >> J.log2_min_cb)) : 0;
    fo[j].scaling = J.scaling ? P.take(sizeof(b265_scaling)) : 0;
    fo[j].pcms = J.n_pcms ? P.take(sizeof(b265_pcm) * (size_t)J.n_pcms) : 0;
    fo[j].pcm_samples = J.n_pcm_samples ? P.take((size_t)J.n_pcm_samples) : 0;
  }
  b->upload_bytes = P.off;
  // device-only scratch after the uploaded part
  const size_t o_ticket = P.take(256);
  const size_t o_zero = P.off;
  struct SOff { size_t done[3], bs, tmp[3], motion; int dp[3]; };
  std::vector<SOff> so(n_jobs);
  for (int j = 0; j < n_jobs; j++) {
    const b265_frame_job& J = jobs[j];
    const int cw = (J.width + 1) >> 1, ch = (J.height + 1) >> 1;
    so[j].dp[0] = (J.width + 3) >> 2;
    so[j].dp[1] = so[j].dp[2] = (cw + 3) >> 2;
    so[j].done[0] = P.take((size_t)so[j].dp[0] * ((J.height + 3) >> 2));
    so[j].done[1] = P.take((size_t)so[j].dp[1] * ((ch + 3) >> 2));
    so[j].done[2] = P.take((size_t)so[j].dp[2] * ((ch + 3) >> 2));
    so[j].bs = P.take((size_t)((J.width + 3) >> 2) * ((J.height + 3) >> 2));
  }
  b->zero_bytes = P.off - o_zero;
  for (int j = 0; j < n_jobs; j++) {  // SAO needs the deblocked picture kept aside
    const b265_frame_job& J = jobs[j];
    so[j].tmp[0] = so[j].tmp[1] = so[j].tmp[2] = 0;
    so[j].motion = (J.n_pus > 0 && (J.flags & B265_FRAME_DEBLOCK)) ? P.take((size_t)16 * ((J.width + 3) >> 2) * ((J.height + 3) >> 2)) : 0;
    if (J.flags & B265_FRAME_SAO) {
      const int pitch = (J.width + 255) & ~255, cpitch = (((J.width + 1) >> 1) + 127) & ~127;
      so[j].tmp[0] = P.take((size_t)pitch * J.height);
      if (J.chroma_format_idc) {
        so[j].tmp[1] = P.take((size_t)cpitch * ((J.height + 1) >> 1));
        so[j].tmp[2] = P.take((size_t)cpitch * ((J.height + 1) >> 1));
      }
    }
  }
  b->dev_bytes = P.off;
  CK(cudaMalloc((void**)&b->dev, b->dev_bytes));
  CK(cudaMallocHost((void**)&b->host, b->upload_bytes));
  uint8_t* H = b->host;
  uint8_t* D = b->dev;
  auto cp = [&](size_t off, const void* src, size_t bytes) { if (bytes && src) memcpy(H + off, src, bytes); };
  cp(o_mc, mc.data(), 4 * mc.size());
  for (int s = 0; s < 4; s++) cp(o_res[s], res[s].data(), 4 * res[s].size());
  cp(o_order, order.data(), 4 * order.size());
  cp(o_groups, groups.data(), sizeof(TuGroup) * groups.size());
  b->d_frames = (FrameDev*)(D + o_frames);
  b->d_mc_items = (uint32_t*)(D + o_mc);
  for (int s = 0; s < 4; s++) b->d_res_order[s] = (uint32_t*)(D + o_res[s]);
  b->d_intra_order = (uint32_t*)(D + o_order);
  b->d_groups = (TuGroup*)(D + o_groups);
  b->d_ticket = (unsigned*)(D + o_ticket);
  b->d_zero_begin = D + o_zero;
  static const b265_slice default_slice = {0, 2, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0, 0, {}, {}};
  for (int j = 0; j < n_jobs; j++) {
    const b265_frame_job& J = jobs[j];
    const int nctb_w = (J.width + (1 << J.log2_ctb) - 1) >> J.log2_ctb, nctb_h = (J.height + (1 << J.log2_ctb) - 1) >> J.log2_ctb;
    const int nctb = nctb_w * nctb_h;
    cp(fo[j].tus, J.tus, sizeof(b265_tu) * (size_t)J.n_tus);
    cp(fo[j].coeffs, J.coeffs, sizeof(b265_coeff) * (size_t)J.n_coeffs);
    cp(fo[j].pus, J.pus, sizeof(b265_pu) * (size_t)J.n_pus);
    if (J.n_slices > 0) cp(fo[j].slices, J.slices, sizeof(b265_slice) * (size_t)J.n_slices);
    else cp(fo[j].slices, &default_slice, sizeof(b265_slice));
    cp(fo[j].ctbs, J.ctbs, sizeof(b265_ctb) * (size_t)nctb);
    cp(fo[j].sao, J.sao, sizeof(b265_sao) * (size_t)nctb);
    cp(fo[j].unit, J.unit_info, (size_t)(J.width >> 2) * (J.height >> 2));
    cp(fo[j].qp, J.qp_y, (size_t)(J.width >> J.log2_min_cb) * (J.height >> J.log2_min_cb));
    cp(fo[j].scaling, J.scaling, sizeof(b265_scaling));
    cp(fo[j].pcms, J.pcms, sizeof(b265_pcm) * (size_t)J.n_pcms);
    cp(fo[j].pcm_samples, J.pcm_samples, (size_t)J.n_pcm_samples);
    FrameDev& f = b->frames[j];
    if (J.dst_surface < 0 || J.dst_surface >= (int)ctx->surfaces.size() || !ctx->surfaces[J.dst_surface].used)
      return B265_ERR_INVALID_ARG;
    const Surface& S = ctx->surfaces[J.dst_surface];
    if (S.width != J.width || S.height != J.height) return B265_ERR_INVALID_ARG;
    for (int c = 0; c < 3; c++) {
      f.out[c] = S.plane[c];
      f.pitch[c] = S.pitch[c];
      f.plane[c] = (so[j].tmp[c] && (J.flags & B265_FRAME_SAO)) ? D + so[j].tmp[c] : S.plane[c];
      f.done[c] = D + so[j].done[c];
      f.done_pitch[c] = so[j].dp[c];
    }
    f.width = J.width; f.height = J.height;
    f.cw = J.chroma_format_idc ? (J.width + 1) >> 1 : 0;
    f.ch = J.chroma_format_idc ? (J.height + 1) >> 1 : 0;
    f.log2_ctb = J.log2_ctb; f.log2_min_cb = J.log2_min_cb; f.log2_min_tb = J.log2_min_tb;
    f.ctbs_w = nctb_w; f.ctbs_h = nctb_h;
    f.flags = J.flags;
    f.cb_qp_offset = J.cb_qp_offset; f.cr_qp_offset = J.cr_qp_offset;
    f.tus = (const b265_tu*)(D + fo[j].tus);
    f.coeffs = (const b265_coeff*)(D + fo[j].coeffs);
    f.pus = (const b265_pu*)(D + fo[j].pus);
    f.slices = (const b265_slice*)(D + fo[j].slices);
    f.ctbs = J.ctbs ? (const b265_ctb*)(D + fo[j].ctbs) : nullptr;
    f.sao = J.sao ? (const b265_sao*)(D + fo[j].sao) : nullptr;
    f.unit_info = J.unit_info ? D + fo[j].unit : nullptr;
    f.qp_y = J.qp_y ? (const int8_t*)(D + fo[j].qp) : nullptr;
    f.scaling = J.scaling ? (const b265_scaling*)(D + fo[j].scaling) : nullptr;
    f.bs = D + so[j].bs;
    f.motion = so[j].motion ? (uint32_t*)(D + so[j].motion) : nullptr;
    f.n_refs = std::min(J.n_refs, 16);
    for (int r = 0; r < f.n_refs; r++) {
      const int sid = J.ref_surfaces[r];
      if (sid < 0 || sid >= (int)ctx->surfaces.size() || !ctx->surfaces[sid].used) return B265_ERR_INVALID_ARG;
      for (int c = 0; c < 3; c++) f.ref_plane[r][c] = ctx->surfaces[sid].plane[c];
      f.ref_poc[r] = J.ref_poc[r];
    }
  }
  memcpy(H + o_frames, b->frames.data(), sizeof(FrameDev) * n_jobs);
  CK(cudaMemcpyAsync(D, H, b->upload_bytes, cudaMemcpyHostToDevice, ctx->stream));
  return B265_OK;
}

// ------------------------------------------------------------------------------------------
// Device work for one batch, all on ctx->stream:
//   clear done/bs maps -> [pcm store] -> MC (all PUs) -> inter residual (4 sizes) -> intra wavefront
//   -> boundary strength -> deblock V -> deblock H -> SAO
// ------------------------------------------------------------------------------------------
static int run_batch(b265_ctx* ctx, b265_batch* b) {
  cudaStream_t st = ctx->stream;
  const bool tm = ctx->stage_timing;
  if (tm) cudaEventRecord(ctx->stage_ev[0], st);
  if (b->zero_bytes) CK(cudaMemsetAsync(b->d_zero_begin, 0, b->zero_bytes, st));
  if (b->n_mc) {
    launch_mc(st, b->d_frames, b->d_mc_items, b->n_mc, ctx->sm_count);
    ctx->launches++;
    CK_LAUNCH("k_mc");
  }
  if (tm) cudaEventRecord(ctx->stage_ev[1], st);
  for (int s = 0; s < 4; s++)
    if (b->n_res[s]) {
      launch_residual_sparse(st, 2 + s, b->d_frames, b->d_res_order[s], b->n_res[s], ctx->sm_count);
      ctx->launches++;
      CK_LAUNCH("k_residual_sparse");
    }
  if (tm) cudaEventRecord(ctx->stage_ev[2], st);
  if (b->n_groups) {
    static int max_ctas = 0;
    if (!max_ctas) max_ctas = intra_max_resident_ctas(ctx->sm_count);
    launch_intra(st, b->d_frames, b->d_intra_order, b->d_groups, b->n_groups, b->d_ticket, max_ctas);
    ctx->launches++;
    CK_LAUNCH("k_intra_wavefront");
  }
  if (tm) cudaEventRecord(ctx->stage_ev[3], st);
  int r = launch_postfilters(ctx, b->d_frames, b->frames.data(), b->n_frames, tm ? ctx->stage_ev[4] : nullptr,
                             b->d_mc_items, b->n_mc);
  if (r) return r;
  if (tm) {
    cudaEventRecord(ctx->stage_ev[5], st);
    CK(cudaEventSynchronize(ctx->stage_ev[5]));
    for (int i = 0; i < 5; i++) cudaEventElapsedTime(&ctx->stage_ms[i], ctx->stage_ev[i], ctx->stage_ev[i + 1]);
    cudaEventElapsedTime(&ctx->stage_ms[5], ctx->stage_ev[0], ctx->stage_ev[5]);
  }
  return B265_OK;
}

extern "C" {

int b265_upload_frames(b265_ctx* ctx, const b265_frame_job* jobs, int n_jobs, b265_batch** out) {
  if (!ctx || !jobs || n_jobs <= 0 || !out) return B265_ERR_INVALID_ARG;
  CK(cudaSetDevice(ctx->device));
  b265_batch* b = new b265_batch();
  int r = build_batch(ctx, jobs, n_jobs, b);
  if (r) { b265_free_batch(ctx, b); return r; }
  *out = b;
  return B265_OK;
}

int b265_run_batch(b265_ctx* ctx, b265_batch* batch) {
  if (!ctx || !batch) return B265_ERR_INVALID_ARG;
  CK(cudaSetDevice(ctx->device));
  return run_batch(ctx, batch);
}

int b265_submit_frames(b265_ctx* ctx, const b265_frame_job* jobs, int n_jobs) {
  if (!ctx || !jobs || n_jobs <= 0) return B265_ERR_INVALID_ARG;
  CK(cudaSetDevice(ctx->device));
  // the previous submission's staging must be drained before it is freed
  if (ctx->submit_batch) {
    CK(cudaStreamSynchronize(ctx->stream));
    b265_free_batch(ctx, ctx->submit_batch);
    ctx->submit_batch = nullptr;
  }
  b265_batch* b = new b265_batch();
  int r = build_batch(ctx, jobs, n_jobs, b);
  if (r) { b265_free_batch(ctx, b); return r; }
  ctx->submit_batch = b;
  return run_batch(ctx, b);
}

int b265_set_stage_timing(b265_ctx* ctx, int enable) {
  if (!ctx) return B265_ERR_INVALID_ARG;
  ctx->stage_timing = enable != 0;
  return B265_OK;
}

int b265_get_stage_times(b265_ctx* ctx, float ms[6]) {
  if (!ctx || !ms) return B265_ERR_INVALID_ARG;
  for (int i = 0; i < 6; i++) ms[i] = ctx->stage_ms[i];
  return B265_OK;
}

}  // extern "C"
