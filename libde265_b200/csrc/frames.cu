// frames.cu — frame-level C ABI: b265_upload_frames / b265_run_batch / b265_submit_frames.
//
// Host work per batch (cheap, linear in the number of TUs): split TUs into the inter-residual lists
// (by size) and note, per CTB, its TU range and which planes hold intra TBs (CtbRec; the wavefront's
// dependency counters and ready queue are built on the device); pack every list of every frame into
// ONE pinned staging buffer and ship it with ONE cudaMemcpyAsync.  Device work: see b265_run_batch.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <vector>
#include "backend.h"
#include "kernels.h"

// A batch may hold pictures that reference earlier pictures of the same batch (P/B pictures in decoding order): it is
// cut into LAUNCH GROUPS of consecutive pictures that do not reference each other; every kernel stage runs once per
// group, the groups follow each other on the stream without a host round trip.  All-intra batches are one group.
struct LaunchGroup {
  int first = 0, n = 0;               // pictures [first, first + n) of the batch; list items carry frame indices relative to `first`
  size_t mc0 = 0, mc1 = 0;            // ranges of the batch-wide lists that belong to this group
  size_t res0[4] = {}, res1[4] = {}, ires0[4] = {}, ires1[4] = {};
  int n_res_plain[4] = {0, 0, 0, 0}, n_ires_plain[4] = {0, 0, 0, 0};   // leading part of each range: plain DCT TUs (sizes 16, 32)
  int n_tickets = 0;                  // (plane, CTB) units with intra TBs
  int n_runs = 0;                     // runs of horizontally adjacent such units: the wavefront's queue entries
  int max_slots = 0;                  // largest planes * CTBs of a picture (grid of the scheduling pre-pass)
  int max_rows = 0;                   // largest CTB row count of a picture
  int max_pcms = 0, max_tus = 0;
  int log2_ctb = 4;                   // largest CTB (sizes the wavefront's shared-memory tiles)
};

struct b265_batch {
  int n_frames = 0;
  uint8_t* dev = nullptr;      // one allocation holding everything below
  size_t dev_bytes = 0;
  uint8_t* host = nullptr;     // pinned staging mirror of the uploaded part
  size_t upload_bytes = 0;
  // device pointers into dev
  FrameDev* d_frames = nullptr;
  uint32_t* d_mc_items = nullptr;
  uint32_t* d_res_order[4] = {};      // inter TUs (residual added in place)
  uint32_t* d_ires_order[4] = {};     // coded intra TUs (residual to scratch)
  std::vector<LaunchGroup> groups;
  int n_tickets = 0;
  IntraSched* d_sched = nullptr;      // wavefront control block and row queue: start of the zeroed region, re-cleared per group
  unsigned long long* d_queue = nullptr;
  unsigned queue_entries = 0;         // power of two
  size_t sched_bytes = 0;             // control block + queue
  uint8_t* d_zero_begin = nullptr;    size_t zero_bytes = 0;   // control block, queue, border mailboxes: cleared per run
  std::vector<FrameDev> frames;       // host copy (for post-filter launches)
  bool pooled = false;                // dev/host belong to the context's grow-only staging (b265_submit_frames)
};

static inline size_t al(size_t v) { return (v + 255) & ~(size_t)255; }

namespace {
struct Packer {
  size_t off = 0;
  size_t take(size_t bytes) { size_t o = off; off = al(off + bytes); return o; }
};
}  // namespace

extern "C" void b265_free_batch(b265_ctx* ctx, b265_batch* b) {
  (void)ctx;
  if (!b) return;
  if (b->dev && !b->pooled) cudaFree(b->dev);
  if (b->host && !b->pooled) cudaFreeHost(b->host);
  delete b;
}

// argument validation failures are reported with their line when B265_DEBUG is set
static int reject(int line, int code) {
  static const bool verbose = getenv("B265_DEBUG") != nullptr;
  if (verbose) fprintf(stderr, "b265: frames.cu:%d rejects the job (error %d)\n", line, code);
  return code;
}

#include <chrono>
static double t_now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
static const bool g_build_timing = getenv("B265_BUILD_TIMING") != nullptr;   // developer aid: host time of build_batch by section

static int build_batch(b265_ctx* ctx, const b265_frame_job* jobs_in, int n_jobs, b265_batch* b) {
  const double tb0 = t_now();
  if (n_jobs > 4096) return reject(__LINE__, B265_ERR_INVALID_ARG);
  // Pictures are reordered by DEPENDENCY LEVEL inside the batch (level = 1 + the highest level among the pictures of this
  // batch it references; stable otherwise): all pictures of a level form one launch group, so the I pictures of a batch -
  // and every picture whose references were finished by earlier submits - share their launches instead of paying the
  // latency of the intra wavefront one by one.  Destination surfaces of a batch are distinct and none is recycled inside
  // it, so the order among independent pictures is free.
  std::vector<b265_frame_job> ordered(jobs_in, jobs_in + n_jobs);
  {
    std::vector<int> level(n_jobs, 0), idx(n_jobs);
    for (int j = 0; j < n_jobs; j++) {
      idx[j] = j;
      const b265_frame_job& J = jobs_in[j];
      for (int r = 0; r < J.n_refs && r < 16; r++)
        for (int i = 0; i < j; i++)
          if (J.ref_surfaces[r] == jobs_in[i].dst_surface) level[j] = std::max(level[j], level[i] + 1);
    }
    std::stable_sort(idx.begin(), idx.end(), [&](int a, int c) { return level[a] < level[c]; });
    for (int j = 0; j < n_jobs; j++) ordered[j] = jobs_in[idx[j]];
  }
  const b265_frame_job* jobs = ordered.data();
  b->n_frames = n_jobs;
  b->frames.assign(n_jobs, FrameDev());
  std::vector<uint32_t> res[4], ires[4], mc;
  std::vector<uint32_t> n_pieces(n_jobs, 0);
  std::vector<std::vector<CtbRec>> ctb_recs(n_jobs);   // per frame, raster order
  std::vector<uint32_t> ctb_area;   // samples covered by intra TBs, per plane and CTB
  std::vector<uint8_t> need_resid(n_jobs, 0);
  std::vector<LaunchGroup>& groups = b->groups;
  groups.clear();
  auto close_group = [&](LaunchGroup& g) {
    g.mc1 = mc.size();
    for (int s = 0; s < 4; s++) { g.res1[s] = res[s].size(); g.ires1[s] = ires[s].size(); }
  };
  auto open_group = [&](int first) {
    LaunchGroup g;
    g.first = first;
    g.mc0 = mc.size();
    for (int s = 0; s < 4; s++) { g.res0[s] = res[s].size(); g.ires0[s] = ires[s].size(); }
    groups.push_back(g);
  };
  // ---- pass 1: classify; TUs arrive in decoding order, so the TUs of one CTB are contiguous ----
  for (int j = 0; j < n_jobs; j++) {
    const b265_frame_job& J = jobs[j];
    if (J.bit_depth != 8 || J.chroma_format_idc > 1 || J.n_tus >= (1 << 24) || J.n_pus >= (1 << 24))
      return reject(__LINE__, B265_ERR_UNSUPPORTED);
    if (J.n_refs < 0 || J.n_refs > 16) return reject(__LINE__, B265_ERR_INVALID_ARG);
    {  // a picture that reads a picture of the current group starts the next group; so does the 256th picture of a group
      // (list items name their picture in 8 bits, relative to the group)
      bool dependent = !groups.empty() && j - groups.back().first >= 255;
      if (!groups.empty())
        for (int r = 0; r < J.n_refs && !dependent; r++)
          for (int i = groups.back().first; i < j; i++)
            if (J.ref_surfaces[r] == jobs[i].dst_surface) { dependent = true; break; }
      if (groups.empty() || dependent) {
        if (!groups.empty()) close_group(groups.back());
        open_group(j);
      }
    }
    LaunchGroup& G = groups.back();
    G.n = j - G.first + 1;
    const uint32_t jrel = (uint32_t)(j - G.first);
    const int cw_ = (J.width + (1 << J.log2_ctb) - 1) >> J.log2_ctb, ch_ = (J.height + (1 << J.log2_ctb) - 1) >> J.log2_ctb;
    const int nctb = cw_ * ch_;
    std::vector<CtbRec>& recs = ctb_recs[j];
    recs.assign(nctb, CtbRec{0, 0, 0, 0});
    ctb_area.assign((size_t)3 * nctb, 0);
    if (cw_ > 1023 || ch_ > 1023) return reject(__LINE__, B265_ERR_UNSUPPORTED);   // queue entries: 10 bits per CTB coordinate
    if (J.log2_ctb < 4 || J.log2_ctb > 6) return reject(__LINE__, B265_ERR_INVALID_ARG);
    G.max_slots = std::max(G.max_slots, (J.chroma_format_idc ? 3 : 1) * nctb);
    G.max_rows = std::max(G.max_rows, ch_);
    G.log2_ctb = std::max(G.log2_ctb, (int)J.log2_ctb);
    G.max_tus = std::max(G.max_tus, J.n_tus);
    G.max_pcms = std::max(G.max_pcms, J.n_pcms);
    // FrameDev holds 16 reference slots (a DPB of 16 pictures is the HEVC maximum, A.4.2); slice indices travel in 8 bits
    if (J.n_slices > 256) return reject(__LINE__, B265_ERR_INVALID_ARG);
    int cur = -1;
    for (int i = 0; i < J.n_tus; i++) {
      const b265_tu& t = J.tus[i];
      const int c = t.cidx_log2 & 3, lg = (t.cidx_log2 >> 2) & 7;
      if (lg < 2 || lg > 5 || c > 2) return reject(__LINE__, B265_ERR_INVALID_ARG);
      if ((uint64_t)t.coeff_off + t.ncoeff > (uint64_t)J.n_coeffs) return reject(__LINE__, B265_ERR_INVALID_ARG);
      const int sh = (c && J.chroma_format_idc) ? 1 : 0;
      const int bx = (t.x << sh) >> J.log2_ctb, by = (t.y << sh) >> J.log2_ctb;
      if (bx >= cw_ || by >= ch_) return reject(__LINE__, B265_ERR_INVALID_ARG);
      const int ctb = by * cw_ + bx;
      if (ctb != cur) {
        if (recs[ctb].tu_count) return reject(__LINE__, B265_ERR_INVALID_ARG);  // TUs of a CTB must be contiguous (decoding order)
        recs[ctb].tu_first = (uint32_t)i;
        cur = ctb;
      }
      if (recs[ctb].tu_count == 0xffff) return reject(__LINE__, B265_ERR_INVALID_ARG);
      recs[ctb].tu_count++;
      if (t.intra_mode != B265_INTRA_NONE) {
        if (c && !J.chroma_format_idc) return reject(__LINE__, B265_ERR_INVALID_ARG);
        recs[ctb].has |= (uint8_t)(1 << c);
        ctb_area[(size_t)c * nctb + ctb] += 1u << (2 * lg);
        if (t.ncoeff > 0) { ires[lg - 2].push_back((jrel << 24) | (uint32_t)i); need_resid[j] = 1; }
      } else if (t.ncoeff > 0) res[lg - 2].push_back((jrel << 24) | (uint32_t)i);
    }
    for (int ctb = 0; ctb < nctb; ctb++) {
      if (!recs[ctb].has) continue;
      const int S = 1 << J.log2_ctb;
      const int bw = std::min(S, J.width - (ctb % cw_) * S), bh = std::min(S, J.height - (ctb / cw_) * S);
      for (int c = 0; c < 3; c++)
        if (recs[ctb].has & (1 << c)) {
          G.n_tickets++;
          if (ctb % cw_ == 0 || !(recs[ctb - 1].has & (1 << c))) G.n_runs++;
          // a CTB plane not fully covered by intra TBs also holds inter / PCM samples written by earlier launches
          const uint32_t full = (c && J.chroma_format_idc) ? (uint32_t)(((bw + 1) >> 1) * ((bh + 1) >> 1)) : (uint32_t)(bw * bh);
          if (ctb_area[(size_t)c * nctb + ctb] != full) recs[ctb].mixed |= (uint8_t)(1 << c);
        }
    }
    {  // one MC item per <= 16 x 8 piece of a prediction block (pu_expand)
      uint32_t n = 0;
      const int n_sl = std::max(1, J.n_slices);
      for (int i = 0; i < J.n_pus; i++) {
        const b265_pu& pu = J.pus[i];
        // a slot or slice index outside the job's tables would become a wild device pointer in k_mc
        if (pu.slice >= n_sl || !(pu.flags & (B265_PU_L0 | B265_PU_L1)) || pu.w == 0 || pu.h == 0 || pu.w > 64 || pu.h > 64 ||
            ((pu.flags & B265_PU_L0) && (pu.ref_slot[0] >= J.n_refs || (uint8_t)pu.ref_idx[0] >= 16)) ||
            ((pu.flags & B265_PU_L1) && (pu.ref_slot[1] >= J.n_refs || (uint8_t)pu.ref_idx[1] >= 16)))
          return reject(__LINE__, B265_ERR_INVALID_ARG);
        for (int k = pu_pieces(pu); k > 0; k--) mc.push_back((jrel << 24) | n++);
      }
      if (n >= (1u << 24)) return reject(__LINE__, B265_ERR_UNSUPPORTED);
      n_pieces[j] = n;
    }
  }
  const double tb1 = t_now();
  close_group(groups.back());
  int n_tickets = 0, max_runs = 0;
  for (LaunchGroup& g : groups) {
    n_tickets += g.n_tickets;
    max_runs = std::max(max_runs, g.n_runs);
    // 16x16 / 32x32 lists: plain DCT TUs first (packed dp2a kernels), transform-skip / bypass TUs after them (generic kernel)
    auto plain = [&](uint32_t o) { return !(jobs[g.first + (o >> 24)].tus[o & 0xffffff].flags & (B265_TU_SKIP | B265_TU_BYPASS)); };
    for (int s = 2; s < 4; s++) {
      g.n_res_plain[s] = (int)(std::stable_partition(res[s].begin() + g.res0[s], res[s].begin() + g.res1[s], plain) - (res[s].begin() + g.res0[s]));
      g.n_ires_plain[s] = (int)(std::stable_partition(ires[s].begin() + g.ires0[s], ires[s].begin() + g.ires1[s], plain) - (ires[s].begin() + g.ires0[s]));
    }
  }
  b->n_tickets = n_tickets;

  const double tb2 = t_now();
  // ---- layout ----
  Packer P;
  const size_t o_frames = P.take(sizeof(FrameDev) * n_jobs);
  const size_t o_mc = P.take(4 * mc.size());
  size_t o_res[4];
  for (int s = 0; s < 4; s++) o_res[s] = P.take(4 * res[s].size());
  size_t o_ires[4];
  for (int s = 0; s < 4; s++) o_ires[s] = P.take(4 * ires[s].size());
  struct FOff { size_t tus, coeffs, pus, slices, ctbs, ctb_rec, sao, sao_rec, unit, qp, scaling, pcms, pcm_samples; };
  std::vector<FOff> fo(n_jobs);
  for (int j = 0; j < n_jobs; j++) {
    const b265_frame_job& J = jobs[j];
    const int nctb = ((J.width + (1 << J.log2_ctb) - 1) >> J.log2_ctb) * ((J.height + (1 << J.log2_ctb) - 1) >> J.log2_ctb);
    fo[j].tus = P.take(sizeof(b265_tu) * (size_t)J.n_tus);
    fo[j].coeffs = P.take(sizeof(b265_coeff) * (size_t)J.n_coeffs);
    fo[j].pus = P.take(sizeof(b265_pu) * (size_t)n_pieces[j]);
    fo[j].slices = P.take(sizeof(b265_slice) * (size_t)std::max(1, J.n_slices));
    fo[j].ctbs = J.ctbs ? P.take(sizeof(b265_ctb) * (size_t)nctb) : 0;
    fo[j].ctb_rec = P.take(sizeof(CtbRec) * (size_t)nctb);
    fo[j].sao = J.sao ? P.take(sizeof(b265_sao) * (size_t)nctb) : 0;
    fo[j].sao_rec = (J.sao && (J.flags & B265_FRAME_SAO)) ? P.take(sizeof(uint2) * 3 * (size_t)nctb) : 0;
    fo[j].unit = J.unit_info ? P.take((size_t)(J.width >> 2) * (J.height >> 2)) : 0;
    fo[j].qp = J.qp_y ? P.take((size_t)(J.width >> J.log2_min_cb) * (J.height >> J.log2_min_cb)) : 0;
    fo[j].scaling = J.scaling ? P.take(sizeof(b265_scaling)) : 0;
    fo[j].pcms = J.n_pcms ? P.take(sizeof(b265_pcm) * (size_t)J.n_pcms) : 0;
    fo[j].pcm_samples = J.n_pcm_samples ? P.take((size_t)J.n_pcm_samples) : 0;
  }
  b->upload_bytes = P.off;
  // device-only scratch after the uploaded part
  const size_t o_zero = P.off;
  static int queue_slack = 0;
  if (!queue_slack) queue_slack = intra_queue_slack(ctx->sm_count);
  const size_t o_sched = P.take(sizeof(IntraSched));
  unsigned queue_entries = 1024;
  while ((size_t)queue_entries < (size_t)max_runs + (size_t)queue_slack) queue_entries <<= 1;
  b->queue_entries = queue_entries;
  const size_t o_queue = P.take(sizeof(unsigned long long) * (size_t)queue_entries);
  b->sched_bytes = P.off - o_zero;
  struct SOff { size_t nbinfo, tmp[3], motion, resid[3], tb_mask, mbox[3]; int slot_base; };
  std::vector<SOff> so(n_jobs);
  {
    int slots = 0;
    for (int j = 0; j < n_jobs; j++) {
      const b265_frame_job& J = jobs[j];
      const int cw_ = (J.width + (1 << J.log2_ctb) - 1) >> J.log2_ctb, ch_ = (J.height + (1 << J.log2_ctb) - 1) >> J.log2_ctb;
      so[j].slot_base = slots;
      slots += 3 * cw_ * ch_;
      for (int c = 0; c < 3; c++) {
        // border mailbox of the intra wavefront: per CTB 2*S samples in flag-in-data form (8 B = 4 samples + tag)
        const int S = (1 << J.log2_ctb) >> ((c && J.chroma_format_idc) ? 1 : 0);
        so[j].mbox[c] = (n_tickets && (c == 0 || J.chroma_format_idc)) ? P.take((size_t)cw_ * ch_ * 4 * S) : 0;
      }
    }
  }
  b->zero_bytes = P.off - o_zero;
  for (int j = 0; j < n_jobs; j++) {   // usable-neighbour masks: written by k_intra_sched_prep, no clearing needed
    const b265_frame_job& J = jobs[j];
    const int cw_ = (J.width + (1 << J.log2_ctb) - 1) >> J.log2_ctb, ch_ = (J.height + (1 << J.log2_ctb) - 1) >> J.log2_ctb;
    so[j].nbinfo = P.take(3 * (size_t)cw_ * ch_);
  }
  for (int j = 0; j < n_jobs; j++) {  // SAO needs the deblocked picture kept aside
    const b265_frame_job& J = jobs[j];
    so[j].tmp[0] = so[j].tmp[1] = so[j].tmp[2] = 0;
    so[j].resid[0] = so[j].resid[1] = so[j].resid[2] = 0;
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
  {  // int16 residual planes of the coded intra TBs (same geometry as the destination surface)
    const std::vector<uint8_t>& need = need_resid;
    for (int j = 0; j < n_jobs; j++) {
      so[j].tb_mask = P.take((size_t)8 * jobs[j].n_tus);
      if (!need[j]) continue;
      const b265_frame_job& J = jobs[j];
      if (J.dst_surface < 0 || J.dst_surface >= (int)ctx->surfaces.size() || !ctx->surfaces[J.dst_surface].used)
        return reject(__LINE__, B265_ERR_INVALID_ARG);
      const Surface& S = ctx->surfaces[J.dst_surface];
      // block-linear (common.cuh resid_block): whole CTBs, S*S values each
      const size_t nctb = (size_t)((J.width + (1 << J.log2_ctb) - 1) >> J.log2_ctb) * ((J.height + (1 << J.log2_ctb) - 1) >> J.log2_ctb);
      (void)S;
      so[j].resid[0] = P.take((size_t)2 * (nctb << (2 * J.log2_ctb)));
      if (J.chroma_format_idc) {
        so[j].resid[1] = P.take((size_t)2 * (nctb << (2 * J.log2_ctb - 2)));
        so[j].resid[2] = P.take((size_t)2 * (nctb << (2 * J.log2_ctb - 2)));
      }
    }
  }
  b->dev_bytes = P.off;
  if (b->pooled) {  // reuse the context's staging (the previous submission has been drained by the caller)
    int r = devbuf_reserve(ctx, &ctx->d_submit, b->dev_bytes, false);
    if (r) return r;
    r = devbuf_reserve(ctx, &ctx->h_submit, b->upload_bytes, true);
    if (r) return r;
    b->dev = (uint8_t*)ctx->d_submit.p;
    b->host = (uint8_t*)ctx->h_submit.p;
  } else {
    CK(cudaMalloc((void**)&b->dev, b->dev_bytes));
    CK(cudaMallocHost((void**)&b->host, b->upload_bytes));
  }
  const double tb3 = t_now();
  uint8_t* H = b->host;
  uint8_t* D = b->dev;
  auto cp = [&](size_t off, const void* src, size_t bytes) { if (bytes && src) memcpy(H + off, src, bytes); };
  cp(o_mc, mc.data(), 4 * mc.size());
  for (int s = 0; s < 4; s++) cp(o_res[s], res[s].data(), 4 * res[s].size());
  for (int s = 0; s < 4; s++) cp(o_ires[s], ires[s].data(), 4 * ires[s].size());
  b->d_frames = (FrameDev*)(D + o_frames);
  b->d_mc_items = (uint32_t*)(D + o_mc);
  for (int s = 0; s < 4; s++) b->d_res_order[s] = (uint32_t*)(D + o_res[s]);
  for (int s = 0; s < 4; s++) b->d_ires_order[s] = (uint32_t*)(D + o_ires[s]);
  b->d_sched = (IntraSched*)(D + o_sched);
  b->d_queue = (unsigned long long*)(D + o_queue);
  b->d_zero_begin = D + o_zero;
  static const b265_slice default_slice = {0, 2, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0, 0, {}, {}};
  for (int j = 0; j < n_jobs; j++) {
    const b265_frame_job& J = jobs[j];
    const int nctb_w = (J.width + (1 << J.log2_ctb) - 1) >> J.log2_ctb, nctb_h = (J.height + (1 << J.log2_ctb) - 1) >> J.log2_ctb;
    const int nctb = nctb_w * nctb_h;
    cp(fo[j].tus, J.tus, sizeof(b265_tu) * (size_t)J.n_tus);
    cp(fo[j].coeffs, J.coeffs, sizeof(b265_coeff) * (size_t)J.n_coeffs);
    {
      b265_pu* out = reinterpret_cast<b265_pu*>(H + fo[j].pus);
      for (int i = 0; i < J.n_pus; i++) out += pu_expand(J.pus[i], out);
    }
    if (J.n_slices > 0) cp(fo[j].slices, J.slices, sizeof(b265_slice) * (size_t)J.n_slices);
    else cp(fo[j].slices, &default_slice, sizeof(b265_slice));
    cp(fo[j].ctbs, J.ctbs, sizeof(b265_ctb) * (size_t)nctb);
    cp(fo[j].ctb_rec, ctb_recs[j].data(), sizeof(CtbRec) * (size_t)nctb);
    cp(fo[j].sao, J.sao, sizeof(b265_sao) * (size_t)nctb);
    if (J.sao && (J.flags & B265_FRAME_SAO)) sao_records(J, nctb_w, nctb_h, false, reinterpret_cast<uint2*>(H + fo[j].sao_rec));
    cp(fo[j].unit, J.unit_info, (size_t)(J.width >> 2) * (J.height >> 2));
    cp(fo[j].qp, J.qp_y, (size_t)(J.width >> J.log2_min_cb) * (J.height >> J.log2_min_cb));
    cp(fo[j].scaling, J.scaling, sizeof(b265_scaling));
    cp(fo[j].pcms, J.pcms, sizeof(b265_pcm) * (size_t)J.n_pcms);
    cp(fo[j].pcm_samples, J.pcm_samples, (size_t)J.n_pcm_samples);
    FrameDev& f = b->frames[j];
    if (J.dst_surface < 0 || J.dst_surface >= (int)ctx->surfaces.size() || !ctx->surfaces[J.dst_surface].used)
      return reject(__LINE__, B265_ERR_INVALID_ARG);
    const Surface& S = ctx->surfaces[J.dst_surface];
    if (S.width != J.width || S.height != J.height) return reject(__LINE__, B265_ERR_INVALID_ARG);
    for (int c = 0; c < 3; c++) {
      f.out[c] = S.plane[c];
      f.pitch[c] = S.pitch[c];
      f.plane[c] = (so[j].tmp[c] && (J.flags & B265_FRAME_SAO)) ? D + so[j].tmp[c] : S.plane[c];
      f.mbox[c] = so[j].mbox[c] ? D + so[j].mbox[c] : nullptr;
      f.resid[c] = so[j].resid[c] ? (int16_t*)(D + so[j].resid[c]) : nullptr;
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
    f.ctb_rec = (const CtbRec*)(D + fo[j].ctb_rec);
    f.nbinfo = D + so[j].nbinfo;
    f.slot_base = so[j].slot_base;
    f.sao = J.sao ? (const b265_sao*)(D + fo[j].sao) : nullptr;
    f.sao_rec = (J.sao && (J.flags & B265_FRAME_SAO)) ? (const uint2*)(D + fo[j].sao_rec) : nullptr;
    f.unit_info = J.unit_info ? D + fo[j].unit : nullptr;
    f.qp_y = J.qp_y ? (const int8_t*)(D + fo[j].qp) : nullptr;
    f.scaling = J.scaling ? (const b265_scaling*)(D + fo[j].scaling) : nullptr;
    f.pcms = J.n_pcms ? (const b265_pcm*)(D + fo[j].pcms) : nullptr;
    f.pcm_samples = J.n_pcm_samples ? D + fo[j].pcm_samples : nullptr;
    f.n_pcms = J.n_pcms;
    f.n_tus = J.n_tus;
    f.tb_mask = (uint64_t*)(D + so[j].tb_mask);
    f.bs = nullptr;   // boundary strengths are derived inside the deblocking kernels
    f.motion = so[j].motion ? (uint32_t*)(D + so[j].motion) : nullptr;
    f.n_refs = J.n_refs;
    for (int r = 0; r < f.n_refs; r++) {
      const int sid = J.ref_surfaces[r];
      if (sid < 0 || sid >= (int)ctx->surfaces.size() || !ctx->surfaces[sid].used) return reject(__LINE__, B265_ERR_INVALID_ARG);
      for (int c = 0; c < 3; c++) f.ref_plane[r][c] = ctx->surfaces[sid].plane[c];
      f.ref_poc[r] = J.ref_poc[r];
    }
  }
  memcpy(H + o_frames, b->frames.data(), sizeof(FrameDev) * n_jobs);
  const double tb4 = t_now();
  CK(cudaMemcpyAsync(D, H, b->upload_bytes, cudaMemcpyHostToDevice, ctx->stream));
  if (g_build_timing)
    fprintf(stderr, "b265 build_batch n=%d (%zu launch groups): classify %.2f ms, list partition %.2f, layout+staging %.2f, copy %.2f, memcpyAsync call %.2f (%.1f MB)\n",
            n_jobs, groups.size(), (tb1 - tb0) * 1e3, (tb2 - tb1) * 1e3, (tb3 - tb2) * 1e3, (tb4 - tb3) * 1e3, (t_now() - tb4) * 1e3, b->upload_bytes / 1e6);
  return B265_OK;
}

// ------------------------------------------------------------------------------------------
// Device work for one batch, all on ctx->stream; per launch group (pictures that do not reference each other):
//   clear border mailboxes -> [pcm store] -> MC (all PUs) -> inter residual (4 sizes, in place)
//   -> intra residual (4 sizes, to the int16 scratch planes) -> intra wavefront
//   -> deblock V -> deblock H -> SAO
// ------------------------------------------------------------------------------------------
static int run_batch(b265_ctx* ctx, b265_batch* b) {
  cudaStream_t st = ctx->stream;
  const bool tm = ctx->stage_timing;
  if (tm) for (int i = 0; i < 6; i++) ctx->stage_ms[i] = 0;
  if (b->zero_bytes) CK(cudaMemsetAsync(b->d_zero_begin, 0, b->zero_bytes, st));
  const int busy_peers = b265_busy_peers(ctx);   // other contexts with a batch on the device right now
  for (size_t gi = 0; gi < b->groups.size(); gi++) {
    const LaunchGroup& g = b->groups[gi];
    const FrameDev* frames = b->d_frames + g.first;
    const uint32_t* mc_items = b->d_mc_items + g.mc0;
    const int n_mc = (int)(g.mc1 - g.mc0);
    if (tm) cudaEventRecord(ctx->stage_ev[0], st);
    if (g.max_pcms) {
      launch_pcm(st, frames, g.n, g.max_pcms);
      ctx->launches++;
      CK_LAUNCH("k_pcm");
    }
    if (n_mc) {
      launch_mc(st, frames, mc_items, n_mc, ctx->sm_count);
      ctx->launches++;
      CK_LAUNCH("k_mc");
    }
    if (tm) cudaEventRecord(ctx->stage_ev[1], st);
    for (int pass = 0; pass < 2; pass++) {   // inter TUs in place, then intra TBs to the int16 scratch planes
      const bool scratch = pass == 1;
      for (int s = 0; s < 4; s++) {
        const uint32_t* order = scratch ? b->d_ires_order[s] + g.ires0[s] : b->d_res_order[s] + g.res0[s];
        const int n = (int)(scratch ? g.ires1[s] - g.ires0[s] : g.res1[s] - g.res0[s]);
        const int n_plain = scratch ? g.n_ires_plain[s] : g.n_res_plain[s];
        if (n_plain) {
          launch_residual_sparse_fast(st, 2 + s, frames, order, n_plain, ctx->sm_count, scratch);
          ctx->launches++;
          CK_LAUNCH("k_residual_sparse_fast");
        }
        if (n > n_plain) {
          launch_residual_sparse(st, 2 + s, frames, order + n_plain, n - n_plain, ctx->sm_count, scratch);
          ctx->launches++;
          CK_LAUNCH("k_residual_sparse");
        }
      }
    }
    if (g.n_tickets) {
      if (gi > 0) CK(cudaMemsetAsync(b->d_sched, 0, b->sched_bytes, st));   // control block + queue are shared by the groups
      launch_intra_prep(st, frames, g.n, g.max_tus, g.max_slots, b->d_sched);
      ctx->launches++;
      CK_LAUNCH("k_intra_prep");
    }
    if (tm) cudaEventRecord(ctx->stage_ev[2], st);   // stage 1 = all residual launches + availability masks
    if (g.n_tickets) {
      launch_intra(st, frames, g.n, g.max_slots, g.max_rows, g.n_runs, b->d_sched, b->d_queue, b->queue_entries,
                   ctx->sm_count, g.log2_ctb, busy_peers);
      ctx->launches += 2;
      CK_LAUNCH("k_intra_sched_prep + k_intra_wavefront");
    }
    if (tm) cudaEventRecord(ctx->stage_ev[3], st);
    int r = launch_postfilters(ctx, frames, b->frames.data() + g.first, g.n, tm ? ctx->stage_ev[4] : nullptr, mc_items, n_mc);
    if (r) return r;
    if (tm) {   // per-stage times are summed over the groups (a host sync per group: timing runs only)
      cudaEventRecord(ctx->stage_ev[5], st);
      CK(cudaEventSynchronize(ctx->stage_ev[5]));
      float ms = 0;
      for (int i = 0; i < 5; i++) { cudaEventElapsedTime(&ms, ctx->stage_ev[i], ctx->stage_ev[i + 1]); ctx->stage_ms[i] += ms; }
      cudaEventElapsedTime(&ms, ctx->stage_ev[0], ctx->stage_ev[5]);
      ctx->stage_ms[5] += ms;
    }
  }
  CK(cudaEventRecord(ctx->ev_busy, st));
  ctx->last_batch_ns.store(b265_now_ns(), std::memory_order_relaxed);
  ctx->busy_recorded.store(true, std::memory_order_release);
  return B265_OK;
}

extern "C" {

int b265_upload_frames(b265_ctx* ctx, const b265_frame_job* jobs, int n_jobs, b265_batch** out) {
  if (!ctx || !jobs || n_jobs <= 0 || !out) return reject(__LINE__, B265_ERR_INVALID_ARG);
  CK(cudaSetDevice(ctx->device));
  b265_batch* b = new b265_batch();
  int r = build_batch(ctx, jobs, n_jobs, b);
  if (r) { b265_free_batch(ctx, b); return r; }
  *out = b;
  return B265_OK;
}

int b265_run_batch(b265_ctx* ctx, b265_batch* batch) {
  if (!ctx || !batch) return reject(__LINE__, B265_ERR_INVALID_ARG);
  CK(cudaSetDevice(ctx->device));
  return run_batch(ctx, batch);
}

int b265_submit_frames(b265_ctx* ctx, const b265_frame_job* jobs, int n_jobs) {
  if (!ctx || !jobs || n_jobs <= 0) return reject(__LINE__, B265_ERR_INVALID_ARG);
  CK(cudaSetDevice(ctx->device));
  // the previous submission's staging must be drained before it is freed
  if (ctx->submit_batch) {
    { int rs = b265_sync(ctx); if (rs) return rs; }
    b265_free_batch(ctx, ctx->submit_batch);
    ctx->submit_batch = nullptr;
  }
  b265_batch* b = new b265_batch();
  b->pooled = true;
  int r = build_batch(ctx, jobs, n_jobs, b);
  if (r) { b265_free_batch(ctx, b); return r; }
  ctx->submit_batch = b;
  return run_batch(ctx, b);
}

int b265_set_stage_timing(b265_ctx* ctx, int enable) {
  if (!ctx) return reject(__LINE__, B265_ERR_INVALID_ARG);
  ctx->stage_timing = enable != 0;
  return B265_OK;
}

int b265_get_stage_times(b265_ctx* ctx, float ms[6]) {
  if (!ctx || !ms) return reject(__LINE__, B265_ERR_INVALID_ARG);
  for (int i = 0; i < 6; i++) ms[i] = ctx->stage_ms[i];
  return B265_OK;
}

}  // extern "C"
