// decoder.cc — the libde265 decoder API (include/de265.h) on top of the host parser and the b265
// CUDA back end.  Counterpart of the reference's de265.cc (API shims :171-702), nal-parser.cc
// (start-code splitting :262-446), decctx.cc (decode_NAL :1228-1330, process_slice_segment_header
// :1940-2110, POC :1532-1580, output queue) and dpb.cc — re-designed as a pipeline:
//
//   API thread:   bytes -> NAL units -> parameter sets / slice headers -> pictures (slice units)
//   parse pool:   one picture per worker: CABAC + syntax -> work lists (slice_decoder.cc)
//   recon thread: consecutive parsed pictures -> one b265_submit_frames() batch -> D2H into pinned
//                 output planes -> output queue (POC order)
//
// so the CPU parses picture N+k while the GPU reconstructs picture N.  There is no CPU pixel path:
// if the CUDA context cannot be created every decode call fails with
// DE265_ERROR_LIBRARY_INITIALIZATION_FAILED.
#include <stdio.h>
#include <string.h>
#include <stdlib.h>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>
#include <unordered_map>
#include "../../../include/b265.h"
#include "../../../include/de265.h"
#include "picture.h"
#include "dpb.h"

using namespace b265h;

struct de265_image {
  Picture pic;
  enum State { ASSEMBLING, QUEUED, PARSED, TAKEN, DONE } state = ASSEMBLING;
  double parsed_at = 0;
  int conf_x[3] = {0, 0, 0}, conf_y[3] = {0, 0, 0}, conf_w[3] = {0, 0, 0}, conf_h[3] = {0, 0, 0};
  size_t host_bytes = 0;
  uint8_t* host_base = nullptr;
};

namespace {

int g_device = -1;
bool g_parse_only = false;  // test/bench aid: build work lists, never touch the GPU, keep the lists
int g_init_count = 0;
std::mutex g_init_mutex;

double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

struct NalIn {
  std::vector<uint8_t> data;  // NAL unit bytes incl. 2-byte header, still escaped
  int64_t pts;
  void* user;
};

struct Decoder {
  std::vector<b265_ctx*> gpus;   // one CUDA context (own stream + staging) per reconstruction thread
  b265_ctx* gpu = nullptr;       // gpus[0]
  int gpu_status = 0;
  // ---- input ----
  std::vector<uint8_t> inbuf;
  int64_t in_pts = 0;
  void* in_user = nullptr;
  std::deque<NalIn> nals;
  bool end_of_stream = false;
  ParamSets ps;
  // ---- picture assembly ----
  std::unique_ptr<de265_image> cur;
  SliceHeader prev_indep;
  bool have_prev_indep = false;
  int prev_tid0_poc_lsb = 0, prev_tid0_poc_msb = 0;
  bool first_picture = true, after_eos = true;
  uint64_t next_decode_index = 0;
  // ---- reference pictures ----
  Dpb dpb;
  uint64_t next_uid = 1, gop_id = 0;
  bool last_no_rasl_output = false;              // the last IRAP picture emptied the DPB
  bool skip_rasl = false;
  std::unordered_map<uint64_t, int> gop_ctx;     // reconstruction context the pictures of a GOP are pinned to (mu)
  std::mutex rel_mu;
  std::vector<std::pair<int, int>> to_release;   // (context, surface) of reference pictures nobody needs any more
  static void recycle_cb(void* owner, int ctx, int surface) {
    Decoder* d = static_cast<Decoder*>(owner);
    std::lock_guard<std::mutex> l(d->rel_mu);
    d->to_release.emplace_back(ctx, surface);
  }
  // ---- pipeline ----
  std::mutex mu;
  std::condition_variable cv_parse, cv_recon, cv_done;
  std::deque<de265_image*> parse_queue;
  std::deque<std::unique_ptr<de265_image>> inflight;  // decoding order; front = oldest not yet reconstructed
  std::vector<std::thread> workers;
  std::vector<std::thread> recon_threads;
  bool draining = false;         // the API thread waits for the pipeline to empty: do not hold back partial batches
  double batch_wait_s = 0.004;
  bool stop = false;
  bool parse_only = false;
  int n_threads = 0;
  int max_batch = 16;
  // ---- output ----
  std::deque<std::unique_ptr<de265_image>> reorder;  // reconstructed, waiting for output order
  std::deque<std::unique_ptr<de265_image>> output;
  std::unique_ptr<de265_image> last_returned;
  std::vector<std::pair<size_t, uint8_t*>> host_pool;
  std::deque<int> warnings;
  int fatal_error = 0;
  // ---- parameters ----
  bool disable_deblocking = false, disable_sao = false, check_hash = false, suppress_faulty = false;
  int limit_tid = 6;
  de265_image_allocation alloc = {nullptr, nullptr};
  void* alloc_user = nullptr;
  // ---- stats ----
  double stat_parse_s = 0, stat_gpu_s = 0, stat_h2d = 0, stat_d2h = 0;
  uint64_t stat_pictures = 0, stat_batches = 0;
  double stat_submit_s = 0;  // part of stat_gpu_s spent inside b265_submit_frames (host list building + enqueue)
  uint64_t stat_hashed = 0;  // pictures whose SEI hash was verified on the device
  uint8_t* hash_host[4] = {nullptr, nullptr, nullptr, nullptr};   // per reconstruction context: pinned digests
  size_t hash_bytes[4] = {0, 0, 0, 0};

  Decoder() {
    int dev = g_device;
    if (dev < 0) {
      const char* e = getenv("B265_DEVICE");
      dev = e ? atoi(e) : 0;
    }
    parse_only = g_parse_only;
    int n_ctx = 3;   // measured on 16 cores, 4K: 2 -> 808, 3 -> 859, 4 -> 654 pictures/s (host list building overlaps the GPU)
    {   // few host cores: the reconstruction threads must not crowd out the parse threads
      const unsigned hw = std::thread::hardware_concurrency();
      if (hw && hw <= 4) n_ctx = 1; else if (hw && hw <= 8) n_ctx = 2;
    }
    if (const char* e = getenv("B265_RECON_CONTEXTS")) n_ctx = std::max(1, std::min(4, atoi(e)));
    if (parse_only) n_ctx = 1;
    for (int i = 0; i < n_ctx; i++) {
      b265_ctx* c = nullptr;
      if (!parse_only) {
        int st = b265_create(dev, &c);
        if (st != B265_OK) { gpu_status = st; c = nullptr; }
      }
      gpus.push_back(c);
    }
    gpu = gpus[0];
    dpb.make_unavailable = [this](int poc, bool lt) {
      auto r = std::make_shared<RefPic>();
      r->poc = poc;
      r->is_lt = lt;
      r->unavailable = true;
      r->uid = next_uid++;
      r->owner = this;
      r->recycle = &Decoder::recycle_cb;
      if (cur) { r->u_w = cur->pic.u_w; r->u_h = cur->pic.u_h; r->ctbs_w = cur->pic.sps->ctbs_w; r->log2_ctb = cur->pic.sps->log2_ctb; }
      r->parsed.store(1, std::memory_order_release);   // empty motion field = all intra (fill_pred_mode(MODE_INTRA))
      return r;
    };
    for (int i = 0; i < n_ctx; i++) recon_threads.emplace_back([this, i] { recon_loop(i); });
  }
  ~Decoder() {
    {
      std::lock_guard<std::mutex> l(mu);
      stop = true;
    }
    cv_parse.notify_all();
    cv_recon.notify_all();
    for (auto& t : workers) t.join();
    for (auto& t : recon_threads) t.join();
    dpb.clear();
    auto drop = [this](std::unique_ptr<de265_image>& i) { if (i) { release_image(*i); i->pic.ref_pics.clear(); i->pic.self.reset(); } };
    for (auto& i : inflight) drop(i);
    for (auto& i : reorder) drop(i);
    for (auto& i : output) drop(i);
    drop(last_returned);
    for (auto& p : host_pool) { if (gpu) b265_host_free(gpu, p.second); }
    for (size_t i = 0; i < gpus.size() && i < 4; i++) if (hash_host[i] && gpus[i]) b265_host_free(gpus[i], hash_host[i]);
    for (b265_ctx* c : gpus) if (c) b265_destroy(c);
  }

  void warn(int w) {
    std::lock_guard<std::mutex> l(mu);
    if (warnings.size() < 20) warnings.push_back(w);
  }

  // ---- host output buffers (pinned, pooled) ----
  uint8_t* host_get(size_t bytes) {
    {
      std::lock_guard<std::mutex> l(mu);
      for (size_t i = 0; i < host_pool.size(); i++)
        if (host_pool[i].first == bytes) {
          uint8_t* p = host_pool[i].second;
          host_pool.erase(host_pool.begin() + i);
          return p;
        }
    }
    void* p = nullptr;
    if (b265_host_alloc(gpu, bytes, &p) != B265_OK) return nullptr;
    return (uint8_t*)p;
  }
  void release_image(de265_image& img) {
    if (img.pic.planes_external) {
      if (alloc.release_buffer) alloc.release_buffer(this, &img, alloc_user);
    } else if (img.host_base) {
      std::lock_guard<std::mutex> l(mu);
      host_pool.emplace_back(img.host_bytes, img.host_base);
    }
    img.host_base = nullptr;
  }

  void start_workers(int n) {
    std::lock_guard<std::mutex> l(mu);
    for (int i = (int)workers.size(); i < n; i++) workers.emplace_back([this] { parse_loop(); });
    n_threads = (int)workers.size();
  }

  void parse_one(de265_image* img) {
    const double t0 = now_s();
    int err = decode_picture_slices(img->pic);
    if (err && getenv("B265_DEBUG")) fprintf(stderr, "b265: slice data of picture %llu: error %d\n", (unsigned long long)img->pic.decode_index, err);
    img->pic.error = err;
    img->pic.parse_seconds = now_s() - t0;
    img->pic.release_parse_state();
  }

  void parse_loop() {
    for (;;) {
      de265_image* img = nullptr;
      {
        std::unique_lock<std::mutex> l(mu);
        cv_parse.wait(l, [this] { return stop || !parse_queue.empty(); });
        if (stop) return;
        img = parse_queue.front();
        parse_queue.pop_front();
      }
      parse_one(img);
      {
        std::lock_guard<std::mutex> l(mu);
        img->state = de265_image::PARSED;
        img->parsed_at = now_s();
        stat_parse_s += img->pic.parse_seconds;
      }
      cv_recon.notify_all();
    }
  }

  // hand the assembled picture to the pipeline
  void dispatch_current() {
    if (!cur) return;
    de265_image* img = cur.get();
    img->pic.decode_index = next_decode_index++;
    img->pic.self->decode_index = img->pic.decode_index;
    bool inline_parse;
    {
      std::lock_guard<std::mutex> l(mu);
      inline_parse = workers.empty();
      img->state = de265_image::QUEUED;
      inflight.push_back(std::move(cur));
      if (!inline_parse) parse_queue.push_back(img);
    }
    if (inline_parse) {
      parse_one(img);
      std::lock_guard<std::mutex> l(mu);
      img->state = de265_image::PARSED;
      img->parsed_at = now_s();
      stat_parse_s += img->pic.parse_seconds;
      cv_recon.notify_all();
    } else {
      cv_parse.notify_one();
    }
  }

  void fill_job(de265_image& img, b265_frame_job& J) {
    Picture& p = img.pic;
    const Sps& s = *p.sps;
    const Pps& pp = *p.pps;
    memset(&J, 0, sizeof(J));
    J.width = s.width; J.height = s.height;
    J.chroma_format_idc = (uint8_t)s.chroma_array_type;
    J.bit_depth = 8;
    J.log2_ctb = (uint8_t)s.log2_ctb; J.log2_min_cb = (uint8_t)s.log2_min_cb; J.log2_min_tb = (uint8_t)s.log2_min_tb;
    J.cb_qp_offset = (int8_t)pp.cb_qp_offset; J.cr_qp_offset = (int8_t)pp.cr_qp_offset;
    uint32_t fl = 0;
    if (s.strong_intra_smoothing) fl |= B265_FRAME_STRONG_INTRA_SMOOTHING;
    if (pp.constrained_intra_pred) fl |= B265_FRAME_CONSTRAINED_INTRA;
    if (pp.lf_across_tiles) fl |= B265_FRAME_LF_ACROSS_TILES;
    if (p.any_deblock && !disable_deblocking) fl |= B265_FRAME_DEBLOCK;
    if (p.any_sao && !disable_sao) fl |= B265_FRAME_SAO;
    if (s.scaling_list_enabled) fl |= B265_FRAME_SCALING_LIST;
    if (s.pcm_lf_disabled) fl |= B265_FRAME_PCM_LF_DISABLE;
    J.flags = fl;
    J.dst_surface = p.surface;
    J.n_refs = (int)p.ref_pics.size();
    for (int i = 0; i < J.n_refs; i++) {
      // parse-only clients get the pictures' unique ids instead of surfaces (tests reconstruct with the CPU oracle)
      J.ref_surfaces[i] = parse_only ? (int32_t)p.ref_pics[i]->uid : p.ref_pics[i]->surface;
      J.ref_poc[i] = p.ref_pics[i]->poc;
    }
    if (parse_only) J.dst_surface = (int32_t)p.self->uid;
    J.tus = p.tus.data(); J.n_tus = (int)p.tus.size();
    J.coeffs = p.coeffs.data(); J.n_coeffs = (int)p.coeffs.size();
    J.pus = p.pus.data(); J.n_pus = (int)p.pus.size();
    J.pcms = p.pcms.data(); J.n_pcms = (int)p.pcms.size();
    J.pcm_samples = p.pcm_samples.data(); J.n_pcm_samples = (int)p.pcm_samples.size();
    J.slices = p.slices.data(); J.n_slices = (int)p.slices.size();
    J.ctbs = p.ctbs.data();
    J.sao = (fl & B265_FRAME_SAO) ? p.sao.data() : nullptr;
    J.unit_info = p.unit_info.data();
    J.qp_y = p.qp_y.data();
    J.scaling = p.scaling ? p.scaling.get() : nullptr;
  }

  bool prepare_output_planes(de265_image& img) {
    Picture& p = img.pic;
    const Sps& s = *p.sps;
    const int nplanes = s.chroma_array_type ? 3 : 1;
    const int sub = s.chroma_array_type == 1 ? 2 : 1;
    for (int c = 0; c < 3; c++) {
      const int sw = c ? sub : 1;
      img.conf_x[c] = s.conf_win[0] * (c ? 1 : sub) ;
      img.conf_y[c] = s.conf_win[2] * (c ? 1 : sub);
      const int pw = c ? (s.width + sw - 1) / sw : s.width, ph = c ? (s.height + sw - 1) / sw : s.height;
      img.conf_w[c] = pw - (s.conf_win[0] + s.conf_win[1]) * (c ? 1 : sub);
      img.conf_h[c] = ph - (s.conf_win[2] + s.conf_win[3]) * (c ? 1 : sub);
      if (c >= nplanes) img.conf_w[c] = img.conf_h[c] = 0;
    }
    if (alloc.get_buffer) {
      de265_image_spec spec;
      memset(&spec, 0, sizeof(spec));
      spec.format = s.chroma_array_type ? de265_image_format_YUV420P8 : de265_image_format_mono8;
      spec.width = s.width; spec.height = s.height; spec.alignment = 16;
      spec.crop_left = img.conf_x[0]; spec.crop_top = img.conf_y[0];
      spec.crop_right = s.width - img.conf_x[0] - img.conf_w[0];
      spec.crop_bottom = s.height - img.conf_y[0] - img.conf_h[0];
      spec.visible_width = img.conf_w[0]; spec.visible_height = img.conf_h[0];
      p.planes_external = true;
      if (!alloc.get_buffer(this, &spec, &img, alloc_user)) return false;
      return p.host_plane[0] != nullptr;
    }
    const int ystride = (s.width + 15) & ~15, cstride = ((s.width + 1) / 2 + 15) & ~15;
    const size_t ysz = (size_t)ystride * s.height, csz = nplanes == 3 ? (size_t)cstride * ((s.height + 1) / 2) : 0;
    img.host_bytes = ysz + 2 * csz + 64;
    img.host_base = host_get(img.host_bytes);
    if (!img.host_base) return false;
    p.host_plane[0] = img.host_base; p.host_stride[0] = ystride;
    if (nplanes == 3) {
      p.host_plane[1] = img.host_base + ysz; p.host_plane[2] = img.host_base + ysz + csz;
      p.host_stride[1] = p.host_stride[2] = cstride;
    }
    return true;
  }

  // ---- reconstruction threads (one per CUDA context/stream): batches of consecutive parsed pictures ->
  //      GPU -> host planes; pictures retire to the reorder buffer strictly in decoding order ----
  // A batch is taken when it is full, when the API thread is draining (end of stream / reset / back
  // pressure), or when the oldest parsed picture has waited longer than `batch_wait_s`: the intra
  // wavefront kernel is latency-bound, so reconstructing a few pictures costs as much as a full batch.
  void recon_loop(int ci) {
    b265_ctx* gpu = gpus[ci];
    std::vector<de265_image*> batch;
    std::vector<b265_frame_job> jobs;
    for (;;) {
      batch.clear();
      {
        std::unique_lock<std::mutex> l(mu);
        for (;;) {
          if (stop) return;
          // first picture not yet taken by a recon thread that this context may take: pictures of a GOP pinned to another
          // context are skipped - no reference crosses a GOP boundary, so the next GOP can be reconstructed on this
          // context's stream while the other context is still busy with its own (pictures retire in decoding order anyway)
          size_t first = 0;
          while (first < inflight.size() &&
                 (inflight[first]->state >= de265_image::TAKEN ||
                  [&] { auto it = gop_ctx.find(inflight[first]->pic.gop); return it != gop_ctx.end() && it->second != ci; }()))
            first++;
          size_t n = 0;
          // Reference pictures live in the surfaces of ONE context: the pictures between two DPB-emptying pictures
          // (a "GOP") are pinned to the context that takes the first of them.  A batch may hold a picture together with
          // its references: b265_submit_frames cuts it into launch groups that follow each other on the stream, so a
          // run of P/B pictures costs one list upload and one host synchronisation, not one per picture.
          auto pinned_elsewhere = [&](const de265_image& im) {
            auto it = gop_ctx.find(im.pic.gop);
            return it != gop_ctx.end() && it->second != ci;
          };
          bool wrong_ctx = false;
          while (first + n < inflight.size() && (int)n < max_batch && inflight[first + n]->state == de265_image::PARSED &&
                 inflight[first + n]->pic.width == inflight[first]->pic.width &&
                 inflight[first + n]->pic.height == inflight[first]->pic.height) {
            if (pinned_elsewhere(*inflight[first + n])) { wrong_ctx = true; break; }
            n++;
          }
          if (wrong_ctx && n == 0) { cv_recon.wait(l); continue; }   // the other context's pictures come first
          const bool blocked = wrong_ctx ||
                               (first + n < inflight.size() && inflight[first + n]->state == de265_image::PARSED);  // geometry change
          bool take = n > 0 && ((int)n >= max_batch || draining || blocked || parse_only);
          if (!take && n > 0) {
            const double waited = now_s() - inflight[first]->parsed_at;
            if (waited >= batch_wait_s) take = true;
            else { cv_recon.wait_for(l, std::chrono::duration<double>(batch_wait_s - waited)); continue; }
          }
          if (take) {
            for (size_t i = 0; i < n; i++) {
              gop_ctx.emplace(inflight[first + i]->pic.gop, ci);
              inflight[first + i]->state = de265_image::TAKEN;
              batch.push_back(inflight[first + i].get());
            }
            break;
          }
          cv_recon.wait(l);
        }
      }
      const double t0 = now_s();
      jobs.resize(batch.size());
      std::vector<de265_image*> good;
      int njobs = 0;
      double d2h = 0, h2d = 0, t_submit = 0;
      if (!parse_only && gpu) {   // surfaces of reference pictures that left the DPB
        std::vector<std::pair<int, int>> rel;
        {
          std::lock_guard<std::mutex> rl(rel_mu);
          for (size_t i = 0; i < to_release.size();)
            if (to_release[i].first == ci) { rel.push_back(to_release[i]); to_release.erase(to_release.begin() + i); }
            else i++;
        }
        for (auto& r : rel) b265_surface_release(gpu, r.second);
      }
      for (de265_image* img : batch) {
        Picture& p = img->pic;
        if (parse_only) continue;
        if (p.error || gpu_status != B265_OK) {
          if (!p.error) p.error = DE265_ERROR_LIBRARY_INITIALIZATION_FAILED;
          continue;
        }
        if (b265_surface_alloc(gpu, p.width, p.height, p.sps->chroma_array_type ? 1 : 0, &p.surface) != B265_OK) {
          p.error = DE265_ERROR_OUT_OF_MEMORY;
          continue;
        }
        // a reference that failed (or was never reconstructed) cannot be read
        bool refs_ok = true;
        for (const auto& r : p.ref_pics) {
          if (r && r->unavailable && r->surface < 0) {   // generate_unavailable_reference_picture: 1 << (bit depth - 1) everywhere
            int sid = -1;
            if (b265_surface_alloc(gpu, p.width, p.height, p.sps->chroma_array_type ? 1 : 0, &sid) == B265_OK &&
                b265_surface_fill(gpu, sid, 128, 128) == B265_OK) {
              r->ctx = ci;
              r->surface = sid;
            } else if (sid >= 0) b265_surface_release(gpu, sid);
          }
          refs_ok &= r && r->surface >= 0 && r->ctx == ci;
        }
        if (!refs_ok) {
          b265_surface_release(gpu, p.surface);
          p.surface = -1;
          p.error = DE265_ERROR_UNSPECIFIED_DECODING_ERROR;
          continue;
        }
        p.self->ctx = ci;
        p.self->surface = p.surface;   // the picture in HBM now belongs to the DPB entry (recycled when it dies)
        fill_job(*img, jobs[njobs++]);
        good.push_back(img);
      }
      if (njobs) {
        const double ts0 = now_s();
        int r = b265_submit_frames(gpu, jobs.data(), njobs);
        t_submit = now_s() - ts0;
        for (de265_image* img : good) {
          Picture& p = img->pic;
          if (r == B265_OK && prepare_output_planes(*img)) {
            const int np = p.sps->chroma_array_type ? 3 : 1;
            for (int c = 0; c < np && r == B265_OK; c++)
              r = b265_surface_download(gpu, p.surface, c, p.host_plane[c], p.host_stride[c], 0);
            d2h += (double)p.width * p.height * (np == 3 ? 1.5 : 1.0);
          } else if (r == B265_OK) {
            p.error = DE265_ERROR_OUT_OF_MEMORY;
          }
          h2d += (double)(p.tus.size() * sizeof(b265_tu) + p.coeffs.size() * sizeof(b265_coeff) + p.unit_info.size() +
                          p.qp_y.size() + p.sao.size() * sizeof(b265_sao) + p.ctbs.size() * sizeof(b265_ctb));
        }
        // decoded picture hash (DE265_DECODER_PARAM_BOOL_SEI_CHECK_HASH): computed on the device from the picture in HBM,
        // 48 bytes per picture come back; pictures that are not output are not checked (sei.cc:277-283)
        std::vector<de265_image*> hashed[3];
        if (r == B265_OK && check_hash) {
          for (de265_image* img : good)
            if (img->pic.hash_type >= 0 && img->pic.pic_output && !img->pic.error) hashed[img->pic.hash_type].push_back(img);
          size_t total = hashed[0].size() + hashed[1].size() + hashed[2].size(), at = 0;
          if (total * 48 > hash_bytes[ci]) {
            if (hash_host[ci]) b265_host_free(gpu, hash_host[ci]);
            void* hp = nullptr;
            hash_bytes[ci] = b265_host_alloc(gpu, total * 48 + 48 * 64, &hp) == B265_OK ? total * 48 + 48 * 64 : 0;
            hash_host[ci] = (uint8_t*)hp;
          }
          for (int t = 0; t < 3 && r == B265_OK && hash_host[ci]; t++) {
            if (hashed[t].empty()) continue;
            std::vector<int> ids;
            for (de265_image* img : hashed[t]) ids.push_back(img->pic.surface);
            r = b265_surface_hash(gpu, ids.data(), (int)ids.size(), t, hash_host[ci] + at * 48);
            at += ids.size();
          }
        }
        if (r == B265_OK) r = b265_sync(gpu);
        if (r == B265_OK && check_hash && hash_host[ci]) {
          size_t at = 0;
          for (int t = 0; t < 3; t++)
            for (de265_image* img : hashed[t]) {
              const uint8_t* got = hash_host[ci] + 48 * at++;
              const int planes = img->pic.sps->chroma_format_idc == 0 ? 1 : 3, per = t == 0 ? 16 : (t == 1 ? 2 : 4);
              for (int c = 0; c < planes; c++)
                if (memcmp(got + 16 * c, img->pic.hash[c], per)) {
                  fprintf(stderr, "SEI decoded picture hash mismatch (POC=%d, plane %d)\n", img->pic.poc, c);
                  img->pic.error = DE265_ERROR_CHECKSUM_MISMATCH;
                  break;
                }
              std::lock_guard<std::mutex> l(mu);
              stat_hashed++;
            }
        }
        for (de265_image* img : good) {
          if (r != B265_OK && getenv("B265_DEBUG")) fprintf(stderr, "b265: back end status %d: %s\n", r, b265_last_error(gpu));
          if (r != B265_OK) img->pic.error = r == B265_ERR_OUT_OF_MEMORY ? DE265_ERROR_OUT_OF_MEMORY : DE265_ERROR_UNSPECIFIED_DECODING_ERROR;
          // (a hash mismatch keeps the picture: it is delivered and the error is reported by de265_decode, like decctx.cc:770-779)
          img->pic.surface = -1;
          img->pic.ref_pics.clear();
          img->pic.self.reset();
        }
      }
      {
        std::lock_guard<std::mutex> l(mu);
        stat_gpu_s += now_s() - t0;
        stat_d2h += d2h;
        stat_h2d += h2d;
        stat_batches++;
        stat_submit_s += t_submit;
        for (de265_image* img : batch) {
          img->state = de265_image::DONE;
          if (!parse_only) {  // work lists are no longer needed
            img->pic.tus.release();
            img->pic.coeffs.release();
            std::vector<uint8_t>().swap(img->pic.unit_info);
          }
        }
        // retire in decoding order
        while (!inflight.empty() && inflight.front()->state == de265_image::DONE) {
          std::unique_ptr<de265_image> up = std::move(inflight.front());
          inflight.pop_front();
          stat_pictures++;
          if (up->pic.error && up->pic.error < 1000 && !fatal_error) fatal_error = up->pic.error;
          const uint64_t gop_done = up->pic.gop;
          push_reorder(std::move(up));
          // context pinning of GOPs nobody can belong to any more (pictures retire in decoding order, GOP ids grow)
          const uint64_t oldest = inflight.empty() ? gop_done : inflight.front()->pic.gop;
          for (auto it = gop_ctx.begin(); it != gop_ctx.end();)
            if (it->first < oldest) it = gop_ctx.erase(it); else ++it;
        }
      }
      cv_done.notify_all();
      cv_recon.notify_all();
    }
  }

  // output order (C.5.2.2 "bumping"): called with mu held
  void push_reorder(std::unique_ptr<de265_image> img) {
    if (img->pic.irap_no_rasl) flush_reorder();   // every IRAP picture with NoRaslOutputFlag (decctx.cc:1357-1365), CRA included
    const int max_reorder = img->pic.sps->max_num_reorder[img->pic.sps->max_sub_layers - 1];
    if (!img->pic.pic_output || (img->pic.error && suppress_faulty)) {
      release_locked(*img);
      return;
    }
    if (parse_only) {   // parse-only clients (tests) get the pictures in DECODING order: they reconstruct with the CPU
      output.push_back(std::move(img));   // oracle, which needs every reference before the picture that uses it
      return;
    }
    reorder.push_back(std::move(img));
    while ((int)reorder.size() > max_reorder) bump();
  }
  void bump() {
    size_t best = 0;
    for (size_t i = 1; i < reorder.size(); i++)
      if (reorder[i]->pic.poc < reorder[best]->pic.poc) best = i;
    output.push_back(std::move(reorder[best]));
    reorder.erase(reorder.begin() + best);
  }
  void flush_reorder() {
    while (!reorder.empty()) bump();
  }
  void release_locked(de265_image& img) {  // mu held
    if (img.pic.planes_external) {
      if (alloc.release_buffer) alloc.release_buffer(this, &img, alloc_user);
    } else if (img.host_base) {
      host_pool.emplace_back(img.host_bytes, img.host_base);
    }
    img.host_base = nullptr;
  }

  // ---- NAL handling ----
  void split_input(bool flush) {
    // find start codes 00 00 01 in inbuf; everything between two start codes is one NAL unit
    size_t n = inbuf.size(), i = 0, nal_start = (size_t)-1;
    size_t consumed = 0;
    const uint8_t* d = inbuf.data();
    while (i + 2 < n) {
      if (d[i] == 0 && d[i + 1] == 0 && d[i + 2] == 1) {
        if (nal_start != (size_t)-1) {
          size_t end = i;
          while (end > nal_start && d[end - 1] == 0) end--;
          push_nal_bytes(d + nal_start, end - nal_start);
        }
        nal_start = i + 3;
        consumed = i;
        i += 3;
      } else {
        i++;
      }
    }
    if (flush) {
      if (nal_start != (size_t)-1 && nal_start < n) {
        size_t end = n;
        while (end > nal_start && d[end - 1] == 0) end--;
        push_nal_bytes(d + nal_start, end - nal_start);
      }
      inbuf.clear();
    } else if (consumed > 0) {
      inbuf.erase(inbuf.begin(), inbuf.begin() + consumed);
    } else if (nal_start == (size_t)-1 && n > 2) {
      // no start code at all yet: keep only the tail that could begin one
      inbuf.erase(inbuf.begin(), inbuf.end() - 2);
    }
  }
  void push_nal_bytes(const uint8_t* p, size_t n) {
    if (n < 2) return;
    NalIn nal;
    nal.data.assign(p, p + n);
    nal.pts = in_pts;
    nal.user = in_user;
    nals.push_back(std::move(nal));
  }

  int compute_poc(const SliceHeader& sh, const Sps& sps, int nal_type, int tid) {
    const bool irap = nal_type >= NAL_BLA_W_LP && nal_type <= NAL_RSV_IRAP_VCL23;
    const bool idr = nal_type == NAL_IDR_W_RADL || nal_type == NAL_IDR_N_LP;
    const bool bla = nal_type >= NAL_BLA_W_LP && nal_type <= NAL_BLA_N_LP;
    const bool no_rasl_output = idr || bla || (irap && (first_picture || after_eos));
    int msb;
    const int max_lsb = 1 << sps.log2_max_poc_lsb;
    last_no_rasl_output = irap && no_rasl_output;
    if (irap && no_rasl_output) msb = 0;
    else {
      if (sh.poc_lsb < prev_tid0_poc_lsb && (prev_tid0_poc_lsb - sh.poc_lsb) >= max_lsb / 2) msb = prev_tid0_poc_msb + max_lsb;
      else if (sh.poc_lsb > prev_tid0_poc_lsb && (sh.poc_lsb - prev_tid0_poc_lsb) > max_lsb / 2) msb = prev_tid0_poc_msb - max_lsb;
      else msb = prev_tid0_poc_msb;
    }
    const int poc = msb + sh.poc_lsb;
    const bool sub_layer_non_ref = nal_type < 16 && (nal_type & 1) == 0;
    const bool rasl_radl = nal_type >= 6 && nal_type <= 9;
    if (tid == 0 && !sub_layer_non_ref && !rasl_radl) {
      prev_tid0_poc_lsb = sh.poc_lsb;
      prev_tid0_poc_msb = msb;
    }
    first_picture = false;
    after_eos = false;
    return poc;
  }

  int decode_nal(NalIn& nal) {
    const uint8_t* d = nal.data.data();
    const int type = (d[0] >> 1) & 0x3f;
    const int layer = ((d[0] & 1) << 5) | (d[1] >> 3);
    const int tid = (d[1] & 7) - 1;
    if (layer != 0 || tid > limit_tid) return DE265_OK;
    std::vector<uint8_t> rbsp;
    std::vector<uint32_t> removed;
    unescape_nal(d + 2, nal.data.size() - 2, rbsp, removed);
    if (type == NAL_VPS || type == NAL_SPS || type == NAL_PPS) {
      // parameter sets start a new access unit: the previous picture is complete
      if (cur && !cur->pic.slice_units.empty()) dispatch_current();
    }
    if (type == NAL_VPS) return DE265_OK;
    if (type == NAL_SPS) {
      BitReader br(rbsp.data(), rbsp.size());
      std::shared_ptr<Sps> sps;
      int r = parse_sps(br, sps);
      if (r) { warn(DE265_WARNING_SPS_HEADER_INVALID); return DE265_OK; }
      if (sps->bit_depth_luma != 8 || sps->bit_depth_chroma != 8 || sps->chroma_array_type > 1 || sps->persistent_rice ||
          sps->cabac_bypass_alignment || sps->extended_precision)
        return DE265_ERROR_NOT_IMPLEMENTED_YET;
      ps.sps[sps->id] = sps;
      return DE265_OK;
    }
    if (type == NAL_PPS) {
      BitReader br(rbsp.data(), rbsp.size());
      std::shared_ptr<Pps> pps;
      int r = parse_pps(br, ps, pps);
      if (r) { warn(r >= 1000 ? r : DE265_WARNING_PPS_HEADER_INVALID); return DE265_OK; }
      if (pps->cross_component_prediction) return DE265_ERROR_NOT_IMPLEMENTED_YET;
      ps.pps[pps->id] = pps;
      return DE265_OK;
    }
    if (type == NAL_EOS || type == NAL_EOB) {
      dispatch_current();
      after_eos = true;
      return DE265_OK;
    }
    if (type == NAL_AUD || type == NAL_PREFIX_SEI) {
      // a new access unit starts: the previous picture is complete
      if (cur && !cur->pic.slice_units.empty()) dispatch_current();
      return DE265_OK;
    }
    if (type == NAL_SUFFIX_SEI) {
      // decoded picture hash of the picture being assembled (read_sei_NAL decctx.cc:584-606: the first message of the NAL,
      // kept for the last image unit; sei.cc:32-61).  Other payloads are ignored like the reference ignores them.
      if (!cur || cur->pic.slice_units.empty() || rbsp.size() < 3) return DE265_OK;
      size_t pos = 0;
      int ptype = 0, psize = 0;
      while (pos < rbsp.size()) { const int b = rbsp[pos++]; ptype += b; if (b != 0xff) break; }
      while (pos < rbsp.size()) { const int b = rbsp[pos++]; psize += b; if (b != 0xff) break; }
      if (ptype != 132 || pos >= rbsp.size()) return DE265_OK;
      const int htype = rbsp[pos++];
      const int planes = cur->pic.sps->chroma_format_idc == 0 ? 1 : 3;
      const int per = htype == 0 ? 16 : (htype == 1 ? 2 : 4);
      if (htype > 2 || pos + (size_t)planes * per > rbsp.size() || psize < 1 + planes * per) { warn(DE265_ERROR_CANNOT_PROCESS_SEI); return DE265_OK; }
      Picture& p = cur->pic;
      memset(p.hash, 0, sizeof(p.hash));
      for (int c = 0; c < planes; c++) {
        if (htype == 0) memcpy(p.hash[c], &rbsp[pos], 16);
        else for (int k = 0; k < per; k++) p.hash[c][per - 1 - k] = rbsp[pos + k];   // big-endian in the stream
        pos += per;
      }
      p.hash_type = htype;
      return DE265_OK;
    }
    if (type > NAL_RSV_IRAP_VCL23 || (type > NAL_RASL_R && type < NAL_BLA_W_LP)) return DE265_OK;  // non-VCL / reserved

    // ---- slice segment ----
    BitReader br(rbsp.data(), rbsp.size());
    SliceHeader sh;
    std::shared_ptr<const Pps> pps;
    int r = parse_slice_header(br, type, ps, have_prev_indep ? &prev_indep : nullptr, removed, sh, pps);
    if (r) { warn(r >= 1000 ? r : DE265_WARNING_SLICEHEADER_INVALID); return DE265_OK; }
    // RASL pictures that follow an IRAP picture with NoRaslOutputFlag are not decoded at all (decctx.cc:2038-2045, :2069):
    // no picture, no POC / reference-picture-set processing, no DPB entry, no error
    if ((type == NAL_RASL_N || type == NAL_RASL_R) && skip_rasl) {
      if (sh.first_slice_segment_in_pic) dispatch_current();
      return DE265_OK;
    }
    if (sh.first_slice_segment_in_pic) {
      dispatch_current();
      cur.reset(new de265_image());
      Picture& p = cur->pic;
      p.sps = pps->sps;
      p.pps = pps;
      p.nal_type = type;
      p.temporal_id = tid;
      p.pts = nal.pts;
      p.user_data = nal.user;
      p.pic_output = sh.pic_output;
      p.init_maps();
      p.poc = compute_poc(sh, *pps->sps, type, tid);
      if (type >= NAL_BLA_W_LP && type <= NAL_RSV_IRAP_VCL23) {
        skip_rasl = last_no_rasl_output;
        p.irap_no_rasl = last_no_rasl_output;
      }
      dpb.apply_rps(sh, *pps->sps, p.poc, last_no_rasl_output);
      if (last_no_rasl_output) gop_id++;
      p.gop = gop_id;
      p.self->poc = p.poc;
      p.self->uid = next_uid++;
      p.self->owner = this;
      p.self->recycle = &Decoder::recycle_cb;
      dpb.pics.push_back(p.self);
      const Sps& s = *pps->sps;
      if (s.scaling_list_enabled) {
        p.scaling = std::make_shared<b265_scaling>();
        const ScalingList& sl = pps->scaling_list_present ? pps->scaling : s.scaling;
        static_assert(sizeof(ScalingList) == sizeof(b265_scaling), "layout");
        memcpy(p.scaling.get(), &sl, sizeof(b265_scaling));
      }
    } else if (!cur) {
      warn(DE265_ERROR_NO_INITIAL_SLICE_HEADER);
      return DE265_OK;
    }
    Picture& p = cur->pic;
    if (!sh.dependent) {
      b265_slice bs;
      memset(&bs, 0, sizeof(bs));
      bs.slice_addr_rs = sh.slice_addr_rs;
      bs.slice_type = (int8_t)sh.slice_type;
      bs.beta_offset = (int8_t)sh.beta_offset;
      bs.tc_offset = (int8_t)sh.tc_offset;
      bs.deblocking_disabled = (uint8_t)sh.deblocking_disabled;
      bs.lf_across_slices = (uint8_t)sh.lf_across_slices;
      bs.sao_luma = (uint8_t)sh.sao_luma;
      bs.sao_chroma = (uint8_t)sh.sao_chroma;
      bs.weighted = (uint8_t)((sh.slice_type == SLICE_P && pps->weighted_pred) || (sh.slice_type == SLICE_B && pps->weighted_bipred));
      bs.pps_weighted_pred = (uint8_t)pps->weighted_pred;
      bs.luma_log2_denom = (uint8_t)sh.pw.luma_log2_denom;
      bs.chroma_log2_denom = (uint8_t)sh.pw.chroma_log2_denom;
      memcpy(bs.weight, sh.pw.weight, sizeof(bs.weight));
      memcpy(bs.offset, sh.pw.offset, sizeof(bs.offset));
      sh.slice_index = (int)p.slices.size();
      p.slices.push_back(bs);
      SliceRefs sr;
      dpb.build_lists(sh, p, sr);   // on failure the lists stay empty and the slice data is rejected
      p.slice_refs.push_back(sr);
      prev_indep = sh;
      have_prev_indep = true;
    }
    if (p.slices.empty()) { warn(DE265_WARNING_SLICEHEADER_INVALID); return DE265_OK; }
    SliceUnit u;
    u.hdr = sh;
    u.pps = pps;
    u.rbsp = std::move(rbsp);
    p.slice_units.push_back(std::move(u));
    return DE265_OK;
  }

  size_t pending_pictures() {
    std::lock_guard<std::mutex> l(mu);
    return inflight.size();
  }

  de265_error decode(int* more) {
    if (gpu_status != B265_OK && !parse_only) {
      if (more) *more = 0;
      return DE265_ERROR_LIBRARY_INITIALIZATION_FAILED;
    }
    if (nals.empty() && end_of_stream) {
      dispatch_current();
      {
        std::unique_lock<std::mutex> l(mu);
        draining = true;
        cv_recon.notify_all();
        cv_done.wait(l, [this] { return inflight.empty(); });
        draining = false;
        flush_reorder();
        if (more) *more = !output.empty();
        if (fatal_error) {
          int e = fatal_error;
          fatal_error = 0;
          if (more) *more = !output.empty();
          return (de265_error)e;
        }
      }
      return DE265_OK;
    }
    if (nals.empty()) {
      if (more) *more = 1;
      return DE265_ERROR_WAITING_FOR_INPUT_DATA;
    }
    {
      // bound the amount of undelivered work
      std::unique_lock<std::mutex> l(mu);
      if (output.size() > 64) {
        if (more) *more = 1;
        return DE265_ERROR_IMAGE_BUFFER_FULL;
      }
      const size_t limit = (size_t)((gpus.size() + 1) * max_batch + 2 * n_threads + 2);
      cv_done.wait(l, [&] { return inflight.size() < limit; });
    }
    NalIn nal = std::move(nals.front());
    nals.pop_front();
    int err = decode_nal(nal);
    {
      std::lock_guard<std::mutex> l(mu);
      if (!err && fatal_error) { err = fatal_error; fatal_error = 0; }
    }
    if (more) *more = (err == DE265_OK);
    return (de265_error)err;
  }
};

const char* nal_name(int t) {
  static const char* names[] = {"TRAIL_N", "TRAIL_R", "TSA_N", "TSA_R", "STSA_N", "STSA_R", "RADL_N", "RADL_R", "RASL_N", "RASL_R"};
  if (t >= 0 && t < 10) return names[t];
  switch (t) {
    case 16: return "BLA_W_LP"; case 17: return "BLA_W_RADL"; case 18: return "BLA_N_LP";
    case 19: return "IDR_W_RADL"; case 20: return "IDR_N_LP"; case 21: return "CRA_NUT";
    case 32: return "VPS_NUT"; case 33: return "SPS_NUT"; case 34: return "PPS_NUT";
    default: return "unknown";
  }
}

}  // namespace

// ================================================================================================
// C API (include/de265.h)
// ================================================================================================
#define DEC(p) (reinterpret_cast<Decoder*>(p))

extern "C" {

LIBDE265_API const char* de265_get_version(void) { return "1.0.0-b265"; }
LIBDE265_API uint32_t de265_get_version_number(void) { return LIBDE265_NUMERIC_VERSION; }
LIBDE265_API int de265_get_version_number_major(void) { return (LIBDE265_NUMERIC_VERSION >> 24) & 0xff; }
LIBDE265_API int de265_get_version_number_minor(void) { return (LIBDE265_NUMERIC_VERSION >> 16) & 0xff; }
LIBDE265_API int de265_get_version_number_maintenance(void) { return (LIBDE265_NUMERIC_VERSION >> 8) & 0xff; }

LIBDE265_API const char* de265_get_error_text(de265_error err) {
  switch ((int)err) {
    case DE265_OK: return "no error";
    case DE265_ERROR_NO_SUCH_FILE: return "no such file";
    case DE265_ERROR_COEFFICIENT_OUT_OF_IMAGE_BOUNDS: return "coefficient out of image bounds";
    case DE265_ERROR_CHECKSUM_MISMATCH: return "image checksum mismatch";
    case DE265_ERROR_CTB_OUTSIDE_IMAGE_AREA: return "CTB outside of image area";
    case DE265_ERROR_OUT_OF_MEMORY: return "out of memory";
    case DE265_ERROR_CODED_PARAMETER_OUT_OF_RANGE: return "coded parameter out of range";
    case DE265_ERROR_IMAGE_BUFFER_FULL: return "DPB/output queue full";
    case DE265_ERROR_CANNOT_START_THREADPOOL: return "cannot start decoding threads";
    case DE265_ERROR_LIBRARY_INITIALIZATION_FAILED: return "global library initialization failed (no usable CUDA device)";
    case DE265_ERROR_LIBRARY_NOT_INITIALIZED: return "cannot free library data (not initialized)";
    case DE265_ERROR_WAITING_FOR_INPUT_DATA: return "no more input data, decoder stalled";
    case DE265_ERROR_CANNOT_PROCESS_SEI: return "SEI data cannot be processed";
    case DE265_ERROR_PARAMETER_PARSING: return "command-line parameter error";
    case DE265_ERROR_NO_INITIAL_SLICE_HEADER: return "first slice missing, cannot decode dependent slice";
    case DE265_ERROR_PREMATURE_END_OF_SLICE: return "premature end of slice data";
    case DE265_ERROR_UNSPECIFIED_DECODING_ERROR: return "unspecified decoding error";
    case DE265_ERROR_NOT_IMPLEMENTED_YET: return "unimplemented decoder feature";
    case DE265_WARNING_SPS_HEADER_INVALID: return "sps header invalid";
    case DE265_WARNING_PPS_HEADER_INVALID: return "pps header invalid";
    case DE265_WARNING_SLICEHEADER_INVALID: return "slice header invalid";
    case DE265_WARNING_NONEXISTING_PPS_REFERENCED: return "non-existing PPS referenced";
    case DE265_WARNING_NONEXISTING_SPS_REFERENCED: return "non-existing SPS referenced";
    case DE265_WARNING_EOSS_BIT_NOT_SET: return "end-of-sub-stream-one-bit not set to 1 when it should be";
    case DE265_WARNING_B265_ACCELERATION_UNAVAILABLE: return "requested CPU acceleration code is not available: this library only has the CUDA pixel path";
    default: return err >= 1000 ? "warning" : "unknown error";
  }
}
LIBDE265_API int de265_isOK(de265_error err) { return err == DE265_OK || err >= 1000; }
LIBDE265_API void de265_disable_logging(void) {}
LIBDE265_API void de265_set_verbosity(int) {}

LIBDE265_API de265_error de265_init(void) {
  std::lock_guard<std::mutex> l(g_init_mutex);
  g_init_count++;
  return DE265_OK;
}
LIBDE265_API de265_error de265_free(void) {
  std::lock_guard<std::mutex> l(g_init_mutex);
  if (g_init_count <= 0) return DE265_ERROR_LIBRARY_NOT_INITIALIZED;
  g_init_count--;
  return DE265_OK;
}

LIBDE265_API void b265_de265_set_device(int device) { g_device = device; }

LIBDE265_API de265_decoder_context* de265_new_decoder(void) {
  if (de265_init() != DE265_OK) return nullptr;
  return new Decoder();
}
LIBDE265_API de265_error de265_free_decoder(de265_decoder_context* c) {
  delete DEC(c);
  return de265_free();
}
LIBDE265_API de265_error de265_start_worker_threads(de265_decoder_context* c, int n) {
  if (n > 64) n = 64;
  if (n > 0) DEC(c)->start_workers(n);
  return DE265_OK;
}

LIBDE265_API de265_error de265_push_data(de265_decoder_context* c, const void* data, int len, de265_PTS pts, void* user) {
  Decoder* d = DEC(c);
  if (len <= 0) return DE265_OK;
  d->in_pts = pts;
  d->in_user = user;
  d->end_of_stream = false;
  const uint8_t* p = (const uint8_t*)data;
  d->inbuf.insert(d->inbuf.end(), p, p + len);
  d->split_input(false);
  return DE265_OK;
}
LIBDE265_API void de265_push_end_of_NAL(de265_decoder_context* c) { DEC(c)->split_input(true); }
LIBDE265_API void de265_push_end_of_frame(de265_decoder_context* c) {
  Decoder* d = DEC(c);
  d->split_input(true);
}
LIBDE265_API de265_error de265_push_NAL(de265_decoder_context* c, const void* data, int len, de265_PTS pts, void* user) {
  Decoder* d = DEC(c);
  d->in_pts = pts;
  d->in_user = user;
  d->end_of_stream = false;
  d->push_nal_bytes((const uint8_t*)data, (size_t)len);
  return DE265_OK;
}
LIBDE265_API de265_error de265_flush_data(de265_decoder_context* c) {
  Decoder* d = DEC(c);
  d->split_input(true);
  d->end_of_stream = true;
  return DE265_OK;
}
LIBDE265_API int de265_get_number_of_input_bytes_pending(de265_decoder_context* c) { return (int)DEC(c)->inbuf.size(); }
LIBDE265_API int de265_get_number_of_NAL_units_pending(de265_decoder_context* c) { return (int)DEC(c)->nals.size(); }

LIBDE265_API de265_error de265_decode(de265_decoder_context* c, int* more) { return DEC(c)->decode(more); }

LIBDE265_API de265_error de265_decode_data(de265_decoder_context* c, const void* data, int len) {
  de265_error err = len > 0 ? de265_push_data(c, data, len, 0, nullptr) : de265_flush_data(c);
  if (err != DE265_OK) return err;
  int more = 0;
  do {
    err = de265_decode(c, &more);
    if (err != DE265_OK) more = 0;
    if (err == DE265_ERROR_WAITING_FOR_INPUT_DATA) err = DE265_OK;   // not an error for this call (de265.cc:283-288)
  } while (more);
  return err;
}

LIBDE265_API void de265_reset(de265_decoder_context* c) {
  Decoder* d = DEC(c);
  d->dispatch_current();
  std::unique_lock<std::mutex> l(d->mu);
  d->draining = true;
  d->cv_recon.notify_all();
  d->cv_done.wait(l, [d] { return d->inflight.empty(); });
  d->draining = false;
  for (auto& i : d->reorder) d->release_locked(*i);
  for (auto& i : d->output) d->release_locked(*i);
  d->reorder.clear();
  d->output.clear();
  d->nals.clear();
  d->inbuf.clear();
  d->end_of_stream = false;
  d->first_picture = true;
  d->after_eos = true;
  d->have_prev_indep = false;
  d->fatal_error = 0;
  d->dpb.clear();
  d->gop_ctx.clear();
}

LIBDE265_API const struct de265_image* de265_peek_next_picture(de265_decoder_context* c) {
  Decoder* d = DEC(c);
  std::lock_guard<std::mutex> l(d->mu);
  return d->output.empty() ? nullptr : d->output.front().get();
}
LIBDE265_API void de265_release_next_picture(de265_decoder_context* c) {
  Decoder* d = DEC(c);
  std::lock_guard<std::mutex> l(d->mu);
  if (d->output.empty()) return;
  if (d->last_returned) d->release_locked(*d->last_returned);
  d->last_returned = std::move(d->output.front());
  d->output.pop_front();
}
LIBDE265_API const struct de265_image* de265_get_next_picture(de265_decoder_context* c) {
  Decoder* d = DEC(c);
  std::lock_guard<std::mutex> l(d->mu);
  if (d->output.empty()) return nullptr;
  if (d->last_returned) d->release_locked(*d->last_returned);
  d->last_returned = std::move(d->output.front());
  d->output.pop_front();
  return d->last_returned.get();
}
LIBDE265_API de265_error de265_get_warning(de265_decoder_context* c) {
  Decoder* d = DEC(c);
  std::lock_guard<std::mutex> l(d->mu);
  if (d->warnings.empty()) return DE265_OK;
  int w = d->warnings.front();
  d->warnings.pop_front();
  return (de265_error)w;
}

LIBDE265_API int de265_get_image_width(const struct de265_image* img, int ch) { return (ch >= 0 && ch < 3) ? img->conf_w[ch] : 0; }
LIBDE265_API int de265_get_image_height(const struct de265_image* img, int ch) { return (ch >= 0 && ch < 3) ? img->conf_h[ch] : 0; }
LIBDE265_API enum de265_chroma de265_get_chroma_format(const struct de265_image* img) {
  return (enum de265_chroma)img->pic.sps->chroma_format_idc;
}
LIBDE265_API int de265_get_bits_per_pixel(const struct de265_image* img, int ch) {
  return ch == 0 ? img->pic.sps->bit_depth_luma : (ch < 3 ? img->pic.sps->bit_depth_chroma : 0);
}
LIBDE265_API const uint8_t* de265_get_image_plane(const struct de265_image* img, int ch, int* stride) {
  if (ch < 0 || ch > 2 || !img->pic.host_plane[ch]) return nullptr;
  if (stride) *stride = img->pic.host_stride[ch];
  return img->pic.host_plane[ch] + (size_t)img->conf_y[ch] * img->pic.host_stride[ch] + img->conf_x[ch];
}
LIBDE265_API void* de265_get_image_plane_user_data(const struct de265_image* img, int ch) {
  return (ch >= 0 && ch < 3) ? img->pic.plane_user[ch] : nullptr;
}
LIBDE265_API void de265_set_image_plane(struct de265_image* img, int c, void* mem, int stride, void* userdata) {
  if (c < 0 || c > 2) return;
  img->pic.host_plane[c] = (uint8_t*)mem;
  img->pic.host_stride[c] = stride;
  img->pic.plane_user[c] = userdata;
}
LIBDE265_API de265_PTS de265_get_image_PTS(const struct de265_image* img) { return img->pic.pts; }
LIBDE265_API void* de265_get_image_user_data(const struct de265_image* img) { return img->pic.user_data; }
LIBDE265_API void de265_set_image_user_data(struct de265_image* img, void* u) { img->pic.user_data = u; }
LIBDE265_API void de265_get_image_NAL_header(const struct de265_image* img, int* type, const char** name, int* layer, int* tid) {
  if (type) *type = img->pic.nal_type;
  if (name) *name = nal_name(img->pic.nal_type);
  if (layer) *layer = 0;
  if (tid) *tid = img->pic.temporal_id;
}

LIBDE265_API void de265_set_image_allocation_functions(de265_decoder_context* c, struct de265_image_allocation* a, void* user) {
  Decoder* d = DEC(c);
  if (a) d->alloc = *a; else d->alloc = de265_image_allocation{nullptr, nullptr};
  d->alloc_user = user;
}
LIBDE265_API const struct de265_image_allocation* de265_get_default_image_allocation_functions(void) {
  static const de265_image_allocation def = {nullptr, nullptr};  // NULL hooks = the built-in pinned-buffer pool
  return &def;
}

LIBDE265_API int de265_get_highest_TID(de265_decoder_context* c) { return DEC(c)->limit_tid; }
LIBDE265_API int de265_get_current_TID(de265_decoder_context* c) { return DEC(c)->limit_tid; }
LIBDE265_API void de265_set_limit_TID(de265_decoder_context* c, int t) { DEC(c)->limit_tid = t < 0 ? 0 : (t > 6 ? 6 : t); }
LIBDE265_API void de265_set_framerate_ratio(de265_decoder_context*, int) {}
LIBDE265_API int de265_change_framerate(de265_decoder_context*, int) { return 100; }

LIBDE265_API void de265_set_parameter_bool(de265_decoder_context* c, enum de265_param p, int v) {
  Decoder* d = DEC(c);
  switch (p) {
    case DE265_DECODER_PARAM_BOOL_SEI_CHECK_HASH: d->check_hash = v != 0; break;
    case DE265_DECODER_PARAM_SUPPRESS_FAULTY_PICTURES: d->suppress_faulty = v != 0; break;
    case DE265_DECODER_PARAM_DISABLE_DEBLOCKING: d->disable_deblocking = v != 0; break;
    case DE265_DECODER_PARAM_DISABLE_SAO: d->disable_sao = v != 0; break;
    default: break;
  }
}
LIBDE265_API void de265_set_parameter_int(de265_decoder_context* c, enum de265_param p, int v) {
  // DUMP_*_HEADERS are debugging aids.  ACCELERATION_CODE (de265.cc:603-612 -> decctx.cc:430-449): the only
  // table this library can install is the CUDA one; a request for a CPU table is refused with a warning
  if (p == DE265_DECODER_PARAM_ACCELERATION_CODE && v != de265_acceleration_AUTO && v != de265_acceleration_CUDA)
    DEC(c)->warn(DE265_WARNING_B265_ACCELERATION_UNAVAILABLE);
}
LIBDE265_API int b265_de265_get_acceleration(de265_decoder_context*) { return de265_acceleration_CUDA; }
LIBDE265_API int de265_get_parameter_bool(de265_decoder_context* c, enum de265_param p) {
  Decoder* d = DEC(c);
  switch (p) {
    case DE265_DECODER_PARAM_BOOL_SEI_CHECK_HASH: return d->check_hash;
    case DE265_DECODER_PARAM_SUPPRESS_FAULTY_PICTURES: return d->suppress_faulty;
    case DE265_DECODER_PARAM_DISABLE_DEBLOCKING: return d->disable_deblocking;
    case DE265_DECODER_PARAM_DISABLE_SAO: return d->disable_sao;
    default: return 0;
  }
}

LIBDE265_API void b265_de265_set_parse_only(int on) { g_parse_only = on != 0; }
LIBDE265_API int b265_de265_image_job(de265_decoder_context* c, const struct de265_image* img, void* out) {
  if (!c || !img || !out) return B265_ERR_INVALID_ARG;
  DEC(c)->fill_job(*const_cast<de265_image*>(img), *static_cast<b265_frame_job*>(out));
  return img->pic.error;
}
LIBDE265_API int b265_de265_image_poc(const struct de265_image* img) { return img->pic.poc; }
LIBDE265_API uint64_t b265_de265_hashes_checked(de265_decoder_context* c) {
  Decoder* d = DEC(c);
  std::lock_guard<std::mutex> l(d->mu);
  return d->stat_hashed;
}
LIBDE265_API int b265_de265_image_starts_cvs(const struct de265_image* img) { return img->pic.irap_no_rasl ? 1 : 0; }
LIBDE265_API void b265_de265_set_batch(de265_decoder_context* c, int n) {
  Decoder* d = DEC(c);
  std::lock_guard<std::mutex> l(d->mu);
  d->max_batch = n < 1 ? 1 : (n > 64 ? 64 : n);
}
LIBDE265_API void b265_de265_get_stats(de265_decoder_context* c, double out[8]) {
  Decoder* d = DEC(c);
  std::lock_guard<std::mutex> l(d->mu);
  out[0] = (double)d->stat_pictures;
  out[1] = d->stat_parse_s;
  out[2] = d->stat_gpu_s;
  out[3] = d->stat_h2d;
  out[4] = d->stat_d2h;
  out[6] = (double)d->stat_batches;
  out[7] = d->stat_submit_s;
  out[5] = 0;
  for (b265_ctx* g : d->gpus) if (g) out[5] += (double)b265_launch_count(g);
}

}  // extern "C"
