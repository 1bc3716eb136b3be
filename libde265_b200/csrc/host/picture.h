// picture.h — one coded picture on the host side: the parse-time metadata the CABAC/syntax decoder
// needs for context selection and prediction-mode derivation, and the WORK LISTS (include/b265.h
// part A.2) that replace the reference's in-place reconstruction calls S1..S5.
// Counterpart of the metadata arrays of the reference's de265_image (image.h:411-417); the pixel
// planes themselves live in HBM (b265 surfaces) and never on the host until output.
#pragma once
#include <stdint.h>
#include <stdlib.h>
#include <atomic>
#include <memory>
#include <vector>
#include "../../../include/b265.h"
#include "params.h"

namespace b265h {

// Growable array of PODs without value-initialisation (the work lists are written once, sequentially, by the
// syntax decoder's inner loops: no zero fill, no per-element capacity check when the caller reserves first).
template <class T>
struct PodBuf {
  T* p = nullptr;
  size_t n = 0, cap = 0;
  PodBuf() {}
  PodBuf(const PodBuf&) = delete;
  PodBuf& operator=(const PodBuf&) = delete;
  PodBuf(PodBuf&& o) noexcept : p(o.p), n(o.n), cap(o.cap) { o.p = nullptr; o.n = o.cap = 0; }
  PodBuf& operator=(PodBuf&& o) noexcept {
    if (this != &o) { free(p); p = o.p; n = o.n; cap = o.cap; o.p = nullptr; o.n = o.cap = 0; }
    return *this;
  }
  ~PodBuf() { free(p); }
  void reserve(size_t c) {
    if (c <= cap) return;
    T* q = (T*)realloc(p, c * sizeof(T));
    if (!q) abort();
    p = q; cap = c;
  }
  // room for `extra` more elements; returns the write position (the caller advances n)
  T* room(size_t extra) {
    if (n + extra > cap) reserve(n + extra > 2 * cap ? n + extra : 2 * cap);
    return p + n;
  }
  void push_back(const T& v) { *room(1) = v; n++; }
  void clear() { n = 0; }
  void release() { free(p); p = nullptr; n = cap = 0; }
  size_t size() const { return n; }
  bool empty() const { return n == 0; }
  T* data() { return p; }
  const T* data() const { return p; }
  T& operator[](size_t i) { return p[i]; }
  const T& operator[](size_t i) const { return p[i]; }
  T& back() { return p[n - 1]; }
};

enum PredMode : uint8_t { MODE_NONE = 0, MODE_INTRA = 1, MODE_INTER = 2, MODE_SKIP = 3 };

// motion of one 4x4 luma unit (the reference's PBMotion, motion.h:36-44); pred_flags 0 = intra / not coded
struct MotionInfo {
  int16_t mv[2][2];     // [list][x,y] quarter samples
  int8_t ref_idx[2];
  uint8_t pred_flags;   // bit 0: L0, bit 1: L1
  uint8_t pad;
};

// reference picture lists of one (independent) slice, resolved against the DPB (slice.h:237 RefPicList +
// the POC / long-term marking the motion vector prediction compares, motion.cc:1020-1180)
struct SliceRefs {
  int n[2] = {0, 0};
  int32_t poc[2][16];
  uint8_t is_lt[2][16];
  uint8_t slot[2][16];     // index into Picture::ref_pics (-> b265_frame_job.ref_surfaces)
};

// What later pictures need from a decoded picture: its identity in the DPB, where its samples live, and its
// motion field for temporal motion vector prediction (8.5.3.2.8).  Shared between the API thread (DPB
// bookkeeping), the parse workers (collocated motion) and the reconstruction threads (surfaces).
struct RefPic {
  int poc = 0;
  uint64_t decode_index = 0;
  uint64_t uid = 0;                // unique per decoder, never 0: identifies the picture to parse-only clients
  bool is_ref = true, is_lt = false, needed_for_output = false;
  bool unavailable = false;        // generated in place of a missing reference (8.3.3.2): mid-grey samples, intra motion field;
                                   // its surface is allocated and filled by the reconstruction thread that first needs it
  int ctx = -1, surface = -1;      // reconstruction context / HBM surface (set by the reconstruction thread)
  void* owner = nullptr;           // decoder that recycles the surface when the last reference drops
  void (*recycle)(void* owner, int ctx, int surface) = nullptr;
  // motion field complete: release-stored by the parse worker after motion / slice_refs / ctb_slice are final,
  // acquire-loaded by the workers of later pictures that use this one as their collocated picture
  std::atomic<int> parsed{0};
  std::atomic<int> failed{0};
  int u_w = 0, u_h = 0, ctbs_w = 0, log2_ctb = 4;
  std::vector<MotionInfo> motion;  // per 4x4 unit
  std::vector<uint16_t> ctb_slice; // per CTB: index into slice_refs
  std::vector<SliceRefs> slice_refs;
  ~RefPic() { if (recycle && surface >= 0) recycle(owner, ctx, surface); }
};

struct SliceUnit {                 // one slice segment NAL of the picture
  SliceHeader hdr;
  std::shared_ptr<const Pps> pps;
  std::vector<uint8_t> rbsp;       // de-escaped NAL payload (after the 2-byte NAL header)
};

struct Picture {
  std::shared_ptr<const Sps> sps;
  std::shared_ptr<const Pps> pps;  // PPS of the first slice
  int width = 0, height = 0;
  int poc = 0, nal_type = 0, temporal_id = 0;
  int pic_output = 1;
  // decoded picture hash of the picture's suffix SEI (sei.cc:32-61): type 0 MD5, 1 CRC, 2 checksum; -1 = none
  int hash_type = -1;
  uint8_t hash[3][16] = {};         // per colour plane: MD5 digest bytes, or the CRC / checksum as a little-endian word
  bool irap_no_rasl = false;       // IRAP picture with NoRaslOutputFlag: pictures waiting for output leave first (C.5.2.2)
  int64_t pts = 0;
  void* user_data = nullptr;
  uint64_t decode_index = 0;       // position in decoding order
  uint64_t gop = 0;                // counts the pictures that empty the DPB (IDR / BLA / first CRA): no reference crosses it

  std::vector<SliceUnit> slice_units;

  // ---- parse-time maps ----
  int cb_w = 0, cb_h = 0;          // in min-CB units
  int u_w = 0, u_h = 0;            // in 4x4 units
  std::vector<uint8_t> ct_depth;   // per min-CB
  std::vector<uint8_t> pred_mode;  // per min-CB (PredMode; MODE_NONE = not decoded)
  std::vector<uint8_t> cu_flags;   // per min-CB: bit0 pcm, bit1 transquant bypass
  std::vector<int8_t>  qp_y;       // per min-CB (-> b265_frame_job.qp_y)
  std::vector<uint8_t> ipm;        // intra luma pred mode per 4x4 unit
  std::vector<uint8_t> unit_info;  // per 4x4 unit, B265_UNIT_* (-> b265_frame_job.unit_info)
  std::vector<int32_t> ctb_slice_addr;  // per CTB (raster): SliceAddrRS of the owning slice, -1 = not decoded
  std::vector<b265_ctb> ctbs;      // per CTB (raster)
  // ---- inter prediction ----
  std::shared_ptr<RefPic> self;                  // this picture as later pictures will see it
  std::vector<std::shared_ptr<RefPic>> ref_pics; // distinct reference pictures of all slices (slot order)
  std::vector<SliceRefs> slice_refs;             // per independent slice (same index as `slices`)
  bool has_inter = false;

  // ---- work lists ----
  PodBuf<b265_tu> tus;
  PodBuf<b265_coeff> coeffs;
  std::vector<b265_pu> pus;
  std::vector<b265_pcm> pcms;
  std::vector<uint8_t> pcm_samples;
  std::vector<b265_slice> slices;  // one per independent slice
  std::vector<b265_sao> sao;       // per CTB (raster)
  bool any_sao = false, any_deblock = false;
  std::shared_ptr<b265_scaling> scaling;

  // ---- reconstruction / output state ----
  int surface = -1;                // HBM surface id
  int error = 0;                   // de265_error of the parse (0 = ok)
  uint8_t* host_plane[3] = {nullptr, nullptr, nullptr};  // output planes after D2H
  int host_stride[3] = {0, 0, 0};
  void* plane_user[3] = {nullptr, nullptr, nullptr};
  bool planes_external = false;    // allocated through de265_image_allocation
  double parse_seconds = 0;

  void init_maps() {
    const Sps& s = *sps;
    width = s.width; height = s.height;
    cb_w = s.min_cb_w; cb_h = s.min_cb_h;
    u_w = width >> 2; u_h = height >> 2;
    ct_depth.assign((size_t)cb_w * cb_h, 0);
    pred_mode.assign((size_t)cb_w * cb_h, MODE_NONE);
    cu_flags.assign((size_t)cb_w * cb_h, 0);
    qp_y.assign((size_t)cb_w * cb_h, 0);
    ipm.assign((size_t)u_w * u_h, 1);
    unit_info.assign((size_t)u_w * u_h, 0);
    ctb_slice_addr.assign((size_t)s.ctbs_w * s.ctbs_h, -1);
    ctbs.assign((size_t)s.ctbs_w * s.ctbs_h, b265_ctb{0, 0, 0});
    sao.assign((size_t)s.ctbs_w * s.ctbs_h, b265_sao{});
    self = std::make_shared<RefPic>();
    self->u_w = u_w; self->u_h = u_h; self->ctbs_w = s.ctbs_w; self->log2_ctb = s.log2_ctb;
    self->motion.clear();   // allocated by the first P/B slice (an empty field reads as "all intra")
    self->ctb_slice.assign((size_t)s.ctbs_w * s.ctbs_h, 0);
    ref_pics.clear(); slice_refs.clear(); has_inter = false;
    tus.clear(); coeffs.clear(); pus.clear(); pcms.clear(); pcm_samples.clear(); slices.clear();
    tus.reserve((size_t)u_w * u_h / 4);
    coeffs.reserve((size_t)width * height / 8);
  }
  void release_parse_state() {
    std::vector<uint8_t>().swap(ct_depth);
    std::vector<uint8_t>().swap(ipm);
    slice_units.clear();
  }
};

// Parses all slice segments of `pic` (CABAC + syntax) and fills its work lists.
// Returns 0 or a de265_error value. Thread-safe across different pictures.
int decode_picture_slices(Picture& pic);

}  // namespace b265h
