// streamgen.cc — random-syntax HEVC test-stream generator.  TEST INFRASTRUCTURE, not part of the product.
//
// The reference's encoder (enc265) only produces intra streams with flat quantisation and without SAO, PCM,
// transform skip, cu_qp_delta, scaling lists, tiles, wavefronts or multiple slices, so those
// tools cannot be pinned at stream level with its output.  This program writes VALID bitstreams whose syntax
// elements are drawn at random: it runs the parser's own syntax walk (libde265_b200/csrc/host/slice_syntax.h)
// over an arithmetic-coding engine that, instead of decoding a bin, draws it, ENCODES it (H.265 9.3.4.4, the
// informative encoder flow charts) and returns it.  The unmodified reference decoder (oracle/_ref/dec265) then
// decodes the stream and its per-plane MD5 is the golden value (tools/make_gen_streams.py); any difference
// between the parser's reading of the syntax and the reference's shows up as an MD5 mismatch or a decode error.
//
// usage: b265_streamgen out.bin key=value ...   (keys: see struct Config)
#include <stdio.h>
#include <stdlib.h>
#include <string>
#include <map>
#include "../../libde265_b200/csrc/host/slice_syntax.h"
#include "../../libde265_b200/csrc/host/dpb.h"

using namespace b265h;

// ---- deterministic random source ------------------------------------------------------------------------
struct Rng {
  uint64_t s;
  explicit Rng(uint64_t seed) : s(seed * 0x9E3779B97F4A7C15ull + 0x1234567ull) {}
  uint64_t next() {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
  }
  uint32_t below(uint32_t n) { return n ? (uint32_t)(next() % n) : 0; }
  int range(int lo, int hi) { return lo + (int)below((uint32_t)(hi - lo + 1)); }   // inclusive
  bool chance(double p) { return (double)(next() >> 11) * (1.0 / 9007199254740992.0) < p; }
};

// ---- bit writer (RBSP, before emulation prevention) ---------------------------------------------------------
struct BitWriter {
  std::vector<uint8_t> b;
  int fill = 0;   // bits used in the last byte (0 = byte aligned)
  void u1(int v) {
    if (fill == 0) b.push_back(0);
    if (v) b.back() |= (uint8_t)(0x80 >> fill);
    fill = (fill + 1) & 7;
  }
  void u(int n, uint32_t v) { for (int i = n - 1; i >= 0; i--) u1((v >> i) & 1); }
  void ue(uint32_t v) {
    const uint64_t x = (uint64_t)v + 1;
    int len = 0;
    while ((x >> len) > 1) len++;
    for (int i = 0; i < len; i++) u1(0);
    for (int i = len; i >= 0; i--) u1((int)((x >> i) & 1));
  }
  void se(int v) { ue(v > 0 ? (uint32_t)(2 * v - 1) : (uint32_t)(-2 * v)); }
  void trailing() { u1(1); while (fill) u1(0); }   // rbsp_trailing_bits / byte_alignment
  void align_zero() { while (fill) u1(0); }
};

// emulation prevention (7.4.2): positions `marks` (indices into `in`) are translated to escaped indices
static std::vector<uint8_t> escape(const std::vector<uint8_t>& in, std::vector<size_t>* marks = nullptr) {
  std::vector<uint8_t> out;
  size_t mi = 0;
  int zeros = 0;
  for (size_t i = 0; i < in.size(); i++) {
    while (marks && mi < marks->size() && (*marks)[mi] == i) (*marks)[mi++] = out.size() + ((zeros >= 2 && in[i] <= 3) ? 1 : 0);
    if (zeros >= 2 && in[i] <= 3) { out.push_back(3); zeros = 0; }
    out.push_back(in[i]);
    zeros = in[i] == 0 ? zeros + 1 : 0;
  }
  while (marks && mi < marks->size()) (*marks)[mi++] = out.size();
  return out;
}

static void put_nal(std::vector<uint8_t>& stream, int nal_type, const std::vector<uint8_t>& escaped_payload) {
  static const uint8_t sc[4] = {0, 0, 0, 1};
  stream.insert(stream.end(), sc, sc + 4);
  stream.push_back((uint8_t)(nal_type << 1));
  stream.push_back(1);   // nuh_layer_id 0, temporal id 0
  stream.insert(stream.end(), escaped_payload.begin(), escaped_payload.end());
}

// ---- the generating arithmetic-coding engine -------------------------------------------------------------------
struct GenSetup {
  Rng* rng = nullptr;
  const double* prob = nullptr;          // P(bin = 1) per context index
  double p_pcm = 0.1, p_big_coeff = 0.01;
  int last_ts = 0;                       // tile-scan address of the last CTB of the segment
  std::vector<uint8_t>* out = nullptr;   // slice segment data
  std::vector<size_t>* substreams = nullptr;   // byte offsets of sub-streams 1..n
};

struct GenEngine {
  typedef GenSetup Setup;
  Setup s;
  const Ctx* base = nullptr;
  uint32_t low = 0, range = 510;
  int outstanding = 0;
  bool first_bit = true;
  int acc = 0, nacc = 0;
  int pending = -1, pending_arg = 0;
  std::vector<int> script;   // forced bypass bins (front first)
  size_t script_pos = 0;
  bool started = false;

  struct Local {
    GenEngine& e;
    explicit Local(GenEngine& g) : e(g) {}
    int bin(Ctx& c) { return e.bin(c); }
    int bypass() { return e.bypass(); }
    uint32_t bypass_bits(int n) { return e.bypass_bits(n); }
    void hint(int k, int a) { e.hint(k, a); }
    void commit(GenEngine&) {}
  };

  void setup(const Setup& st) { s = st; }
  void set_ctx_base(const Ctx* b) { base = b; }
  void set_data(const uint8_t*, const uint8_t*) {}

  void write_bit(int b) {
    acc = (acc << 1) | b;
    if (++nacc == 8) { s.out->push_back((uint8_t)acc); acc = 0; nacc = 0; }
  }
  void put_bit(int b) {
    if (first_bit) first_bit = false; else write_bit(b);
    while (outstanding > 0) { write_bit(1 - b); outstanding--; }
  }
  void renorm() {
    while (range < 256) {
      if (low < 256) put_bit(0);
      else if (low >= 512) { low -= 512; put_bit(1); }
      else { low -= 256; outstanding++; }
      range <<= 1;
      low <<= 1;
    }
  }
  void start() {
    if (started && s.substreams) s.substreams->push_back(s.out->size());
    started = true;
    reinit();
  }
  void reinit() { low = 0; range = 510; outstanding = 0; first_bit = true; }
  void encode(Ctx& c, int b) {
    const uint32_t st = c;
    const uint32_t lps = g_cabac_lps[st >> 1][(range >> 6) & 3];
    range -= lps;
    const int is_lps = b != (int)(st & 1);
    if (is_lps) { low += range; range = lps; }
    c = g_cabac_next[(st << 1) | (uint32_t)is_lps];
    renorm();
  }
  int bin(Ctx& c) {
    const int b = s.rng->chance(s.prob[&c - base]) ? 1 : 0;
    encode(c, b);
    return b;
  }
  int bypass() {
    int b;
    if (script_pos < script.size()) {
      b = script[script_pos++];
      if (script_pos == script.size()) { script.clear(); script_pos = 0; }
    } else b = (int)(s.rng->next() & 1);
    low <<= 1;
    if (b) low += range;
    if (low >= 1024) { put_bit(1); low -= 1024; }
    else if (low < 512) put_bit(0);
    else { low -= 512; outstanding++; }
    return b;
  }
  uint32_t bypass_bits(int n) {
    uint32_t v = 0;
    while (n-- > 0) v = (v << 1) | (uint32_t)bypass();
    return v;
  }
  void hint(int kind, int arg) {
    switch (kind) {
      case HINT_QP_DELTA_SUFFIX: {   // EG0 suffix of cu_qp_delta_abs: keep |CuQpDeltaVal| <= 19
        const int k = s.rng->range(0, 3);
        for (int i = 0; i < k; i++) script.push_back(1);
        script.push_back(0);
        break;
      }
      case HINT_MVD_SUFFIX: {        // abs_mvd_minus2 (EG1): keep motion vector differences moderate
        const int k = s.rng->range(0, s.rng->chance(0.05) ? 9 : 4);
        for (int i = 0; i < k; i++) script.push_back(1);
        script.push_back(0);
        break;
      }
      case HINT_COEFF_REMAINING: {   // prefix of coeff_abs_level_remaining: levels stay inside 16 bits (7.4.9.11)
        int prefix = 0;
        if (s.rng->chance(s.p_big_coeff)) prefix = s.rng->range(7, 16 - arg);
        else while (prefix < 6 && (s.rng->next() & 1)) prefix++;
        for (int i = 0; i < prefix; i++) script.push_back(1);
        script.push_back(0);
        break;
      }
      default:
        pending = kind;
        pending_arg = arg;
    }
  }
  void flush() {   // 9.3.4.4.5 EncodeFlush, then zero bits up to the byte boundary
    range = 2;
    renorm();
    put_bit((low >> 9) & 1);
    write_bit((low >> 8) & 1);
    write_bit(1);
    while (nacc) write_bit(0);
  }
  int terminate() {
    int b;
    switch (pending) {
      case HINT_PCM_FLAG: b = s.rng->chance(s.p_pcm); break;
      case HINT_END_OF_SEGMENT: b = pending_arg == s.last_ts; break;
      case HINT_END_OF_SUBSET: b = 1; break;
      default: fprintf(stderr, "streamgen: terminating bin without a hint\n"); abort();
    }
    pending = -1;
    range -= 2;
    if (b) { low += range; flush(); }
    else renorm();
    return b;
  }
  template <class Vec>
  bool pcm_samples(int n_luma, int bits_luma, int shift_luma, int n_chroma, int bits_chroma, int shift_chroma, Vec& out) {
    for (int i = 0; i < n_luma + n_chroma; i++) {
      const int bits = i < n_luma ? bits_luma : bits_chroma, sh = i < n_luma ? shift_luma : shift_chroma;
      const uint32_t v = s.rng->below(1u << bits);
      for (int k = bits - 1; k >= 0; k--) write_bit((v >> k) & 1);
      out.push_back((uint8_t)(v << sh));
    }
    while (nacc) write_bit(0);
    reinit();
    return true;
  }
};

// ---- configuration ------------------------------------------------------------------------------------------------
struct Config {
  std::map<std::string, double> kv;
  double get(const char* k, double d) const { auto it = kv.find(k); return it == kv.end() ? d : it->second; }
  int geti(const char* k, int d) const { return (int)get(k, d); }
};

static void profile_tier_level(BitWriter& bw, int profile_idc) {
  bw.u(2, 0); bw.u1(0); bw.u(5, profile_idc);
  for (int i = 0; i < 32; i++) bw.u1(i == profile_idc || (profile_idc == 1 && i == 2));
  bw.u1(1); bw.u1(0); bw.u1(0); bw.u1(1);   // progressive, not interlaced, not packed, frame only
  bw.u(32, 0); bw.u(11, 0);                 // 43 reserved zero bits
  bw.u1(0);
  bw.u(8, 153);                             // level 5.1
}

// scaling_list_data() (7.3.4) with random but legal content
static void scaling_list_data(BitWriter& bw, Rng& rng) {
  for (int size_id = 0; size_id < 4; size_id++)
    for (int matrix_id = 0; matrix_id < 6; matrix_id += (size_id == 3 ? 3 : 1)) {
      const bool explicit_list = rng.chance(0.6);
      bw.u1(explicit_list);
      if (!explicit_list) {
        // 0 = the default list, d = copy of list matrix_id - d (for 32x32: matrix_id - 3 d)
        const int max_delta = size_id == 3 ? matrix_id / 3 : matrix_id;
        bw.ue((uint32_t)rng.range(0, max_delta));
        continue;
      }
      int next = 8;
      const int n = size_id == 0 ? 16 : 64;
      if (size_id > 1) {
        const int dc = rng.range(1, 255);
        bw.se(dc - 8);
        next = dc;
      }
      const int centre = rng.range(8, 60);
      for (int i = 0; i < n; i++) {
        int target = centre + (i * rng.range(0, 3)) / 2 + rng.range(-4, 4);
        if (rng.chance(0.03)) target = rng.range(1, 255);
        target = target < 1 ? 1 : (target > 255 ? 255 : target);
        int delta = target - next;
        if (delta > 127) delta -= 256;
        if (delta < -128) delta += 256;
        bw.se(delta);
        next = (next + delta + 256) % 256;
      }
    }
}

struct StreamParams {
  int w, h, log2_ctb, log2_min_cb, log2_min_tb, log2_max_tb, th_depth_intra;
  int sao, pcm, pcm_bits_luma, pcm_bits_chroma, log2_min_pcm, log2_max_pcm, pcm_lf_disabled;
  int scaling;   // 0 off, 1 default lists, 2 SPS lists, 3 PPS lists
  int strong_intra, log2_max_poc_lsb;
  int rext_rotation, rext_context, rext_implicit_rdpcm;
  // PPS
  int dep_slices, sign_hiding, init_qp, cip, tskip, log2_max_tskip, cu_qp_delta, diff_cu_qp_delta_depth;
  int cb_qp_offset, cr_qp_offset, slice_chroma_offsets, tq_bypass, tile_cols, tile_rows, tiles_uniform, wpp;
  int lf_across_tiles, lf_across_slices, dbk_control, dbk_override_enabled, dbk_disabled, beta_div2, tc_div2;
  int chroma_qp_offset_list, diff_cu_chroma_qp_offset_depth;
  // inter
  int inter, th_depth_inter, amp, tmvp, weighted_pred, weighted_bipred, lists_modification, par_merge_level, cabac_init_present;
  int num_ref_default[2];
  int reorder, long_term;
  std::vector<int> col_w, row_h;   // explicit tile sizes (non-uniform)
};

static std::vector<uint8_t> write_vps(int dpb_minus1 = 1, int reorder = 0) {
  BitWriter bw;
  bw.u(4, 0); bw.u1(1); bw.u1(1); bw.u(6, 0); bw.u(3, 0); bw.u1(1); bw.u(16, 0xffff);
  profile_tier_level(bw, 1);
  bw.u1(1); bw.ue(dpb_minus1); bw.ue(reorder); bw.ue(0);
  bw.u(6, 0); bw.ue(0); bw.u1(0); bw.u1(0);
  bw.trailing();
  return bw.b;
}

static std::vector<uint8_t> write_sps(const StreamParams& p, Rng& rng) {
  BitWriter bw;
  const bool rext = p.rext_rotation || p.rext_context || p.rext_implicit_rdpcm;
  bw.u(4, 0); bw.u(3, 0); bw.u1(1);
  profile_tier_level(bw, rext ? 4 : 1);
  bw.ue(0);            // sps id
  bw.ue(1);            // 4:2:0
  bw.ue(p.w); bw.ue(p.h);
  bw.u1(0);            // no conformance window
  bw.ue(0); bw.ue(0);  // 8 bit
  bw.ue(p.log2_max_poc_lsb - 4);
  bw.u1(1); bw.ue(p.inter ? 6 : 1); bw.ue(p.reorder ? 3 : 0); bw.ue(0);
  bw.ue(p.log2_min_cb - 3); bw.ue(p.log2_ctb - p.log2_min_cb);
  bw.ue(p.log2_min_tb - 2); bw.ue(p.log2_max_tb - p.log2_min_tb);
  bw.ue(p.th_depth_inter);
  bw.ue(p.th_depth_intra);
  bw.u1(p.scaling != 0);
  if (p.scaling) {
    bw.u1(p.scaling == 2);
    if (p.scaling == 2) scaling_list_data(bw, rng);
  }
  bw.u1(p.amp);
  bw.u1(p.sao);
  bw.u1(p.pcm);
  if (p.pcm) {
    bw.u(4, p.pcm_bits_luma - 1); bw.u(4, p.pcm_bits_chroma - 1);
    bw.ue(p.log2_min_pcm - 3); bw.ue(p.log2_max_pcm - p.log2_min_pcm);
    bw.u1(p.pcm_lf_disabled);
  }
  bw.ue(0);            // no short-term RPS in the SPS
  bw.u1(p.long_term);
  if (p.long_term) bw.ue(0);   // none listed in the SPS
  bw.u1(p.tmvp);
  bw.u1(p.strong_intra);
  bw.u1(0);            // VUI
  bw.u1(rext);
  if (rext) {
    bw.u1(1); bw.u1(0); bw.u1(0); bw.u(5, 0);
    bw.u1(p.rext_rotation); bw.u1(p.rext_context); bw.u1(p.rext_implicit_rdpcm);
    bw.u1(0); bw.u1(0); bw.u1(0); bw.u1(0); bw.u1(0); bw.u1(0);
  }
  bw.trailing();
  return bw.b;
}

static std::vector<uint8_t> write_pps(const StreamParams& p, Rng& rng) {
  BitWriter bw;
  const bool tiles = p.tile_cols > 1 || p.tile_rows > 1;
  bw.ue(0); bw.ue(0);
  bw.u1(p.dep_slices); bw.u1(0); bw.u(3, 0); bw.u1(p.sign_hiding); bw.u1(p.cabac_init_present);
  bw.ue(p.num_ref_default[0] - 1); bw.ue(p.num_ref_default[1] - 1);
  bw.se(p.init_qp - 26);
  bw.u1(p.cip); bw.u1(p.tskip);
  bw.u1(p.cu_qp_delta);
  if (p.cu_qp_delta) bw.ue(p.diff_cu_qp_delta_depth);
  bw.se(p.cb_qp_offset); bw.se(p.cr_qp_offset);
  bw.u1(p.slice_chroma_offsets);
  bw.u1(p.weighted_pred); bw.u1(p.weighted_bipred);
  bw.u1(p.tq_bypass);
  bw.u1(tiles); bw.u1(p.wpp);
  if (tiles) {
    bw.ue(p.tile_cols - 1); bw.ue(p.tile_rows - 1);
    bw.u1(p.tiles_uniform);
    if (!p.tiles_uniform) {
      for (int i = 0; i + 1 < p.tile_cols; i++) bw.ue(p.col_w[i] - 1);
      for (int i = 0; i + 1 < p.tile_rows; i++) bw.ue(p.row_h[i] - 1);
    }
    bw.u1(p.lf_across_tiles);
  }
  bw.u1(p.lf_across_slices);
  bw.u1(p.dbk_control);
  if (p.dbk_control) {
    bw.u1(p.dbk_override_enabled);
    bw.u1(p.dbk_disabled);
    if (!p.dbk_disabled) { bw.se(p.beta_div2); bw.se(p.tc_div2); }
  }
  bw.u1(p.scaling == 3);
  if (p.scaling == 3) scaling_list_data(bw, rng);
  bw.u1(p.lists_modification);
  bw.ue(p.par_merge_level - 2);
  bw.u1(0);            // slice header extension
  const bool rext = (p.tskip && p.log2_max_tskip > 2) || p.chroma_qp_offset_list;
  bw.u1(rext);
  if (rext) {
    bw.u1(1); bw.u1(0); bw.u1(0); bw.u(5, 0);
    if (p.tskip) bw.ue(p.log2_max_tskip - 2);
    bw.u1(0);          // cross component prediction (4:4:4 only)
    bw.u1(p.chroma_qp_offset_list);
    if (p.chroma_qp_offset_list) {
      bw.ue(p.diff_cu_chroma_qp_offset_depth);
      const int n = rng.range(1, 6);
      bw.ue(n - 1);
      for (int i = 0; i < n; i++) { bw.se(rng.range(-12, 12)); bw.se(rng.range(-12, 12)); }
    }
    bw.ue(0); bw.ue(0);   // SAO offset scaling (bit depth <= 10: 0)
  }
  bw.trailing();
  return bw.b;
}

struct SegmentParams {
  int nal_type = NAL_IDR_W_RADL;
  bool first = true, dependent = false;
  int address = 0;
  int poc_lsb = 0;
  bool sao_luma = false, sao_chroma = false;
  int qp_delta = 0, cb_off = 0, cr_off = 0;
  bool cu_chroma_qp_offset = false;
  bool dbk_override = false, dbk_disabled = false;
  int beta_div2 = 0, tc_div2 = 0;
  bool lf_across = false;
  // inter
  int slice_type = SLICE_I;
  int n_neg = 0, delta_poc_minus1[16], used[16];   // short-term RPS of the picture: negative pictures ...
  int n_pos = 0, delta_pos_minus1[16], used_pos[16]; // ... and positive ones (streams with reordering)
  int n_lt = 0, lt_lsb[16], lt_used[16], lt_msb_present[16], lt_msb_delta[16];   // long-term pictures
  bool tmvp = false, override_num_ref = false, mvd_l1_zero = false, cabac_init = false, col_from_l0 = true;
  int num_ref[2] = {1, 1}, max_merge_cand = 5;
  bool list_mod[2] = {false, false};
  int list_entry[2][16];
  // prediction weights
  int luma_denom = 0, delta_chroma_denom = 0;
  int lw_flag[2][16], cw_flag[2][16], d_lw[2][16], l_off[2][16], d_cw[2][16][2], d_coff[2][16][2];
};

static int ceil_log2(uint32_t v) { int n = 0; while ((1u << n) < v) n++; return n; }

// slice_segment_header() (7.3.6.1) of an I slice; entry_sizes = bytes of every sub-stream but the last
static std::vector<uint8_t> write_slice_header(const Sps& sps, const Pps& pps, const SegmentParams& s,
                                               const std::vector<uint32_t>& entry_sizes) {
  BitWriter bw;
  bw.u1(s.first);
  if (s.nal_type >= NAL_BLA_W_LP && s.nal_type <= NAL_RSV_IRAP_VCL23) bw.u1(0);
  bw.ue(0);
  if (!s.first) {
    if (pps.dependent_slices_enabled) bw.u1(s.dependent);
    bw.u(ceil_log2((uint32_t)(sps.ctbs_w * sps.ctbs_h)), (uint32_t)s.address);
  }
  if (!s.dependent) {
    bw.ue((uint32_t)s.slice_type);
    int n_curr = 0;
    if (s.nal_type != NAL_IDR_W_RADL && s.nal_type != NAL_IDR_N_LP) {
      bw.u(sps.log2_max_poc_lsb, (uint32_t)s.poc_lsb);
      bw.u1(0);               // RPS coded in the slice header (the SPS holds no sets, so no inter-RPS flag)
      bw.ue((uint32_t)s.n_neg); bw.ue((uint32_t)s.n_pos);
      for (int i = 0; i < s.n_neg; i++) { bw.ue((uint32_t)s.delta_poc_minus1[i]); bw.u1(s.used[i]); n_curr += s.used[i]; }
      for (int i = 0; i < s.n_pos; i++) { bw.ue((uint32_t)s.delta_pos_minus1[i]); bw.u1(s.used_pos[i]); n_curr += s.used_pos[i]; }
      if (sps.long_term_ref_pics_present) {
        bw.ue((uint32_t)s.n_lt);
        for (int i = 0; i < s.n_lt; i++) {
          bw.u(sps.log2_max_poc_lsb, (uint32_t)s.lt_lsb[i]);
          bw.u1(s.lt_used[i]);
          bw.u1(s.lt_msb_present[i]);
          if (s.lt_msb_present[i]) bw.ue((uint32_t)s.lt_msb_delta[i]);
          n_curr += s.lt_used[i];
        }
      }
      if (sps.temporal_mvp_enabled) bw.u1(s.tmvp);
    }
    if (sps.sao_enabled) { bw.u1(s.sao_luma); bw.u1(s.sao_chroma); }
    if (s.slice_type != SLICE_I) {
      const int nl = s.slice_type == SLICE_B ? 2 : 1;
      bw.u1(s.override_num_ref);
      if (s.override_num_ref) for (int l = 0; l < nl; l++) bw.ue((uint32_t)s.num_ref[l] - 1);
      if (pps.lists_modification_present && n_curr > 1) {
        int nb = 0;
        while ((1 << nb) < n_curr) nb++;
        for (int l = 0; l < nl; l++) {
          bw.u1(s.list_mod[l]);
          if (s.list_mod[l]) for (int i = 0; i < s.num_ref[l]; i++) bw.u(nb, (uint32_t)s.list_entry[l][i]);
        }
      }
      if (s.slice_type == SLICE_B) bw.u1(s.mvd_l1_zero);
      if (pps.cabac_init_present) bw.u1(s.cabac_init);
      if (s.tmvp) {
        if (s.slice_type == SLICE_B) bw.u1(s.col_from_l0);
        if ((s.col_from_l0 && s.num_ref[0] > 1) || (!s.col_from_l0 && s.num_ref[1] > 1)) bw.ue(0);   // collocated_ref_idx
      }
      if ((pps.weighted_pred && s.slice_type == SLICE_P) || (pps.weighted_bipred && s.slice_type == SLICE_B)) {
        bw.ue((uint32_t)s.luma_denom);
        bw.se(s.delta_chroma_denom);
        for (int l = 0; l < nl; l++) {
          for (int i = 0; i < s.num_ref[l]; i++) bw.u1(s.lw_flag[l][i]);
          for (int i = 0; i < s.num_ref[l]; i++) bw.u1(s.cw_flag[l][i]);
          for (int i = 0; i < s.num_ref[l]; i++) {
            if (s.lw_flag[l][i]) { bw.se(s.d_lw[l][i]); bw.se(s.l_off[l][i]); }
            if (s.cw_flag[l][i]) for (int j = 0; j < 2; j++) { bw.se(s.d_cw[l][i][j]); bw.se(s.d_coff[l][i][j]); }
          }
        }
      }
      bw.ue((uint32_t)(5 - s.max_merge_cand));
    }
    bw.se(s.qp_delta);
    if (pps.slice_chroma_qp_offsets_present) { bw.se(s.cb_off); bw.se(s.cr_off); }
    if (pps.chroma_qp_offset_list_enabled) bw.u1(s.cu_chroma_qp_offset);
    if (pps.deblocking_override_enabled) bw.u1(s.dbk_override);
    bool disabled = pps.deblocking_disabled;
    if (s.dbk_override) {
      bw.u1(s.dbk_disabled);
      disabled = s.dbk_disabled;
      if (!s.dbk_disabled) { bw.se(s.beta_div2); bw.se(s.tc_div2); }
    }
    if (pps.lf_across_slices && (s.sao_luma || s.sao_chroma || !disabled)) bw.u1(s.lf_across);
  }
  if (pps.tiles_enabled || pps.entropy_coding_sync) {
    bw.ue((uint32_t)entry_sizes.size());
    if (!entry_sizes.empty()) {
      uint32_t mx = 0;
      for (uint32_t e : entry_sizes) mx = std::max(mx, e - 1);
      int len = 1;
      while (len < 32 && (mx >> len)) len++;
      bw.ue(len - 1);
      for (uint32_t e : entry_sizes) bw.u(len, e - 1);
    }
  }
  bw.trailing();   // byte_alignment()
  return bw.b;
}

int main(int argc, char** argv) {
  if (argc < 2) { fprintf(stderr, "usage: %s out.bin key=value ...\n", argv[0]); return 2; }
  Config cfg;
  for (int i = 2; i < argc; i++) {
    const char* eq = strchr(argv[i], '=');
    if (!eq) { fprintf(stderr, "bad argument %s\n", argv[i]); return 2; }
    cfg.kv[std::string(argv[i], eq - argv[i])] = atof(eq + 1);
  }
  Rng rng((uint64_t)cfg.geti("seed", 1));
  StreamParams p = {};
  p.w = cfg.geti("w", 416); p.h = cfg.geti("h", 240);
  p.log2_ctb = cfg.geti("ctb", 5); p.log2_min_cb = cfg.geti("mincb", 3);
  p.log2_min_tb = cfg.geti("mintb", 2); p.log2_max_tb = cfg.geti("maxtb", std::min(5, p.log2_ctb));
  p.th_depth_intra = cfg.geti("thdepth", 2);
  p.sao = cfg.geti("sao", 0);
  p.pcm = cfg.geti("pcm", 0);
  p.pcm_bits_luma = cfg.geti("pcm_bits_luma", 8); p.pcm_bits_chroma = cfg.geti("pcm_bits_chroma", 8);
  p.log2_min_pcm = cfg.geti("minpcm", 3); p.log2_max_pcm = cfg.geti("maxpcm", std::min(5, p.log2_ctb));
  p.pcm_lf_disabled = cfg.geti("pcm_lf_disabled", 0);
  p.scaling = cfg.geti("scaling", 0);
  p.strong_intra = cfg.geti("strong_intra", 1);
  p.log2_max_poc_lsb = cfg.geti("poc_bits", 8);
  p.rext_rotation = cfg.geti("rext_rotation", 0); p.rext_context = cfg.geti("rext_context", 0);
  p.rext_implicit_rdpcm = cfg.geti("rext_implicit_rdpcm", 0);
  p.dep_slices = cfg.geti("dep_slices", 0);
  p.sign_hiding = cfg.geti("sign_hiding", 0);
  p.init_qp = cfg.geti("init_qp", 26);
  p.cip = cfg.geti("cip", 0);
  p.tskip = cfg.geti("tskip", 0); p.log2_max_tskip = cfg.geti("max_tskip", 2);
  p.cu_qp_delta = cfg.geti("cu_qp_delta", 0); p.diff_cu_qp_delta_depth = cfg.geti("qp_delta_depth", 1);
  p.cb_qp_offset = cfg.geti("cb_qp_offset", 0); p.cr_qp_offset = cfg.geti("cr_qp_offset", 0);
  p.slice_chroma_offsets = cfg.geti("slice_chroma_offsets", 0);
  p.tq_bypass = cfg.geti("tq_bypass", 0);
  p.tile_cols = cfg.geti("tile_cols", 1); p.tile_rows = cfg.geti("tile_rows", 1);
  p.tiles_uniform = cfg.geti("tiles_uniform", 1);
  p.wpp = cfg.geti("wpp", 0);
  p.lf_across_tiles = cfg.geti("lf_across_tiles", 1); p.lf_across_slices = cfg.geti("lf_across_slices", 1);
  p.dbk_control = cfg.geti("dbk_control", 1); p.dbk_override_enabled = cfg.geti("dbk_override", 0);
  p.dbk_disabled = cfg.geti("dbk_disabled", 0);
  p.beta_div2 = cfg.geti("beta_div2", 0); p.tc_div2 = cfg.geti("tc_div2", 0);
  p.chroma_qp_offset_list = cfg.geti("chroma_qp_offset_list", 0);
  p.diff_cu_chroma_qp_offset_depth = cfg.geti("chroma_qp_offset_depth", 1);
  p.inter = cfg.geti("inter", 0);
  p.reorder = cfg.geti("reorder", 0);
  p.long_term = cfg.geti("long_term", 0);
  p.th_depth_inter = cfg.geti("thdepth_inter", 1);
  p.amp = cfg.geti("amp", 0);
  p.tmvp = cfg.geti("tmvp", 0);
  p.weighted_pred = cfg.geti("weighted_pred", 0); p.weighted_bipred = cfg.geti("weighted_bipred", 0);
  p.lists_modification = cfg.geti("lists_modification", 0);
  p.par_merge_level = cfg.geti("par_merge_level", 2);
  p.cabac_init_present = cfg.geti("cabac_init_present", 0);
  p.num_ref_default[0] = cfg.geti("num_ref_default_l0", 1); p.num_ref_default[1] = cfg.geti("num_ref_default_l1", 1);
  const int max_refs = cfg.geti("max_refs", 3);
  const double p_bslice = cfg.get("p_bslice", 0.5), p_islice = cfg.get("p_islice", 0.15);
  const int frames = cfg.geti("frames", 2);
  const int slice_ctbs = cfg.geti("slice_ctbs", 0);   // mean slice-segment length in CTBs (0 = one per picture)
  const int qp_lo = cfg.geti("qp_lo", 22), qp_hi = cfg.geti("qp_hi", 38);
  const double density = cfg.get("density", 1.0);

  const int ctbs_w = (p.w + (1 << p.log2_ctb) - 1) >> p.log2_ctb, ctbs_h = (p.h + (1 << p.log2_ctb) - 1) >> p.log2_ctb;
  if (!p.tiles_uniform) {   // random explicit tile sizes
    auto split = [&](int total, int parts, std::vector<int>& out) {
      out.assign(parts, 1);
      for (int left = total - parts; left > 0; left--) out[rng.below(parts)]++;
    };
    split(ctbs_w, p.tile_cols, p.col_w);
    split(ctbs_h, p.tile_rows, p.row_h);
  }

  std::vector<uint8_t> stream;
  put_nal(stream, NAL_VPS, escape(write_vps(p.inter ? 6 : 1, p.reorder ? 3 : 0)));
  const std::vector<uint8_t> sps_rbsp = write_sps(p, rng), pps_rbsp = write_pps(p, rng);
  put_nal(stream, NAL_SPS, escape(sps_rbsp));
  put_nal(stream, NAL_PPS, escape(pps_rbsp));

  // parameter sets as the PARSER understands them (derived tables included)
  ParamSets ps;
  {
    BitReader br(sps_rbsp.data(), sps_rbsp.size());
    std::shared_ptr<Sps> sps;
    if (parse_sps(br, sps)) { fprintf(stderr, "streamgen: own SPS does not parse\n"); return 1; }
    ps.sps[sps->id] = sps;
    BitReader bp(pps_rbsp.data(), pps_rbsp.size());
    std::shared_ptr<Pps> pps;
    if (parse_pps(bp, ps, pps)) { fprintf(stderr, "streamgen: own PPS does not parse\n"); return 1; }
    ps.pps[pps->id] = pps;
  }
  std::shared_ptr<const Sps> sps = ps.sps[0];
  std::shared_ptr<const Pps> pps = ps.pps[0];
  const int n_ctbs = sps->ctbs_w * sps->ctbs_h;

  // bin probabilities per context
  double prob[CTX_COUNT];
  for (int i = 0; i < CTX_COUNT; i++) prob[i] = 0.5;
  auto setp = [&](int off, int n, double v) { for (int i = 0; i < n; i++) prob[off + i] = v; };
  setp(CTX_SAO_MERGE, 1, cfg.get("p_sao_merge", 0.3));
  setp(CTX_SAO_TYPE, 1, cfg.get("p_sao_on", 0.7));
  setp(CTX_SPLIT_CU, 3, cfg.get("p_split_cu", 0.5));
  setp(CTX_CU_TQ_BYPASS, 1, cfg.get("p_tq_bypass", 0.15));
  setp(CTX_PART_MODE, 4, cfg.get("p_2nx2n", 0.6));
  setp(CTX_PREV_INTRA_LUMA, 1, 0.4);
  setp(CTX_INTRA_CHROMA, 1, 0.5);
  setp(CTX_SPLIT_TRANSFORM, 3, cfg.get("p_split_tu", 0.4));
  setp(CTX_CBF_LUMA, 2, std::min(0.95, 0.6 * density));
  setp(CTX_CBF_CHROMA, 5, std::min(0.95, 0.4 * density));
  prob[CTX_CU_QP_DELTA] = 0.6; prob[CTX_CU_QP_DELTA + 1] = 0.6;
  setp(CTX_TRANSFORM_SKIP, 2, cfg.get("p_tskip", 0.35));
  setp(CTX_LAST_X, 36, cfg.get("p_last", 0.55));
  setp(CTX_CODED_SUB_BLOCK, 4, std::min(0.9, 0.4 * density));
  setp(CTX_SIG_COEFF, 44, std::min(0.9, 0.25 * density));
  setp(CTX_GT1, 24, 0.3);
  setp(CTX_GT2, 6, 0.3);
  setp(CTX_CU_SKIP, 3, cfg.get("p_skip", 0.3));
  setp(CTX_PRED_MODE, 1, cfg.get("p_intra_cu", 0.25));
  setp(CTX_RQT_ROOT_CBF, 1, 0.6);
  setp(CTX_MERGE_FLAG, 1, cfg.get("p_merge", 0.5));
  setp(CTX_ABS_MVD_GT0, 1, 0.6);
  setp(CTX_CHROMA_QP_OFFSET_FLAG, 1, 0.6);
  setp(CTX_CHROMA_QP_OFFSET_IDX, 1, 0.5);

  int poc = 0;
  size_t total_pus = 0;
  int anchor = 0, mini = 0;
  const bool open_gop = cfg.geti("open_gop", 0) != 0;
  bool rasl_group = false;
  Dpb dpb;
  for (int f = 0; f < frames; f++) {
    Picture pic;
    pic.sps = sps;
    pic.pps = pps;
    pic.init_maps();
    PicDecodeState st;
    st.wpp.assign((size_t)sps->ctbs_h * CTX_COUNT, 0);
    st.wpp_valid.assign(sps->ctbs_h, 0);
    int nal_type = NAL_IDR_W_RADL;
    if (f > 0) {
      const int r = (int)rng.below(p.inter ? 16 : 8);
      nal_type = r == 0 ? NAL_IDR_N_LP : (r == 1 ? NAL_CRA : (r == 2 && !p.inter ? NAL_TRAIL_N : NAL_TRAIL_R));
    }
    const bool idr = nal_type == NAL_IDR_W_RADL || nal_type == NAL_IDR_N_LP;
    if (p.reorder && !idr && nal_type != NAL_TRAIL_R) nal_type = NAL_TRAIL_R;   // (no CRA inside a reordered group)
    // open GOP: the anchor of a reordered group becomes a CRA picture (intra, but its reference picture set keeps the
    // earlier pictures) and the three pictures that precede it in output order become its RASL pictures.  Decoded from
    // the start they are ordinary pictures; a decoder that starts at the CRA (or meets it after an end-of-sequence NAL)
    // must skip them.
    if (idr) rasl_group = false;
    if (open_gop && p.reorder && p.inter && !idr && f > 0) {
      if (mini == 0) {
        rasl_group = rng.chance(cfg.get("p_cra", 0.4));
        if (rasl_group) nal_type = NAL_CRA;
      } else if (rasl_group) nal_type = NAL_RASL_R;
    }
    if (idr) { poc = 0; anchor = 0; mini = 0; rasl_group = false; }
    else if (!p.reorder) poc++;
    else {   // groups of four pictures decoded as anchor+4, +2, +1, +3 (output order differs from decoding order)
      static const int step[4] = {4, 2, 1, 3};
      poc = anchor + step[mini];
      if (++mini == 4) { mini = 0; anchor += 4; }
    }
    pic.poc = poc;
    pic.self->poc = poc;
    // ---- reference picture set of the picture (negative pictures only: decoding order = output order) ----
    SegmentParams rps_params;
    bool pic_tmvp = false;
    const bool cra = nal_type == NAL_CRA;
    if (p.inter && (nal_type == NAL_TRAIL_R || nal_type == NAL_RASL_R || (cra && open_gop))) {
      std::vector<int> pocs;
      for (const auto& r : dpb.pics) pocs.push_back(r->poc);
      std::sort(pocs.begin(), pocs.end(), [](int a, int b) { return a > b; });   // closest first
      int n_after = 0;
      for (int q : pocs) n_after += q > poc;
      if (p.long_term) {
        const int max_lsb = 1 << sps->log2_max_poc_lsb;
        std::vector<int> lts;
        for (const auto& r : dpb.pics) {
          if (r->poc > poc) continue;   // (pictures still to be output after this one stay short-term)
          if (r->is_lt ? rng.chance(0.9) : rng.chance(0.2)) lts.push_back(r->poc);
          else if (r->is_lt) pocs.erase(std::find(pocs.begin(), pocs.end(), r->poc));   // once long-term, never short-term again
        }
        std::sort(lts.begin(), lts.end(), [](int a, int b) { return a > b; });   // MSB cycles must not decrease
        int acc = 0;
        for (int q : lts) {
          pocs.erase(std::find(pocs.begin(), pocs.end(), q));
          if (rps_params.n_lt >= 2) continue;   // dropped
          const int i = rps_params.n_lt++;
          int same_lsb = 0;
          for (const auto& r : dpb.pics) same_lsb += ((r->poc ^ q) & (max_lsb - 1)) == 0;
          const int cycle = ((poc & ~(max_lsb - 1)) - (q & ~(max_lsb - 1))) / max_lsb;
          rps_params.lt_lsb[i] = q & (max_lsb - 1);
          rps_params.lt_used[i] = !cra && rng.chance(0.8);
          rps_params.lt_msb_present[i] = same_lsb > 1 || rng.chance(0.4);
          rps_params.lt_msb_delta[i] = cycle - acc;
          if (rps_params.lt_msb_present[i]) acc = cycle;
        }
      }
      int prev = poc;
      for (int q : pocs) {   // pictures before this one in output order, closest first
        if (q > poc) continue;
        // in a reordered group the anchors and the group's earlier pictures must stay until the group is complete
        const bool must_keep = p.reorder && q >= anchor - 4;
        if (!must_keep && (rps_params.n_neg >= max_refs + 1 || !rng.chance(0.85))) continue;    // dropped from the DPB
        if (rps_params.n_neg + rps_params.n_lt + n_after >= 6) continue;   // sps_max_dec_pic_buffering_minus1
        rps_params.delta_poc_minus1[rps_params.n_neg] = prev - q - 1;
        rps_params.used[rps_params.n_neg] = !cra && rng.chance(0.8);
        rps_params.n_neg++;
        prev = q;
      }
      prev = poc;
      for (auto it = pocs.rbegin(); it != pocs.rend(); ++it) {   // pictures after this one in output order, closest first
        const int q = *it;
        if (q < poc) continue;
        rps_params.delta_pos_minus1[rps_params.n_pos] = q - prev - 1;
        rps_params.used_pos[rps_params.n_pos] = !cra && rng.chance(0.8);
        rps_params.n_pos++;
        prev = q;
      }
      pic_tmvp = !cra && p.tmvp && rng.chance(0.8);
    }
    int n_curr = 0;
    for (int i = 0; i < rps_params.n_neg; i++) n_curr += rps_params.used[i];
    for (int i = 0; i < rps_params.n_pos; i++) n_curr += rps_params.used_pos[i];
    for (int i = 0; i < rps_params.n_lt; i++) n_curr += rps_params.lt_used[i];
    bool rps_applied = false;

    // ---- slice segment plan in tile-scan order: (first, last) addresses + dependent flag -------------------
    struct Seg { int first_ts, last_ts; bool dependent; };
    std::vector<Seg> segs;
    {
      const bool tiles = pps->tiles_enabled;
      const bool whole_tiles = tiles && rng.chance(0.5);
      int ts = 0;
      bool slice_started_midrow = false;
      while (ts < n_ctbs) {
        const int rs = pps->ts_to_rs[ts];
        const int tile = pps->tile_id_rs[rs];
        int tile_end = ts;   // last ts address of this tile
        while (tile_end + 1 < n_ctbs && pps->tile_id_rs[pps->ts_to_rs[tile_end + 1]] == tile) tile_end++;
        int last;
        bool force_independent = segs.empty();
        if (slice_ctbs <= 0) last = n_ctbs - 1;
        else if (whole_tiles) {
          last = tile_end;
          // a few more whole tiles
          while (last + 1 < n_ctbs && rng.chance(0.4)) {
            const int t2 = pps->tile_id_rs[pps->ts_to_rs[last + 1]];
            while (last + 1 < n_ctbs && pps->tile_id_rs[pps->ts_to_rs[last + 1]] == t2) last++;
          }
        } else {
          last = ts + (int)rng.below((uint32_t)(2 * slice_ctbs));
          if (tiles) {
            last = std::min(last, tile_end);
            // a slice lies inside one tile: the first segment of every tile starts a new slice
            const bool tile_start = ts == 0 || pps->tile_id_rs[pps->ts_to_rs[ts - 1]] != tile;
            if (tile_start) force_independent = true;
          }
          if (pps->entropy_coding_sync) {
            const int x = rs % sps->ctbs_w;
            const int row_end = ts + (sps->ctbs_w - 1 - x);
            if (x != 0) {
              last = std::min(last, row_end);   // a segment that starts inside a row ends in that row
            } else if (slice_started_midrow) {
              force_independent = true;         // ... and so does the slice it belongs to
            }
          }
          last = std::min(last, n_ctbs - 1);
        }
        const bool dependent = !force_independent && pps->dependent_slices_enabled && rng.chance(0.5);
        if (!dependent) slice_started_midrow = pps->entropy_coding_sync && (rs % sps->ctbs_w) != 0;
        segs.push_back({ts, last, dependent});
        ts = last + 1;
      }
    }

    SliceHeader prev_independent;
    int n_slices = 0;
    for (size_t si = 0; si < segs.size(); si++) {
      SegmentParams sp;
      sp.nal_type = nal_type;
      sp.first = si == 0;
      sp.dependent = segs[si].dependent;
      sp.address = pps->ts_to_rs[segs[si].first_ts];
      sp.poc_lsb = poc & ((1 << sps->log2_max_poc_lsb) - 1);
      sp.n_neg = rps_params.n_neg;
      memcpy(sp.delta_poc_minus1, rps_params.delta_poc_minus1, sizeof(sp.delta_poc_minus1));
      memcpy(sp.used, rps_params.used, sizeof(sp.used));
      sp.n_lt = rps_params.n_lt;
      memcpy(sp.lt_lsb, rps_params.lt_lsb, sizeof(sp.lt_lsb)); memcpy(sp.lt_used, rps_params.lt_used, sizeof(sp.lt_used));
      memcpy(sp.lt_msb_present, rps_params.lt_msb_present, sizeof(sp.lt_msb_present));
      memcpy(sp.lt_msb_delta, rps_params.lt_msb_delta, sizeof(sp.lt_msb_delta));
      sp.n_pos = rps_params.n_pos;
      memcpy(sp.delta_pos_minus1, rps_params.delta_pos_minus1, sizeof(sp.delta_pos_minus1));
      memcpy(sp.used_pos, rps_params.used_pos, sizeof(sp.used_pos));
      sp.tmvp = pic_tmvp;
      if (!sp.dependent && n_curr > 0 && !rng.chance(p_islice)) {
        sp.slice_type = rng.chance(p_bslice) ? SLICE_B : SLICE_P;
        // Reference quirk kept out of the golden set: for a P slice the reference's collocated-MV rule walks
        // RefPicList[1] (motion.cc:1189-1197), which a P slice never fills - it reads stale DPB indices, and with output
        // reordering those can name a picture after the current one, flipping the list the collocated vector is taken
        // from.  The spec (8.5.3.2.9) and this repo's parser only look at the lists the slice has.
        if (open_gop && pic_tmvp) sp.slice_type = SLICE_B;
        const int nl = sp.slice_type == SLICE_B ? 2 : 1;
        sp.override_num_ref = rng.chance(0.6);
        for (int l = 0; l < 2; l++) sp.num_ref[l] = sp.override_num_ref ? rng.range(1, std::min(4, n_curr + 1)) : pps->num_ref_idx_default[l];
        if (pps->lists_modification_present && n_curr > 1)
          for (int l = 0; l < nl; l++) {
            sp.list_mod[l] = rng.chance(0.5);
            // entry 0 stays the first picture of the initial list: every slice of a picture must name the same
            // collocated picture (collocated_ref_idx is always 0 here)
            for (int i = 0; i < sp.num_ref[l]; i++) sp.list_entry[l][i] = i == 0 ? 0 : rng.range(0, n_curr - 1);
          }
        sp.mvd_l1_zero = rng.chance(0.3);
        sp.cabac_init = rng.chance(0.5);
        sp.col_from_l0 = sp.slice_type == SLICE_P || rng.chance(0.5);
        sp.max_merge_cand = rng.range(1, 5);
        sp.luma_denom = rng.range(0, 7);
        sp.delta_chroma_denom = rng.range(-sp.luma_denom, 7 - sp.luma_denom);
        for (int l = 0; l < 2; l++)
          for (int i = 0; i < 16; i++) {
            sp.lw_flag[l][i] = rng.chance(0.6); sp.cw_flag[l][i] = rng.chance(0.6);
            sp.d_lw[l][i] = rng.range(-128, 127); sp.l_off[l][i] = rng.range(-128, 127);
            for (int j = 0; j < 2; j++) { sp.d_cw[l][i][j] = rng.range(-128, 127); sp.d_coff[l][i][j] = rng.range(-512, 511); }
            if (rng.chance(0.5)) {   // moderate weights: the picture stays inside the sample range more often
              sp.d_lw[l][i] = rng.range(-8, 8); sp.l_off[l][i] = rng.range(-10, 10);
              for (int j = 0; j < 2; j++) { sp.d_cw[l][i][j] = rng.range(-8, 8); sp.d_coff[l][i][j] = rng.range(-20, 20); }
            }
          }
      }
      if (!sp.dependent) {
        sp.sao_luma = sps->sao_enabled && rng.chance(0.8);
        sp.sao_chroma = sps->sao_enabled && rng.chance(0.8);
        sp.qp_delta = rng.range(qp_lo, qp_hi) - pps->init_qp;
        sp.cb_off = rng.range(-6, 6); sp.cr_off = rng.range(-6, 6);
        // pps + slice offsets must stay inside [-12, 12]
        sp.cb_off = std::max(-12 - pps->cb_qp_offset, std::min(12 - pps->cb_qp_offset, sp.cb_off));
        sp.cr_off = std::max(-12 - pps->cr_qp_offset, std::min(12 - pps->cr_qp_offset, sp.cr_off));
        sp.cu_chroma_qp_offset = rng.chance(0.8);
        sp.dbk_override = pps->deblocking_override_enabled && rng.chance(0.6);
        sp.dbk_disabled = rng.chance(0.3);
        sp.beta_div2 = rng.range(-6, 6); sp.tc_div2 = rng.range(-6, 6);
        sp.lf_across = rng.chance(0.5);
      }
      // draft header (no entry points) -> the parser's view of it, derived fields included
      SliceUnit unit;
      unit.pps = pps;
      {
        const std::vector<uint8_t> draft = write_slice_header(*sps, *pps, sp, {});
        BitReader br(draft.data(), draft.size());
        std::shared_ptr<const Pps> pp;
        const int r = parse_slice_header(br, nal_type, ps, n_slices ? &prev_independent : nullptr, {}, unit.hdr, pp);
        if (r) { fprintf(stderr, "streamgen: own slice header does not parse (%d)\n", r); return 1; }
      }
      if (!rps_applied) {   // once per picture, from its first slice segment header
        dpb.apply_rps(unit.hdr, *sps, poc, idr || f == 0);
        dpb.pics.push_back(pic.self);
        rps_applied = true;
      }
      if (!sp.dependent) {
        unit.hdr.slice_index = n_slices++;
        prev_independent = unit.hdr;
        SliceRefs sr;
        if (!dpb.build_lists(unit.hdr, pic, sr)) { fprintf(stderr, "streamgen: reference picture missing (frame %d poc %d nal %d, %zu+%zu+%zu in the set, missing %d)\n", f, poc, nal_type, dpb.st_before.size(), dpb.st_after.size(), dpb.lt_curr.size(), (int)dpb.missing); return 1; }
        pic.slice_refs.push_back(sr);
      }
      unit.hdr.data_offset = 0;
      unit.hdr.entry_points.clear();

      std::vector<uint8_t> data;
      std::vector<size_t> subs;
      GenSetup gs;
      gs.rng = &rng;
      gs.prob = prob;
      gs.p_pcm = cfg.get("p_pcm", 0.15);
      gs.p_big_coeff = cfg.get("p_big_coeff", 0.01);
      gs.last_ts = segs[si].last_ts;
      gs.out = &data;
      gs.substreams = &subs;
      const int r = decode_slice_unit<GenEngine>(pic, unit, st, gs);
      if (r) { fprintf(stderr, "streamgen: syntax walk failed (%d) in frame %d segment %zu\n", r, f, si); return 1; }

      std::vector<size_t> marks = subs;
      const std::vector<uint8_t> esc_data = escape(data, &marks);
      std::vector<uint32_t> sizes;
      size_t prev = 0;
      for (size_t m : marks) { sizes.push_back((uint32_t)(m - prev)); prev = m; }
      std::vector<uint8_t> payload = escape(write_slice_header(*sps, *pps, sp, sizes));
      payload.insert(payload.end(), esc_data.begin(), esc_data.end());
      put_nal(stream, nal_type, payload);
    }
    pic.self->slice_refs = pic.slice_refs;
    pic.self->parsed.store(1, std::memory_order_release);
    total_pus += pic.pus.size();
  }
  if (cfg.geti("verbose", 0)) fprintf(stderr, "streamgen: %d pictures, %zu prediction units\n", frames, total_pus);
  FILE* fo = fopen(argv[1], "wb");
  if (!fo) { perror(argv[1]); return 1; }
  fwrite(stream.data(), 1, stream.size(), fo);
  fclose(fo);
  return 0;
}
