// params.cc — SPS / PPS / slice-segment-header parsing (H.265 7.3.2.2, 7.3.2.3, 7.3.6, E.2).
// Functional counterpart of the reference's sps.cc (seq_parameter_set::read :195-520, scaling lists
// :790-940), pps.cc (pic_parameter_set::read :275-620, set_derived_values :622-760), vui.cc,
// refpic.cc (read_short_term_ref_pic_set :60-260) and slice.cc (slice_segment_header::read :360-880).
#include "params.h"
#include <string.h>
#include <algorithm>

namespace b265h {

enum { E_OK = 0, E_RANGE = 8, E_PARSE = 15, E_NOSPS = 1010, E_NOPPS = 1009, E_SLICEHDR = 1007 };

static int ceil_log2(uint32_t v) {
  int b = 0;
  while ((1u << b) < v) b++;
  return b;
}

// ---- scans -------------------------------------------------------------------------------------
static void diag_scan(int n, uint8_t* xs, uint8_t* ys) {
  int i = 0, x = 0, y = 0;
  for (;;) {
    while (y >= 0) {
      if (x < n && y < n) { xs[i] = (uint8_t)x; ys[i] = (uint8_t)y; i++; }
      y--; x++;
    }
    y = x; x = 0;
    if (i >= n * n) break;
  }
}

static const uint8_t kDefault4[16] = {16, 16, 16, 16, 16, 16, 16, 16, 16, 16, 16, 16, 16, 16, 16, 16};
static const uint8_t kDefaultIntra[64] = {16, 16, 16, 16, 16, 16, 16, 16, 16, 16, 17, 16, 17, 16, 17, 18, 17, 18, 18, 17, 18, 21,
                                          19, 20, 21, 20, 19, 21, 24, 22, 22, 24, 24, 22, 22, 24, 25, 25, 27, 30, 27, 25, 25, 29,
                                          31, 35, 35, 31, 29, 36, 41, 44, 41, 36, 47, 54, 54, 47, 65, 70, 65, 88, 88, 115};
static const uint8_t kDefaultInter[64] = {16, 16, 16, 16, 16, 16, 16, 16, 16, 16, 17, 17, 17, 17, 17, 18, 18, 18, 18, 18, 18, 20,
                                          20, 20, 20, 20, 20, 20, 24, 24, 24, 24, 24, 24, 24, 24, 25, 25, 25, 25, 25, 25, 25, 28,
                                          28, 28, 28, 28, 28, 33, 33, 33, 33, 33, 41, 41, 41, 41, 54, 54, 54, 71, 71, 91};

// expand a coded list (diagonal order) into ScalingFactor[y*n+x] for size id 0..3 (7.4.5)
static void fill_factor(ScalingList& sl, int size_id, int mid, const uint8_t* list, int dc) {
  uint8_t xs[64], ys[64];
  if (size_id == 0) {
    diag_scan(4, xs, ys);
    for (int i = 0; i < 16; i++) sl.s4[mid][ys[i] * 4 + xs[i]] = list[i];
    return;
  }
  diag_scan(8, xs, ys);
  const int rep = 1 << (size_id - 1), n = 8 * rep;
  uint8_t* dst = size_id == 1 ? sl.s8[mid] : size_id == 2 ? sl.s16[mid] : sl.s32[mid];
  for (int i = 0; i < 64; i++)
    for (int dy = 0; dy < rep; dy++)
      for (int dx = 0; dx < rep; dx++) dst[(ys[i] * rep + dy) * n + xs[i] * rep + dx] = list[i];
  if (size_id >= 2) dst[0] = (uint8_t)dc;
}

static void default_scaling(ScalingList& sl) {
  for (int m = 0; m < 6; m++) {
    fill_factor(sl, 0, m, kDefault4, 16);
    fill_factor(sl, 1, m, m < 3 ? kDefaultIntra : kDefaultInter, 16);
    fill_factor(sl, 2, m, m < 3 ? kDefaultIntra : kDefaultInter, 16);
  }
  fill_factor(sl, 3, 0, kDefaultIntra, 16);
  fill_factor(sl, 3, 1, kDefaultInter, 16);
}

static int parse_scaling_list(BitReader& br, ScalingList& sl) {
  uint8_t lists[6][64];
  int dcs[6];
  for (int size_id = 0; size_id < 4; size_id++) {
    const int n_mat = size_id == 3 ? 2 : 6, len = size_id == 0 ? 16 : 64;
    for (int m = 0; m < n_mat; m++) {
      int dc = 16;
      if (!br.u1()) {
        const uint32_t delta = br.ue();
        if (br.err || (int)delta > m) return E_RANGE;
        if (delta == 0) {
          const bool intra = size_id == 3 ? (m == 0) : (m < 3);
          memcpy(lists[m], size_id == 0 ? kDefault4 : intra ? kDefaultIntra : kDefaultInter, len);
        } else {
          memcpy(lists[m], lists[m - delta], len);
          dc = dcs[m - delta];
        }
      } else {
        int next = 8;
        if (size_id > 1) {
          const int d = br.se();
          if (d < -7 || d > 247) return E_RANGE;
          dc = next = d + 8;
        }
        for (int i = 0; i < len; i++) {
          const int d = br.se();
          if (d < -128 || d > 127) return E_RANGE;
          next = (next + d + 256) % 256;
          lists[m][i] = (uint8_t)next;
        }
      }
      dcs[m] = dc;
      fill_factor(sl, size_id, m, lists[m], dc);
    }
  }
  return br.err ? E_PARSE : E_OK;
}

// ---- profile_tier_level / VUI / HRD: parsed only to stay in sync ---------------------------------
static void skip_profile_tier_level(BitReader& br, int max_sub_layers_minus1) {
  br.skip(88);  // general profile space/tier/idc, 32 compat flags, 4 source flags, 43+1 reserved
  br.skip(8);   // general_level_idc
  uint8_t prof[8] = {0}, lev[8] = {0};
  for (int i = 0; i < max_sub_layers_minus1; i++) { prof[i] = br.u1(); lev[i] = br.u1(); }
  if (max_sub_layers_minus1 > 0)
    for (int i = max_sub_layers_minus1; i < 8; i++) br.skip(2);
  for (int i = 0; i < max_sub_layers_minus1; i++) {
    if (prof[i]) br.skip(88);
    if (lev[i]) br.skip(8);
  }
}

static void skip_sub_layer_hrd(BitReader& br, int cpb_cnt, int sub_pic) {
  for (int j = 0; j < cpb_cnt; j++) {
    br.ue(); br.ue();
    if (sub_pic) { br.ue(); br.ue(); }
    br.u1();
  }
}
static void skip_hrd(BitReader& br, int common, int max_sub_layers_minus1) {
  int nal = 0, vcl = 0, sub_pic = 0;
  if (common) {
    nal = br.u1(); vcl = br.u1();
    if (nal || vcl) {
      sub_pic = br.u1();
      if (sub_pic) br.skip(8 + 5 + 1 + 5);
      br.skip(8);
      if (sub_pic) br.skip(4);
      br.skip(15);
    }
  }
  for (int i = 0; i <= max_sub_layers_minus1 && !br.err; i++) {
    int fixed_general = br.u1(), fixed_cvs = 1, low_delay = 0, cpb_cnt = 1;
    if (!fixed_general) fixed_cvs = br.u1();
    if (fixed_cvs) br.ue();
    else low_delay = br.u1();
    if (!low_delay) cpb_cnt = br.ue() + 1;
    if (cpb_cnt > 32) { br.err = true; return; }
    if (nal) skip_sub_layer_hrd(br, cpb_cnt, sub_pic);
    if (vcl) skip_sub_layer_hrd(br, cpb_cnt, sub_pic);
  }
}
static void skip_vui(BitReader& br, int max_sub_layers_minus1) {
  if (br.u1()) { if (br.u(8) == 255) br.skip(32); }
  if (br.u1()) br.u1();
  if (br.u1()) { br.skip(4); if (br.u1()) br.skip(24); }
  if (br.u1()) { br.ue(); br.ue(); }
  br.skip(3);
  if (br.u1()) { br.ue(); br.ue(); br.ue(); br.ue(); }
  if (br.u1()) {
    br.skip(64);
    if (br.u1()) br.ue();
    if (br.u1()) skip_hrd(br, 1, max_sub_layers_minus1);
  }
  if (br.u1()) { br.skip(3); br.ue(); br.ue(); br.ue(); br.ue(); br.ue(); }
}

// ---- short-term RPS (7.3.7, 7.4.8) -----------------------------------------------------------------
static int parse_st_rps(BitReader& br, ShortTermRps& out, int idx, int num_sets, const std::vector<ShortTermRps>& sets,
                        bool in_slice_header) {
  out = ShortTermRps();
  int inter = 0;
  if (idx != 0) inter = br.u1();
  if (inter) {
    int delta_idx = 1;
    if (in_slice_header) delta_idx = br.ue() + 1;
    if (delta_idx > idx) return E_RANGE;
    const ShortTermRps& ref = sets[idx - delta_idx];
    const int sign = br.u1(), abs_m1 = br.ue();
    const int delta_rps = (1 - 2 * sign) * (abs_m1 + 1);
    const int nref = ref.num_delta_pocs();
    uint8_t used[33], use_delta[33];
    for (int j = 0; j <= nref; j++) {
      used[j] = br.u1();
      use_delta[j] = 1;
      if (!used[j]) use_delta[j] = br.u1();
    }
    int i = 0;
    for (int j = ref.num_positive - 1; j >= 0; j--) {
      const int d = ref.delta_poc_s1[j] + delta_rps;
      if (d < 0 && use_delta[ref.num_negative + j]) { if (i >= 16) return E_RANGE; out.delta_poc_s0[i] = d; out.used_s0[i++] = used[ref.num_negative + j]; }
    }
    if (delta_rps < 0 && use_delta[nref]) { if (i >= 16) return E_RANGE; out.delta_poc_s0[i] = delta_rps; out.used_s0[i++] = used[nref]; }
    for (int j = 0; j < ref.num_negative; j++) {
      const int d = ref.delta_poc_s0[j] + delta_rps;
      if (d < 0 && use_delta[j]) { if (i >= 16) return E_RANGE; out.delta_poc_s0[i] = d; out.used_s0[i++] = used[j]; }
    }
    out.num_negative = i;
    i = 0;
    for (int j = ref.num_negative - 1; j >= 0; j--) {
      const int d = ref.delta_poc_s0[j] + delta_rps;
      if (d > 0 && use_delta[j]) { if (i >= 16) return E_RANGE; out.delta_poc_s1[i] = d; out.used_s1[i++] = used[j]; }
    }
    if (delta_rps > 0 && use_delta[nref]) { if (i >= 16) return E_RANGE; out.delta_poc_s1[i] = delta_rps; out.used_s1[i++] = used[nref]; }
    for (int j = 0; j < ref.num_positive; j++) {
      const int d = ref.delta_poc_s1[j] + delta_rps;
      if (d > 0 && use_delta[ref.num_negative + j]) { if (i >= 16) return E_RANGE; out.delta_poc_s1[i] = d; out.used_s1[i++] = used[ref.num_negative + j]; }
    }
    out.num_positive = i;
  } else {
    const uint32_t nn = br.ue(), np = br.ue();
    if (br.err || nn > 16 || np > 16 || nn + np > 16) return E_RANGE;
    out.num_negative = nn; out.num_positive = np;
    int poc = 0;
    for (uint32_t i = 0; i < nn; i++) { poc -= (int)br.ue() + 1; out.delta_poc_s0[i] = (int16_t)poc; out.used_s0[i] = br.u1(); }
    poc = 0;
    for (uint32_t i = 0; i < np; i++) { poc += (int)br.ue() + 1; out.delta_poc_s1[i] = (int16_t)poc; out.used_s1[i] = br.u1(); }
  }
  (void)num_sets;
  return br.err ? E_PARSE : E_OK;
}

// ---- SPS -----------------------------------------------------------------------------------------
int parse_sps(BitReader& br, std::shared_ptr<Sps>& out) {
  std::shared_ptr<Sps> s = std::make_shared<Sps>();
  s->vps_id = br.u(4);
  const int msl1 = br.u(3);
  s->max_sub_layers = msl1 + 1;
  br.u1();
  skip_profile_tier_level(br, msl1);
  s->id = br.ue();
  if (br.err || s->id > 15) return E_RANGE;
  s->chroma_format_idc = br.ue();
  if (s->chroma_format_idc > 3) return E_RANGE;
  if (s->chroma_format_idc == 3) s->separate_colour_plane = br.u1();
  s->chroma_array_type = s->separate_colour_plane ? 0 : s->chroma_format_idc;
  s->width = br.ue();
  s->height = br.ue();
  if (br.u1()) for (int i = 0; i < 4; i++) s->conf_win[i] = br.ue();
  s->bit_depth_luma = br.ue() + 8;
  s->bit_depth_chroma = br.ue() + 8;
  s->log2_max_poc_lsb = br.ue() + 4;
  if (br.err || s->bit_depth_luma > 16 || s->bit_depth_chroma > 16 || s->log2_max_poc_lsb > 16) return E_RANGE;
  const int ordering_present = br.u1();
  for (int i = ordering_present ? 0 : msl1; i <= msl1; i++) {
    s->max_dec_pic_buffering[i] = br.ue() + 1;
    s->max_num_reorder[i] = br.ue();
    s->max_latency_increase[i] = br.ue();
  }
  if (!ordering_present)
    for (int i = 0; i < msl1; i++) {
      s->max_dec_pic_buffering[i] = s->max_dec_pic_buffering[msl1];
      s->max_num_reorder[i] = s->max_num_reorder[msl1];
      s->max_latency_increase[i] = s->max_latency_increase[msl1];
    }
  s->log2_min_cb = br.ue() + 3;
  s->log2_ctb = s->log2_min_cb + br.ue();
  s->log2_min_tb = br.ue() + 2;
  s->log2_max_tb = s->log2_min_tb + br.ue();
  s->max_th_depth_inter = br.ue();
  s->max_th_depth_intra = br.ue();
  if (br.err || s->log2_ctb > 6 || s->log2_ctb < 4 || s->log2_max_tb > 5 || s->log2_min_tb >= s->log2_min_cb ||
      s->log2_max_tb > s->log2_ctb || s->width <= 0 || s->height <= 0 || s->width > 16384 || s->height > 16384)
    return E_RANGE;
  if ((s->width & ((1 << s->log2_min_cb) - 1)) || (s->height & ((1 << s->log2_min_cb) - 1))) return E_RANGE;
  {  // the conformance window must leave at least one sample (7.4.3.2.1), in chroma units of the format
    const int sub_w = (s->chroma_format_idc == 1 || s->chroma_format_idc == 2) ? 2 : 1, sub_h = s->chroma_format_idc == 1 ? 2 : 1;
    for (int i = 0; i < 4; i++) if (s->conf_win[i] < 0 || s->conf_win[i] > 16384) return E_RANGE;
    if (sub_w * (s->conf_win[0] + s->conf_win[1]) >= s->width || sub_h * (s->conf_win[2] + s->conf_win[3]) >= s->height) return E_RANGE;
  }
  for (int i = 0; i <= msl1; i++)   // A.4.2: the DPB holds at most 16 pictures
    if (s->max_dec_pic_buffering[i] < 1 || s->max_dec_pic_buffering[i] > 16 || s->max_num_reorder[i] < 0 ||
        s->max_num_reorder[i] > s->max_dec_pic_buffering[i] - 1) return E_RANGE;
  s->scaling_list_enabled = br.u1();
  if (s->scaling_list_enabled) {
    default_scaling(s->scaling);
    s->scaling_list_present = br.u1();
    if (s->scaling_list_present) {
      int r = parse_scaling_list(br, s->scaling);
      if (r) return r;
    }
  }
  s->amp_enabled = br.u1();
  s->sao_enabled = br.u1();
  s->pcm_enabled = br.u1();
  if (s->pcm_enabled) {
    s->pcm_bits_luma = br.u(4) + 1;
    s->pcm_bits_chroma = br.u(4) + 1;
    s->log2_min_pcm = br.ue() + 3;
    s->log2_max_pcm = s->log2_min_pcm + br.ue();
    s->pcm_lf_disabled = br.u1();
    if (br.err || s->pcm_bits_luma > s->bit_depth_luma || s->pcm_bits_chroma > s->bit_depth_chroma ||
        s->log2_min_pcm > 5 || s->log2_max_pcm > 5 || s->log2_max_pcm > s->log2_ctb || s->log2_min_pcm < (s->log2_min_cb < 5 ? s->log2_min_cb : 5))
      return E_RANGE;
  }
  const uint32_t n_rps = br.ue();
  if (br.err || n_rps > 64) return E_RANGE;
  s->st_rps.resize(n_rps);
  for (uint32_t i = 0; i < n_rps; i++) {
    int r = parse_st_rps(br, s->st_rps[i], (int)i, (int)n_rps, s->st_rps, false);
    if (r) return r;
  }
  s->long_term_ref_pics_present = br.u1();
  if (s->long_term_ref_pics_present) {
    s->num_lt_ref_pics_sps = br.ue();
    if (s->num_lt_ref_pics_sps > 32) return E_RANGE;
    for (int i = 0; i < s->num_lt_ref_pics_sps; i++) {
      s->lt_ref_pic_poc_lsb[i] = br.u(s->log2_max_poc_lsb);
      s->lt_used_by_curr[i] = br.u1();
    }
  }
  s->temporal_mvp_enabled = br.u1();
  s->strong_intra_smoothing = br.u1();
  if (br.u1()) skip_vui(br, msl1);
  if (br.u1()) {  // sps_extension_present_flag
    const int range_ext = br.u1();
    br.skip(7);
    if (range_ext) {
      s->transform_skip_rotation = br.u1();
      s->transform_skip_context = br.u1();
      s->implicit_rdpcm = br.u1();
      s->explicit_rdpcm = br.u1();
      s->extended_precision = br.u1();
      s->intra_smoothing_disabled = br.u1();
      s->high_precision_offsets = br.u1();
      s->persistent_rice = br.u1();
      s->cabac_bypass_alignment = br.u1();
    }
  }
  if (br.err) return E_PARSE;
  s->ctbs_w = (s->width + (1 << s->log2_ctb) - 1) >> s->log2_ctb;
  s->ctbs_h = (s->height + (1 << s->log2_ctb) - 1) >> s->log2_ctb;
  s->min_cb_w = s->width >> s->log2_min_cb;
  s->min_cb_h = s->height >> s->log2_min_cb;
  s->qp_bd_offset_y = 6 * (s->bit_depth_luma - 8);
  s->qp_bd_offset_c = 6 * (s->bit_depth_chroma - 8);
  out = s;
  return E_OK;
}

// ---- PPS -----------------------------------------------------------------------------------------
int parse_pps(BitReader& br, const ParamSets& ps, std::shared_ptr<Pps>& out) {
  std::shared_ptr<Pps> p = std::make_shared<Pps>();
  p->id = br.ue();
  p->sps_id = br.ue();
  if (br.err || p->id > 63 || p->sps_id > 15) return E_RANGE;
  std::shared_ptr<Sps> sps = ps.sps[p->sps_id];
  if (!sps) return E_NOSPS;
  p->sps = sps;
  p->dependent_slices_enabled = br.u1();
  p->output_flag_present = br.u1();
  p->num_extra_slice_header_bits = br.u(3);
  p->sign_data_hiding = br.u1();
  p->cabac_init_present = br.u1();
  p->num_ref_idx_default[0] = br.ue() + 1;
  p->num_ref_idx_default[1] = br.ue() + 1;
  p->init_qp = 26 + br.se();
  p->constrained_intra_pred = br.u1();
  p->transform_skip_enabled = br.u1();
  p->cu_qp_delta_enabled = br.u1();
  if (p->cu_qp_delta_enabled) p->diff_cu_qp_delta_depth = br.ue();
  p->cb_qp_offset = br.se();
  p->cr_qp_offset = br.se();
  p->slice_chroma_qp_offsets_present = br.u1();
  p->weighted_pred = br.u1();
  p->weighted_bipred = br.u1();
  p->transquant_bypass_enabled = br.u1();
  p->tiles_enabled = br.u1();
  p->entropy_coding_sync = br.u1();
  if (br.err || p->num_ref_idx_default[0] > 16 || p->num_ref_idx_default[1] > 16 ||
      p->diff_cu_qp_delta_depth > sps->log2_ctb - sps->log2_min_cb || p->diff_cu_qp_delta_depth < 0 ||
      p->init_qp < -(6 * (sps->bit_depth_luma - 8)) || p->init_qp > 51 ||
      p->cb_qp_offset < -12 || p->cb_qp_offset > 12 || p->cr_qp_offset < -12 || p->cr_qp_offset > 12)
    return E_RANGE;
  std::vector<int> col_w, row_h;
  if (p->tiles_enabled) {
    p->num_tile_cols = br.ue() + 1;
    p->num_tile_rows = br.ue() + 1;
    if (br.err || p->num_tile_cols > sps->ctbs_w || p->num_tile_rows > sps->ctbs_h || p->num_tile_cols > 20 || p->num_tile_rows > 22)
      return E_RANGE;
    p->uniform_spacing = br.u1();
    if (!p->uniform_spacing) {
      int rem = sps->ctbs_w;
      for (int i = 0; i < p->num_tile_cols - 1; i++) { col_w.push_back(br.ue() + 1); rem -= col_w.back(); }
      if (rem <= 0) return E_RANGE;
      col_w.push_back(rem);
      rem = sps->ctbs_h;
      for (int i = 0; i < p->num_tile_rows - 1; i++) { row_h.push_back(br.ue() + 1); rem -= row_h.back(); }
      if (rem <= 0) return E_RANGE;
      row_h.push_back(rem);
    }
    p->lf_across_tiles = br.u1();
  }
  if (col_w.empty())
    for (int i = 0; i < p->num_tile_cols; i++)
      col_w.push_back(((i + 1) * sps->ctbs_w) / p->num_tile_cols - (i * sps->ctbs_w) / p->num_tile_cols);
  if (row_h.empty())
    for (int i = 0; i < p->num_tile_rows; i++)
      row_h.push_back(((i + 1) * sps->ctbs_h) / p->num_tile_rows - (i * sps->ctbs_h) / p->num_tile_rows);
  p->lf_across_slices = br.u1();
  p->deblocking_control_present = br.u1();
  if (p->deblocking_control_present) {
    p->deblocking_override_enabled = br.u1();
    p->deblocking_disabled = br.u1();
    if (!p->deblocking_disabled) {
      p->beta_offset = 2 * br.se();
      p->tc_offset = 2 * br.se();
    }
  }
  p->scaling_list_present = br.u1();
  if (p->scaling_list_present) {
    default_scaling(p->scaling);
    int r = parse_scaling_list(br, p->scaling);
    if (r) return r;
  }
  p->lists_modification_present = br.u1();
  p->log2_parallel_merge_level = br.ue() + 2;
  if (br.err || p->log2_parallel_merge_level > sps->log2_ctb) return E_RANGE;
  p->slice_header_extension_present = br.u1();
  if (br.u1()) {  // pps_extension_present_flag
    const int range_ext = br.u1();
    br.skip(7);
    if (range_ext) {
      if (p->transform_skip_enabled) p->log2_max_transform_skip = br.ue() + 2;
      p->cross_component_prediction = br.u1();
      p->chroma_qp_offset_list_enabled = br.u1();
      if (p->chroma_qp_offset_list_enabled) {
        p->diff_cu_chroma_qp_offset_depth = br.ue();
        if (br.err || p->diff_cu_chroma_qp_offset_depth < 0 || p->diff_cu_chroma_qp_offset_depth > sps->log2_ctb - sps->log2_min_cb) return E_RANGE;
        p->chroma_qp_offset_list_len = br.ue() + 1;
        if (p->chroma_qp_offset_list_len > 6) return E_RANGE;
        for (int i = 0; i < p->chroma_qp_offset_list_len; i++) {
          p->cb_qp_offset_list[i] = br.se();
          p->cr_qp_offset_list[i] = br.se();
        }
      }
      p->log2_sao_offset_scale_luma = br.ue();
      p->log2_sao_offset_scale_chroma = br.ue();
    }
  }
  if (br.err) return E_PARSE;
  // derived: tile geometry and CTB address conversion (6.5.1)
  p->col_bd.assign(1, 0);
  for (int w : col_w) p->col_bd.push_back(p->col_bd.back() + w);
  p->row_bd.assign(1, 0);
  for (int h : row_h) p->row_bd.push_back(p->row_bd.back() + h);
  const int n = sps->ctbs_w * sps->ctbs_h;
  p->rs_to_ts.assign(n, 0);
  p->ts_to_rs.assign(n, 0);
  p->tile_id_rs.assign(n, 0);
  for (int rs = 0; rs < n; rs++) {
    const int tbx = rs % sps->ctbs_w, tby = rs / sps->ctbs_w;
    int tx = 0, ty = 0;
    for (int i = 0; i < p->num_tile_cols; i++) if (tbx >= p->col_bd[i]) tx = i;
    for (int j = 0; j < p->num_tile_rows; j++) if (tby >= p->row_bd[j]) ty = j;
    int v = 0;
    for (int i = 0; i < tx; i++) v += row_h[ty] * col_w[i];
    for (int j = 0; j < ty; j++) v += sps->ctbs_w * row_h[j];
    v += (tby - p->row_bd[ty]) * col_w[tx] + tbx - p->col_bd[tx];
    p->rs_to_ts[rs] = v;
    p->ts_to_rs[v] = rs;
    p->tile_id_rs[rs] = ty * p->num_tile_cols + tx;
  }
  p->log2_min_cu_qp_delta = sps->log2_ctb - p->diff_cu_qp_delta_depth;
  p->log2_min_cu_chroma_qp_offset = sps->log2_ctb - p->diff_cu_chroma_qp_offset_depth;
  out = p;
  return E_OK;
}

// ---- slice segment header ---------------------------------------------------------------------------
static int parse_pred_weight_table(BitReader& br, const Sps& sps, SliceHeader& sh) {
  PredWeight& pw = sh.pw;
  pw.luma_log2_denom = br.ue();
  if (pw.luma_log2_denom > 7) return E_RANGE;
  pw.chroma_log2_denom = pw.luma_log2_denom;
  if (sps.chroma_array_type != 0) {
    pw.chroma_log2_denom += br.se();
    if (pw.chroma_log2_denom < 0 || pw.chroma_log2_denom > 7) return E_RANGE;
  }
  const int half = sps.high_precision_offsets ? 1 << (sps.bit_depth_chroma - 1) : 128;
  for (int l = 0; l < (sh.slice_type == SLICE_B ? 2 : 1); l++) {
    uint8_t lf[16], cf[16];
    for (int i = 0; i < sh.num_ref_idx[l]; i++) lf[i] = br.u1();
    for (int i = 0; i < sh.num_ref_idx[l]; i++) cf[i] = sps.chroma_array_type != 0 ? br.u1() : 0;
    for (int i = 0; i < sh.num_ref_idx[l]; i++) {
      pw.weight[l][i][0] = (int16_t)(1 << pw.luma_log2_denom);
      pw.offset[l][i][0] = 0;
      pw.weight[l][i][1] = pw.weight[l][i][2] = (int16_t)(1 << pw.chroma_log2_denom);
      pw.offset[l][i][1] = pw.offset[l][i][2] = 0;
      if (lf[i]) {
        const int dw = br.se();
        if (dw < -128 || dw > 127) return E_RANGE;
        pw.weight[l][i][0] += dw;
        pw.offset[l][i][0] = (int16_t)br.se();
      }
      if (cf[i])
        for (int j = 1; j < 3; j++) {
          const int dw = br.se(), dof = br.se();
          if (dw < -128 || dw > 127) return E_RANGE;
          pw.weight[l][i][j] += dw;
          int o = half + dof - ((half * pw.weight[l][i][j]) >> pw.chroma_log2_denom);
          pw.offset[l][i][j] = (int16_t)std::min(half - 1, std::max(-half, o));
        }
    }
  }
  return br.err ? E_PARSE : E_OK;
}

int parse_slice_header(BitReader& br, int nal_type, const ParamSets& ps, const SliceHeader* prev_independent,
                       const std::vector<uint32_t>& removed, SliceHeader& sh, std::shared_ptr<const Pps>& pps_out) {
  sh = SliceHeader();
  sh.nal_type = nal_type;
  sh.first_slice_segment_in_pic = br.u1();
  if (nal_type >= NAL_BLA_W_LP && nal_type <= NAL_RSV_IRAP_VCL23) sh.no_output_of_prior_pics = br.u1();
  sh.pps_id = br.ue();
  if (br.err || sh.pps_id > 63) return E_SLICEHDR;
  std::shared_ptr<Pps> pps = ps.pps[sh.pps_id];
  if (!pps) return E_NOPPS;
  const Sps& sps = *pps->sps;
  pps_out = pps;
  const int n_ctbs = sps.ctbs_w * sps.ctbs_h;
  if (!sh.first_slice_segment_in_pic) {
    if (pps->dependent_slices_enabled) sh.dependent = br.u1();
    sh.segment_address = br.u(ceil_log2(n_ctbs));
    if (sh.segment_address >= n_ctbs || sh.segment_address == 0) return E_SLICEHDR;
  }
  if (sh.dependent) {
    if (!prev_independent) return E_SLICEHDR;
    // inherit everything of the slice header proper (7.4.7.1), keep the segment's own fields
    SliceHeader own = sh;
    sh = *prev_independent;
    sh.nal_type = own.nal_type;
    sh.first_slice_segment_in_pic = own.first_slice_segment_in_pic;
    sh.no_output_of_prior_pics = own.no_output_of_prior_pics;
    sh.pps_id = own.pps_id;
    sh.dependent = 1;
    sh.segment_address = own.segment_address;
    sh.entry_points.clear();
  } else {
    br.skip(pps->num_extra_slice_header_bits);
    sh.slice_type = br.ue();
    if (sh.slice_type > 2) return E_SLICEHDR;
    if (pps->output_flag_present) sh.pic_output = br.u1();
    if (sps.separate_colour_plane) sh.colour_plane_id = br.u(2);
    if (nal_type != NAL_IDR_W_RADL && nal_type != NAL_IDR_N_LP) {
      sh.poc_lsb = br.u(sps.log2_max_poc_lsb);
      sh.short_term_ref_pic_set_sps = br.u1();
      if (!sh.short_term_ref_pic_set_sps) {
        const size_t p0 = br.pos;
        int r = parse_st_rps(br, sh.st_rps, (int)sps.st_rps.size(), (int)sps.st_rps.size(), sps.st_rps, true);
        if (r) return r;
        sh.st_rps_bits = (int)(br.pos - p0);
      } else {
        const int nb = ceil_log2((uint32_t)sps.st_rps.size());
        sh.st_rps_idx = nb ? br.u(nb) : 0;
        if (sh.st_rps_idx >= (int)sps.st_rps.size()) return E_SLICEHDR;
        sh.st_rps = sps.st_rps[sh.st_rps_idx];
      }
      if (sps.long_term_ref_pics_present) {
        if (sps.num_lt_ref_pics_sps > 0) sh.num_long_term_sps = br.ue();
        sh.num_long_term_pics = br.ue();
        if (br.err || sh.num_long_term_sps + sh.num_long_term_pics > 32) return E_SLICEHDR;
        for (int i = 0; i < sh.num_long_term_sps + sh.num_long_term_pics; i++) {
          if (i < sh.num_long_term_sps) {
            const int nb = ceil_log2((uint32_t)sps.num_lt_ref_pics_sps);
            sh.lt_idx_sps[i] = nb ? br.u(nb) : 0;
            if (sh.lt_idx_sps[i] >= sps.num_lt_ref_pics_sps) return E_SLICEHDR;
            sh.poc_lsb_lt[i] = sps.lt_ref_pic_poc_lsb[sh.lt_idx_sps[i]];
            sh.used_by_curr_lt[i] = sps.lt_used_by_curr[sh.lt_idx_sps[i]];
          } else {
            sh.poc_lsb_lt[i] = br.u(sps.log2_max_poc_lsb);
            sh.used_by_curr_lt[i] = br.u1();
          }
          sh.delta_poc_msb_present[i] = br.u1();
          sh.delta_poc_msb_cycle[i] = sh.delta_poc_msb_present[i] ? br.ue() : 0;
          if (i != 0 && i != sh.num_long_term_sps) sh.delta_poc_msb_cycle[i] += sh.delta_poc_msb_cycle[i - 1];
        }
      }
      if (sps.temporal_mvp_enabled) sh.temporal_mvp_enabled = br.u1();
    }
    if (sps.sao_enabled) {
      sh.sao_luma = br.u1();
      if (sps.chroma_array_type != 0) sh.sao_chroma = br.u1();
    }
    if (sh.slice_type != SLICE_I) {
      sh.num_ref_idx[0] = pps->num_ref_idx_default[0];
      sh.num_ref_idx[1] = sh.slice_type == SLICE_B ? pps->num_ref_idx_default[1] : 0;
      if (br.u1()) {
        sh.num_ref_idx[0] = br.ue() + 1;
        if (sh.slice_type == SLICE_B) sh.num_ref_idx[1] = br.ue() + 1;
      }
      if (sh.num_ref_idx[0] > 16 || sh.num_ref_idx[1] > 16) return E_SLICEHDR;
      int num_pic_total = 0;
      for (int i = 0; i < sh.st_rps.num_negative; i++) num_pic_total += sh.st_rps.used_s0[i];
      for (int i = 0; i < sh.st_rps.num_positive; i++) num_pic_total += sh.st_rps.used_s1[i];
      for (int i = 0; i < sh.num_long_term_sps + sh.num_long_term_pics; i++) num_pic_total += sh.used_by_curr_lt[i];
      if (pps->lists_modification_present && num_pic_total > 1) {
        const int nb = ceil_log2((uint32_t)num_pic_total);
        for (int l = 0; l < (sh.slice_type == SLICE_B ? 2 : 1); l++) {
          sh.ref_pic_list_modification[l] = br.u1();
          if (sh.ref_pic_list_modification[l])
            for (int i = 0; i < sh.num_ref_idx[l]; i++) sh.list_entry[l][i] = br.u(nb);
        }
      }
      if (sh.slice_type == SLICE_B) sh.mvd_l1_zero = br.u1();
      if (pps->cabac_init_present) sh.cabac_init_flag = br.u1();
      if (sh.temporal_mvp_enabled) {
        if (sh.slice_type == SLICE_B) sh.collocated_from_l0 = br.u1();
        if ((sh.collocated_from_l0 && sh.num_ref_idx[0] > 1) || (!sh.collocated_from_l0 && sh.num_ref_idx[1] > 1))
          sh.collocated_ref_idx = br.ue();
      }
      if ((pps->weighted_pred && sh.slice_type == SLICE_P) || (pps->weighted_bipred && sh.slice_type == SLICE_B)) {
        int r = parse_pred_weight_table(br, sps, sh);
        if (r) return r;
      }
      sh.max_num_merge_cand = 5 - (int)br.ue();
      if (sh.max_num_merge_cand < 1 || sh.max_num_merge_cand > 5) return E_SLICEHDR;
    }
    sh.qp_delta = br.se();
    if (pps->slice_chroma_qp_offsets_present) {
      sh.cb_qp_offset = br.se();
      sh.cr_qp_offset = br.se();
    }
    if (pps->chroma_qp_offset_list_enabled) sh.cu_chroma_qp_offset_enabled = br.u1();
    if (pps->deblocking_override_enabled) sh.deblocking_override = br.u1();
    sh.deblocking_disabled = pps->deblocking_disabled;
    sh.beta_offset = pps->beta_offset;
    sh.tc_offset = pps->tc_offset;
    if (sh.deblocking_override) {
      sh.deblocking_disabled = br.u1();
      if (!sh.deblocking_disabled) {
        sh.beta_offset = 2 * br.se();
        sh.tc_offset = 2 * br.se();
      }
    }
    sh.lf_across_slices = pps->lf_across_slices;
    if (pps->lf_across_slices && (sh.sao_luma || sh.sao_chroma || !sh.deblocking_disabled)) sh.lf_across_slices = br.u1();
    sh.slice_qp = pps->init_qp + sh.qp_delta;
    if (sh.slice_qp < -(6 * (pps->sps->bit_depth_luma - 8)) || sh.slice_qp > 51) return E_SLICEHDR;   // SliceQpY range, 7.4.7.1
    sh.slice_addr_rs = sh.segment_address;
  }
  if (pps->tiles_enabled || pps->entropy_coding_sync) {
    const uint32_t n = br.ue();
    if (br.err || n > (uint32_t)n_ctbs) return E_SLICEHDR;
    if (n > 0) {
      const int len = br.ue() + 1;
      if (len > 32) return E_SLICEHDR;
      uint32_t acc = 0;
      for (uint32_t i = 0; i < n; i++) {
        acc += br.u(len) + 1;
        sh.entry_points.push_back(acc);
      }
    }
  }
  if (pps->slice_header_extension_present) {
    const uint32_t n = br.ue();
    if (br.err || n > 256) return E_SLICEHDR;
    br.skip(8 * n);
  }
  // byte_alignment(): a one bit followed by zero bits
  br.u1();
  br.align();
  if (br.err) return E_SLICEHDR;
  sh.data_offset = br.pos >> 3;
  // entry points are given in bytes of the escaped NAL; translate to de-escaped offsets (slice.cc:815-850)
  if (!sh.entry_points.empty() && !removed.empty()) {
    // `removed` holds output indices (relative to the NAL payload start, header included) of dropped bytes
    for (uint32_t& ep : sh.entry_points) {
      // position in escaped stream relative to slice data start = ep; count removed bytes before it
      uint32_t esc_abs = 0, k = 0;
      // escaped absolute position of data start = data_offset + number of removed bytes before data_offset
      uint32_t removed_before_data = 0;
      for (uint32_t r : removed) if (r <= sh.data_offset) removed_before_data++;
      esc_abs = (uint32_t)sh.data_offset + removed_before_data + ep;
      // an emulation byte dropped at output index r sat at escaped index r + (its rank)
      for (uint32_t i = 0; i < removed.size(); i++) if (removed[i] + i < esc_abs) k = i + 1;
      ep = esc_abs - k - (uint32_t)sh.data_offset;
    }
  }
  return E_OK;
}

}  // namespace b265h
