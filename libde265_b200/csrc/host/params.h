// params.h — HEVC parameter sets and slice segment header as the host parser needs them.
// Syntax per H.265 7.3.2.2 (SPS), 7.3.2.3 (PPS), 7.3.6 (slice segment header); mirrors what the
// reference keeps in sps.h / pps.h / slice.h (seq_parameter_set, pic_parameter_set,
// slice_segment_header), restricted to fields that influence parsing or the pixel path.
#pragma once
#include <stdint.h>
#include <memory>
#include <vector>
#include "bits.h"

namespace b265h {

enum NalType {
  NAL_TRAIL_N = 0, NAL_TRAIL_R = 1, NAL_RASL_N = 8, NAL_RASL_R = 9,
  NAL_BLA_W_LP = 16, NAL_BLA_W_RADL = 17, NAL_BLA_N_LP = 18, NAL_IDR_W_RADL = 19, NAL_IDR_N_LP = 20,
  NAL_CRA = 21, NAL_RSV_IRAP_VCL23 = 23,
  NAL_VPS = 32, NAL_SPS = 33, NAL_PPS = 34, NAL_AUD = 35, NAL_EOS = 36, NAL_EOB = 37, NAL_FD = 38,
  NAL_PREFIX_SEI = 39, NAL_SUFFIX_SEI = 40
};
enum SliceType { SLICE_B = 0, SLICE_P = 1, SLICE_I = 2 };

struct ShortTermRps {
  int num_negative = 0, num_positive = 0;
  int16_t delta_poc_s0[16], delta_poc_s1[16];
  uint8_t used_s0[16], used_s1[16];
  int num_delta_pocs() const { return num_negative + num_positive; }
};

struct ScalingList {
  // ScalingFactor per size id / matrix id in raster order [y*n+x] (7.4.5), as b265_scaling expects
  uint8_t s4[6][16], s8[6][64], s16[6][256], s32[6][1024];
};

struct Sps {
  int id = 0, vps_id = 0;
  int max_sub_layers = 1;
  int chroma_format_idc = 1, separate_colour_plane = 0, chroma_array_type = 1;
  int width = 0, height = 0;
  int conf_win[4] = {0, 0, 0, 0};  // left right top bottom, in chroma units
  int bit_depth_luma = 8, bit_depth_chroma = 8;
  int log2_max_poc_lsb = 4;
  int max_dec_pic_buffering[8], max_num_reorder[8], max_latency_increase[8];
  int log2_min_cb = 3, log2_ctb = 4, log2_min_tb = 2, log2_max_tb = 5;
  int max_th_depth_inter = 0, max_th_depth_intra = 0;
  int scaling_list_enabled = 0, scaling_list_present = 0;
  ScalingList scaling;
  int amp_enabled = 0, sao_enabled = 0;
  int pcm_enabled = 0, pcm_bits_luma = 8, pcm_bits_chroma = 8, log2_min_pcm = 3, log2_max_pcm = 5, pcm_lf_disabled = 0;
  std::vector<ShortTermRps> st_rps;
  int long_term_ref_pics_present = 0, num_lt_ref_pics_sps = 0;
  int lt_ref_pic_poc_lsb[32], lt_used_by_curr[32];
  int temporal_mvp_enabled = 0, strong_intra_smoothing = 0;
  // range extension
  int transform_skip_rotation = 0, transform_skip_context = 0, implicit_rdpcm = 0, explicit_rdpcm = 0;
  int extended_precision = 0, intra_smoothing_disabled = 0, high_precision_offsets = 0;
  int persistent_rice = 0, cabac_bypass_alignment = 0;
  // derived
  int ctbs_w = 0, ctbs_h = 0, min_cb_w = 0, min_cb_h = 0;
  int qp_bd_offset_y = 0, qp_bd_offset_c = 0;
};

struct Pps {
  int id = 0, sps_id = 0;
  int dependent_slices_enabled = 0, output_flag_present = 0, num_extra_slice_header_bits = 0;
  int sign_data_hiding = 0, cabac_init_present = 0;
  int num_ref_idx_default[2] = {1, 1};
  int init_qp = 26;
  int constrained_intra_pred = 0, transform_skip_enabled = 0;
  int cu_qp_delta_enabled = 0, diff_cu_qp_delta_depth = 0;
  int cb_qp_offset = 0, cr_qp_offset = 0, slice_chroma_qp_offsets_present = 0;
  int weighted_pred = 0, weighted_bipred = 0;
  int transquant_bypass_enabled = 0;
  int tiles_enabled = 0, entropy_coding_sync = 0;
  int num_tile_cols = 1, num_tile_rows = 1, uniform_spacing = 1;
  int lf_across_tiles = 1, lf_across_slices = 0;
  int deblocking_control_present = 0, deblocking_override_enabled = 0, deblocking_disabled = 0;
  int beta_offset = 0, tc_offset = 0;  // already *2
  int scaling_list_present = 0;
  ScalingList scaling;
  int lists_modification_present = 0, log2_parallel_merge_level = 2;
  int slice_header_extension_present = 0;
  // range extension
  int log2_max_transform_skip = 2, cross_component_prediction = 0;
  int chroma_qp_offset_list_enabled = 0, diff_cu_chroma_qp_offset_depth = 0, chroma_qp_offset_list_len = 0;
  int cb_qp_offset_list[6], cr_qp_offset_list[6];
  int log2_sao_offset_scale_luma = 0, log2_sao_offset_scale_chroma = 0;
  // derived (6.5.1) — valid for the SPS they were derived against
  std::shared_ptr<const Sps> sps;
  std::vector<int> col_bd, row_bd;             // tile boundaries in CTBs (size cols+1 / rows+1)
  std::vector<int> rs_to_ts, ts_to_rs, tile_id_rs;
  int log2_min_cu_qp_delta = 0, log2_min_cu_chroma_qp_offset = 0;
};

struct PredWeight {
  int luma_log2_denom = 0, chroma_log2_denom = 0;
  int16_t weight[2][16][3], offset[2][16][3];
};

struct SliceHeader {
  int nal_type = 0;
  int first_slice_segment_in_pic = 0, no_output_of_prior_pics = 0;
  int pps_id = 0;
  int dependent = 0, segment_address = 0;
  int slice_type = SLICE_I, pic_output = 1, colour_plane_id = 0;
  int poc_lsb = 0;
  int short_term_ref_pic_set_sps = 0, st_rps_idx = 0, st_rps_bits = 0;
  ShortTermRps st_rps;
  int num_long_term_sps = 0, num_long_term_pics = 0;
  int lt_idx_sps[32], poc_lsb_lt[32], used_by_curr_lt[32], delta_poc_msb_present[32], delta_poc_msb_cycle[32];
  int temporal_mvp_enabled = 0;
  int sao_luma = 0, sao_chroma = 0;
  int num_ref_idx[2] = {0, 0};
  int ref_pic_list_modification[2] = {0, 0};
  int list_entry[2][16];
  int mvd_l1_zero = 0, cabac_init_flag = 0, collocated_from_l0 = 1, collocated_ref_idx = 0;
  PredWeight pw;
  int max_num_merge_cand = 5;
  int qp_delta = 0, cb_qp_offset = 0, cr_qp_offset = 0, cu_chroma_qp_offset_enabled = 0;
  int deblocking_override = 0, deblocking_disabled = 0, beta_offset = 0, tc_offset = 0;  // offsets already *2
  int lf_across_slices = 0;
  std::vector<uint32_t> entry_points;  // byte offsets (in de-escaped slice data) of sub-streams 1..n
  // derived
  int slice_qp = 26;
  int slice_addr_rs = 0;      // address of the first CTB of the (independent) slice this segment belongs to
  size_t data_offset = 0;     // byte offset of slice_segment_data() inside the de-escaped NAL payload
  int slice_index = 0;        // index into Picture::slices (one per independent slice)
};

struct ParamSets {
  std::shared_ptr<Sps> sps[16];
  std::shared_ptr<Pps> pps[64];
};

// all return 0 on success, a de265_error-compatible code otherwise
int parse_sps(BitReader& br, std::shared_ptr<Sps>& out);
int parse_pps(BitReader& br, const ParamSets& ps, std::shared_ptr<Pps>& out);
int parse_slice_header(BitReader& br, int nal_type, const ParamSets& ps, const SliceHeader* prev_independent,
                       const std::vector<uint32_t>& removed, SliceHeader& sh, std::shared_ptr<const Pps>& pps_out);

}  // namespace b265h
