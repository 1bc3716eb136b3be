// cabac.cc — CABAC state-transition / range tables (H.265 tables 9-46, 9-47) and the initValue tables
// of the context variables (tables 9-5 .. 9-37), laid out per initType for cabac_init_contexts().
#include "cabac.h"

namespace b265h {

// rangeTabLps[pStateIdx][qRangeIdx]
const uint8_t g_cabac_lps[64][4] = {
    {128, 176, 208, 240}, {128, 167, 197, 227}, {128, 158, 187, 216}, {123, 150, 178, 205}, {116, 142, 169, 195},
    {111, 135, 160, 185}, {105, 128, 152, 175}, {100, 122, 144, 166}, {95, 116, 137, 158},  {90, 110, 130, 150},
    {85, 104, 123, 142},  {81, 99, 117, 135},   {77, 94, 111, 128},   {73, 89, 105, 122},   {69, 85, 100, 116},
    {66, 80, 95, 110},    {62, 76, 90, 104},    {59, 72, 86, 99},     {56, 69, 81, 94},     {53, 65, 77, 89},
    {51, 62, 73, 85},     {48, 59, 69, 80},     {46, 56, 66, 76},     {43, 53, 63, 72},     {41, 50, 59, 69},
    {39, 48, 56, 65},     {37, 45, 54, 62},     {35, 43, 51, 59},     {33, 41, 48, 56},     {32, 39, 46, 53},
    {30, 37, 43, 50},     {29, 35, 41, 48},     {27, 33, 39, 45},     {26, 31, 37, 43},     {24, 30, 35, 41},
    {23, 28, 33, 39},     {22, 27, 32, 37},     {21, 26, 30, 35},     {20, 24, 29, 33},     {19, 23, 27, 31},
    {18, 22, 26, 30},     {17, 21, 25, 28},     {16, 20, 23, 27},     {15, 19, 22, 25},     {14, 18, 21, 24},
    {14, 17, 20, 23},     {13, 16, 19, 22},     {12, 15, 18, 21},     {12, 14, 17, 20},     {11, 14, 16, 19},
    {11, 13, 15, 18},     {10, 12, 15, 17},     {10, 12, 14, 16},     {9, 11, 13, 15},      {9, 11, 12, 14},
    {8, 10, 12, 14},      {8, 9, 11, 13},       {7, 9, 11, 12},       {7, 9, 10, 12},       {7, 8, 10, 11},
    {6, 8, 9, 11},        {6, 7, 9, 10},        {6, 7, 8, 9},         {2, 2, 2, 2}};

const uint8_t g_cabac_next_mps[64] = {1,  2,  3,  4,  5,  6,  7,  8,  9,  10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 21, 22,
                                      23, 24, 25, 26, 27, 28, 29, 30, 31, 32, 33, 34, 35, 36, 37, 38, 39, 40, 41, 42, 43, 44,
                                      45, 46, 47, 48, 49, 50, 51, 52, 53, 54, 55, 56, 57, 58, 59, 60, 61, 62, 62, 63};

const uint8_t g_cabac_next_lps[64] = {0,  0,  1,  2,  2,  4,  4,  5,  6,  7,  8,  9,  9,  11, 11, 12, 13, 13, 15, 15, 16, 16,
                                      18, 18, 19, 19, 21, 21, 22, 22, 23, 24, 24, 25, 26, 26, 27, 27, 28, 29, 29, 30, 30, 30,
                                      31, 32, 32, 33, 33, 33, 34, 34, 35, 35, 35, 36, 36, 36, 37, 37, 37, 38, 38, 63};

// number of renormalisation shifts after an LPS, indexed by rangeLps >> 3
const uint8_t g_cabac_renorm[32] = {6, 5, 4, 4, 3, 3, 3, 3, 2, 2, 2, 2, 2, 2, 2, 2,
                                    1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1};

// next context byte for (ctx << 1) | decodedLps, built from the two transition tables above
uint8_t g_cabac_next[256];
namespace {
struct NextTabInit {
  NextTabInit() {
    for (int c = 0; c < 128; c++) {
      const int state = c >> 1, mps = c & 1;
      g_cabac_next[(c << 1) | 0] = (uint8_t)((g_cabac_next_mps[state] << 1) | mps);
      g_cabac_next[(c << 1) | 1] = (uint8_t)((g_cabac_next_lps[state] << 1) | (state == 0 ? mps ^ 1 : mps));
    }
  }
};
const NextTabInit g_next_init;
}  // namespace

namespace {

struct InitRow {
  int offset, count;
  const uint8_t* v[3];  // per initType; nullptr => 154 (equiprobable-ish default, never used)
};

const uint8_t k154[8] = {154, 154, 154, 154, 154, 154, 154, 154};
const uint8_t k139[2] = {139, 139};
const uint8_t sao_merge[1] = {153};
const uint8_t sao_type[3][1] = {{200}, {185}, {160}};
const uint8_t split_cu[3][3] = {{139, 141, 157}, {107, 139, 126}, {107, 139, 126}};
const uint8_t cu_skip[3] = {197, 185, 201};
const uint8_t pred_mode[3][1] = {{154}, {149}, {134}};
const uint8_t part_mode[3][4] = {{184, 154, 139, 154}, {154, 139, 154, 154}, {154, 139, 154, 154}};
const uint8_t prev_intra[3][1] = {{184}, {154}, {183}};
const uint8_t intra_chroma[3][1] = {{63}, {152}, {152}};
const uint8_t rqt_root[1] = {79};
const uint8_t merge_flag[3][1] = {{154}, {110}, {154}};
const uint8_t merge_idx[3][1] = {{154}, {122}, {137}};
const uint8_t inter_pred_idc[5] = {95, 79, 63, 31, 31};
const uint8_t ref_idx[2] = {153, 153};
const uint8_t mvp_flag[1] = {168};
const uint8_t split_tr[3][3] = {{153, 138, 138}, {124, 138, 94}, {224, 167, 122}};
const uint8_t cbf_luma[3][2] = {{111, 141}, {153, 111}, {153, 111}};
const uint8_t cbf_chroma[3][5] = {{94, 138, 182, 154, 154}, {149, 107, 167, 154, 154}, {149, 92, 167, 154, 154}};
const uint8_t mvd_gt0[3][1] = {{154}, {140}, {169}};
const uint8_t mvd_gt1[3][1] = {{154}, {198}, {198}};
const uint8_t last_prefix[3][18] = {
    {110, 110, 124, 125, 140, 153, 125, 127, 140, 109, 111, 143, 127, 111, 79, 108, 123, 63},
    {125, 110, 94, 110, 95, 79, 125, 111, 110, 78, 110, 111, 111, 95, 94, 108, 123, 108},
    {125, 110, 124, 110, 95, 94, 125, 111, 111, 79, 125, 126, 111, 111, 79, 108, 123, 93}};
const uint8_t csbf[3][4] = {{91, 171, 134, 141}, {121, 140, 61, 154}, {121, 140, 61, 154}};
const uint8_t sig[3][44] = {
    {111, 111, 125, 110, 110, 94,  124, 108, 124, 107, 125, 141, 179, 153, 125, 107, 125, 141, 179, 153, 125, 107,
     125, 141, 179, 153, 125, 140, 139, 182, 182, 152, 136, 152, 136, 153, 136, 139, 111, 136, 139, 111, 141, 111},
    {155, 154, 139, 153, 139, 123, 123, 63,  153, 166, 183, 140, 136, 153, 154, 166, 183, 140, 136, 153, 154, 166,
     183, 140, 136, 153, 154, 170, 153, 123, 123, 107, 121, 107, 121, 167, 151, 183, 140, 151, 183, 140, 140, 140},
    {170, 154, 139, 153, 139, 123, 123, 63,  124, 166, 183, 140, 136, 153, 154, 166, 183, 140, 136, 153, 154, 166,
     183, 140, 136, 153, 154, 170, 153, 138, 138, 122, 121, 122, 121, 167, 151, 183, 140, 151, 183, 140, 140, 140}};
const uint8_t gt1[3][24] = {
    {140, 92, 137, 138, 140, 152, 138, 139, 153, 74, 149, 92, 139, 107, 122, 152, 140, 179, 166, 182, 140, 227, 122, 197},
    {154, 196, 196, 167, 154, 152, 167, 182, 182, 134, 149, 136, 153, 121, 136, 137, 169, 194, 166, 167, 154, 167, 137, 182},
    {154, 196, 167, 167, 154, 152, 167, 182, 182, 134, 149, 136, 153, 121, 136, 122, 169, 208, 166, 167, 154, 152, 167, 182}};
const uint8_t gt2[3][6] = {{138, 153, 136, 167, 152, 152}, {107, 167, 91, 122, 107, 167}, {107, 167, 91, 107, 107, 167}};

#define ROW1(off, n, a) {off, n, {a, a, a}}
#define ROW3(off, n, a) {off, n, {a[0], a[1], a[2]}}
const InitRow kRows[] = {
    ROW1(CTX_SAO_MERGE, 1, sao_merge),
    ROW3(CTX_SAO_TYPE, 1, sao_type),
    ROW3(CTX_SPLIT_CU, 3, split_cu),
    ROW1(CTX_CU_TQ_BYPASS, 1, k154),
    ROW1(CTX_CU_SKIP, 3, cu_skip),
    ROW3(CTX_PRED_MODE, 1, pred_mode),
    ROW3(CTX_PART_MODE, 4, part_mode),
    ROW3(CTX_PREV_INTRA_LUMA, 1, prev_intra),
    ROW3(CTX_INTRA_CHROMA, 1, intra_chroma),
    ROW1(CTX_RQT_ROOT_CBF, 1, rqt_root),
    ROW3(CTX_MERGE_FLAG, 1, merge_flag),
    ROW3(CTX_MERGE_IDX, 1, merge_idx),
    ROW1(CTX_INTER_PRED_IDC, 5, inter_pred_idc),
    ROW1(CTX_REF_IDX, 2, ref_idx),
    ROW1(CTX_MVP_FLAG, 1, mvp_flag),
    ROW3(CTX_SPLIT_TRANSFORM, 3, split_tr),
    ROW3(CTX_CBF_LUMA, 2, cbf_luma),
    ROW3(CTX_CBF_CHROMA, 5, cbf_chroma),
    ROW3(CTX_ABS_MVD_GT0, 1, mvd_gt0),
    ROW3(CTX_ABS_MVD_GT1, 1, mvd_gt1),
    ROW1(CTX_CU_QP_DELTA, 2, k154),
    ROW1(CTX_TRANSFORM_SKIP, 2, k139),
    ROW3(CTX_LAST_X, 18, last_prefix),
    ROW3(CTX_LAST_Y, 18, last_prefix),
    ROW3(CTX_CODED_SUB_BLOCK, 4, csbf),
    ROW3(CTX_SIG_COEFF, 44, sig),
    ROW3(CTX_GT1, 24, gt1),
    ROW3(CTX_GT2, 6, gt2),
    ROW1(CTX_EXPLICIT_RDPCM, 2, k139),
    ROW1(CTX_EXPLICIT_RDPCM_DIR, 2, k139),
    ROW1(CTX_CHROMA_QP_OFFSET_FLAG, 1, k154),
    ROW1(CTX_CHROMA_QP_OFFSET_IDX, 1, k154),
    ROW1(CTX_LOG2_RES_SCALE, 8, k154),
    ROW1(CTX_RES_SCALE_SIGN, 2, k154),
};

}  // namespace

void cabac_init_contexts(Ctx* ctx, int init_type, int slice_qp) {
  const int qp = slice_qp < 0 ? 0 : (slice_qp > 51 ? 51 : slice_qp);
  for (const InitRow& r : kRows) {
    for (int i = 0; i < r.count; i++) {
      const int iv = r.v[init_type][i];
      const int m = (iv >> 4) * 5 - 45, n = ((iv & 15) << 3) - 16;
      int pre = ((m * qp) >> 4) + n;
      pre = pre < 1 ? 1 : (pre > 126 ? 126 : pre);
      const int mps = pre > 63;
      const int state = mps ? pre - 64 : 63 - pre;
      ctx[r.offset + i] = (Ctx)((state << 1) | mps);
    }
  }
}

}  // namespace b265h
