// cabac.h — CABAC arithmetic decoding engine (H.265 9.3.4.3) and context-variable initialisation
// (9.3.2.2, tables 9-5 .. 9-37) for the host syntax parser.  Replaces the reference's cabac.cc
// (decode_CABAC_bit :159-260, decode_CABAC_bypass :346-374, decode_CABAC_term_bit :277-315) and
// contextmodel.cc (initialize_CABAC_models :279-347).  The register layout (offset scaled by 7 bits,
// one byte fetched every 8 shifts) makes the byte position after a terminating bin identical to the
// reference's, which PCM sample reading and sub-stream switching rely on (slice.cc:4186-4213,4824).
#pragma once
#include <stdint.h>
#include <string.h>
#include "bits.h"

namespace b265h {

// context variable: bit 0 = valMps, bits 1..6 = pStateIdx
typedef uint8_t Ctx;

enum CtxOffset {
  CTX_SAO_MERGE = 0,                          // 1
  CTX_SAO_TYPE = CTX_SAO_MERGE + 1,           // 1
  CTX_SPLIT_CU = CTX_SAO_TYPE + 1,            // 3
  CTX_CU_TQ_BYPASS = CTX_SPLIT_CU + 3,        // 1
  CTX_CU_SKIP = CTX_CU_TQ_BYPASS + 1,         // 3
  CTX_PRED_MODE = CTX_CU_SKIP + 3,            // 1
  CTX_PART_MODE = CTX_PRED_MODE + 1,          // 4
  CTX_PREV_INTRA_LUMA = CTX_PART_MODE + 4,    // 1
  CTX_INTRA_CHROMA = CTX_PREV_INTRA_LUMA + 1, // 1
  CTX_RQT_ROOT_CBF = CTX_INTRA_CHROMA + 1,    // 1
  CTX_MERGE_FLAG = CTX_RQT_ROOT_CBF + 1,      // 1
  CTX_MERGE_IDX = CTX_MERGE_FLAG + 1,         // 1
  CTX_INTER_PRED_IDC = CTX_MERGE_IDX + 1,     // 5
  CTX_REF_IDX = CTX_INTER_PRED_IDC + 5,       // 2
  CTX_MVP_FLAG = CTX_REF_IDX + 2,             // 1
  CTX_SPLIT_TRANSFORM = CTX_MVP_FLAG + 1,     // 3
  CTX_CBF_LUMA = CTX_SPLIT_TRANSFORM + 3,     // 2
  CTX_CBF_CHROMA = CTX_CBF_LUMA + 2,          // 5 (4 + one for 4:2:2 depth 4)
  CTX_ABS_MVD_GT0 = CTX_CBF_CHROMA + 5,       // 1
  CTX_ABS_MVD_GT1 = CTX_ABS_MVD_GT0 + 1,      // 1
  CTX_CU_QP_DELTA = CTX_ABS_MVD_GT1 + 1,      // 2
  CTX_TRANSFORM_SKIP = CTX_CU_QP_DELTA + 2,   // 2 (luma, chroma)
  CTX_LAST_X = CTX_TRANSFORM_SKIP + 2,        // 18
  CTX_LAST_Y = CTX_LAST_X + 18,               // 18
  CTX_CODED_SUB_BLOCK = CTX_LAST_Y + 18,      // 4
  CTX_SIG_COEFF = CTX_CODED_SUB_BLOCK + 4,    // 42 + 2 (transform_skip_context)
  CTX_GT1 = CTX_SIG_COEFF + 44,               // 24
  CTX_GT2 = CTX_GT1 + 24,                     // 6
  CTX_EXPLICIT_RDPCM = CTX_GT2 + 6,           // 2
  CTX_EXPLICIT_RDPCM_DIR = CTX_EXPLICIT_RDPCM + 2,     // 2
  CTX_CHROMA_QP_OFFSET_FLAG = CTX_EXPLICIT_RDPCM_DIR + 2,  // 1
  CTX_CHROMA_QP_OFFSET_IDX = CTX_CHROMA_QP_OFFSET_FLAG + 1, // 1
  CTX_LOG2_RES_SCALE = CTX_CHROMA_QP_OFFSET_IDX + 1,   // 8
  CTX_RES_SCALE_SIGN = CTX_LOG2_RES_SCALE + 8,         // 2
  CTX_COUNT = CTX_RES_SCALE_SIGN + 2
};

void cabac_init_contexts(Ctx* ctx, int init_type, int slice_qp);

extern const uint8_t g_cabac_lps[64][4];
extern const uint8_t g_cabac_next_mps[64];
extern const uint8_t g_cabac_next_lps[64];
extern const uint8_t g_cabac_renorm[32];
extern uint8_t g_cabac_next[256];   // next context byte, indexed by (ctx << 1) | decodedLps

// what the syntax walk tells the engine about the bin(s) that follow; the decoder ignores it, the
// test-stream generator uses it to keep randomly drawn values legal
enum Hint { HINT_PCM_FLAG, HINT_QP_DELTA_SUFFIX, HINT_COEFF_REMAINING, HINT_END_OF_SEGMENT, HINT_END_OF_SUBSET, HINT_MVD_SUFFIX };

struct Cabac {
  typedef Cabac Local;            // residual_coding works on a register copy of the engine
  struct Setup {};
  void setup(const Setup&) {}
  void commit(Cabac& dst) const { dst = *this; }
  static inline void hint(int, int) {}
  static inline void set_ctx_base(const Ctx*) {}
  const uint8_t* cur = nullptr;
  const uint8_t* end = nullptr;
  uint32_t range = 510;   // after start(): kept in [256, 510] between decisions
  uint32_t value = 0;
  int bits_needed = 8;

  void set_data(const uint8_t* p, const uint8_t* e) { cur = p; end = e; }
  // 9.3.2.5 initialisation at the current byte position
  void start() {
    range = 510;
    bits_needed = 8;
    value = 0;
    if (cur < end) { value = (uint32_t)(*cur++) << 8; bits_needed -= 8; }
    if (cur < end) { value |= *cur++; bits_needed -= 8; }
  }
  // 9.3.4.3.2 DecodeDecision without data-dependent branches: the MPS/LPS outcome is a mask, the state
  // transition one table look-up (g_cabac_next[ctx*2 + lps]), the renormalisation shift a leading-zero count
  // (0 or 1 after an MPS, 1..6 after an LPS).  The only branch left is the byte refill (every 8th shift).
  inline int bin(Ctx& c) {
    const uint32_t s = c;
    const uint32_t lps = g_cabac_lps[s >> 1][(range >> 6) & 3];
    uint32_t r = range - lps;
    const uint32_t scaled = r << 7;
    const uint32_t m = (uint32_t)((int32_t)(scaled - value - 1) >> 31);   // ~0 iff value >= scaled (LPS)
    value -= scaled & m;
    r += (lps - r) & m;
    c = g_cabac_next[(s << 1) | (m & 1)];
    const int sh = __builtin_clz(r) - 23;   // r in [6, 510]
    range = r << sh;
    value <<= sh;
    bits_needed += sh;
    if (bits_needed >= 0) {
      if (cur < end) value |= (uint32_t)(*cur++) << bits_needed;
      bits_needed -= 8;
    }
    return (int)((s ^ m) & 1);
  }
  inline int bypass() {
    value <<= 1;
    if (++bits_needed >= 0) {
      bits_needed = -8;
      if (cur < end) value |= *cur++;
    }
    const uint32_t scaled = range << 7;
    const uint32_t m = (uint32_t)((int32_t)(scaled - value - 1) >> 31);
    value -= scaled & m;
    return (int)(m & 1);
  }
  // n <= 8 bypass bins at once: n bypass decisions are the long division of the scaled offset by the scaled
  // range (same register contents afterwards as n calls of bypass(); cabac.cc:377-410 of the reference
  // uses the same identity)
  inline uint32_t bypass_chunk(int n) {
    value <<= n;
    bits_needed += n;
    if (bits_needed >= 0) {
      if (cur < end) value |= (uint32_t)(*cur++) << bits_needed;
      bits_needed -= 8;
    }
    const uint32_t scaled = range << 7;
    uint32_t v = value / scaled;
    const uint32_t lim = (1u << n) - 1;
    if (v > lim) v = lim;  // only on corrupt data
    value -= v * scaled;
    return v;
  }
  inline uint32_t bypass_bits(int n) {
    uint32_t v = 0;
    while (n > 8) { v = (v << 8) | bypass_chunk(8); n -= 8; }
    if (n >= 3) return (v << n) | bypass_chunk(n);
    while (n-- > 0) v = (v << 1) | (uint32_t)bypass();
    return v;
  }
  // PCM sample data (7.3.8.7): byte-aligned raw bits at the current position, then re-initialisation.
  // Appends the samples (shifted to the full bit depth) to `out`; false when the data ends early.
  template <class Vec>
  bool pcm_samples(int n_luma, int bits_luma, int shift_luma, int n_chroma, int bits_chroma, int shift_chroma, Vec& out) {
    BitReader br(cur, (size_t)(end - cur));
    for (int i = 0; i < n_luma; i++) out.push_back((uint8_t)(br.u(bits_luma) << shift_luma));
    for (int i = 0; i < n_chroma; i++) out.push_back((uint8_t)(br.u(bits_chroma) << shift_chroma));
    if (br.err) return false;
    br.align();
    cur += br.pos >> 3;
    start();
    return true;
  }
  inline int terminate() {
    range -= 2;
    const uint32_t scaled = range << 7;
    if (value >= scaled) return 1;
    if (scaled < (256u << 7)) {
      range = scaled >> 6;
      value <<= 1;
      if (++bits_needed == 0) {
        bits_needed = -8;
        if (cur < end) value |= *cur++;
      }
    }
    return 0;
  }
};

}  // namespace b265h
