/*
 * oracle.h — CPU restatement (plain C) of the reference's pixel path, used ONLY as the checker in
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg.  TEST INFRASTRUCTURE: nothing
 * in libde265_b200/ links, loads or calls this.  Every function cites the reference lines it restates.
 *
 * Parity status: PINNED — each function here is compared against the compiled, unmodified reference
 * (oracle/_ref/libde265_ref.so, built by oracle/Makefile from /root/reference) in
 * tests/test_oracle_vs_ref.py, and the frame-level path against `dec265 -o` MD5s committed under
 * tests/golden/streams/.
 */
#ifndef B265_ORACLE_H
#define B265_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#include "../include/b265.h"

/* --- transforms (fallback-dct.cc, transform.cc) --- */
void orc_dequant_flat(int16_t* out, const int16_t* in, int log2, int qp);           /* transform.cc:700-729 */
void orc_idct_add(uint8_t* dst, ptrdiff_t stride, int nT, const int16_t* coeffs);     /* fallback-dct.cc:552-692 */
void orc_idst4_add(uint8_t* dst, ptrdiff_t stride, const int16_t* coeffs);            /* fallback-dct.cc:270-337 */
void orc_skip_add(uint8_t* dst, ptrdiff_t stride, int nT, const int16_t* coeffs, int rdpcm);   /* :80-90,186-213 */
void orc_bypass_add(uint8_t* dst, ptrdiff_t stride, int nT, const int16_t* coeffs, int rdpcm); /* :160-183,216-224 */
/* dense batch with the layout of b265_residual_batch */
void orc_residual_batch(int kind, int log2, int rdpcm, int n_blocks, const int16_t* coeffs,
                        const uint8_t* qp, uint8_t* dst);
/* one work-list TU onto a plane (scale_coefficients_internal, transform.cc:353-625) */
void orc_residual_tu(const b265_tu* tu, const b265_coeff* coeffs, const b265_scaling* sl,
                     uint8_t* plane, int pitch);

/* --- motion compensation (motion.cc:49-661, fallback-motion.cc) --- */
void orc_mc_batch(const b265_pu* pus, int n_pus, const b265_slice* slices,
                  const uint8_t* const* ref_planes, int n_refs, uint8_t* const* dst_planes,
                  int width, int height, int pitch, int pitch_c);
void orc_interp_block(int luma, int w, int h, const uint8_t* src, ptrdiff_t stride, int fx, int fy,
                      int16_t* out);

/* --- intra prediction (intrapred.cc:344-1139) --- */
void orc_intra_from_border(int log2, int cidx, int mode, const uint8_t* border_centre, int strong,
                           int no_edge_filter, uint8_t* out, int out_stride);
/* prediction of one TU inside a picture under reconstruction (border from the plane + availability) */
void orc_intra_tu(const b265_frame_job* job, const b265_tu* tu, uint8_t* const* planes,
                  const int* pitch);

/* --- whole-frame reconstruction from a work list (the six call sites of SURVEY §8b) --- */
/* planes: recon target; ref planes looked up through ref_planes[3*slot + c] */
void orc_reconstruct_frame(const b265_frame_job* job, uint8_t* const* planes, const int* pitch,
                           const uint8_t* const* ref_planes);

/* --- post filters (deblock.cc, sao.cc) --- */
void orc_deblock(const b265_frame_job* job, uint8_t* const* planes, const int* pitch);
void orc_sao(const b265_frame_job* job, const uint8_t* const* in_planes, uint8_t* const* out_planes,
             const int* pitch);
/* bS derivation for tests: out[(W/4)*(H/4)], bits 0-1 vertical edge bS, bits 2-3 horizontal */
void orc_boundary_strength(const b265_frame_job* job, uint8_t* bs_out);

#endif
