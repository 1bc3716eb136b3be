/*
 * b265_acceleration.h — the reference's kernel-level plugin surface, filled with CUDA entry points.
 *
 * libde265 selects its pixel primitives through one table of function pointers,
 * `struct acceleration_functions` (libde265/acceleration.h:29-201), installed per decoder context in
 * decctx.cc:430-449: first init_acceleration_functions_fallback() (fallback.cc:26-125), then - when
 * the CPU supports it - init_acceleration_functions_sse() overrides some slots (x86/sse.cc:37-103).
 * init_acceleration_functions_cuda() is a third initialiser of the same kind: call it AFTER the
 * fallback initialiser and it overrides the 8-bit decoder slots listed below with functions that run
 * the b265 CUDA kernels (the same device code the frame-level path of b265.h uses).
 *
 * `struct b265_acceleration_functions` restates the DATA members of the reference struct in their
 * declaration order (85 function pointers; the reference's non-virtual member functions take no
 * space), so a `struct acceleration_functions*` of the reference can be passed as is.
 *
 * The slots work on HOST memory, one block per call, exactly like the CPU slots they replace: every
 * call copies the block to the device, launches a kernel and copies the result back.  This keeps the
 * plugin contract but is slow by construction - it exists for drop-in compatibility and for slot-by-
 * slot parity checks (the `acceleration-speed --check` pattern); a decoder that wants throughput uses
 * the batched work-list path of b265.h, which is what de265_decode() of this library does.
 *
 * Slots installed (8-bit, decoder side):
 *   put_unweighted_pred_8, put_weighted_pred_avg_8, put_weighted_pred_8, put_weighted_bipred_8
 *                                                             (fallback-motion.cc:33-158)
 *   put_hevc_epel_8, put_hevc_epel_h_8, put_hevc_epel_v_8, put_hevc_epel_hv_8
 *                                                             (fallback-motion.cc:257-407)
 *   put_hevc_qpel_8[4][4]                                     (fallback-motion.cc:482-625)
 *   transform_skip_8, transform_4x4_dst_add_8, transform_add_8[4]
 *                                                             (fallback-dct.cc:160-177, 270-337, 552-692, 865-883)
 * Left as installed by the caller (no 8-bit pixel work in them, or not on the Main-profile path):
 *   the 16-bit slots, transform_bypass*, transform_skip_rdpcm_*, rotate_coefficients, the int32 residual
 *   slots (transform_idst_4x4 ... transform_skip_residual), and the encoder's forward transforms.
 */
#ifndef B265_ACCELERATION_H
#define B265_ACCELERATION_H

#include <stddef.h>
#include <stdint.h>
#include "b265.h"

#ifdef __cplusplus
extern "C" {
#endif

struct b265_acceleration_functions {
  /* acceleration.h:31-48 */
  void (*put_weighted_pred_avg_8)(uint8_t* dst, ptrdiff_t dststride, const int16_t* src1, const int16_t* src2,
                                  ptrdiff_t srcstride, int width, int height);
  void (*put_unweighted_pred_8)(uint8_t* dst, ptrdiff_t dststride, const int16_t* src, ptrdiff_t srcstride, int width,
                                int height);
  void (*put_weighted_pred_8)(uint8_t* dst, ptrdiff_t dststride, const int16_t* src, ptrdiff_t srcstride, int width,
                              int height, int w, int o, int log2WD);
  void (*put_weighted_bipred_8)(uint8_t* dst, ptrdiff_t dststride, const int16_t* src1, const int16_t* src2,
                                ptrdiff_t srcstride, int width, int height, int w1, int o1, int w2, int o2, int log2WD);
  /* acceleration.h:51-67 (9-16 bit writers: not installed) */
  void (*put_weighted_pred_avg_16)(uint16_t*, ptrdiff_t, const int16_t*, const int16_t*, ptrdiff_t, int, int, int);
  void (*put_unweighted_pred_16)(uint16_t*, ptrdiff_t, const int16_t*, ptrdiff_t, int, int, int);
  void (*put_weighted_pred_16)(uint16_t*, ptrdiff_t, const int16_t*, ptrdiff_t, int, int, int, int, int, int);
  void (*put_weighted_bipred_16)(uint16_t*, ptrdiff_t, const int16_t*, const int16_t*, ptrdiff_t, int, int, int, int, int,
                                 int, int, int);
  /* acceleration.h:89-104 */
  void (*put_hevc_epel_8)(int16_t* dst, ptrdiff_t dststride, const uint8_t* src, ptrdiff_t srcstride, int width, int height,
                          int mx, int my, int16_t* mcbuffer);
  void (*put_hevc_epel_h_8)(int16_t* dst, ptrdiff_t dststride, const uint8_t* src, ptrdiff_t srcstride, int width, int height,
                            int mx, int my, int16_t* mcbuffer, int bit_depth);
  void (*put_hevc_epel_v_8)(int16_t* dst, ptrdiff_t dststride, const uint8_t* src, ptrdiff_t srcstride, int width, int height,
                            int mx, int my, int16_t* mcbuffer, int bit_depth);
  void (*put_hevc_epel_hv_8)(int16_t* dst, ptrdiff_t dststride, const uint8_t* src, ptrdiff_t srcstride, int width, int height,
                             int mx, int my, int16_t* mcbuffer, int bit_depth);
  void (*put_hevc_qpel_8[4][4])(int16_t* dst, ptrdiff_t dststride, const uint8_t* src, ptrdiff_t srcstride, int width,
                                int height, int16_t* mcbuffer);
  /* acceleration.h:107-122 (9-16 bit interpolation: not installed) */
  void (*put_hevc_epel_16)(int16_t*, ptrdiff_t, const uint16_t*, ptrdiff_t, int, int, int, int, int16_t*, int);
  void (*put_hevc_epel_h_16)(int16_t*, ptrdiff_t, const uint16_t*, ptrdiff_t, int, int, int, int, int16_t*, int);
  void (*put_hevc_epel_v_16)(int16_t*, ptrdiff_t, const uint16_t*, ptrdiff_t, int, int, int, int, int16_t*, int);
  void (*put_hevc_epel_hv_16)(int16_t*, ptrdiff_t, const uint16_t*, ptrdiff_t, int, int, int, int, int16_t*, int);
  void (*put_hevc_qpel_16[4][4])(int16_t*, ptrdiff_t, const uint16_t*, ptrdiff_t, int, int, int16_t*, int);
  /* acceleration.h:143-145 (int32 residual of bypass blocks: not installed) */
  void (*transform_bypass)(int32_t* residual, const int16_t* coeffs, int nT);
  void (*transform_bypass_rdpcm_v)(int32_t* r, const int16_t* coeffs, int nT);
  void (*transform_bypass_rdpcm_h)(int32_t* r, const int16_t* coeffs, int nT);
  /* acceleration.h:149-153 */
  void (*transform_skip_8)(uint8_t* dst, const int16_t* coeffs, ptrdiff_t stride);
  void (*transform_skip_rdpcm_v_8)(uint8_t* dst, const int16_t* coeffs, int nT, ptrdiff_t stride);
  void (*transform_skip_rdpcm_h_8)(uint8_t* dst, const int16_t* coeffs, int nT, ptrdiff_t stride);
  void (*transform_4x4_dst_add_8)(uint8_t* dst, const int16_t* coeffs, ptrdiff_t stride);
  void (*transform_add_8[4])(uint8_t* dst, const int16_t* coeffs, ptrdiff_t stride); /* index log2(nT) - 2 */
  /* acceleration.h:157-159 (9-16 bit: not installed) */
  void (*transform_skip_16)(uint16_t*, const int16_t*, ptrdiff_t, int);
  void (*transform_4x4_dst_add_16)(uint16_t*, const int16_t*, ptrdiff_t, int);
  void (*transform_add_16[4])(uint16_t*, const int16_t*, ptrdiff_t, int);
  /* acceleration.h:162-178 (int32 residual path of the range extensions: not installed) */
  void (*rotate_coefficients)(int16_t* coeff, int nT);
  void (*transform_idst_4x4)(int32_t* dst, const int16_t* coeffs, int bdShift, int max_coeff_bits);
  void (*transform_idct_4x4)(int32_t* dst, const int16_t* coeffs, int bdShift, int max_coeff_bits);
  void (*transform_idct_8x8)(int32_t* dst, const int16_t* coeffs, int bdShift, int max_coeff_bits);
  void (*transform_idct_16x16)(int32_t* dst, const int16_t* coeffs, int bdShift, int max_coeff_bits);
  void (*transform_idct_32x32)(int32_t* dst, const int16_t* coeffs, int bdShift, int max_coeff_bits);
  void (*add_residual_8)(uint8_t* dst, ptrdiff_t stride, const int32_t* r, int nT, int bit_depth);
  void (*add_residual_16)(uint16_t* dst, ptrdiff_t stride, const int32_t* r, int nT, int bit_depth);
  void (*rdpcm_v)(int32_t* residual, const int16_t* coeffs, int nT, int tsShift, int bdShift);
  void (*rdpcm_h)(int32_t* residual, const int16_t* coeffs, int nT, int tsShift, int bdShift);
  void (*transform_skip_residual)(int32_t* residual, const int16_t* coeffs, int nT, int tsShift, int bdShift);
  /* acceleration.h:191-200 (encoder: not installed) */
  void (*fwd_transform_4x4_dst_8)(int16_t* coeffs, const int16_t* src, ptrdiff_t stride);
  void (*fwd_transform_8[4])(int16_t* coeffs, const int16_t* src, ptrdiff_t stride);
  void (*hadamard_transform_8[4])(int16_t* coeffs, const int16_t* src, ptrdiff_t stride);
};

/* Counterpart of init_acceleration_functions_sse(): overrides the slots listed at the top of this file.  `accel` may be the
 * reference's `struct acceleration_functions*` (same layout).  The CUDA context behind the slots is created on first use on
 * the device set by b265_acceleration_set_device() (default 0 or env B265_DEVICE); returns B265_OK or B265_ERR_NO_DEVICE, in
 * which case no slot is touched (there is no CPU code in this library to fall back to: the caller's table stays as it was). */
B265_API int init_acceleration_functions_cuda(struct b265_acceleration_functions* accel);
B265_API void b265_acceleration_set_device(int device);
/* kernels launched through the slots so far (tests: proves the GPU did the work) */
B265_API uint64_t b265_acceleration_launch_count(void);

/* The four prediction writers on dense blocks (put_*_pred_8 slots, fallback-motion.cc:33-158): block i = w x h samples at
 * src1/src2 + i*w*h (row stride w), dst + i*w*h.  mode 0: unweighted, 1: average of two, 2: weighted (w1, o1),
 * 3: weighted bi-prediction (w1, o1, w2, o2); log2WD as the slots define it. */
B265_API int b265_pred_write_batch(b265_ctx* ctx, int mem, int mode, int n_blocks, int w, int h, const int16_t* src1,
                                   const int16_t* src2, int w1, int o1, int w2, int o2, int log2WD, uint8_t* dst);

#ifdef __cplusplus
}
#endif
#endif /* B265_ACCELERATION_H */
