/*
 * b265.h — C ABI of the B200 (sm_100a) reconstruction back end for libde265's HEVC decoder.
 *
 * This is the drop-in boundary for the *pixel path* (SURVEY.md §8): everything the reference does
 * after CABAC/syntax parsing — dequantisation + inverse DST/DCT, motion compensation, intra
 * prediction, deblocking, SAO — executed by hand-written CUDA kernels on HBM-resident pictures.
 *
 * Two levels, both plain C (pointers + sizes, no C++/torch types):
 *
 *  1. FRAME LEVEL (this file, part A): the host parser emits per-frame *work lists* instead of
 *     calling the reference's six reconstruction call sites
 *         S1 decode_intra_prediction            libde265/slice.cc:3453   (intrapred.cc:1115)
 *         S2 scale_coefficients                 libde265/slice.cc:3471,3485 (transform.cc:353)
 *         S3 generate_inter_prediction_samples  libde265/motion.cc:2161  (motion.cc:279)
 *         S4 read_pcm_samples                   libde265/slice.cc:4356
 *         S5 run_postprocessing_filters_*       libde265/decctx.cc:757-766 (deblock.cc:1020, sao.cc:318)
 *         S6 picture lifecycle                  libde265/decctx.cc:2000, dpb.cc:204, image.cc:477
 *     and hands them to b265_submit_frames().
 *
 *  2. KERNEL LEVEL (part B): batched forms of the `acceleration_functions` slots
 *     (libde265/acceleration.h:29-201) operating on host or device buffers; used by the
 *     slot table in b265_acceleration.h, by the parity tests and by bench.py.
 *
 * All functions return 0 (B265_OK) on success or a negative b265_status. There is NO CPU fallback:
 * if no CUDA device is usable b265_create() fails with B265_ERR_NO_DEVICE.
 */
#ifndef B265_H
#define B265_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define B265_API __attribute__((visibility("default")))
#else
#define B265_API
#endif

typedef enum {
  B265_OK = 0,
  B265_ERR_NO_DEVICE = -1,      /* maps to DE265_ERROR_LIBRARY_INITIALIZATION_FAILED (de265.h:82-139) */
  B265_ERR_OUT_OF_MEMORY = -2,  /* maps to DE265_ERROR_OUT_OF_MEMORY */
  B265_ERR_INVALID_ARG = -3,
  B265_ERR_CUDA = -4,           /* maps to DE265_ERROR_UNSPECIFIED_DECODING_ERROR */
  B265_ERR_UNSUPPORTED = -5
} b265_status;

typedef struct b265_ctx b265_ctx; /* one per decoder instance == one CUDA device + one stream */

B265_API const char* b265_version(void);
B265_API const char* b265_status_string(int status);
/* last CUDA error text recorded on this context (never NULL) */
B265_API const char* b265_last_error(const b265_ctx* ctx);

B265_API int  b265_create(int device, b265_ctx** out);
B265_API void b265_destroy(b265_ctx* ctx);
B265_API int  b265_sync(b265_ctx* ctx);
/* Work submitted to `ctx` after this call starts only when everything submitted to `other` so far has finished (an event
 * between the two contexts' CUDA streams, no host wait).  Both contexts on the same device.  No reference counterpart:
 * the reference's thread pool orders its tasks on the host (threads.cc:150-260).  Contexts of one process that have
 * batches on the device at the same time share the SMs: the intra wavefront sizes its grid by the number of busy peers. */
B265_API int  b265_stream_wait(b265_ctx* ctx, b265_ctx* other);
/* number of kernels launched by this context since creation (bench.py's gpu_launches claim) */
B265_API uint64_t b265_launch_count(const b265_ctx* ctx);

/* ------------------------------------------------------------------------------------------
 * A.1  Surfaces: decoded pictures resident in HBM (replaces de265_image planes, image.cc:232-459).
 *      8-bit 4:2:0/4:0:0 planes; luma pitch = width rounded up to 256 bytes, chroma pitch = half of
 *      it rounded up to 128, so every row start is sector- and TMA-aligned. A surface id is the
 *      DPB slot handle the host keeps in its de265_image replacement.
 * ------------------------------------------------------------------------------------------ */
B265_API int b265_surface_alloc(b265_ctx* ctx, int width, int height, int chroma_format_idc,
                                int* surface_id);
B265_API int b265_surface_release(b265_ctx* ctx, int surface_id);
/* memset all planes (generate_unavailable_reference_picture, decctx.cc:1408-1434) */
B265_API int b265_surface_fill(b265_ctx* ctx, int surface_id, int luma_value, int chroma_value);
/* host <-> HBM copies of one plane; stride in bytes; asynchronous on the context stream.
 * download is followed by a stream sync only if `sync` != 0. */
B265_API int b265_surface_upload(b265_ctx* ctx, int surface_id, int plane, const uint8_t* src,
                                 ptrdiff_t stride);
B265_API int b265_surface_download(b265_ctx* ctx, int surface_id, int plane, uint8_t* dst,
                                   ptrdiff_t stride, int sync);
/* device pointer + pitch of a plane (for consumers that stay on the GPU) */
B265_API int b265_surface_plane(b265_ctx* ctx, int surface_id, int plane, void** dev_ptr,
                                int* pitch, int* width, int* height);

/* Decoded picture hash (SEI payload 132) of n surfaces, computed on the device from the pictures in HBM the way the
 * reference computes it from its host planes (sei.cc:160-270; verification :273-354): hash_type 0 = MD5, 1 = CRC,
 * 2 = checksum.  out receives n * 3 slots of 16 bytes (per colour plane: the MD5 digest, or the CRC / checksum as a
 * little-endian 32-bit word; absent planes: zeros), copied asynchronously on the context stream - valid after
 * b265_sync(); `out` should be pinned (b265_host_alloc). */
B265_API int b265_surface_hash(b265_ctx* ctx, const int* surface_ids, int n, int hash_type, uint8_t* out);

/* CRC (the hash_type 1 value of b265_surface_hash) of ONE plane, for callers that cannot allocate: code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc.  NO COUNTERPART IN THE REFERENCE.
 * The call allocates nothing (no cudaMalloc / cudaFree, no growable context buffer): its per-row
 * scratch is a static device array of B265_CRC_NOALLOC_MAX_ROWS 16-bit entries (16 KB, held by every GPU that loads the
 * library until exit).  dev_plane = device pointer to `height` rows of `width` 8-bit samples, `pitch` bytes apart
 * (address, pitch and width multiples of 4); dev_crc = device pointer to one 32-bit word that receives the CRC.
 * Asynchronous on the context stream (valid after b265_sync or any later work on that stream).
 * height > B265_CRC_NOALLOC_MAX_ROWS -> B265_ERR_UNSUPPORTED with the reason in b265_last_error().
 * All calls share the scratch: calls that overlap on the GPU (other streams, contexts or threads) RACE - serialise them. */
#define B265_CRC_NOALLOC_MAX_ROWS 8192
B265_API int b265_plane_crc_noalloc(b265_ctx* ctx, const uint8_t* dev_plane, int pitch, int width, int height,
                                    uint32_t* dev_crc);

/* ------------------------------------------------------------------------------------------
 * A.2  Work-list records (all little-endian PODs, written by the host parser into pinned memory)
 * ------------------------------------------------------------------------------------------ */

/* one non-zero transform coefficient as the parser produces it (slice.cc:3408-3409:
 * tctx->coeffList / tctx->coeffPos) */
typedef struct {
  int16_t  level;  /* quantised level, NOT dequantised (dequant runs on the GPU) */
  uint16_t pos;    /* x + y*nT inside the transform block */
} b265_coeff;

/* b265_tu.flags */
#define B265_TU_DST          0x01 /* 4x4 luma intra: inverse DST-VII (transform.cc:584-589) */
#define B265_TU_SKIP         0x02 /* transform_skip_flag (transform.cc:530-579) */
#define B265_TU_BYPASS       0x04 /* cu_transquant_bypass (transform.cc:399-439) */
#define B265_TU_RDPCM_H      0x08 /* rdpcmMode==1 */
#define B265_TU_RDPCM_V      0x10 /* rdpcmMode==2 */
#define B265_TU_ROTATE       0x20 /* transform_skip_rotation (4x4, intra) */
#define B265_TU_NO_EDGE_FILT 0x40 /* intra boundary filters off (intrapred.cc:1102-1104) */
#define B265_TU_INTRA_CU     0x80 /* CU pred mode is MODE_INTRA (scaling-list matrixID, transform.cc:481-485) */

#define B265_INTRA_NONE 0xFF

/* one transform block of one colour component: optional intra prediction (S1) followed by an
 * optional residual (S2). Emitted in decoding order. 16 bytes. */
typedef struct {
  uint16_t x, y;        /* top-left, in samples of component cIdx */
  uint8_t  cidx_log2;   /* bits 0-1: cIdx, bits 2-4: log2(nT) in 2..5 */
  uint8_t  flags;       /* B265_TU_* */
  uint8_t  qp;          /* qP'(Y|Cb|Cr) incl. QpBdOffset (transform.cc:362-368) */
  uint8_t  intra_mode;  /* 0..34 => predict first; B265_INTRA_NONE => residual only */
  uint32_t coeff_off;   /* first entry in the frame's coefficient pool */
  uint16_t ncoeff;      /* entries (0 => prediction only) */
  uint16_t reserved;
} b265_tu;

/* b265_pu.flags */
#define B265_PU_L0       0x01
#define B265_PU_L1       0x02

/* one prediction block (S3: motion.cc:279-661). 24 bytes. */
typedef struct {
  uint16_t x, y;        /* luma position of the PB */
  uint8_t  w, h;        /* luma size (4..64) */
  uint8_t  flags;       /* B265_PU_L0 | B265_PU_L1 (identical-MV collapse motion.cc:326-335 is done
                           on the device, it depends on weighted_pred) */
  uint8_t  slice;       /* index into b265_frame_job.slices (weights, slice type) */
  int16_t  mv[2][2];    /* [list][x,y] quarter-sample units */
  int8_t   ref_idx[2];  /* refIdx per list (weight-table row) */
  uint8_t  ref_slot[2]; /* index into b265_frame_job.ref_surfaces per list (RefPicList resolved on the
                           host, motion.cc:348) */
  uint32_t reserved;
} b265_pu;

/* PCM coding block (S4: slice.cc:4143-4183); samples already shifted to bit depth, stored
 * Y (n*n), then Cb, Cr ((n/2)^2 each) in the frame's pcm pool. */
typedef struct {
  uint16_t x, y;       /* luma position */
  uint8_t  log2_size;  /* log2 of the luma block size */
  uint8_t  reserved[3];
  uint32_t sample_off; /* offset into pcm pool */
} b265_pcm;

/* per-CTB SAO parameters (slice.h:265-273 sao_info, after merge resolution slice.cc:2825-2831) */
typedef struct {
  uint8_t type[3];      /* SaoTypeIdx per component: 0 off, 1 band, 2 edge */
  uint8_t band_eo[3];   /* band position (type 1) or edge class 0..3 (type 2) */
  int8_t  offset[3][4]; /* final signed, scaled offsets (slice.cc:2790-2816) */
} b265_sao;

/* per-CTB addressing/ownership (pps CtbAddrRStoTS / TileIdRS, image.h ctb_info.SliceAddrRS) */
typedef struct {
  int32_t  addr_ts;    /* CTB address in tile scan (decoding order) */
  uint16_t slice;      /* index into b265_frame_job.slices (one entry per *independent* slice) */
  uint16_t tile;       /* TileIdRS */
} b265_ctb;

/* b265_frame_job.unit_info bits, one byte per 4x4 luma unit (image.h:70-74 deblk_info + cb_info) */
#define B265_UNIT_EDGE_V    0x01 /* left edge of the unit is a filtered TU or PU edge (deblock.cc:130-225) */
#define B265_UNIT_EDGE_H    0x02 /* top edge likewise */
#define B265_UNIT_TU_EDGE_V 0x04 /* the left edge is a transform-block edge (bS=1 by cbf rule) */
#define B265_UNIT_TU_EDGE_H 0x08
#define B265_UNIT_INTRA     0x10 /* pred_mode == MODE_INTRA */
#define B265_UNIT_CBF       0x20 /* luma TB containing the unit has non-zero coefficients */
#define B265_UNIT_NOFILT    0x40 /* (pcm && pcm_loop_filter_disable) || cu_transquant_bypass
                                    (deblock.cc:577-590, sao.cc:112-117) */
#define B265_UNIT_DONE      0x80 /* reconstructed before the intra wavefront starts (inter/PCM) */

/* per independent slice: the header fields the pixel path reads (slice.h:136-247) */
typedef struct {
  int32_t slice_addr_rs;
  int8_t  slice_type;                 /* 0 B, 1 P, 2 I (slice.h SLICE_TYPE_*) */
  int8_t  beta_offset, tc_offset;     /* slice_beta_offset, slice_tc_offset (already *2) */
  uint8_t deblocking_disabled;        /* slice_deblocking_filter_disabled_flag */
  uint8_t lf_across_slices;           /* slice_loop_filter_across_slices_enabled_flag */
  uint8_t sao_luma, sao_chroma;       /* slice_sao_luma_flag / slice_sao_chroma_flag */
  uint8_t weighted;                   /* 1 => explicit weighted prediction applies to this slice
                                         (P: pps weighted_pred_flag, B: weighted_bipred_flag) */
  uint8_t luma_log2_denom, chroma_log2_denom;
  uint8_t pps_weighted_pred;          /* pps weighted_pred_flag: the identical-MV bi->uni collapse is
                                         keyed on THIS flag for P and B slices alike (motion.cc:326) */
  uint8_t pad;
  int16_t weight[2][16][3];           /* LumaWeight / ChromaWeight[.][.][0..1] (slice.h:195-203) */
  int16_t offset[2][16][3];           /* luma_offset / ChromaOffset, NOT yet scaled by bit depth */
} b265_slice;

/* b265_frame_job.flags */
#define B265_FRAME_STRONG_INTRA_SMOOTHING 0x0001
#define B265_FRAME_CONSTRAINED_INTRA      0x0002
#define B265_FRAME_LF_ACROSS_TILES        0x0004 /* pps loop_filter_across_tiles_enabled_flag */
#define B265_FRAME_DEBLOCK                0x0008 /* run the deblocking pass for this picture */
#define B265_FRAME_SAO                    0x0010 /* run SAO for this picture */
#define B265_FRAME_SCALING_LIST           0x0020 /* sps scaling_list_enable_flag */
#define B265_FRAME_PCM_LF_DISABLE         0x0040
#define B265_FRAME_SIGN_UNUSED            0x0080

/* scaling factors (pps.scaling_list ScalingFactor_Size0..3, transform.cc:489-495), [matrixID][y*nT+x] */
typedef struct {
  uint8_t s4[6][16];
  uint8_t s8[6][64];
  uint8_t s16[6][256];
  uint8_t s32[6][1024]; /* matrixID 0..5 as the reference indexes it (only 0,1 / 0,3 used in 4:2:0) */
} b265_scaling;

/* One picture's complete reconstruction job. Pointers are HOST pointers (ideally pinned, see
 * b265_host_alloc); the call copies them to the device asynchronously. */
typedef struct {
  int32_t width, height;            /* luma samples (pic_{width,height}_in_luma_samples) */
  uint8_t chroma_format_idc;        /* 1 (4:2:0) or 0 (4:0:0) */
  uint8_t bit_depth;                /* 8 */
  uint8_t log2_ctb, log2_min_cb, log2_min_tb;
  int8_t  cb_qp_offset, cr_qp_offset; /* pps pic_cb/cr_qp_offset (deblock.cc:809-819) */
  uint8_t log2_sao_offset_scale_unused;
  uint32_t flags;                   /* B265_FRAME_* */
  int32_t dst_surface;              /* output / DPB surface */
  int32_t ref_surfaces[32];         /* surfaces referenced by b265_pu.ref_slot */
  int32_t ref_poc[32];              /* an identity per ref_slot used by the bS motion compare
                                       (deblock.cc:292-357 compares pictures, not indices) */
  int32_t n_refs;

  const b265_tu*    tus;    int32_t n_tus;     /* decoding order */
  const b265_coeff* coeffs; int32_t n_coeffs;
  const b265_pu*    pus;    int32_t n_pus;
  const b265_pcm*   pcms;   int32_t n_pcms;
  const uint8_t*    pcm_samples; int32_t n_pcm_samples;
  const b265_slice* slices; int32_t n_slices;
  const b265_ctb*   ctbs;              /* PicSizeInCtbsY entries, raster order; NULL => single slice,
                                          single tile, raster == tile scan */
  const b265_sao*   sao;               /* per CTB; NULL if !B265_FRAME_SAO */
  const uint8_t*    unit_info;         /* (W/4)*(H/4) B265_UNIT_* bytes; NULL => all-intra, no edges */
  const int8_t*     qp_y;              /* QP_Y per min-CB (W>>log2_min_cb)*(H>>log2_min_cb); deblock */
  const b265_scaling* scaling;         /* NULL unless B265_FRAME_SCALING_LIST */
} b265_frame_job;

/* pinned host memory for work lists and output planes */
B265_API int  b265_host_alloc(b265_ctx* ctx, size_t bytes, void** out);
B265_API void b265_host_free(b265_ctx* ctx, void* p);

/* Reconstruct n_jobs (<= 4096) pictures, given in decoding order.  A picture may reference pictures of the same
 * call: the call is cut into launch groups by dependency level (pictures that reference nothing of this call
 * share the first group, and so on; at most 255 pictures per group), each group runs MC -> inter residual ->
 * intra wavefront (+residual) -> deblock V -> deblock H -> SAO on the context stream, groups one after the other.
 * Destination surfaces of one call are distinct.  Returns when everything is enqueued; the job arrays may be
 * reused after b265_sync() or after a later download with sync. */
B265_API int b265_submit_frames(b265_ctx* ctx, const b265_frame_job* jobs, int n_jobs);

/* Same, but with work lists already resident on the device from a previous b265_upload_frames();
 * used by bench.py to time the kernels with inputs in HBM. */
typedef struct b265_batch b265_batch;
B265_API int  b265_upload_frames(b265_ctx* ctx, const b265_frame_job* jobs, int n_jobs,
                                 b265_batch** out);
B265_API int  b265_run_batch(b265_ctx* ctx, b265_batch* batch);
B265_API void b265_free_batch(b265_ctx* ctx, b265_batch* batch);
/* per-stage device time of the last b265_run_batch with timing enabled (ms):
 * [0] pcm+mc [1] residual (inter in place, intra to scratch) + availability masks [2] intra wavefront
 * [3] deblock (bs+V+H) [4] sao [5] total */
B265_API int  b265_set_stage_timing(b265_ctx* ctx, int enable);
B265_API int  b265_get_stage_times(b265_ctx* ctx, float ms[6]);

/* ------------------------------------------------------------------------------------------
 * B.  Kernel-level batched entry points (dense forms of the acceleration.h slots).
 *     `mem` selects whether the buffers are host (0, copied in/out inside the call) or device (1)
 *     pointers. These launch the same device code the frame-level path uses.
 * ------------------------------------------------------------------------------------------ */
#define B265_MEM_HOST   0
#define B265_MEM_DEVICE 1

/* b265_residual_batch kinds */
#define B265_TR_DCT    0 /* transform_add_8[log2-2]        acceleration.h:153  (fallback-dct.cc:552-692) */
#define B265_TR_DST    1 /* transform_4x4_dst_add_8        acceleration.h:152  (fallback-dct.cc:270-337) */
#define B265_TR_SKIP   2 /* transform_skip_residual + add  acceleration.h:177  (fallback-dct.cc:80-90)   */
#define B265_TR_BYPASS 3 /* transform_bypass + add         acceleration.h:143  (fallback-dct.cc:216-224) */

/* n blocks of size nT = 1<<log2, block i: coeffs[i*nT*nT .. ] dense row-major int16 (already
 * dequantised unless qp != NULL, in which case the flat dequantiser of dequant_coefficients,
 * transform.cc:700-729, runs first with qP = qp[i]); dst block i = nT x nT bytes at
 * dst + i*nT*nT (row stride nT): dst = Clip1(dst + residual). rdpcm: 0 none, 1 H, 2 V (skip/bypass only). */
B265_API int b265_residual_batch(b265_ctx* ctx, int mem, int kind, int log2, int rdpcm, int n_blocks,
                                 const int16_t* coeffs, const uint8_t* qp, uint8_t* dst);

/* Same arithmetic through the sparse work-list form the frame path uses: tus[i] addresses plane
 * `dst` (pitch in bytes) of a single component; coefficient levels are dequantised on the device. */
B265_API int b265_residual_sparse(b265_ctx* ctx, int mem, const b265_tu* tus, int n_tus,
                                  const b265_coeff* coeffs, int n_coeffs, uint8_t* dst, int pitch,
                                  int width, int height);

/* Motion compensation of a PU list onto a destination picture from reference pictures
 * (generate_inter_prediction_samples, motion.cc:279-661, including mc_luma/mc_chroma edge
 * clamping, put_hevc_qpel/epel and the four put_*_pred writers). Planes are 4:2:0:
 * ref_planes[3*r + c] / dst_planes[c], luma pitch `pitch`, chroma pitch `pitch_c`. */
B265_API int b265_mc_batch(b265_ctx* ctx, int mem, const b265_pu* pus, int n_pus,
                           const b265_slice* slices, int n_slices,
                           const uint8_t* const* ref_planes, int n_refs,
                           uint8_t* const* dst_planes, int width, int height, int pitch, int pitch_c);

/* Raw interpolation slots (put_hevc_qpel_8[xFrac][yFrac], put_hevc_epel*_8; acceleration.h:87-102):
 * n blocks of w x h from src (+ i*src_block_stride, row stride src_stride, with the 3/4 (luma) or
 * 1/2 (chroma) sample halo present around each block) to int16 out (w*h each, row stride w). */
B265_API int b265_interp_batch(b265_ctx* ctx, int mem, int luma, int n_blocks, int w, int h,
                               const uint8_t* src, ptrdiff_t src_stride, ptrdiff_t src_block_stride,
                               const uint8_t* frac_xy, int16_t* out);

/* Intra prediction of isolated blocks from explicit border arrays
 * (intra_prediction_{angular,planar,DC} + intra_prediction_sample_filtering, intrapred.cc:816-1069):
 * border[i] has 4*nT+1 samples p[-2nT..2nT] (centre = corner). out block i = nT*nT bytes. */
B265_API int b265_intra_batch(b265_ctx* ctx, int mem, int log2, int cidx, int n_blocks,
                              const uint8_t* modes, const uint8_t* borders, int strong_filter,
                              int no_edge_filter, uint8_t* out);

/* Whole-plane post filters on raw 4:2:0 planes with explicit metadata (apply_deblocking_filter
 * deblock.cc:1020-1058, apply_sample_adaptive_offset_sequential sao.cc:318-370). The job supplies
 * width/height/log2 sizes/flags/slices/ctbs/unit_info/qp_y/sao; planes are host or device. */
B265_API int b265_deblock_planes(b265_ctx* ctx, int mem, const b265_frame_job* meta,
                                 uint8_t* const* planes, int pitch, int pitch_c);
B265_API int b265_sao_planes(b265_ctx* ctx, int mem, const b265_frame_job* meta,
                             const uint8_t* const* in_planes, uint8_t* const* out_planes,
                             int pitch, int pitch_c);

/* raw device memory helpers for MEM_DEVICE callers (bench.py, torch-free) */
B265_API int  b265_dev_alloc(b265_ctx* ctx, size_t bytes, void** out);
B265_API void b265_dev_free(b265_ctx* ctx, void* p);
B265_API int  b265_memcpy_h2d(b265_ctx* ctx, void* dst, const void* src, size_t bytes);
B265_API int  b265_memcpy_d2h(b265_ctx* ctx, void* dst, const void* src, size_t bytes);
/* CUDA-event timing on the context stream */
B265_API int  b265_timer_start(b265_ctx* ctx);
B265_API int  b265_timer_stop(b265_ctx* ctx, float* ms);
/* write `bytes` of scratch (> L2) to evict cached inputs between timed iterations */
B265_API int  b265_flush_l2(b265_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* B265_H */
