/*
 * de265.h — the public decoder API of libde265 as exported by libb265.so (B200 back end).
 *
 * Source-compatible with the reference's libde265/de265.h:72-430 for the calls a decoder client
 * uses (dec265/dec265.cc:620-800): same function names, argument meaning, enum values and error
 * behaviour, so an application written against libde265 links against libb265.so unchanged.
 * Differences, all inside what the reference API permits:
 *   - pictures are reconstructed on the GPU; de265_get_image_plane() returns pinned host memory
 *     filled by a device->host copy of the finished picture (de265.h:174, valid until the next
 *     de265_* call exactly as in the reference);
 *   - de265_start_worker_threads(n) sets the number of host CABAC/syntax parsing threads
 *     (pictures are parsed concurrently; the reference uses the pool for WPP rows/tiles);
 *   - streams needing tools this back end does not implement yet (inter prediction parsing)
 *     fail with DE265_ERROR_NOT_IMPLEMENTED_YET instead of decoding on the CPU: there is no CPU
 *     pixel path in this library.
 */
#ifndef B265_DE265_H
#define B265_DE265_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define LIBDE265_API __attribute__((visibility("default")))
#else
#define LIBDE265_API
#endif

#define LIBDE265_NUMERIC_VERSION 0x01000000

LIBDE265_API const char* de265_get_version(void);
LIBDE265_API uint32_t de265_get_version_number(void);
LIBDE265_API int de265_get_version_number_major(void);
LIBDE265_API int de265_get_version_number_minor(void);
LIBDE265_API int de265_get_version_number_maintenance(void);

/* numeric values identical to the reference's enum (de265.h:82-139) */
typedef enum {
  DE265_OK = 0,
  DE265_ERROR_NO_SUCH_FILE = 1,
  DE265_ERROR_COEFFICIENT_OUT_OF_IMAGE_BOUNDS = 4,
  DE265_ERROR_CHECKSUM_MISMATCH = 5,
  DE265_ERROR_CTB_OUTSIDE_IMAGE_AREA = 6,
  DE265_ERROR_OUT_OF_MEMORY = 7,
  DE265_ERROR_CODED_PARAMETER_OUT_OF_RANGE = 8,
  DE265_ERROR_IMAGE_BUFFER_FULL = 9,
  DE265_ERROR_CANNOT_START_THREADPOOL = 10,
  DE265_ERROR_LIBRARY_INITIALIZATION_FAILED = 11,
  DE265_ERROR_LIBRARY_NOT_INITIALIZED = 12,
  DE265_ERROR_WAITING_FOR_INPUT_DATA = 13,
  DE265_ERROR_CANNOT_PROCESS_SEI = 14,
  DE265_ERROR_PARAMETER_PARSING = 15,
  DE265_ERROR_NO_INITIAL_SLICE_HEADER = 16,
  DE265_ERROR_PREMATURE_END_OF_SLICE = 17,
  DE265_ERROR_UNSPECIFIED_DECODING_ERROR = 18,
  DE265_ERROR_NOT_IMPLEMENTED_YET = 502,
  DE265_WARNING_NO_WPP_CANNOT_USE_MULTITHREADING = 1000,
  DE265_WARNING_WARNING_BUFFER_FULL = 1001,
  DE265_WARNING_PREMATURE_END_OF_SLICE_SEGMENT = 1002,
  DE265_WARNING_INCORRECT_ENTRY_POINT_OFFSET = 1003,
  DE265_WARNING_CTB_OUTSIDE_IMAGE_AREA = 1004,
  DE265_WARNING_SPS_HEADER_INVALID = 1005,
  DE265_WARNING_PPS_HEADER_INVALID = 1006,
  DE265_WARNING_SLICEHEADER_INVALID = 1007,
  DE265_WARNING_INCORRECT_MOTION_VECTOR_SCALING = 1008,
  DE265_WARNING_NONEXISTING_PPS_REFERENCED = 1009,
  DE265_WARNING_NONEXISTING_SPS_REFERENCED = 1010,
  DE265_WARNING_BOTH_PREDFLAGS_ZERO = 1011,
  DE265_WARNING_NONEXISTING_REFERENCE_PICTURE_ACCESSED = 1012,
  DE265_WARNING_NUMMVP_NOT_EQUAL_TO_NUMMVQ = 1013,
  DE265_WARNING_NUMBER_OF_SHORT_TERM_REF_PIC_SETS_OUT_OF_RANGE = 1014,
  DE265_WARNING_SHORT_TERM_REF_PIC_SET_OUT_OF_RANGE = 1015,
  DE265_WARNING_FAULTY_REFERENCE_PICTURE_LIST = 1016,
  DE265_WARNING_EOSS_BIT_NOT_SET = 1017,
  DE265_WARNING_MAX_NUM_REF_PICS_EXCEEDED = 1018,
  DE265_WARNING_INVALID_CHROMA_FORMAT = 1019,
  DE265_WARNING_SLICE_SEGMENT_ADDRESS_INVALID = 1020,
  DE265_WARNING_DEPENDENT_SLICE_WITH_ADDRESS_ZERO = 1021,
  DE265_WARNING_NUMBER_OF_THREADS_LIMITED_TO_MAXIMUM = 1022,
  DE265_NON_EXISTING_LT_REFERENCE_CANDIDATE_IN_SLICE_HEADER = 1023,
  DE265_WARNING_CANNOT_APPLY_SAO_OUT_OF_MEMORY = 1024,
  DE265_WARNING_SPS_MISSING_CANNOT_DECODE_SEI = 1025,
  DE265_WARNING_COLLOCATED_MOTION_VECTOR_OUTSIDE_IMAGE_AREA = 1026
} de265_error;

LIBDE265_API const char* de265_get_error_text(de265_error err);
LIBDE265_API int de265_isOK(de265_error err); /* OK or a warning (>= 1000) */
LIBDE265_API void de265_disable_logging(void);
LIBDE265_API void de265_set_verbosity(int level);

/* ---- image ---- */
struct de265_image;
enum de265_chroma { de265_chroma_mono = 0, de265_chroma_420 = 1, de265_chroma_422 = 2, de265_chroma_444 = 3 };
typedef int64_t de265_PTS;

LIBDE265_API int de265_get_image_width(const struct de265_image*, int channel);
LIBDE265_API int de265_get_image_height(const struct de265_image*, int channel);
LIBDE265_API enum de265_chroma de265_get_chroma_format(const struct de265_image*);
LIBDE265_API int de265_get_bits_per_pixel(const struct de265_image*, int channel);
/* pointer to the first sample of the conformance window; *out_stride = bytes per line */
LIBDE265_API const uint8_t* de265_get_image_plane(const struct de265_image*, int channel, int* out_stride);
LIBDE265_API void* de265_get_image_plane_user_data(const struct de265_image*, int channel);
LIBDE265_API de265_PTS de265_get_image_PTS(const struct de265_image*);
LIBDE265_API void* de265_get_image_user_data(const struct de265_image*);
LIBDE265_API void de265_set_image_user_data(struct de265_image*, void* user_data);
LIBDE265_API void de265_get_image_NAL_header(const struct de265_image*, int* nal_unit_type,
                                             const char** nal_unit_name, int* nuh_layer_id,
                                             int* nuh_temporal_id);

/* ---- decoder ---- */
typedef void de265_decoder_context;

LIBDE265_API de265_decoder_context* de265_new_decoder(void);
LIBDE265_API de265_error de265_start_worker_threads(de265_decoder_context*, int number_of_threads);
LIBDE265_API de265_error de265_free_decoder(de265_decoder_context*);

/* input: a byte stream with start codes (push_data) or single NAL units without them (push_NAL);
 * the data is copied. flush_data marks the end of the stream. */
LIBDE265_API de265_error de265_push_data(de265_decoder_context*, const void* data, int length,
                                         de265_PTS pts, void* user_data);
LIBDE265_API void de265_push_end_of_NAL(de265_decoder_context*);
LIBDE265_API void de265_push_end_of_frame(de265_decoder_context*);
LIBDE265_API de265_error de265_push_NAL(de265_decoder_context*, const void* data, int length,
                                        de265_PTS pts, void* user_data);
LIBDE265_API de265_error de265_flush_data(de265_decoder_context*);
LIBDE265_API int de265_get_number_of_input_bytes_pending(de265_decoder_context*);
LIBDE265_API int de265_get_number_of_NAL_units_pending(de265_decoder_context*);

/* err == OK && more: call again; err != OK && more: stalled (WAITING_FOR_INPUT_DATA /
 * IMAGE_BUFFER_FULL); err == OK && !more: finished; err != OK && !more: fatal (de265.h:263-276) */
LIBDE265_API de265_error de265_decode(de265_decoder_context*, int* more);
LIBDE265_API void de265_reset(de265_decoder_context*);
/* deprecated one-call form (de265.h:206-219, de265.cc:262-300): push `length` bytes (0 = flush) and decode
 * everything that is complete; WAITING_FOR_INPUT_DATA is not reported as an error */
LIBDE265_API de265_error de265_decode_data(de265_decoder_context*, const void* data, int length);

LIBDE265_API const struct de265_image* de265_peek_next_picture(de265_decoder_context*);
LIBDE265_API const struct de265_image* de265_get_next_picture(de265_decoder_context*);
LIBDE265_API void de265_release_next_picture(de265_decoder_context*);
LIBDE265_API de265_error de265_get_warning(de265_decoder_context*);

enum de265_image_format {
  de265_image_format_mono8 = 1,
  de265_image_format_YUV420P8 = 2,
  de265_image_format_YUV422P8 = 3,
  de265_image_format_YUV444P8 = 4
};
struct de265_image_spec {
  enum de265_image_format format;
  int width, height, alignment;
  int crop_left, crop_right, crop_top, crop_bottom;
  int visible_width, visible_height;
};
/* Output-plane allocation hook (de265.h:325-343). With this back end the DPB itself lives in HBM;
 * the hook supplies the HOST buffers finished pictures are copied into. */
struct de265_image_allocation {
  int (*get_buffer)(de265_decoder_context* ctx, struct de265_image_spec* spec, struct de265_image* img,
                    void* userdata);
  void (*release_buffer)(de265_decoder_context* ctx, struct de265_image* img, void* userdata);
};
LIBDE265_API void de265_set_image_allocation_functions(de265_decoder_context*, struct de265_image_allocation*,
                                                       void* userdata);
LIBDE265_API const struct de265_image_allocation* de265_get_default_image_allocation_functions(void);
LIBDE265_API void de265_set_image_plane(struct de265_image* img, int cIdx, void* mem, int stride, void* userdata);

LIBDE265_API int de265_get_highest_TID(de265_decoder_context*);
LIBDE265_API int de265_get_current_TID(de265_decoder_context*);
LIBDE265_API void de265_set_limit_TID(de265_decoder_context*, int max_tid);
LIBDE265_API void de265_set_framerate_ratio(de265_decoder_context*, int percent);
LIBDE265_API int de265_change_framerate(de265_decoder_context*, int more_vs_less);

enum de265_param {
  DE265_DECODER_PARAM_BOOL_SEI_CHECK_HASH = 0,
  DE265_DECODER_PARAM_DUMP_SPS_HEADERS = 1,
  DE265_DECODER_PARAM_DUMP_VPS_HEADERS = 2,
  DE265_DECODER_PARAM_DUMP_PPS_HEADERS = 3,
  DE265_DECODER_PARAM_DUMP_SLICE_HEADERS = 4,
  DE265_DECODER_PARAM_ACCELERATION_CODE = 5,
  DE265_DECODER_PARAM_SUPPRESS_FAULTY_PICTURES = 6,
  DE265_DECODER_PARAM_DISABLE_DEBLOCKING = 7,
  DE265_DECODER_PARAM_DISABLE_SAO = 8
};
enum de265_acceleration {
  de265_acceleration_SCALAR = 0,
  de265_acceleration_MMX = 10,
  de265_acceleration_SSE = 20,
  de265_acceleration_SSE2 = 30,
  de265_acceleration_SSE4 = 40,
  de265_acceleration_AVX = 50,
  de265_acceleration_AVX2 = 60,
  de265_acceleration_ARM = 70,
  de265_acceleration_NEON = 80,
  de265_acceleration_AUTO = 10000,
  /* extension: the only pixel path this library has.  de265_set_parameter_int(ACCELERATION_CODE, v) accepts
   * AUTO and CUDA; any CPU code (SCALAR .. NEON) cannot be honoured (there is no CPU pixel path) and is
   * reported through de265_get_warning() as DE265_WARNING_B265_ACCELERATION_UNAVAILABLE */
  de265_acceleration_CUDA = 20000,
  de265_acceleration_CUDA_SM100 = 20000
};
/* b265 warnings (>= 1000 like every libde265 warning; outside the reference's range 1000-1026) */
#define DE265_WARNING_B265_ACCELERATION_UNAVAILABLE ((de265_error)1100)
/* current ACCELERATION_CODE (always de265_acceleration_CUDA) */
LIBDE265_API int b265_de265_get_acceleration(de265_decoder_context*);
LIBDE265_API void de265_set_parameter_bool(de265_decoder_context*, enum de265_param param, int value);
LIBDE265_API void de265_set_parameter_int(de265_decoder_context*, enum de265_param param, int value);
LIBDE265_API int de265_get_parameter_bool(de265_decoder_context*, enum de265_param param);

LIBDE265_API de265_error de265_init(void);
LIBDE265_API de265_error de265_free(void);

/* ---- b265 extensions (not in the reference) ---- */
/* CUDA device used by decoders created afterwards (default 0, or env B265_DEVICE) */
LIBDE265_API void b265_de265_set_device(int device);
/* pictures submitted to the GPU per launch batch (default 8); larger batches raise kernel efficiency */
LIBDE265_API void b265_de265_set_batch(de265_decoder_context*, int max_pictures);
/* statistics: [0] pictures decoded [1] host parse seconds (summed over threads) [2] GPU seconds
 * [3] bytes H2D [4] bytes D2H [5] kernel launches [6] GPU batches [7] seconds inside b265_submit_frames */
LIBDE265_API void b265_de265_get_stats(de265_decoder_context*, double out[8]);

/* Test/bench aid: decoders created while this is on only PARSE (CABAC + syntax -> work lists) and never
 * touch the GPU; returned images carry no pixels, their work lists are read with
 * b265_de265_image_job(ctx, img, struct b265_frame_job* out) declared in b265.h terms (pointers stay
 * valid until the image is released).  Such decoders deliver the pictures in DECODING order. */
LIBDE265_API void b265_de265_set_parse_only(int on);
/* `out` is a b265_frame_job (b265.h); void* here so that de265.h stays free of b265.h types */
LIBDE265_API int b265_de265_image_job(de265_decoder_context*, const struct de265_image*, void* out_b265_frame_job);
LIBDE265_API int b265_de265_image_poc(const struct de265_image*);
/* pictures whose decoded-picture-hash SEI was verified (on the device) so far; a mismatch makes de265_decode return
 * DE265_ERROR_CHECKSUM_MISMATCH while the picture is still delivered (decctx.cc:770-779, sei.cc:273-354) */
LIBDE265_API uint64_t b265_de265_hashes_checked(de265_decoder_context*);
/* 1 = IRAP picture with NoRaslOutputFlag (first of a coded video sequence: POCs restart, earlier pictures are output first) */
LIBDE265_API int b265_de265_image_starts_cvs(const struct de265_image*);

#ifdef __cplusplus
}
#endif
#endif
