/*
 * dec_client.c — a plain C decoder client written against include/de265.h only, the way a libde265 user's program is
 * (the decode loop of the reference's dec265/dec265.cc:700-770: de265_push_data in chunks -> de265_decode ->
 * de265_get_next_picture -> write the planes).  Built by tests/test_c_client.py with
 *     gcc -std=c99 -I include tests/c/dec_client.c -L libde265_b200 -lb265
 * and run on a golden stream; the test compares the MD5 of every plane with the reference decoder's.
 *
 * usage: dec_client <in.bin> <out.yuv> [threads] [check_hash]
 */
#include <stdio.h>
#include <stdlib.h>
#include "de265.h"

static int write_picture(FILE* out, const struct de265_image* img) {
  int c;
  for (c = 0; c < 3; c++) {
    int stride, y;
    const uint8_t* p = de265_get_image_plane(img, c, &stride);
    const int w = de265_get_image_width(img, c), h = de265_get_image_height(img, c);
    if (!p || w <= 0 || h <= 0) continue;
    for (y = 0; y < h; y++)
      if (fwrite(p + (size_t)y * stride, 1, (size_t)w, out) != (size_t)w) return -1;
  }
  return 0;
}

int main(int argc, char** argv) {
  FILE *in, *out;
  de265_decoder_context* ctx;
  de265_error err = DE265_OK;
  static uint8_t buf[40960];
  int more = 1, eof = 0, pictures = 0, warnings = 0;
  if (argc < 3) { fprintf(stderr, "usage: %s in.bin out.yuv [threads] [check_hash]\n", argv[0]); return 2; }
  in = fopen(argv[1], "rb");
  out = fopen(argv[2], "wb");
  if (!in || !out) { perror("open"); return 2; }
  ctx = de265_new_decoder();
  if (!ctx) { fprintf(stderr, "de265_new_decoder failed\n"); return 3; }
  if (argc > 3 && atoi(argv[3]) > 0) de265_start_worker_threads(ctx, atoi(argv[3]));
  if (argc > 4 && atoi(argv[4])) de265_set_parameter_bool(ctx, DE265_DECODER_PARAM_BOOL_SEI_CHECK_HASH, 1);
  while (more) {
    const struct de265_image* img;
    more = 0;
    err = de265_decode(ctx, &more);
    if (err == DE265_ERROR_WAITING_FOR_INPUT_DATA && !eof) {
      const size_t n = fread(buf, 1, sizeof(buf), in);
      if (n > 0) err = de265_push_data(ctx, buf, (int)n, 0, NULL);
      else { err = de265_flush_data(ctx); eof = 1; }
      if (!de265_isOK(err)) break;
      more = 1;
    } else if (err == DE265_ERROR_IMAGE_BUFFER_FULL) {
      more = 1;
    } else if (!de265_isOK(err)) {
      break;
    }
    while ((img = de265_get_next_picture(ctx)) != NULL) {
      if (write_picture(out, img)) { fprintf(stderr, "write failed\n"); return 4; }
      pictures++;
    }
    for (;;) {
      const de265_error w = de265_get_warning(ctx);
      if (w == DE265_OK) break;
      fprintf(stderr, "warning: %s\n", de265_get_error_text(w));
      warnings++;
    }
  }
  fclose(in);
  fclose(out);
  de265_free_decoder(ctx);
  if (!de265_isOK(err)) { fprintf(stderr, "decoding error %d: %s\n", (int)err, de265_get_error_text(err)); return 10 + (int)err % 100; }
  printf("pictures %d warnings %d\n", pictures, warnings);
  return 0;
}
