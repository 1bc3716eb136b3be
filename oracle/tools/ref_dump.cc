// ref_dump.cc — debugging aid (test infrastructure): decodes a stream with the UNMODIFIED reference library and
// prints, per output picture and 4x4 unit, prediction mode and motion (image.h:497,847), to diff against the
// host parser's PU lists.  Built by hand against /root/reference headers + oracle/_ref/libde265_ref.so:
//   g++ -std=gnu++11 -I/root/reference -I/root/reference/libde265 -Ioracle/_ref/gen -Ioracle/_ref/gen/libde265 \
//       -DHAVE_STDINT_H oracle/tools/ref_dump.cc -Loracle/_ref -lde265_ref -Wl,-rpath,$PWD/oracle/_ref -o oracle/_ref/ref_dump
#include <stdio.h>
#include <vector>
#include "libde265/de265.h"
#include "libde265/image.h"

int main(int argc, char** argv) {
  FILE* f = fopen(argv[1], "rb");
  std::vector<uint8_t> d;
  int c;
  while ((c = fgetc(f)) != EOF) d.push_back((uint8_t)c);
  de265_decoder_context* ctx = de265_new_decoder();
  de265_push_data(ctx, d.data(), (int)d.size(), 0, nullptr);
  de265_flush_data(ctx);
  int more = 1, n = 0;
  while (more) {
    de265_error e = de265_decode(ctx, &more);
    if (e != DE265_OK && e != DE265_ERROR_WAITING_FOR_INPUT_DATA) break;
    while (const de265_image* img = de265_get_next_picture(ctx)) {
      printf("picture %d poc %d\n", n++, img->PicOrderCntVal);
      for (int y = 0; y < img->get_height(); y += 4)
        for (int x = 0; x < img->get_width(); x += 4) {
          const PBMotion& m = img->get_mv_info(x, y);
          printf("%d %d mode %d pf %d%d ref %d %d mv %d %d %d %d\n", x, y, (int)img->get_pred_mode(x, y), m.predFlag[0], m.predFlag[1],
                 m.predFlag[0] ? m.refIdx[0] : -1, m.predFlag[1] ? m.refIdx[1] : -1, m.predFlag[0] ? m.mv[0].x : 0,
                 m.predFlag[0] ? m.mv[0].y : 0, m.predFlag[1] ? m.mv[1].x : 0, m.predFlag[1] ? m.mv[1].y : 0);
        }
    }
  }
  return 0;
}
