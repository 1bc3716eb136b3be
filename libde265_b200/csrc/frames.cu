// frames.cu — frame-level C ABI: b265_upload_frames / b265_run_batch / b265_submit_frames.
#include "backend.h"
#include "kernels.h"

struct b265_batch {
  int n_frames = 0;
};

extern "C" {
void b265_free_batch(b265_ctx* ctx, b265_batch* batch) {
  (void)ctx;
  delete batch;
}
}
