#include "vxk_common.cuh"
extern "C" int vxk_fp4_gemm(void*, const void*, const void*, const void*, const void*, const void*, int64_t,
                            int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int,
                            int, vxk_stream_t) {
  return vxk::fail(VXK_ERR_UNSUPPORTED, "fp4_gemm not built yet");
}
