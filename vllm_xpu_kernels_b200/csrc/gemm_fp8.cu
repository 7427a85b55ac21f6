#include "vxk_common.cuh"
extern "C" int vxk_fp8_gemm(void*, const void*, const void*, const void*, const void*, const void*, int64_t,
                            int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int, int, int64_t, int64_t,
                            int64_t, int64_t, int64_t, int64_t, int, int, int, int, vxk_stream_t) {
  return vxk::fail(VXK_ERR_UNSUPPORTED, "fp8_gemm not built yet");
}
