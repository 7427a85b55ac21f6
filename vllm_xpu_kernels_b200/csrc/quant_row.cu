// Row-resident quantisation kernels: ops that need a reduction over the whole
// token row (per-token amax and/or the rms-norm variance).
//
//   a1  dynamic_per_token_scaled_fp8_quant        (MODE_TOKEN_DIV)
//   a6  rms_norm_dynamic_per_token_quant          (MODE_TOKEN_MUL / int8)
//   a7  rms_norm_per_block_quant                  (MODE_GROUP)
//   a8  rms_norm_mxfp4_quant                      (MODE_MXFP4)
//   a9  rms_norm_static_fp8_quant (+fused_add)    (MODE_STATIC)
//
// Design (quant_row_impl.cuh): persistent CTAs loop over rows; a row lives in registers
// in its packed input format (up to 16 vectors of 8 elements per thread, all 128-bit
// streaming loads issued before the first use), so HBM sees exactly one read of the
// activation (+ residual read/write) and one write of the codes.  The reference
// kernels re-read the row 2-3 times (csrc/layernorm_quant.cpp:27-142,
// fp8_quant.cpp:393-477).  Small CTAs (32-256 threads) keep the two row reductions
// cheap (a single warp needs no barrier at all).
// Rows whose length / alignment does not fit the vector path go through
// row_quant_generic (scalar, multi-pass, any H / stride).
#include "quant_row_impl.cuh"

namespace vxk {
extern template int row_launch<float>(int, int, const RowParams&, int64_t, bool, cudaStream_t);
extern template int row_launch<__half>(int, int, const RowParams&, int64_t, bool, cudaStream_t);
extern template int row_launch<__nv_bfloat16>(int, int, const RowParams&, int64_t, bool, cudaStream_t);

static int row_dispatch(int in_dtype, int op, int q_dtype, const RowParams& p, int64_t T_rows, bool vec_ok,
                        cudaStream_t st) {
  switch (in_dtype) {
    case VXK_F32: return row_launch<float>(op, q_dtype, p, T_rows, vec_ok, st);
    case VXK_F16: return row_launch<__half>(op, q_dtype, p, T_rows, vec_ok, st);
    case VXK_BF16: return row_launch<__nv_bfloat16>(op, q_dtype, p, T_rows, vec_ok, st);
    default: return fail(VXK_ERR_INVALID, "unsupported input dtype %d", in_dtype);
  }
}

static size_t esize(int in_dtype) { return in_dtype == VXK_F32 ? 4 : 2; }

// vector path: H % 8 == 0, 16-byte aligned base pointers, row strides multiple of 16 B
static bool vec_path_ok(const RowParams& p, int in_dtype, bool packed4 = false) {
  const size_t es = esize(in_dtype);
  if (p.H % kVec) return false;
  if (!aligned16(p.in) || (p.in_row_stride * es) % 16) return false;
  if (p.residual && !aligned16(p.residual)) return false;
  if (p.weight && !aligned16(p.weight)) return false;
  if (p.residual && (p.H * es) % 16) return false;
  const int64_t out_row_bytes = packed4 ? p.H / 2 : p.H;
  if (reinterpret_cast<uintptr_t>(p.out) % (packed4 ? 4 : 8)) return false;
  if (out_row_bytes % (packed4 ? 4 : 8)) return false;
  return true;
}

}  // namespace vxk

using namespace vxk;

extern "C" int vxk_dynamic_per_token_scaled_fp8_quant(void* out, const void* in, float* scales,
                                                      const float* scale_ub, int64_t T_, int64_t H,
                                                      int in_dtype, int q_dtype, vxk_stream_t stream) {
  VXK_REQUIRE(T_ >= 0 && H >= 0, "negative shape");
  VXK_REQUIRE(q_dtype == VXK_Q_E4M3 || q_dtype == VXK_Q_E5M2, "output must be fp8 (e4m3 / e5m2)");
  RowParams p{};
  p.out = out; p.in = in; p.scales = scales; p.scale_in = scale_ub;
  p.H = H; p.in_row_stride = H; p.nvec = static_cast<int>(H / kVec);
  const bool v = vec_path_ok(p, in_dtype);
  return row_dispatch(in_dtype, ROW_A1, q_dtype, p, T_, v, stream);
}

extern "C" int vxk_rms_norm_dynamic_per_token_quant(void* out, const void* in, const void* weight,
                                                    float* scales, double epsilon,
                                                    const float* scale_ub, void* residual, int64_t T_,
                                                    int64_t H, int in_dtype, int q_dtype,
                                                    vxk_stream_t stream) {
  VXK_REQUIRE(T_ >= 0 && H >= 0, "negative shape");
  VXK_REQUIRE(q_dtype == VXK_Q_E4M3 || q_dtype == VXK_Q_INT8, "output must be float8_e4m3fn or int8");
  VXK_REQUIRE(!(scale_ub && q_dtype != VXK_Q_E4M3), "scale_ub is only supported for FP8 output");
  RowParams p{};
  p.out = out; p.in = in; p.weight = weight; p.residual = residual; p.scales = scales;
  p.scale_in = scale_ub; p.H = H; p.in_row_stride = H; p.nvec = static_cast<int>(H / kVec);
  p.eps = static_cast<float>(epsilon);
  const bool v = vec_path_ok(p, in_dtype);
  return row_dispatch(in_dtype, ROW_A6, q_dtype, p, T_, v, stream);
}

extern "C" int vxk_rms_norm_per_block_quant(void* out, const void* in, const void* weight,
                                            float* scales, double epsilon, void* residual, int64_t T_,
                                            int64_t H, int64_t in_row_stride, int64_t group_size,
                                            int64_t ss0, int64_t ss1, int scale_ue8m0, int in_dtype,
                                            int q_dtype, vxk_stream_t stream) {
  VXK_REQUIRE(T_ >= 0 && H >= 0, "negative shape");
  VXK_REQUIRE(q_dtype == VXK_Q_E4M3 || q_dtype == VXK_Q_INT8, "output must be float8_e4m3fn or int8");
  VXK_REQUIRE(group_size > 0 && H % group_size == 0, "hidden_size must be divisible by group_size");
  RowParams p{};
  p.out = out; p.in = in; p.weight = weight; p.residual = residual; p.scales = scales;
  p.H = H; p.in_row_stride = in_row_stride; p.nvec = static_cast<int>(H / kVec);
  p.group = static_cast<int>(group_size); p.ss0 = ss0; p.ss1 = ss1;
  p.ue8m0 = (q_dtype == VXK_Q_E4M3) ? scale_ue8m0 : 0;
  p.eps = static_cast<float>(epsilon);
  const int lanes = static_cast<int>(group_size / kVec);
  const bool v = vec_path_ok(p, in_dtype) && group_size % kVec == 0 && lanes <= 32 &&
                 (lanes & (lanes - 1)) == 0;
  return row_dispatch(in_dtype, ROW_A7, q_dtype, p, T_, v, stream);
}

extern "C" int vxk_rms_norm_mxfp4_quant(void* out, const void* in, const void* weight, float* scales,
                                        double epsilon, void* residual, int64_t T_, int64_t H,
                                        int64_t group_size, int64_t ss0, int64_t ss1, int in_dtype,
                                        vxk_stream_t stream) {
  return vxk_rms_norm_mxfp4_quant_sf(out, in, weight, scales, nullptr, epsilon, residual, T_, H, group_size, ss0, ss1,
                                     in_dtype, stream);
}
extern "C" int vxk_rms_norm_mxfp4_quant_sf(void* out, const void* in, const void* weight, float* scales,
                                           void* sf_packed, double epsilon, void* residual, int64_t T_, int64_t H,
                                           int64_t group_size, int64_t ss0, int64_t ss1, int in_dtype,
                                           vxk_stream_t stream) {
  VXK_REQUIRE(T_ >= 0 && H >= 0, "negative shape");
  VXK_REQUIRE(group_size == 32, "MXFP4 requires group_size == 32, got %lld", (long long)group_size);
  VXK_REQUIRE(H % 32 == 0, "hidden_size must be divisible by group_size");
  RowParams p{};
  p.out = out; p.in = in; p.weight = weight; p.residual = residual; p.scales = scales;
  p.H = H; p.in_row_stride = H; p.nvec = static_cast<int>(H / kVec);
  p.group = 32; p.ss0 = ss0; p.ss1 = ss1; p.eps = static_cast<float>(epsilon);
  p.sf_packed = static_cast<uint8_t*>(sf_packed); p.sf_nblk = sf_pack_blocks(T_, 0);
  const bool v = vec_path_ok(p, in_dtype, /*packed4=*/true);
  return row_dispatch(in_dtype, ROW_A8, VXK_Q_E4M3, p, T_, v, stream);
}

extern "C" int vxk_rms_norm_static_fp8_quant(void* out, const void* in, const void* weight,
                                             const float* scale, double epsilon, void* residual,
                                             int64_t T_, int64_t H, int64_t in_row_stride, int in_dtype,
                                             int q_dtype, vxk_stream_t stream) {
  VXK_REQUIRE(T_ >= 0 && H >= 0, "negative shape");
  VXK_REQUIRE(q_dtype == VXK_Q_E4M3 || q_dtype == VXK_Q_E5M2, "output must be fp8 (e4m3 / e5m2)");
  RowParams p{};
  p.out = out; p.in = in; p.weight = weight; p.residual = residual; p.scale_in = scale;
  p.H = H; p.in_row_stride = in_row_stride; p.nvec = static_cast<int>(H / kVec);
  p.eps = static_cast<float>(epsilon);
  const bool v = vec_path_ok(p, in_dtype);
  return row_dispatch(in_dtype, ROW_A9, q_dtype, p, T_, v, stream);
}
