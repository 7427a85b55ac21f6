// PyTorch operator surface of the quantized-linear hot path, registered under
// the CUDA dispatch key.  Schemas are the reference's, verbatim:
//   library _C      <- /root/reference/csrc/torch_bindings.cpp:38-107,146-184
//   library _xpu_C  <- /root/reference/csrc/xpu/torch_bindings.cpp:14-27
// Each implementation replicates the reference launcher's argument checks
// (TORCH_CHECK -> RuntimeError), takes the current CUDA stream and calls the
// C ABI of libvxk_b200.so (include/vxk_b200.h).  No kernels live here and
// there is no CPU or library fallback: a failing C-ABI call raises.
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <ATen/ATen.h>
#include <torch/library.h>

#include <optional>
#include <string>
#include <tuple>

#include "../../include/vxk_b200.h"

namespace {

using at::Tensor;

inline vxk_stream_t cur_stream() {
  return reinterpret_cast<vxk_stream_t>(at::cuda::getCurrentCUDAStream().stream());
}

inline void check_rc(int rc, const char* op) {
  TORCH_CHECK(rc == 0, op, ": ", vxk_last_error());
}

inline int in_dtype_of(const Tensor& t, const char* op) {
  switch (t.scalar_type()) {
    case at::kFloat: return VXK_F32;
    case at::kHalf: return VXK_F16;
    case at::kBFloat16: return VXK_BF16;
    default: TORCH_CHECK(false, op, ": unsupported input dtype ", t.scalar_type());
  }
}

inline int q_dtype_of(const Tensor& t, const char* op, bool allow_int8) {
  switch (t.scalar_type()) {
    case at::kFloat8_e4m3fn: return VXK_Q_E4M3;
    case at::kFloat8_e5m2: return VXK_Q_E5M2;
    case at::kChar:
      TORCH_CHECK(allow_int8, op, ": int8 output is not supported by this op");
      return VXK_Q_INT8;
    default: TORCH_CHECK(false, op, ": unsupported output dtype ", t.scalar_type());
  }
}

inline void check_cuda(const Tensor& t, const char* name) {
  TORCH_CHECK(t.is_cuda(), name, " must be a CUDA tensor");
}

// every tensor argument of an op must live on the device of its first input: a host or
// foreign-device pointer would otherwise reach the kernel and fault the context
inline void same_device(const Tensor&) {}
template <typename... Rest>
inline void same_device(const Tensor& ref, const std::optional<Tensor>& t, const Rest&... rest);
template <typename... Rest>
inline void same_device(const Tensor& ref, const Tensor& t, const Rest&... rest) {
  // (an empty tensor stands for "absent": the reference tests pass torch.Tensor() as "no bias")
  TORCH_CHECK(!t.defined() || t.numel() == 0 || (t.is_cuda() && t.device() == ref.device()),
              "all tensor arguments must be on the same CUDA device (got ", t.device(), " and ", ref.device(), ")");
  same_device(ref, rest...);
}
template <typename... Rest>
inline void same_device(const Tensor& ref, const std::optional<Tensor>& t, const Rest&... rest) {
  if (t.has_value()) same_device(ref, *t);
  same_device(ref, rest...);
}

inline const float* opt_f32(const std::optional<Tensor>& t, const char* name) {
  if (!t.has_value() || !t->defined()) return nullptr;
  TORCH_CHECK(t->scalar_type() == at::kFloat, name, " must be float32");
  check_cuda(*t, name);
  return t->data_ptr<float>();
}

// ------------------------------------------------------------------- _C ops

// ref: csrc/quantization/fp8/fp8_quant.cpp:842-881
void dynamic_per_token_scaled_fp8_quant(Tensor& out, const Tensor& input, Tensor& scales,
                                        const std::optional<Tensor>& scale_ub) {
  TORCH_CHECK(input.is_contiguous());
  TORCH_CHECK(out.is_contiguous());
  TORCH_CHECK(scales.scalar_type() == at::kFloat, "scales must be float32");
  check_cuda(input, "input");
  same_device(input, out, scales, scale_ub);
  const int64_t H = input.size(-1);
  const int64_t T = H ? input.numel() / H : 0;
  TORCH_CHECK(scales.numel() >= T && scales.is_contiguous(), "scales must hold one fp32 per token");
  const c10::cuda::CUDAGuard guard(input.device());
  check_rc(vxk_dynamic_per_token_scaled_fp8_quant(
               out.data_ptr(), input.data_ptr(), scales.data_ptr<float>(),
               opt_f32(scale_ub, "scale_ub"), T, H, in_dtype_of(input, "dynamic_per_token_scaled_fp8_quant"),
               q_dtype_of(out, "dynamic_per_token_scaled_fp8_quant", false), cur_stream()),
           "dynamic_per_token_scaled_fp8_quant");
}

// ref: fp8_quant.cpp:481-665 (scale-shape logic :502-623)
void static_scaled_fp8_quant(Tensor& out, const Tensor& input, const Tensor& scale,
                             std::optional<std::tuple<int64_t, int64_t>> opt_group_shape) {
  TORCH_CHECK(input.stride(-1) == 1, "last dimension of input must be contiguous");
  TORCH_CHECK(out.stride(-1) == 1, "last dimension of output must be contiguous");
  TORCH_CHECK(scale.scalar_type() == at::kFloat, "scale must be float32");
  check_cuda(input, "input");
  same_device(input, out, scale);
  const int64_t hidden_size = input.size(-1);
  const int64_t num_tokens = hidden_size ? input.numel() / hidden_size : 0;
  int64_t group_m, group_n, ss_i, ss_j;
  if (scale.dim() == 0 || scale.numel() == 1) {
    group_m = num_tokens; group_n = hidden_size; ss_i = 0; ss_j = 0;
  } else if (scale.dim() == 1) {
    TORCH_CHECK(opt_group_shape.has_value(),
                "a 1-D scale needs group_shape: (1, -1) = one scale per token, (-1, 1) = one per channel");
    const auto [ogm, ogn] = opt_group_shape.value();
    group_m = ogm == -1 ? num_tokens : ogm;
    group_n = ogn == -1 ? hidden_size : ogn;
    TORCH_CHECK(group_m > 0 && group_n > 0, "invalid group_shape");
    const int64_t em = num_tokens / group_m, en = hidden_size / group_n;
    TORCH_CHECK(scale.numel() == em * en, "scale has ", scale.numel(), " elements, group_shape (", ogm, ", ", ogn,
                ") over a [", num_tokens, ", ", hidden_size, "] input needs ", em * en);
    if (em == 1) { ss_i = 0; ss_j = scale.stride(0); }
    else if (en == 1) { ss_i = scale.stride(0); ss_j = 0; }
    else TORCH_CHECK(false, "a 1-D scale covers per-token or per-channel groups only; pass a 2-D scale for 2-D groups");
  } else if (scale.dim() == 2) {
    const int64_t s0 = scale.size(0), s1 = scale.size(1);
    TORCH_CHECK(s0 > 0 && num_tokens % s0 == 0, "rows of the input (", num_tokens, ") must be divisible by scale.size(0) = ", s0);
    TORCH_CHECK(s1 > 0 && hidden_size % s1 == 0, "columns of the input (", hidden_size, ") must be divisible by scale.size(1) = ", s1);
    const int64_t igm = num_tokens / s0, ign = hidden_size / s1;
    if (opt_group_shape.has_value()) {
      const auto [ogm, ogn] = opt_group_shape.value();
      group_m = ogm == -1 ? num_tokens : ogm;
      group_n = ogn == -1 ? hidden_size : ogn;
      TORCH_CHECK(group_m == igm && group_n == ign, "group_shape (", ogm, ", ", ogn, ") contradicts the [", s0, ", ", s1,
                  "] scale tensor, which implies groups of (", igm, ", ", ign, ")");
    } else {
      group_m = igm; group_n = ign;
    }
    ss_i = scale.stride(0); ss_j = scale.stride(1);
  } else {
    TORCH_CHECK(false, "scale must have 0, 1 or 2 dimensions, got ", scale.dim());
  }
  const int64_t in_rs = input.dim() >= 2 ? input.stride(-2) : hidden_size;
  const int64_t out_rs = out.dim() >= 2 ? out.stride(-2) : hidden_size;
  const c10::cuda::CUDAGuard guard(input.device());
  check_rc(vxk_static_scaled_fp8_quant(out.data_ptr(), input.data_ptr(), scale.data_ptr<float>(),
                                       num_tokens, hidden_size, in_rs, out_rs, group_m, group_n, ss_i,
                                       ss_j, in_dtype_of(input, "static_scaled_fp8_quant"),
                                       q_dtype_of(out, "static_scaled_fp8_quant", false), cur_stream()),
           "static_scaled_fp8_quant");
}

// ref: fp8_quant.cpp:667-722
void dynamic_scaled_fp8_quant(Tensor& out, const Tensor& input, Tensor& scale) {
  TORCH_CHECK(scale.scalar_type() == at::kFloat, "scale must be a float tensor");
  TORCH_CHECK(scale.numel() == 1, "scale must be a single-element tensor");
  TORCH_CHECK(input.stride(-1) == 1, "last dimension of input must be contiguous");
  TORCH_CHECK(out.stride(-1) == 1, "last dimension of output must be contiguous");
  check_cuda(input, "input");
  same_device(input, out, scale);
  const int64_t H = input.size(-1);
  const int64_t T = H ? input.numel() / H : 0;
  const int64_t in_rs = input.dim() >= 2 ? input.stride(-2) : H;
  const int64_t out_rs = out.dim() >= 2 ? out.stride(-2) : H;
  const c10::cuda::CUDAGuard guard(input.device());
  check_rc(vxk_dynamic_scaled_fp8_quant(out.data_ptr(), input.data_ptr(), scale.data_ptr<float>(), T, H,
                                        in_rs, out_rs, in_dtype_of(input, "dynamic_scaled_fp8_quant"),
                                        q_dtype_of(out, "dynamic_scaled_fp8_quant", false), cur_stream()),
           "dynamic_scaled_fp8_quant");
}

// ref: fp8_quant.cpp:724-840.  fp8_min / fp8_max are ignored like in the reference
// (the type's max is used); the two dummy flags are ignored too: the scale layout
// is read off output_s's strides (:763-765).
void per_token_group_fp8_quant(const Tensor& input, Tensor& output_q, Tensor& output_s, int64_t group_size,
                               double eps, double /*fp8_min*/, double /*fp8_max*/, bool scale_ue8m0,
                               bool /*dummy_is_scale_transposed*/, bool /*dummy_is_tma_aligned*/) {
  TORCH_CHECK(input.is_contiguous());
  TORCH_CHECK(output_q.is_contiguous());
  TORCH_CHECK(group_size > 0 && input.numel() % group_size == 0);
  TORCH_CHECK(output_s.dim() == 2);
  TORCH_CHECK(output_s.scalar_type() == at::kFloat, "output_s must be float32");
  check_cuda(input, "input");
  same_device(input, output_q, output_s);
  const int64_t T = output_s.size(0);
  if (T == 0 || input.numel() == 0) return;
  const int64_t H = input.numel() / T;
  TORCH_CHECK(H % group_size == 0 && output_s.size(1) == H / group_size,
              "output_s must have shape [num_tokens, hidden / group_size]");
  const c10::cuda::CUDAGuard guard(input.device());
  check_rc(vxk_per_token_group_fp8_quant(input.data_ptr(), output_q.data_ptr(), output_s.data_ptr<float>(),
                                         T, H, group_size, eps, scale_ue8m0 ? 1 : 0, output_s.stride(0),
                                         output_s.stride(1), in_dtype_of(input, "per_token_group_fp8_quant"),
                                         q_dtype_of(output_q, "per_token_group_fp8_quant", false),
                                         cur_stream()),
           "per_token_group_fp8_quant");
}

// ref: csrc/quantization/fp4/mxfp4_quant.cpp:24-126
// packed scale-factor buffer of an activation tensor [T, H] (include/vxk_b200.h "(f) next #2"); nullptr when absent
void* sf_packed_ptr(const std::optional<Tensor>& sf, int64_t rows, int64_t K, int operand, const Tensor& like) {
  if (!sf.has_value() || !sf->defined()) return nullptr;
  TORCH_CHECK(sf->scalar_type() == at::kByte && sf->is_contiguous(), "packed scale buffer must be contiguous uint8");
  TORCH_CHECK(sf->device() == like.device(), "packed scale buffer must be on the operands' device");
  TORCH_CHECK(sf->numel() >= vxk_mx_sf_packed_bytes(rows, K, /*kind=*/1, operand),
              "packed scale buffer too small: ", sf->numel(), " < ", vxk_mx_sf_packed_bytes(rows, K, 1, operand));
  TORCH_CHECK(reinterpret_cast<uintptr_t>(sf->data_ptr()) % 16 == 0, "packed scale buffer must be 16-byte aligned");
  return sf->data_ptr();
}

void per_token_group_quant_mxfp4_impl(const Tensor& input, Tensor& output_q, Tensor& output_s, int64_t group_size,
                                      double eps, const std::optional<Tensor>& sf_packed) {
  TORCH_CHECK(input.is_contiguous(), "input must be contiguous");
  TORCH_CHECK(output_q.is_contiguous(), "output_q must be contiguous");
  TORCH_CHECK(group_size == 32, "MXFP4 requires group_size == 32, got ", group_size);
  TORCH_CHECK(input.numel() % group_size == 0, "input numel must be divisible by group_size");
  TORCH_CHECK(output_s.dim() == 2, "output_s must be 2-D");
  TORCH_CHECK(output_q.scalar_type() == at::kByte, "output_q must be uint8");
  TORCH_CHECK(output_s.scalar_type() == at::kFloat, "output_s must be float32");
  if (input.numel() / group_size == 0) {
    TORCH_CHECK(output_q.numel() == 0, "output_q must be empty when input is empty");
    TORCH_CHECK(output_s.numel() == 0, "output_s must be empty when input is empty");
    return;
  }
  check_cuda(input, "input");
  same_device(input, output_q, output_s);
  const int64_t T = output_s.size(0);
  const int64_t H = input.numel() / T;
  TORCH_CHECK(H % group_size == 0 && output_s.size(1) == H / group_size,
              "output_s must have shape [num_tokens, hidden / 32]");
  const c10::cuda::CUDAGuard guard(input.device());
  check_rc(vxk_per_token_group_quant_mxfp4_sf(input.data_ptr(), output_q.data_ptr(), output_s.data_ptr<float>(),
                                              sf_packed_ptr(sf_packed, T, H, 0, input), T, H, group_size, eps,
                                              output_s.stride(0), output_s.stride(1),
                                              in_dtype_of(input, "per_token_group_quant_mxfp4"), cur_stream()),
           "per_token_group_quant_mxfp4");
}
void per_token_group_quant_mxfp4(const Tensor& input, Tensor& output_q, Tensor& output_s, int64_t group_size,
                                 double eps) {
  per_token_group_quant_mxfp4_impl(input, output_q, output_s, group_size, eps, std::nullopt);
}
void per_token_group_quant_mxfp4_sf(const Tensor& input, Tensor& output_q, Tensor& output_s, Tensor& sf_packed,
                                    int64_t group_size, double eps) {
  per_token_group_quant_mxfp4_impl(input, output_q, output_s, group_size, eps, sf_packed);
}

void check_residual(const std::optional<Tensor>& residual, const Tensor& input) {
  if (residual.has_value()) {
    TORCH_CHECK(residual->scalar_type() == input.scalar_type(), "residual and input must have the same dtype");
    TORCH_CHECK(residual->is_contiguous(), "residual must be contiguous");
    TORCH_CHECK(residual->numel() == input.numel(), "residual and input must have the same size");
  }
}

// ref: csrc/layernorm_quant.cpp:988-1045
void rms_norm_dynamic_per_token_quant(Tensor& out, const Tensor& input, const Tensor& weight, Tensor& scales,
                                      double epsilon, std::optional<Tensor> scale_ub,
                                      std::optional<Tensor> residual) {
  TORCH_CHECK(input.is_contiguous(), "input must be contiguous");
  TORCH_CHECK(out.is_contiguous(), "out must be contiguous");
  TORCH_CHECK(weight.dtype() == input.dtype(), "weight and input must have the same dtype");
  TORCH_CHECK(scales.dtype() == at::kFloat, "scales must be float32");
  TORCH_CHECK(out.dtype() == at::kFloat8_e4m3fn || out.dtype() == at::kChar,
              "output must be float8_e4m3fn or int8");
  if (scale_ub.has_value())
    TORCH_CHECK(out.dtype() == at::kFloat8_e4m3fn, "scale_ub is only supported for FP8 output");
  check_residual(residual, input);
  check_cuda(input, "input");
  same_device(input, out, weight, scales, scale_ub, residual);
  const int64_t H = input.size(-1);
  const int64_t T = H ? input.numel() / H : 0;
  TORCH_CHECK(weight.numel() == H && weight.is_contiguous(), "weight must be a contiguous [hidden] tensor");
  TORCH_CHECK(scales.numel() >= T && scales.is_contiguous(), "scales must hold one fp32 per token");
  const c10::cuda::CUDAGuard guard(input.device());
  check_rc(vxk_rms_norm_dynamic_per_token_quant(
               out.data_ptr(), input.data_ptr(), weight.data_ptr(), scales.data_ptr<float>(), epsilon,
               opt_f32(scale_ub, "scale_ub"), residual.has_value() ? residual->data_ptr() : nullptr, T, H,
               in_dtype_of(input, "rms_norm_dynamic_per_token_quant"),
               q_dtype_of(out, "rms_norm_dynamic_per_token_quant", true), cur_stream()),
           "rms_norm_dynamic_per_token_quant");
}

// ref: layernorm_quant.cpp:1047-1114.  scale_ub is accepted and ignored, as in the reference (:1088-1097).
void rms_norm_per_block_quant(Tensor& out, const Tensor& input, const Tensor& weight, Tensor& scales,
                              double epsilon, std::optional<Tensor> /*scale_ub*/, std::optional<Tensor> residual,
                              int64_t group_size, bool is_scale_transposed, bool scale_ue8m0) {
  TORCH_CHECK(input.stride(-1) == 1, "input must be contiguous in the last dimension");
  TORCH_CHECK(out.is_contiguous(), "out must be contiguous");
  TORCH_CHECK(weight.dtype() == input.dtype(), "weight and input must have the same dtype");
  TORCH_CHECK(scales.dtype() == at::kFloat, "scales must be float32");
  TORCH_CHECK(out.dtype() == at::kFloat8_e4m3fn || out.dtype() == at::kChar,
              "output must be float8_e4m3fn or int8");
  TORCH_CHECK(group_size > 0 && input.size(-1) % group_size == 0, "hidden_size must be divisible by group_size");
  check_residual(residual, input);
  if (scale_ue8m0)
    TORCH_CHECK(out.dtype() == at::kFloat8_e4m3fn,
                "scale_ue8m0 (MX FP8) is only supported for float8_e4m3fn output");
  check_cuda(input, "input");
  same_device(input, out, weight, scales, residual);
  const int64_t H = input.size(-1);
  const int64_t T = H ? input.numel() / H : 0;
  const int64_t G = H / group_size;
  TORCH_CHECK(input.dim() <= 2 || input.is_contiguous(), "only 2-D inputs may be row-strided");
  const int64_t in_rs = input.dim() >= 2 ? input.stride(-2) : H;
  TORCH_CHECK(weight.numel() == H && weight.is_contiguous(), "weight must be a contiguous [hidden] tensor");
  TORCH_CHECK(scales.numel() >= T * G, "scales too small");
  // reference addressing: transposed -> tok + g * T, else tok * G + g  (:751-753)
  const int64_t ss0 = is_scale_transposed ? 1 : G;
  const int64_t ss1 = is_scale_transposed ? T : 1;
  const c10::cuda::CUDAGuard guard(input.device());
  check_rc(vxk_rms_norm_per_block_quant(out.data_ptr(), input.data_ptr(), weight.data_ptr(),
                                        scales.data_ptr<float>(), epsilon,
                                        residual.has_value() ? residual->data_ptr() : nullptr, T, H, in_rs,
                                        group_size, ss0, ss1, scale_ue8m0 ? 1 : 0,
                                        in_dtype_of(input, "rms_norm_per_block_quant"),
                                        q_dtype_of(out, "rms_norm_per_block_quant", true), cur_stream()),
           "rms_norm_per_block_quant");
}

// ref: layernorm_quant.cpp:1172-1219
void rms_norm_mxfp4_quant_impl(Tensor& out, const Tensor& input, const Tensor& weight, Tensor& scales,
                               double epsilon, std::optional<Tensor> residual, int64_t group_size,
                               const std::optional<Tensor>& sf_packed) {
  TORCH_CHECK(input.is_contiguous(), "input must be contiguous");
  TORCH_CHECK(out.is_contiguous(), "out must be contiguous");
  TORCH_CHECK(weight.dtype() == input.dtype(), "weight and input must have the same dtype");
  TORCH_CHECK(scales.dtype() == at::kFloat, "scales must be float32");
  TORCH_CHECK(out.scalar_type() == at::ScalarType::Byte, "output must be uint8 (packed FP4)");
  TORCH_CHECK(group_size == 32, "MXFP4 requires group_size == 32, got ", group_size);
  TORCH_CHECK(input.size(-1) % group_size == 0, "hidden_size must be divisible by group_size");
  TORCH_CHECK(input.size(-1) % 2 == 0, "hidden_size must be even for FP4 packing");
  TORCH_CHECK(scales.dim() == 2, "scales must be 2-D");
  check_residual(residual, input);
  check_cuda(input, "input");
  same_device(input, out, weight, scales, residual);
  const int64_t H = input.size(-1);
  const int64_t T = H ? input.numel() / H : 0;
  TORCH_CHECK(scales.size(0) == T && scales.size(1) == H / group_size, "scales must be [num_tokens, hidden/32]");
  TORCH_CHECK(weight.numel() == H && weight.is_contiguous(), "weight must be a contiguous [hidden] tensor");
  const c10::cuda::CUDAGuard guard(input.device());
  check_rc(vxk_rms_norm_mxfp4_quant_sf(out.data_ptr(), input.data_ptr(), weight.data_ptr(), scales.data_ptr<float>(),
                                       sf_packed_ptr(sf_packed, T, H, 0, input), epsilon,
                                       residual.has_value() ? residual->data_ptr() : nullptr, T, H, group_size,
                                       scales.stride(0), scales.stride(1),
                                       in_dtype_of(input, "rms_norm_mxfp4_quant"), cur_stream()),
           "rms_norm_mxfp4_quant");
}
void rms_norm_mxfp4_quant(Tensor& out, const Tensor& input, const Tensor& weight, Tensor& scales,
                          double epsilon, std::optional<Tensor> residual, int64_t group_size) {
  rms_norm_mxfp4_quant_impl(out, input, weight, scales, epsilon, residual, group_size, std::nullopt);
}
void rms_norm_mxfp4_quant_sf(Tensor& out, const Tensor& input, const Tensor& weight, Tensor& scales, Tensor& sf_packed,
                             double epsilon, std::optional<Tensor> residual, int64_t group_size) {
  rms_norm_mxfp4_quant_impl(out, input, weight, scales, epsilon, residual, group_size, sf_packed);
}

void rms_norm_static_impl(Tensor& out, const Tensor& input, Tensor* residual, const Tensor& weight,
                          const Tensor& scale, double epsilon, const char* op) {
  TORCH_CHECK(out.is_contiguous(), "out must be contiguous");
  TORCH_CHECK(weight.scalar_type() == input.scalar_type(), "weight and input must have the same dtype");
  TORCH_CHECK(input.stride(-1) == 1, "input must be contiguous in the last dimension");
  TORCH_CHECK(scale.scalar_type() == at::kFloat && scale.numel() == 1, "scale must be one float32");
  check_cuda(input, "input");
  same_device(input, out, weight, scale); if (residual) same_device(input, *residual);
  const int64_t H = input.size(-1);
  const int64_t T = H ? input.numel() / H : 0;
  TORCH_CHECK(input.dim() <= 2 || input.is_contiguous(), "only 2-D inputs may be row-strided");
  const int64_t in_rs = input.dim() >= 2 ? input.stride(-2) : H;
  TORCH_CHECK(weight.numel() == H && weight.is_contiguous(), "weight must be a contiguous [hidden] tensor");
  const c10::cuda::CUDAGuard guard(input.device());
  check_rc(vxk_rms_norm_static_fp8_quant(out.data_ptr(), input.data_ptr(), weight.data_ptr(),
                                         scale.data_ptr<float>(), epsilon,
                                         residual ? residual->data_ptr() : nullptr, T, H, in_rs,
                                         in_dtype_of(input, op), q_dtype_of(out, op, false), cur_stream()),
           op);
}

// ref: layernorm_quant.cpp:1116-1136
void rms_norm_static_fp8_quant(Tensor& out, const Tensor& input, const Tensor& weight, const Tensor& scale,
                               double epsilon) {
  rms_norm_static_impl(out, input, nullptr, weight, scale, epsilon, "rms_norm_static_fp8_quant");
}

// ref: layernorm_quant.cpp:1138-1170
void fused_add_rms_norm_static_fp8_quant(Tensor& out, const Tensor& input, Tensor& residual, const Tensor& weight,
                                         const Tensor& scale, double epsilon) {
  TORCH_CHECK(residual.is_contiguous(), "residual must be contiguous");
  TORCH_CHECK(residual.scalar_type() == input.scalar_type(), "residual and input must have the same dtype");
  TORCH_CHECK(residual.numel() == input.numel(), "residual and input must have the same size");
  rms_norm_static_impl(out, input, &residual, weight, scale, epsilon, "fused_add_rms_norm_static_fp8_quant");
}

// ref: csrc/activation.cpp:552-566
void silu_and_mul_quant(Tensor& out, const Tensor& input, const Tensor& scale) {
  TORCH_CHECK(out.dtype() == at::kFloat8_e4m3fn || out.dtype() == at::kFloat8_e5m2);
  TORCH_CHECK(input.dtype() == at::kHalf || input.dtype() == at::kBFloat16);
  TORCH_CHECK(input.size(-1) % 2 == 0);
  TORCH_CHECK(input.is_contiguous() && out.is_contiguous(), "input and out must be contiguous");
  TORCH_CHECK(scale.scalar_type() == at::kFloat && scale.numel() >= 1, "scale must be float32");
  const int64_t d = input.size(-1) / 2;
  const int64_t T = input.size(-1) ? input.numel() / input.size(-1) : 0;
  if (T == 0) return;
  check_cuda(input, "input");
  same_device(input, out, scale);
  const c10::cuda::CUDAGuard guard(input.device());
  check_rc(vxk_silu_and_mul_quant(out.data_ptr(), input.data_ptr(), scale.data_ptr<float>(), T, d,
                                  in_dtype_of(input, "silu_and_mul_quant"),
                                  q_dtype_of(out, "silu_and_mul_quant", false), cur_stream()),
           "silu_and_mul_quant");
}

// ref: csrc/quantization/fused_kernels/fused_silu_mul_block_quant.cpp:141-239
void silu_and_mul_per_block_quant(Tensor& out, const Tensor& input, Tensor& scales, int64_t group_size,
                                  std::optional<Tensor> scale_ub, bool is_scale_transposed, bool scale_ue8m0) {
  TORCH_CHECK(out.dtype() == at::kFloat8_e4m3fn || out.dtype() == at::kFloat8_e5m2 ||
                  out.dtype() == at::kChar,
              "out must be float8_e4m3fn, float8_e5m2, or int8");
  TORCH_CHECK(input.dtype() == at::kHalf || input.dtype() == at::kBFloat16,
              "input must be float16 or bfloat16");
  TORCH_CHECK(out.is_contiguous(), "out must be contiguous");
  TORCH_CHECK(input.is_contiguous(), "input must be contiguous");
  TORCH_CHECK(scales.dtype() == at::kFloat, "scales must be float32");
  TORCH_CHECK(group_size == 32 || group_size == 64 || group_size == 128,
              "group_size must be 32, 64, or 128, got ", group_size);
  if (scale_ub.has_value())
    TORCH_CHECK(out.dtype() == at::kFloat8_e4m3fn || out.dtype() == at::kFloat8_e5m2,
                "scale_ub is only supported for FP8 output");
  if (scale_ue8m0)
    TORCH_CHECK(out.dtype() != at::kChar, "scale_ue8m0 is only supported for FP8 output");
  TORCH_CHECK(input.dim() == 2, "input must be 2-dimensional");
  TORCH_CHECK(out.dim() == 2, "out must be 2-dimensional");
  TORCH_CHECK(scales.dim() == 2, "scales must be 2-dimensional");
  const int64_t T = input.size(0);
  const int64_t H = out.size(1);
  const int64_t G = H / group_size;
  TORCH_CHECK(input.size(-1) == H * 2, "input last dim must be 2 * hidden_size");
  TORCH_CHECK(H % group_size == 0, "hidden_size must be divisible by group_size");
  TORCH_CHECK(out.size(0) == T, "out.size(0) must equal num_tokens (input.size(0))");
  TORCH_CHECK(scales.size(0) == T && scales.size(1) == G, "scales must have logical shape [num_tokens, num_groups]");
  if (T == 0) return;
  check_cuda(input, "input");
  same_device(input, out, scales, scale_ub);
  const int64_t ss0 = is_scale_transposed ? 1 : G;
  const int64_t ss1 = is_scale_transposed ? T : 1;
  const c10::cuda::CUDAGuard guard(input.device());
  check_rc(vxk_silu_and_mul_per_block_quant(out.data_ptr(), input.data_ptr(), scales.data_ptr<float>(), T, H,
                                            group_size, opt_f32(scale_ub, "scale_ub"), ss0, ss1,
                                            scale_ue8m0 ? 1 : 0, in_dtype_of(input, "silu_and_mul_per_block_quant"),
                                            q_dtype_of(out, "silu_and_mul_per_block_quant", true), cur_stream()),
           "silu_and_mul_per_block_quant");
}

// ref: fused_silu_mul_mxfp4_quant.cpp:134-209
void silu_and_mul_mxfp4_quant_impl(Tensor& out, const Tensor& input, Tensor& scales, int64_t group_size, double eps,
                                   const std::optional<Tensor>& sf_packed) {
  TORCH_CHECK(input.dtype() == at::kHalf || input.dtype() == at::kBFloat16,
              "input must be float16 or bfloat16");
  TORCH_CHECK(out.scalar_type() == at::ScalarType::Byte, "out must be uint8 (packed FP4)");
  TORCH_CHECK(out.is_contiguous(), "out must be contiguous");
  TORCH_CHECK(input.is_contiguous(), "input must be contiguous");
  TORCH_CHECK(scales.dtype() == at::kFloat, "scales must be float32");
  TORCH_CHECK(group_size == 32, "MXFP4 requires group_size == 32, got ", group_size);
  TORCH_CHECK(input.dim() == 2, "input must be 2-dimensional");
  TORCH_CHECK(out.dim() == 2, "out must be 2-dimensional");
  TORCH_CHECK(scales.dim() == 2, "scales must be 2-dimensional");
  const int64_t T = input.size(0);
  const int64_t H = input.size(-1) / 2;
  const int64_t G = H / group_size;
  TORCH_CHECK(input.size(-1) == H * 2, "input last dim must be 2 * hidden_size");
  TORCH_CHECK(H % group_size == 0, "hidden_size must be divisible by group_size");
  TORCH_CHECK(out.size(0) == T && out.size(1) == H / 2, "out must have shape [num_tokens, hidden_size / 2]");
  TORCH_CHECK(scales.size(0) == T && scales.size(1) == G, "scales must have logical shape [num_tokens, num_groups]");
  if (T == 0) return;
  check_cuda(input, "input");
  same_device(input, out, scales);
  const c10::cuda::CUDAGuard guard(input.device());
  check_rc(vxk_silu_and_mul_mxfp4_quant_sf(out.data_ptr(), input.data_ptr(), scales.data_ptr<float>(),
                                           sf_packed_ptr(sf_packed, T, H, 0, input), T, H, group_size, eps,
                                           scales.stride(0), scales.stride(1),
                                           in_dtype_of(input, "silu_and_mul_mxfp4_quant"), cur_stream()),
           "silu_and_mul_mxfp4_quant");
}
void silu_and_mul_mxfp4_quant(Tensor& out, const Tensor& input, Tensor& scales, int64_t group_size, double eps) {
  silu_and_mul_mxfp4_quant_impl(out, input, scales, group_size, eps, std::nullopt);
}
void silu_and_mul_mxfp4_quant_sf(Tensor& out, const Tensor& input, Tensor& scales, Tensor& sf_packed,
                                 int64_t group_size, double eps) {
  silu_and_mul_mxfp4_quant_impl(out, input, scales, group_size, eps, sf_packed);
}

// --------------------------------------------------------------- _xpu_C ops

// ref: csrc/xpu/onednn/onednn_matmul.cpp:17-52
Tensor check_and_create_output_tensor(const Tensor& A, const Tensor& B, std::optional<c10::ScalarType> out_dtype) {
  TORCH_CHECK(A.dim() == 2 || A.dim() == 3, "Matmul only support 2D and 3D inputs!\n");
  TORCH_CHECK(B.dim() == 2, "Matmul only support 2D weights!\n");
  std::vector<int64_t> result_shape;
  if (A.dim() == 2) result_shape = {A.size(0), B.size(1)};
  else result_shape = {A.size(0), A.size(1), B.size(1)};
  const auto k = A.size(A.dim() - 1);
  const auto n = result_shape.back();
  auto res_stride = A.strides().vec();
  for (size_t i = 0; i + 1 < res_stride.size(); i++) res_stride[i] = res_stride[i] / k * n;
  const auto out_dtype_ = out_dtype.value_or(at::kHalf);  // reference default (:49)
  return at::empty_strided(result_shape, res_stride, A.options().dtype(out_dtype_));
}

struct ARows {  // A viewed as M rows of K contiguous elements with ONE row stride
  int64_t M, lda, ldd;
};
ARows flatten_rows(const Tensor& A, const Tensor& D) {
  TORCH_CHECK(A.stride(-1) == 1, "A must be contiguous in its last dimension");
  if (A.dim() == 2) return {A.size(0), A.stride(0), D.stride(0)};
  // [b, m, k]: rows must be expressible with one stride (contiguous [b,m,k], or b == 1 / m == 1)
  const int64_t b = A.size(0), m = A.size(1);
  if (b == 1) return {m, A.stride(1), D.stride(1)};
  if (m == 1) return {b, A.stride(0), D.stride(0)};
  TORCH_CHECK(A.stride(0) == m * A.stride(1),
              "3-D A must have uniformly strided rows ([b,m,k] contiguous); got strides ", A.strides());
  return {b * m, A.stride(1), D.stride(1)};
}

int bias_dtype_of(const Tensor& b) {
  switch (b.scalar_type()) {
    case at::kFloat: return VXK_F32;
    case at::kHalf: return VXK_F16;
    case at::kBFloat16: return VXK_BF16;
    default: TORCH_CHECK(false, "Unsupported data type for bias in matmul: ", b.scalar_type());
  }
}

// ref: onednn_matmul.cpp:54-79, fp8_gemm_w8a8.h:12-184
Tensor fp8_gemm(const Tensor& A, const Tensor& B, std::optional<c10::ScalarType> out_dtype,
                const std::optional<Tensor>& A_scale_, const std::optional<Tensor>& B_scale_,
                const std::optional<Tensor>& bias_) {
  const c10::cuda::CUDAGuard guard(A.device());
  Tensor result = check_and_create_output_tensor(A, B, out_dtype);
  const auto a_st = A.scalar_type(), b_st = B.scalar_type();
  TORCH_CHECK((a_st == at::kFloat8_e5m2 || a_st == at::kFloat8_e4m3fn) && a_st == b_st,
              "input and weight must be f8_e5m2 or f8_e4m3fn for fp8 matmul");
  TORCH_CHECK(result.scalar_type() == at::kHalf || result.scalar_type() == at::kBFloat16,
              "output must be float16 or bfloat16 for fp8 matmul");
  TORCH_CHECK(B.stride(0) == 1 || B.stride(1) == 1, "B must be contiguous along K or along N");
  const ARows r = flatten_rows(A, result);
  const int64_t K = A.size(-1), N = B.size(1);
  TORCH_CHECK(B.size(0) == K, "A and B have mismatched K: ", K, " vs ", B.size(0));
  if (r.M == 0 || N == 0) return result;

  // missing scales default to 1.0 (onednn_matmul.cpp:75-76)
  Tensor A_scale = A_scale_.has_value() && A_scale_->defined() && A_scale_->numel() > 0
                       ? *A_scale_ : at::ones({1}, A.options().dtype(at::kFloat));
  Tensor B_scale = B_scale_.has_value() && B_scale_->defined() && B_scale_->numel() > 0
                       ? *B_scale_ : at::ones({1}, A.options().dtype(at::kFloat));
  check_cuda(A, "A");
  same_device(A, B, A_scale, B_scale, bias_);

  const bool is_block = A_scale.dim() == 2 && A_scale.size(1) > 1;  // fp8_gemm_w8a8.h:28
  int a_mode, b_mode, scale_dtype = VXK_S_F32;
  int64_t gk = 0, gn = 0, sa0 = 0, sa1 = 0, sb0 = 0, sb1 = 0;
  if (is_block) {
    TORCH_CHECK(B_scale.dim() == 2, "block-quantised weights need a 2-D [K/gk, N/gn] scale");
    TORCH_CHECK(A_scale.size(1) == B_scale.size(0), "Mismatch group size in input and weight.",
                A_scale.size(1), " vs ", B_scale.size(0));
    TORCH_CHECK(A_scale.size(0) == r.M, "A_scale must have one row per row of A");
    TORCH_CHECK(K % B_scale.size(0) == 0 && N % B_scale.size(1) == 0,
                "K and N must be divisible by the scale grid");
    gk = K / B_scale.size(0);
    gn = N / B_scale.size(1);
    a_mode = b_mode = 2;
    sa0 = A_scale.stride(0); sa1 = A_scale.stride(1);
    sb0 = B_scale.stride(0); sb1 = B_scale.stride(1);
    TORCH_CHECK(A_scale.scalar_type() == B_scale.scalar_type(), "A_scale / B_scale dtype mismatch");
    if (A_scale.scalar_type() == at::kFloat8_e8m0fnu) scale_dtype = VXK_S_E8M0;
    else TORCH_CHECK(A_scale.scalar_type() == at::kFloat, "block scales must be float32 or float8_e8m0fnu");
  } else {
    TORCH_CHECK(A_scale.scalar_type() == at::kFloat && B_scale.scalar_type() == at::kFloat,
                "per-tensor / per-token / per-channel scales must be float32");
    a_mode = A_scale.numel() == 1 ? 0 : 1;
    b_mode = B_scale.numel() == 1 ? 0 : 1;
    if (a_mode == 1) TORCH_CHECK(A_scale.numel() == r.M && A_scale.is_contiguous(), "A_scale must be [M] / [M,1]");
    if (b_mode == 1) TORCH_CHECK(B_scale.numel() == N && B_scale.is_contiguous(), "B_scale must be [N] / [N,1]");
  }
  const void* bias = nullptr;
  int bias_dtype = VXK_F32;
  if (bias_.has_value() && bias_->defined() && bias_->numel() > 0) {  // empty tensor == no bias (onednn_ext.h:317-318)
    TORCH_CHECK(bias_->numel() == N && bias_->is_contiguous(), "only a contiguous [N] bias is supported, got ",
                bias_->sizes());
    check_cuda(*bias_, "bias");
    bias = bias_->data_ptr();
    bias_dtype = bias_dtype_of(*bias_);
  }
  // per-call scratch, like the reference's scratchpad tensor (fp8_gemm_w8a8.h:176-179)
  const int64_t ws_bytes = vxk_fp8_gemm_workspace_bytes(r.M, N, K, a_mode, scale_dtype);
  Tensor ws;
  if (ws_bytes > 0) ws = at::empty({ws_bytes}, A.options().dtype(at::kByte));
  check_rc(vxk_fp8_gemm(result.data_ptr(), A.data_ptr(), B.data_ptr(), A_scale.data_ptr(), B_scale.data_ptr(),
                        bias, r.M, N, K, r.lda, B.stride(0), B.stride(1), r.ldd, a_mode, b_mode, gk, gn, sa0, sa1,
                        sb0, sb1, scale_dtype, a_st == at::kFloat8_e4m3fn ? VXK_Q_E4M3 : VXK_Q_E5M2,
                        result.scalar_type() == at::kBFloat16 ? VXK_BF16 : VXK_F16, bias_dtype,
                        ws_bytes > 0 ? ws.data_ptr() : nullptr, ws_bytes, cur_stream()),
           "fp8_gemm");
  return result;
}

// ref: onednn_matmul.cpp:101-123, fp4_gemm_w4a4.h:12-167
Tensor fp4_gemm_impl(const Tensor& A, const Tensor& B, const Tensor& A_scale, const Tensor& B_scale,
                     std::optional<c10::ScalarType> out_dtype, const std::optional<Tensor>& bias_,
                     const std::optional<Tensor>& A_sf_packed, const std::optional<Tensor>& B_sf_packed) {
  const c10::cuda::CUDAGuard guard(A.device());
  Tensor result = check_and_create_output_tensor(A, B, out_dtype);
  const auto a_st = A.scalar_type(), b_st = B.scalar_type();
  TORCH_CHECK(a_st == at::ScalarType::Float4_e2m1fn_x2 && a_st == b_st,
              "input and weight must be f4_e2m1x2 or f4_e2m1x2 for fp4 matmul");
  TORCH_CHECK(result.scalar_type() == at::kHalf || result.scalar_type() == at::kBFloat16,
              "output must be float16 or bfloat16 for fp4 matmul");
  TORCH_CHECK(A_scale.scalar_type() == at::kFloat8_e8m0fnu, "Unsupported scale type for fp4 matmul: ",
              A_scale.scalar_type());
  TORCH_CHECK(B_scale.scalar_type() == at::kFloat8_e8m0fnu, "Mismatched scale data types in mxfp4 matmul: ",
              A_scale.scalar_type(), " vs ", B_scale.scalar_type());
  TORCH_CHECK(B.stride(0) == 1, "FP4 weights must be K-major: pass W[N, K/2].transpose(0, 1)");
  check_cuda(A, "A");
  same_device(A, B, A_scale, B_scale, bias_);
  const ARows r = flatten_rows(A, result);
  const int64_t K = 2 * A.size(-1), N = B.size(1);
  TORCH_CHECK(B.size(0) * 2 == K, "A and B have mismatched K");
  TORCH_CHECK(K % 32 == 0, "K must be a multiple of the MX block (32)");
  TORCH_CHECK(A_scale.dim() == 2 && A_scale.size(0) == r.M && A_scale.size(1) == K / 32,
              "A_scale must be [M, K/32]");
  TORCH_CHECK(B_scale.dim() == 2 && B_scale.size(0) == N && B_scale.size(1) == K / 32,
              "B_scale must be [N, K/32] (tests/test_fp4_gemm_onednn.py:65-71 convention)");
  if (r.M == 0 || N == 0) return result;
  const void* bias = nullptr;
  int bias_dtype = VXK_F32;
  if (bias_.has_value() && bias_->defined() && bias_->numel() > 0) {
    TORCH_CHECK(bias_->numel() == N && bias_->is_contiguous(), "only a contiguous [N] bias is supported");
    bias = bias_->data_ptr();
    bias_dtype = bias_dtype_of(*bias_);
  }
  const void* a_packed = sf_packed_ptr(A_sf_packed, r.M, K, 0, A);
  const void* b_packed = sf_packed_ptr(B_sf_packed, N, K, 1, A);
  // per-call scratch, like the reference's scratchpad tensor (fp4_gemm_w4a4.h:159-162); none when both scale
  // grids arrive in the tensor core's layout
  const int64_t ws_bytes = (a_packed && b_packed) ? 0 : vxk_fp4_gemm_workspace_bytes(r.M, N, K);
  Tensor ws;
  if (ws_bytes > 0) ws = at::empty({ws_bytes}, A.options().dtype(at::kByte));
  check_rc(vxk_fp4_gemm_sf(result.data_ptr(), A.data_ptr(), B.data_ptr(), A_scale.data_ptr(), B_scale.data_ptr(),
                           a_packed, b_packed, bias, r.M, N, K, r.lda, B.stride(1), r.ldd, A_scale.stride(0),
                           A_scale.stride(1), B_scale.stride(0), B_scale.stride(1),
                           result.scalar_type() == at::kBFloat16 ? VXK_BF16 : VXK_F16, bias_dtype,
                           ws_bytes > 0 ? ws.data_ptr() : nullptr, ws_bytes, cur_stream()),
           "fp4_gemm");
  return result;
}
Tensor fp4_gemm(const Tensor& A, const Tensor& B, const Tensor& A_scale, const Tensor& B_scale,
                std::optional<c10::ScalarType> out_dtype, const std::optional<Tensor>& bias_) {
  return fp4_gemm_impl(A, B, A_scale, B_scale, out_dtype, bias_, std::nullopt, std::nullopt);
}
// extension (no reference counterpart): scale grids also handed over in tcgen05.cp order
Tensor fp4_gemm_sf(const Tensor& A, const Tensor& B, const Tensor& A_scale, const Tensor& B_scale,
                   std::optional<c10::ScalarType> out_dtype, const std::optional<Tensor>& bias_,
                   const std::optional<Tensor>& A_sf_packed, const std::optional<Tensor>& B_sf_packed) {
  return fp4_gemm_impl(A, B, A_scale, B_scale, out_dtype, bias_, A_sf_packed, B_sf_packed);
}
int64_t mx_sf_packed_bytes(int64_t rows, int64_t K, bool fp4, bool is_weight) {
  const int64_t n = vxk_mx_sf_packed_bytes(rows, K, fp4 ? 1 : 0, is_weight ? 1 : 0);
  TORCH_CHECK(n >= 0, "mx_sf_packed_bytes: bad arguments");
  return n;
}
// scale: [rows, K/32] float8_e8m0fnu (any strides) -> packed uint8 buffer, every byte written
Tensor mx_sf_pack(const Tensor& scale, bool fp4, bool is_weight) {
  TORCH_CHECK(scale.dim() == 2 && scale.scalar_type() == at::kFloat8_e8m0fnu, "scale must be a 2-D float8_e8m0fnu tensor");
  check_cuda(scale, "scale");
  const c10::cuda::CUDAGuard guard(scale.device());
  const int64_t rows = scale.size(0), K = scale.size(1) * 32;
  Tensor out = at::empty({mx_sf_packed_bytes(rows, K, fp4, is_weight)}, scale.options().dtype(at::kByte));
  check_rc(vxk_mx_sf_pack(out.data_ptr(), scale.data_ptr(), rows, K, scale.stride(0), scale.stride(1), fp4 ? 1 : 0,
                          is_weight ? 1 : 0, cur_stream()),
           "mx_sf_pack");
  return out;
}

// ref: onednn_matmul.cpp:81-99, fp8_gemm_w8a16.h:13-145
Tensor fp8_gemm_w8a16(const Tensor& A, const Tensor& B, const std::optional<Tensor>& B_scale_,
                      const std::optional<Tensor>& bias_) {
  const c10::cuda::CUDAGuard guard(A.device());
  TORCH_CHECK(A.scalar_type() == at::kHalf || A.scalar_type() == at::kBFloat16,
              "Unsupported data type for fp8 matmul: ", A.scalar_type());      // fp8_gemm_w8a16.h:51-53
  Tensor result = check_and_create_output_tensor(A, B, A.scalar_type());
  const auto b_st = B.scalar_type();
  TORCH_CHECK(b_st == at::kFloat8_e5m2 || b_st == at::kFloat8_e4m3fn,
              "weight must be f8_e5m2 or f8_e4m3fn for fp8 matmul");
  check_cuda(A, "A");
  same_device(A, B, B_scale_, bias_);
  const ARows r = flatten_rows(A, result);
  const int64_t K = A.size(-1), N = B.size(1);
  TORCH_CHECK(B.size(0) == K, "A and B have mismatched K: ", K, " vs ", B.size(0));
  if (r.M == 0 || N == 0) return result;

  // missing scale defaults to 1.0 (onednn_matmul.cpp:94-96); any float dtype is taken, the kernels read fp32
  Tensor B_scale = B_scale_.has_value() && B_scale_->defined() && B_scale_->numel() > 0
                       ? *B_scale_ : at::ones({1}, A.options().dtype(at::kFloat));
  check_cuda(B_scale, "B_scale");
  if (B_scale.scalar_type() != at::kFloat) {
    TORCH_CHECK(B_scale.scalar_type() == at::kHalf || B_scale.scalar_type() == at::kBFloat16,
                "weight scale must be float32, float16 or bfloat16");
    B_scale = B_scale.to(at::kFloat);
  }
  int b_mode;
  int64_t gk = 0, gn = 0, sb0 = 0, sb1 = 0;
  const bool is_block = B_scale.dim() == 2 && B_scale.size(0) > 1;   // fp8_gemm_w8a16.h:29-35
  if (is_block) {
    TORCH_CHECK(K % B_scale.size(0) == 0 && N % B_scale.size(1) == 0, "K and N must be divisible by the scale grid");
    gk = K / B_scale.size(0);
    gn = N / B_scale.size(1);
    sb0 = B_scale.stride(0); sb1 = B_scale.stride(1);
    b_mode = 2;
  } else if (B_scale.numel() == 1) {
    b_mode = 0;
  } else {
    TORCH_CHECK(B_scale.numel() == N, "per-channel weight scale must have N elements ([N] or [1, N]), got ",
                B_scale.sizes());
    B_scale = B_scale.contiguous();
    b_mode = 1;
  }
  const void* bias = nullptr;
  int bias_dtype = VXK_F32;
  if (bias_.has_value() && bias_->defined() && bias_->numel() > 0) {
    TORCH_CHECK(bias_->numel() == N && bias_->is_contiguous(), "only a contiguous [N] bias is supported, got ",
                bias_->sizes());
    check_cuda(*bias_, "bias");
    bias = bias_->data_ptr();
    bias_dtype = bias_dtype_of(*bias_);
  }
  check_rc(vxk_fp8_gemm_w8a16(result.data_ptr(), A.data_ptr(), B.data_ptr(), B_scale.data_ptr<float>(), bias, r.M, N,
                              K, r.lda, B.stride(0), B.stride(1), r.ldd, b_mode, gk, gn, sb0, sb1,
                              A.scalar_type() == at::kBFloat16 ? VXK_BF16 : VXK_F16,
                              b_st == at::kFloat8_e4m3fn ? VXK_Q_E4M3 : VXK_Q_E5M2, bias_dtype, 0, cur_stream()),
           "fp8_gemm_w8a16");
  return result;
}

// ref: cutlass_grouped_gemm_interface  csrc/xpu/grouped_gemm/grouped_gemm_interface.cpp:12-59,
//      argument checks csrc/xpu/grouped_gemm/xe_2/grouped_gemm_xe2_interface.hpp:160-246,303-311,339-341
Tensor cutlass_grouped_gemm_interface(Tensor ptr_A, Tensor ptr_B, const std::optional<Tensor>& ptr_scales,
                                      const std::optional<Tensor>& ptr_bias, Tensor ptr_D, Tensor rows_per_expert,
                                      int64_t N, int64_t K, int64_t num_experts) {
  check_cuda(ptr_A, "ptr_A");
  same_device(ptr_A, ptr_B, ptr_scales, ptr_bias);
  TORCH_CHECK(ptr_D.device() == ptr_A.device() && rows_per_expert.device() == ptr_A.device(),
              "ptr_D and rows_per_expert must be on ptr_A's device");
  const c10::cuda::CUDAGuard guard(ptr_A.device());
  const auto a_st = ptr_A.scalar_type(), b_st = ptr_B.scalar_type();
  TORCH_CHECK(a_st == at::kBFloat16 || a_st == at::kHalf, "activations must be bfloat16 or float16, got ", a_st);
  TORCH_CHECK(ptr_D.scalar_type() == a_st, "ptr_D must have the activation dtype");
  const bool has_scales = ptr_scales.has_value() && ptr_scales->defined();
  const bool is_fp8 = b_st == at::kFloat8_e4m3fn || b_st == at::kFloat8_e5m2;
  const bool is_mxfp4 = b_st == at::kFloat4_e2m1fn_x2 && has_scales;
  TORCH_CHECK(!(b_st == at::kChar && has_scales),
              "the int4 recipe of cutlass_grouped_gemm_interface is not implemented in the B200 build");
  TORCH_CHECK(N % 8 == 0, "N must be divisible by 8");
  TORCH_CHECK(ptr_A.dim() == 2, "ptr_A must be 2D [Total_M, K]");
  TORCH_CHECK(ptr_B.dim() == 3, "ptr_B must be 3D [num_experts, K, N]");
  TORCH_CHECK(ptr_D.dim() == 2, "ptr_D must be 2D [Total_M, N]");
  // extension: fp8 weights may be the transposed VIEW of a contiguous [E, N, K] tensor (the widened tile then goes to
  // tensor memory, include/vxk_b200.h VXK_MOE_FP8_*_NK); everything else must be contiguous like the reference
  const bool fp8_nk = is_fp8 && !ptr_B.is_contiguous() && ptr_B.dim() == 3 && ptr_B.transpose(1, 2).is_contiguous();
  TORCH_CHECK(ptr_A.is_contiguous() && (ptr_B.is_contiguous() || fp8_nk) && ptr_D.is_contiguous(),
              "ptr_A, ptr_B and ptr_D must be contiguous");
  const void* bias = nullptr;
  if (ptr_bias.has_value() && ptr_bias->defined()) {
    TORCH_CHECK(ptr_bias->dim() == 2, "ptr_bias must be 2D [num_experts, N]");
    TORCH_CHECK(ptr_bias->is_contiguous(), "ptr_bias must be contiguous");
    TORCH_CHECK(ptr_bias->size(0) == num_experts && ptr_bias->size(1) == N, "ptr_bias must be [num_experts, N]");
    TORCH_CHECK(ptr_bias->scalar_type() == a_st, "ptr_bias must have the activation dtype");
    bias = ptr_bias->data_ptr();
  }
  int64_t B_K = ptr_B.size(1), B_N = ptr_B.size(2);
  if (is_mxfp4) { B_K = ptr_B.size(2) * 2; B_N = ptr_B.size(1); }      // [E, N, K/2]
  TORCH_CHECK(ptr_B.size(0) == num_experts, "ptr_B.size(0) must match num_experts");
  TORCH_CHECK(ptr_A.size(0) == ptr_D.size(0), "ptr_A.size(0) must match ptr_D.size(0)");
  TORCH_CHECK(ptr_A.size(1) == B_K && B_K == K, "ptr_A.size(1) must match ptr_B.size(1)");
  TORCH_CHECK(B_N == ptr_D.size(1) && B_N == N, "ptr_B.size(2) must match ptr_D.size(1)");
  TORCH_CHECK(rows_per_expert.scalar_type() == at::kInt && rows_per_expert.is_contiguous() &&
                  rows_per_expert.numel() >= num_experts,
              "rows_per_expert must be a contiguous int32 tensor with num_experts entries");
  int kind = VXK_MOE_W16;
  const void* scales = nullptr;
  if (is_mxfp4) {
    TORCH_CHECK(ptr_scales->is_contiguous() && ptr_scales->dim() == 3, "ptr_scales must be contiguous [num_experts, N, K / 32]");
    TORCH_CHECK(ptr_scales->size(0) == num_experts && ptr_scales->size(1) == N && ptr_scales->size(2) * 32 == K,
                "mxfp4 ptr_scales must be [num_experts, N, K / 32] (group size 32)");
    TORCH_CHECK(ptr_scales->element_size() == 1, "mxfp4 ptr_scales must be E8M0 bytes (uint8 / float8_e8m0fnu)");
    kind = VXK_MOE_MXFP4;
    scales = ptr_scales->data_ptr();
  } else if (is_fp8) {
    TORCH_CHECK(has_scales, "w8a16 grouped gemm must have scales");
    TORCH_CHECK(ptr_scales->is_contiguous() && ptr_scales->dim() == 1 && ptr_scales->size(0) == num_experts,
                "ptr_scales of fp8 must be 1D [num_experts]");
    TORCH_CHECK(ptr_scales->scalar_type() == at::kFloat, "ptr_scales must be float");
    kind = b_st == at::kFloat8_e4m3fn ? (fp8_nk ? VXK_MOE_FP8_E4M3_NK : VXK_MOE_FP8_E4M3)
                                      : (fp8_nk ? VXK_MOE_FP8_E5M2_NK : VXK_MOE_FP8_E5M2);
    scales = ptr_scales->data_ptr();
  } else {
    TORCH_CHECK(b_st == a_st, "w16a16 grouped gemm: ptr_B must have the activation dtype, got ", b_st);
    TORCH_CHECK(!has_scales, "w16a16 grouped gemm must not have scales");
  }
  check_rc(vxk_moe_grouped_gemm(ptr_D.data_ptr(), ptr_A.data_ptr(), ptr_B.data_ptr(), scales, bias,
                                rows_per_expert.data_ptr<int32_t>(), ptr_A.size(0), N, K, num_experts,
                                a_st == at::kBFloat16 ? VXK_BF16 : VXK_F16, kind, 0, cur_stream()),
           "cutlass_grouped_gemm_interface");
  return ptr_D;
}

// ------------------------------------------------------------ _C_cache_ops
// ref: csrc/cache.cpp:1558-1590
void indexer_k_quant_and_cache(Tensor& k, Tensor& kv_cache, Tensor& slot_mapping, int64_t quant_block_size,
                               const std::string& scale_fmt) {
  check_cuda(k, "k");
  TORCH_CHECK(k.device() == kv_cache.device(), "k and kv_cache must be on the same device");
  TORCH_CHECK(k.device() == slot_mapping.device(), "k and slot_mapping must be on the same device");
  TORCH_CHECK(k.dim() == 2 && kv_cache.dim() == 3, "k must be [num_tokens, head_dim], kv_cache [blocks, block_size, stride]");
  TORCH_CHECK(kv_cache.element_size() == 1, "kv_cache must be a byte tensor (uint8 / fp8)");
  TORCH_CHECK(slot_mapping.scalar_type() == at::kLong && slot_mapping.is_contiguous(), "slot_mapping must be contiguous int64");
  const int64_t num_tokens = k.size(0), head_dim = k.size(1);
  TORCH_CHECK(slot_mapping.numel() >= num_tokens, "slot_mapping must hold one slot per token");
  TORCH_CHECK(quant_block_size > 0 && head_dim % quant_block_size == 0, "head_dim must be divisible by quant_block_size");
  TORCH_CHECK(k.stride(1) == 1 && kv_cache.stride(2) == 1 && kv_cache.stride(1) == kv_cache.size(2),
              "k rows and kv_cache blocks must be dense");
  const c10::cuda::CUDAGuard guard(k.device());
  check_rc(vxk_indexer_k_quant_and_cache(k.data_ptr(), kv_cache.data_ptr(), slot_mapping.data_ptr<int64_t>(), num_tokens,
                                         head_dim, k.stride(0), quant_block_size, kv_cache.size(1), kv_cache.size(2),
                                         kv_cache.stride(0), scale_fmt == "ue8m0" ? 1 : 0,
                                         in_dtype_of(k, "indexer_k_quant_and_cache"), cur_stream()),
           "indexer_k_quant_and_cache");
}

// ref: csrc/cache.cpp:1625-1680
void cp_gather_indexer_k_quant_cache(const Tensor& kv_cache, Tensor& dst_k, Tensor& dst_scale, const Tensor& block_table,
                                     const Tensor& cu_seq_lens) {
  check_cuda(kv_cache, "kv_cache");
  TORCH_CHECK(kv_cache.device() == dst_k.device(), "kv_cache and dst_k must be on the same device");
  TORCH_CHECK(kv_cache.device() == dst_scale.device(), "kv_cache and dst_scale must be on the same device");
  TORCH_CHECK(kv_cache.device() == block_table.device(), "kv_cache and block_table must be on the same device");
  TORCH_CHECK(kv_cache.device() == cu_seq_lens.device(), "kv_cache and cu_seq_lens must be on the same device");
  TORCH_CHECK(kv_cache.dim() == 3 && dst_k.dim() == 2 && dst_scale.dim() == 2 && block_table.dim() == 2,
              "kv_cache [blocks, block_size, stride], dst_k [tokens, head_dim], dst_scale [tokens, 4 * groups], "
              "block_table [batch, blocks]");
  TORCH_CHECK(kv_cache.element_size() == 1 && dst_k.element_size() == 1 && dst_scale.element_size() == 1,
              "kv_cache, dst_k and dst_scale must be byte tensors");
  TORCH_CHECK(block_table.scalar_type() == at::kInt && cu_seq_lens.scalar_type() == at::kInt,
              "block_table and cu_seq_lens must be int32");
  TORCH_CHECK(block_table.stride(1) == 1 && cu_seq_lens.is_contiguous() && dst_k.stride(1) == 1 && dst_scale.stride(1) == 1 &&
                  kv_cache.stride(2) == 1 && kv_cache.stride(1) == kv_cache.size(2),
              "inner dimensions must be dense");
  const int64_t batch = block_table.size(0), num_tokens = dst_k.size(0), head_dim = dst_k.size(1);
  TORCH_CHECK(cu_seq_lens.numel() >= batch + 1, "cu_seq_lens must have batch_size + 1 entries");
  TORCH_CHECK(dst_scale.size(1) > 0 && (head_dim * 4) % dst_scale.size(1) == 0, "dst_scale must be [tokens, 4 * head_dim / quant_block_size]");
  const int64_t quant_block_size = head_dim * 4 / dst_scale.size(1);
  TORCH_CHECK(head_dim % quant_block_size == 0, "head_dim must be divisible by quant_block_size");
  TORCH_CHECK(block_table.stride(0) == block_table.size(1), "block_table rows must be dense");
  const c10::cuda::CUDAGuard guard(kv_cache.device());
  check_rc(vxk_cp_gather_indexer_k_quant_cache(kv_cache.data_ptr(), dst_k.data_ptr(), dst_scale.data_ptr(),
                                               block_table.data_ptr<int32_t>(), cu_seq_lens.data_ptr<int32_t>(), batch,
                                               block_table.size(1), num_tokens, head_dim, quant_block_size,
                                               kv_cache.size(1), kv_cache.stride(0), dst_k.stride(0), dst_scale.stride(0),
                                               cur_stream()),
           "cp_gather_indexer_k_quant_cache");
}

// ref: csrc/cache.cpp:1478-1540
void convert_fp8(Tensor& dst, const Tensor& src, double scale, const std::string& kv_cache_dtype) {
  TORCH_CHECK(src.is_cuda(), "src must be on a CUDA device");
  TORCH_CHECK(dst.is_cuda(), "dst must be on a CUDA device");
  TORCH_CHECK(src.device() == dst.device(), "src and dst must be on the same device");
  int mode;
  if (kv_cache_dtype == "auto") mode = 0;
  else if (kv_cache_dtype == "fp8" || kv_cache_dtype == "fp8_e4m3") mode = 1;
  else TORCH_CHECK(false, "Unsupported data type: ", kv_cache_dtype);
  auto kind = [](const Tensor& t) -> int {
    switch (t.scalar_type()) {
      case at::kFloat: return VXK_F32;
      case at::kHalf: return VXK_F16;
      case at::kBFloat16: return VXK_BF16;
      default: return t.element_size() == 1 ? 3 : -1;
    }
  };
  const int sk = kind(src), dk = kind(dst);
  TORCH_CHECK(sk >= 0 && dk >= 0 && ((sk == 3) != (dk == 3)), "Unsupported data type combination for ", kv_cache_dtype,
              " convert_fp8");
  TORCH_CHECK(src.numel() == dst.numel(), "src and dst must have the same number of elements");
  TORCH_CHECK(src.is_contiguous() && dst.is_contiguous(), "src and dst must be contiguous");
  const c10::cuda::CUDAGuard guard(src.device());
  check_rc(vxk_convert_fp8(dst.data_ptr(), src.data_ptr(), scale, src.numel(), sk, dk, mode, cur_stream()), "convert_fp8");
}

}  // namespace

// Schemas verbatim from /root/reference/csrc/torch_bindings.cpp (line numbers in SURVEY.md 8b).
TORCH_LIBRARY_FRAGMENT(_C, ops) {
  ops.def("rms_norm_dynamic_per_token_quant(Tensor! result, Tensor input, Tensor weight, Tensor! scale, "
          "float epsilon, Tensor? scale_ub, Tensor!? residual) -> ()");
  ops.def("rms_norm_per_block_quant(Tensor! result, Tensor input, Tensor weight, Tensor! scale, float epsilon, "
          "Tensor? scale_ub, Tensor!? residual, int group_size, bool is_scale_transposed, "
          "bool scale_ue8m0=False) -> ()");
  ops.def("rms_norm_mxfp4_quant(Tensor! result, Tensor input, Tensor weight, Tensor! scale, float epsilon, "
          "Tensor!? residual, int group_size) -> ()");
  ops.def("rms_norm_static_fp8_quant(Tensor! result, Tensor input, Tensor weight, Tensor scale, "
          "float epsilon) -> ()");
  ops.def("fused_add_rms_norm_static_fp8_quant(Tensor! result, Tensor input, Tensor! residual, Tensor weight, "
          "Tensor scale, float epsilon) -> ()");
  ops.def("silu_and_mul_quant(Tensor! result, Tensor input, Tensor scale) -> ()");
  ops.def("silu_and_mul_per_block_quant(Tensor! out, Tensor input, Tensor! scales, int group_size, "
          "Tensor? scale_ub=None, bool is_scale_transposed=False, bool scale_ue8m0=False) -> ()");
  ops.def("silu_and_mul_mxfp4_quant(Tensor! out, Tensor input, Tensor! scales, int group_size, "
          "float eps=1e-10) -> ()");
  ops.def("static_scaled_fp8_quant(Tensor! result, Tensor input, Tensor scale, "
          "(int, int)? group_shape=None) -> ()");
  ops.def("dynamic_scaled_fp8_quant(Tensor! result, Tensor input, Tensor! scale) -> ()");
  ops.def("dynamic_per_token_scaled_fp8_quant(Tensor! result, Tensor input, Tensor! scale, "
          "Tensor? scale_ub) -> ()");
  ops.def("per_token_group_fp8_quant(Tensor input, Tensor! output_q, Tensor! output_s, int group_size, "
          "float eps, float fp8_min, float fp8_max, bool scale_ue8m0, bool dummy_is_scale_transposed, "
          "bool dummy_is_tma_aligned ) -> ()");
  ops.def("per_token_group_quant_mxfp4(Tensor input, Tensor! output_q, Tensor! output_s, int group_size, "
          "float eps) -> ()");
}

TORCH_LIBRARY_IMPL(_C, CUDA, ops) {
  ops.impl("rms_norm_dynamic_per_token_quant", &rms_norm_dynamic_per_token_quant);
  ops.impl("rms_norm_per_block_quant", &rms_norm_per_block_quant);
  ops.impl("rms_norm_mxfp4_quant", &rms_norm_mxfp4_quant);
  ops.impl("rms_norm_static_fp8_quant", &rms_norm_static_fp8_quant);
  ops.impl("fused_add_rms_norm_static_fp8_quant", &fused_add_rms_norm_static_fp8_quant);
  ops.impl("silu_and_mul_quant", &silu_and_mul_quant);
  ops.impl("silu_and_mul_per_block_quant", &silu_and_mul_per_block_quant);
  ops.impl("silu_and_mul_mxfp4_quant", &silu_and_mul_mxfp4_quant);
  ops.impl("static_scaled_fp8_quant", &static_scaled_fp8_quant);
  ops.impl("dynamic_scaled_fp8_quant", &dynamic_scaled_fp8_quant);
  ops.impl("dynamic_per_token_scaled_fp8_quant", &dynamic_per_token_scaled_fp8_quant);
  ops.impl("per_token_group_fp8_quant", &per_token_group_fp8_quant);
  ops.impl("per_token_group_quant_mxfp4", &per_token_group_quant_mxfp4);
}

// Schemas verbatim from /root/reference/csrc/xpu/torch_bindings.cpp:14-27.
TORCH_LIBRARY_FRAGMENT(_xpu_C, xpu_ops) {
  xpu_ops.def("fp8_gemm(Tensor A, Tensor B, ScalarType? out_dtype, Tensor? A_scale_, Tensor? B_scale_, "
              "Tensor? bias_) -> Tensor");
  xpu_ops.def("fp8_gemm_w8a16(Tensor A, Tensor B, Tensor? B_scale_, Tensor? bias_) -> Tensor");
  xpu_ops.def("fp4_gemm(Tensor A, Tensor B, Tensor A_scale, Tensor B_scale, ScalarType? out_dtype, "
              "Tensor? bias_) -> Tensor");
  // csrc/xpu/torch_bindings.cpp:41-53
  xpu_ops.def("cutlass_grouped_gemm_interface(Tensor ptr_A, Tensor ptr_B, Tensor? ptr_scales, Tensor? ptr_bias, "
              "Tensor ptr_D, Tensor rows_per_expert, int N, int K, int num_experts) -> Tensor");
}

TORCH_LIBRARY_IMPL(_xpu_C, CUDA, xpu_ops) {
  xpu_ops.impl("fp8_gemm", &fp8_gemm);
  xpu_ops.impl("fp4_gemm", &fp4_gemm);
  xpu_ops.impl("fp8_gemm_w8a16", &fp8_gemm_w8a16);
  xpu_ops.impl("cutlass_grouped_gemm_interface", &cutlass_grouped_gemm_interface);
}

// Extensions without a reference counterpart (scale factors in the tensor core's layout).
// SURVEY 8f #2, second half (no reference counterpart): gate_up projection with silu_and_mul and the 1 x 128 block FP8
// quantisation fused into the GEMM's epilogue.  Equals
//   t = fp8_gemm(A, B, mid_dtype, A_scale, B_scale, bias);  silu_and_mul_per_block_quant(out_q, t, out_s, 128)
// bit for bit; shapes the fused kernel does not take run exactly those two ops.
void fp8_gemm_silu_mul_quant(const Tensor& A, const Tensor& B, const Tensor& A_scale, const Tensor& B_scale,
                             const std::optional<Tensor>& bias_, at::ScalarType mid_dtype, Tensor& out_q, Tensor& out_s) {
  check_cuda(A, "A");
  same_device(A, B, A_scale, B_scale, bias_);
  TORCH_CHECK(out_q.device() == A.device() && out_s.device() == A.device(), "outputs must be on A's device");
  const c10::cuda::CUDAGuard guard(A.device());
  TORCH_CHECK(A.dim() == 2 && B.dim() == 2, "A must be [M, K], B the [K, 2 * inter] view of W[2 * inter, K]");
  TORCH_CHECK(mid_dtype == at::kBFloat16 || mid_dtype == at::kHalf, "mid_dtype must be bfloat16 or float16");
  const auto st = A.scalar_type();
  TORCH_CHECK((st == at::kFloat8_e4m3fn || st == at::kFloat8_e5m2) && B.scalar_type() == st,
              "A and B must be float8_e4m3fn or float8_e5m2, same type");
  const int64_t M = A.size(0), K = A.size(1), N = B.size(1);
  TORCH_CHECK(B.size(0) == K && N % 2 == 0, "B must be [K, 2 * inter]");
  const int64_t inter = N / 2;
  TORCH_CHECK(out_q.scalar_type() == at::kFloat8_e4m3fn && out_q.is_contiguous() && out_q.dim() == 2 &&
                  out_q.size(0) == M && out_q.size(1) == inter, "out_q must be contiguous float8_e4m3fn [M, inter]");
  TORCH_CHECK(inter % 128 == 0, "inter must be a multiple of the group size 128");
  TORCH_CHECK(out_s.scalar_type() == at::kFloat && out_s.dim() == 2 && out_s.size(0) == M && out_s.size(1) == inter / 128,
              "out_s must be float32 [M, inter / 128]");
  TORCH_CHECK(A_scale.scalar_type() == at::kFloat && B_scale.scalar_type() == at::kFloat, "scales must be float32");
  const int a_mode = A_scale.numel() == 1 ? 0 : 1, b_mode = B_scale.numel() == 1 ? 0 : 1;
  if (a_mode == 1) TORCH_CHECK(A_scale.numel() == M && A_scale.is_contiguous(), "A_scale must be [M] / [M,1]");
  if (b_mode == 1) TORCH_CHECK(B_scale.numel() == N && B_scale.is_contiguous(), "B_scale must be [N] / [N,1]");
  if (M == 0 || inter == 0) return;
  const void* bias = nullptr;
  int bias_dtype = VXK_F32;
  if (bias_.has_value() && bias_->defined() && bias_->numel() > 0) {
    TORCH_CHECK(bias_->numel() == N && bias_->is_contiguous(), "only a contiguous [N] bias is supported");
    bias = bias_->data_ptr();
    bias_dtype = bias_dtype_of(*bias_);
  }
  int rc = VXK_ERR_UNSUPPORTED;
  if (B.stride(0) == 1 && A.stride(1) == 1)
    rc = vxk_fp8_gemm_silu_mul_quant(out_q.data_ptr(), out_s.data_ptr<float>(), A.data_ptr(), B.data_ptr(),
                                     A_scale.data_ptr<float>(), B_scale.data_ptr<float>(), bias, M, inter, K, A.stride(0),
                                     B.stride(1), a_mode, b_mode, out_s.stride(0), out_s.stride(1),
                                     st == at::kFloat8_e4m3fn ? VXK_Q_E4M3 : VXK_Q_E5M2,
                                     mid_dtype == at::kBFloat16 ? VXK_BF16 : VXK_F16, bias_dtype, cur_stream());
  if (rc == VXK_ERR_UNSUPPORTED) {
    Tensor t = fp8_gemm(A, B, mid_dtype, A_scale, B_scale, bias_);
    silu_and_mul_per_block_quant(out_q, t, out_s, 128, std::nullopt, out_s.stride(0) == 1 && out_s.size(1) > 1, false);
    return;
  }
  check_rc(rc, "fp8_gemm_silu_mul_quant");
}

TORCH_LIBRARY_FRAGMENT(_vxk, vxk_ops) {
  vxk_ops.def("mx_sf_packed_bytes(int rows, int K, bool fp4, bool is_weight) -> int", &mx_sf_packed_bytes);
  vxk_ops.def("mx_sf_pack(Tensor scale, bool fp4, bool is_weight) -> Tensor");
  vxk_ops.def("per_token_group_quant_mxfp4_sf(Tensor input, Tensor! output_q, Tensor! output_s, Tensor! sf_packed, "
              "int group_size, float eps) -> ()");
  vxk_ops.def("rms_norm_mxfp4_quant_sf(Tensor! result, Tensor input, Tensor weight, Tensor! scale, Tensor! sf_packed, "
              "float epsilon, Tensor!? residual, int group_size) -> ()");
  vxk_ops.def("silu_and_mul_mxfp4_quant_sf(Tensor! out, Tensor input, Tensor! scales, Tensor! sf_packed, int group_size, "
              "float eps=1e-10) -> ()");
  vxk_ops.def("fp4_gemm_sf(Tensor A, Tensor B, Tensor A_scale, Tensor B_scale, ScalarType? out_dtype, Tensor? bias_, "
              "Tensor? A_sf_packed, Tensor? B_sf_packed) -> Tensor");
  vxk_ops.def("fp8_gemm_silu_mul_quant(Tensor A, Tensor B, Tensor A_scale, Tensor B_scale, Tensor? bias_, "
              "ScalarType mid_dtype, Tensor! out_q, Tensor! out_s) -> ()");
}
TORCH_LIBRARY_IMPL(_vxk, CUDA, vxk_ops) {
  vxk_ops.impl("mx_sf_pack", &mx_sf_pack);
  vxk_ops.impl("per_token_group_quant_mxfp4_sf", &per_token_group_quant_mxfp4_sf);
  vxk_ops.impl("rms_norm_mxfp4_quant_sf", &rms_norm_mxfp4_quant_sf);
  vxk_ops.impl("silu_and_mul_mxfp4_quant_sf", &silu_and_mul_mxfp4_quant_sf);
  vxk_ops.impl("fp4_gemm_sf", &fp4_gemm_sf);
  vxk_ops.impl("fp8_gemm_silu_mul_quant", &fp8_gemm_silu_mul_quant);
}

// Schemas verbatim from /root/reference/csrc/torch_bindings.cpp:281-284,298-309 (library _C_cache_ops).
TORCH_LIBRARY_FRAGMENT(_C_cache_ops, cache_ops) {
  cache_ops.def("convert_fp8(Tensor! dst, Tensor src, "
                "            float scale, str kv_cache_dtype) -> ()");
  cache_ops.def("indexer_k_quant_and_cache(Tensor k, Tensor! kv_cache,"
                "Tensor slot_mapping, int quant_block_size, str scale_fmt) -> ()");
  cache_ops.def("cp_gather_indexer_k_quant_cache(Tensor kv_cache, Tensor! dst_k, "
                "Tensor! dst_scale, Tensor block_table, Tensor cu_seq_lens) -> ()");
}

TORCH_LIBRARY_IMPL(_C_cache_ops, CUDA, cache_ops) {
  cache_ops.impl("convert_fp8", &convert_fp8);
  cache_ops.impl("indexer_k_quant_and_cache", &indexer_k_quant_and_cache);
  cache_ops.impl("cp_gather_indexer_k_quant_cache", &cp_gather_indexer_k_quant_cache);
}
