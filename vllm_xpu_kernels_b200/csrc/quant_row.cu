// Row-resident quantisation kernels: ops that need a reduction over the whole
// token row (per-token amax and/or the rms-norm variance).
//
//   a1  dynamic_per_token_scaled_fp8_quant        (MODE_TOKEN_DIV)
//   a6  rms_norm_dynamic_per_token_quant          (MODE_TOKEN_MUL / int8)
//   a7  rms_norm_per_block_quant                  (MODE_GROUP)
//   a8  rms_norm_mxfp4_quant                      (MODE_MXFP4)
//   a9  rms_norm_static_fp8_quant (+fused_add)    (MODE_STATIC)
//
// Design: one CTA per row, the row lives in registers as fp32 (VPT vectors of
// 8 elements per thread, loaded with 128-bit streaming loads), so HBM sees
// exactly one read of the activation (+ residual read/write) and one write of
// the codes.  The reference kernels re-read the row 2-3 times
// (csrc/layernorm_quant.cpp:27-142, fp8_quant.cpp:393-477).
// Rows whose length / alignment does not fit the vector path go through
// row_quant_generic (scalar, multi-pass, any H / stride).
#include "vxk_common.cuh"

namespace vxk {

enum RowMode { MODE_TOKEN_DIV = 0, MODE_TOKEN_MUL = 1, MODE_GROUP = 2, MODE_MXFP4 = 3, MODE_STATIC = 4 };

struct RowParams {
  void* out;            // [T, H] codes (or [T, H/2] packed fp4)
  const void* in;       // [T, H], row stride in_row_stride
  const void* weight;   // [H] (rms only)
  void* residual;       // [T, H] in/out, nullable
  float* scales;        // per-token [T] or per-group (t*ss0 + g*ss1)
  const float* scale_in;  // scale_ub (token modes) or the static scale
  int64_t H, in_row_stride, ss0, ss1;
  int nvec;             // H / 8
  int group;            // group size (MODE_GROUP / MODE_MXFP4)
  int ue8m0;
  float eps;
};

template <typename T, int Q, int MODE, bool RMS, int VPT>
__global__ void __launch_bounds__(512) row_quant_kernel(const RowParams p) {
  __shared__ float scratch[33];
  const int64_t row = blockIdx.x;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const T* __restrict__ xin = static_cast<const T*>(p.in) + row * p.in_row_stride;
  float v[VPT][8];

#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int idx = tid + i * nthr;
    if (idx < p.nvec) load8<T, true>(xin + idx * kVec, v[i]);
    else {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[i][j] = 0.0f;
    }
  }

  if constexpr (RMS) {
    if (p.residual != nullptr) {
      T* res = static_cast<T*>(p.residual) + row * p.H;
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        const int idx = tid + i * nthr;
        if (idx < p.nvec) {
          float r[8];
          load8<T, false>(res + idx * kVec, r);
#pragma unroll
          for (int j = 0; j < 8; ++j) v[i][j] = round_to<T>(v[i][j] + r[j]);
          store8<T>(res + idx * kVec, v[i]);
        }
      }
    }
    float ss = 0.0f;
#pragma unroll
    for (int i = 0; i < VPT; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) ss += v[i][j] * v[i][j];
    ss = block_reduce<false>(ss, scratch);
    const float var = __fdiv_rn(ss, static_cast<float>(p.H));
    const float inv_rms = __fdiv_rn(1.0f, __fsqrt_rn(var + p.eps));
    const T* __restrict__ w = static_cast<const T*>(p.weight);
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int idx = tid + i * nthr;
      if (idx < p.nvec) {
        float wv[8];
        load8<T, false>(w + idx * kVec, wv);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[i][j] = (v[i][j] * inv_rms) * wv[j];
      }
    }
  }

  constexpr float kQMax = QTraits<Q>::kMax;

  if constexpr (MODE == MODE_TOKEN_DIV || MODE == MODE_TOKEN_MUL) {
    float amax = 0.0f;
#pragma unroll
    for (int i = 0; i < VPT; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) amax = fmaxf(amax, fabsf(v[i][j]));
    amax = block_reduce<true>(amax, scratch);
    float s;
    if constexpr (Q == VXK_Q_INT8) {
      s = amax > 0.0f ? __fdiv_rn(amax, 127.0f) : 1.0f;
    } else {
      if (p.scale_in != nullptr) amax = fminf(amax, __ldg(p.scale_in));
      s = fmaxf(__fdiv_rn(amax, kQMax), __fdiv_rn(1.0f, kQMax * 512.0f));
    }
    const float r = __fdiv_rn(1.0f, s);
    if (tid == 0) p.scales[row] = s;
    uint8_t* o = static_cast<uint8_t*>(p.out) + row * p.H;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int idx = tid + i * nthr;
      if (idx < p.nvec) {
        float q[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if constexpr (MODE == MODE_TOKEN_DIV || Q == VXK_Q_INT8) q[j] = div_by(v[i][j], s, r);
          else q[j] = v[i][j] * r;
        }
        stg_stream(o + idx * kVec, pack_q8<Q>(q));
      }
    }
  } else if constexpr (MODE == MODE_STATIC) {
    const float r = __fdiv_rn(1.0f, __ldg(p.scale_in));
    uint8_t* o = static_cast<uint8_t*>(p.out) + row * p.H;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int idx = tid + i * nthr;
      if (idx < p.nvec) {
        float q[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) q[j] = v[i][j] * r;
        stg_stream(o + idx * kVec, pack_q8<Q>(q));
      }
    }
  } else {  // MODE_GROUP / MODE_MXFP4: 1 x G groups, G/8 consecutive lanes per group
    const int lanes = p.group / kVec;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int idx = tid + i * nthr;
      const bool valid = idx < p.nvec;
      float gmax = 0.0f;
#pragma unroll
      for (int j = 0; j < 8; ++j) gmax = fmaxf(gmax, fabsf(v[i][j]));
      gmax = seg_max(gmax, lanes);  // all lanes participate (invalid ones carry 0)
      const int g = idx / lanes;
      const bool leader = valid && (idx % lanes == 0);
      if constexpr (MODE == MODE_MXFP4) {
        const int e = rceil_e8m0(__fdiv_rn(gmax, 6.0f));
        const float d = e8m0_descale(e);
        if (leader) p.scales[row * p.ss0 + g * p.ss1] = e8m0_scale(e);
        if (valid) {
          float q[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float c = fminf(fmaxf(v[i][j] * d, -6.0f), 6.0f);
            q[j] = __bfloat162float(__float2bfloat16_rn(c));
          }
          uint8_t* o = static_cast<uint8_t*>(p.out) + row * (p.H / 2);
          stg_stream(o + idx * 4, pack_e2m1x8(q));
        }
      } else {
        float s;
        if constexpr (Q == VXK_Q_INT8) {
          s = gmax > 0.0f ? __fdiv_rn(gmax, 127.0f) : 1.0f;
        } else {
          s = fmaxf(__fdiv_rn(gmax, kQMax), __fdiv_rn(1.0f, kQMax * 512.0f));
          if (p.ue8m0) s = exp2_ceil_log2(fmaxf(fabsf(s), 1e-10f));
        }
        const float r = __fdiv_rn(1.0f, s);
        if (leader) p.scales[row * p.ss0 + g * p.ss1] = s;
        if (valid) {
          float q[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            if constexpr (Q == VXK_Q_INT8) q[j] = div_by(v[i][j], s, r);
            else q[j] = v[i][j] * r;
          }
          uint8_t* o = static_cast<uint8_t*>(p.out) + row * p.H;
          stg_stream(o + idx * kVec, pack_q8<Q>(q));
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------
// Generic fallback: any H, any alignment.  One CTA per row, scalar accesses,
// the row is re-read from L2 for each pass.  Same arithmetic as above.
// ---------------------------------------------------------------------------
template <typename T, int Q, int MODE, bool RMS>
__global__ void __launch_bounds__(256) row_quant_generic(const RowParams p) {
  __shared__ float scratch[33];
  const int64_t row = blockIdx.x;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const T* __restrict__ xin = static_cast<const T*>(p.in) + row * p.in_row_stride;
  T* res = p.residual ? static_cast<T*>(p.residual) + row * p.H : nullptr;
  const T* __restrict__ w = static_cast<const T*>(p.weight);
  const int H = static_cast<int>(p.H);
  constexpr float kQMax = QTraits<Q>::kMax;

  float inv_rms = 1.0f;
  if constexpr (RMS) {
    float ss = 0.0f;
    for (int h = tid; h < H; h += nthr) {
      float x = to_f32<T>(xin[h]);
      if (res) {
        x = round_to<T>(x + to_f32<T>(res[h]));
        res[h] = from_f32<T>(x);
      }
      ss += x * x;
    }
    ss = block_reduce<false>(ss, scratch);
    inv_rms = __fdiv_rn(1.0f, __fsqrt_rn(__fdiv_rn(ss, static_cast<float>(H)) + p.eps));
  }
  // value to quantise at column h (after the prologue)
  auto value = [&](int h) -> float {
    if constexpr (RMS) {
      const float x = res ? to_f32<T>(res[h]) : to_f32<T>(xin[h]);
      return (x * inv_rms) * to_f32<T>(w[h]);
    } else {
      return to_f32<T>(xin[h]);
    }
  };

  if constexpr (MODE == MODE_TOKEN_DIV || MODE == MODE_TOKEN_MUL) {
    float amax = 0.0f;
    for (int h = tid; h < H; h += nthr) amax = fmaxf(amax, fabsf(value(h)));
    amax = block_reduce<true>(amax, scratch);
    float s;
    if constexpr (Q == VXK_Q_INT8) {
      s = amax > 0.0f ? __fdiv_rn(amax, 127.0f) : 1.0f;
    } else {
      if (p.scale_in != nullptr) amax = fminf(amax, __ldg(p.scale_in));
      s = fmaxf(__fdiv_rn(amax, kQMax), __fdiv_rn(1.0f, kQMax * 512.0f));
    }
    const float r = __fdiv_rn(1.0f, s);
    if (tid == 0) p.scales[row] = s;
    uint8_t* o = static_cast<uint8_t*>(p.out) + row * p.H;
    for (int h = tid; h < H; h += nthr) {
      const float x = value(h);
      const float q = (MODE == MODE_TOKEN_DIV || Q == VXK_Q_INT8) ? div_by(x, s, r) : x * r;
      o[h] = pack_q1<Q>(q);
    }
  } else if constexpr (MODE == MODE_STATIC) {
    const float r = __fdiv_rn(1.0f, __ldg(p.scale_in));
    uint8_t* o = static_cast<uint8_t*>(p.out) + row * p.H;
    for (int h = tid; h < H; h += nthr) o[h] = pack_q1<Q>(value(h) * r);
  } else {
    // one warp per group, round-robin over the row's groups
    const int G = p.group, ngroups = H / G;
    const int lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
    for (int g = warp; g < ngroups; g += nwarp) {
      float gmax = 0.0f;
      for (int h = g * G + lane; h < (g + 1) * G; h += 32) gmax = fmaxf(gmax, fabsf(value(h)));
      gmax = warp_max(gmax);
      if constexpr (MODE == MODE_MXFP4) {
        const int e = rceil_e8m0(__fdiv_rn(gmax, 6.0f));
        const float d = e8m0_descale(e);
        if (lane == 0) p.scales[row * p.ss0 + g * p.ss1] = e8m0_scale(e);
        uint8_t* o = static_cast<uint8_t*>(p.out) + row * (p.H / 2);
        for (int h2 = g * G / 2 + lane; h2 < (g + 1) * G / 2; h2 += 32) {
          float c0 = fminf(fmaxf(value(2 * h2) * d, -6.0f), 6.0f);
          float c1 = fminf(fmaxf(value(2 * h2 + 1) * d, -6.0f), 6.0f);
          c0 = __bfloat162float(__float2bfloat16_rn(c0));
          c1 = __bfloat162float(__float2bfloat16_rn(c1));
          o[h2] = static_cast<uint8_t>(cvt_e2m1x2(c1, c0));
        }
      } else {
        float s;
        if constexpr (Q == VXK_Q_INT8) {
          s = gmax > 0.0f ? __fdiv_rn(gmax, 127.0f) : 1.0f;
        } else {
          s = fmaxf(__fdiv_rn(gmax, kQMax), __fdiv_rn(1.0f, kQMax * 512.0f));
          if (p.ue8m0) s = exp2_ceil_log2(fmaxf(fabsf(s), 1e-10f));
        }
        const float r = __fdiv_rn(1.0f, s);
        if (lane == 0) p.scales[row * p.ss0 + g * p.ss1] = s;
        uint8_t* o = static_cast<uint8_t*>(p.out) + row * p.H;
        for (int h = g * G + lane; h < (g + 1) * G; h += 32) {
          const float x = value(h);
          o[h] = pack_q1<Q>(Q == VXK_Q_INT8 ? div_by(x, s, r) : x * r);
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------
// host dispatch
// ---------------------------------------------------------------------------
template <typename T, int Q, int MODE, bool RMS>
static int launch_row(const RowParams& p, int64_t T_rows, bool vec_ok, cudaStream_t st) {
  if (T_rows == 0 || p.H == 0) return VXK_OK;
  if (T_rows > 0x7fffffffLL) return fail(VXK_ERR_INVALID, "too many rows (%lld)", (long long)T_rows);
  const dim3 grid(static_cast<unsigned>(T_rows));
  const int nvec = p.nvec;
  if (vec_ok && nvec <= 4096) {
    int vpt = nvec <= 128 ? 1 : nvec <= 512 ? 2 : nvec <= 2048 ? 4 : 8;
    int threads = ((nvec + vpt - 1) / vpt + 31) / 32 * 32;
    switch (vpt) {
      case 1: row_quant_kernel<T, Q, MODE, RMS, 1><<<grid, threads, 0, st>>>(p); break;
      case 2: row_quant_kernel<T, Q, MODE, RMS, 2><<<grid, threads, 0, st>>>(p); break;
      case 4: row_quant_kernel<T, Q, MODE, RMS, 4><<<grid, threads, 0, st>>>(p); break;
      default: row_quant_kernel<T, Q, MODE, RMS, 8><<<grid, threads, 0, st>>>(p); break;
    }
  } else {
    row_quant_generic<T, Q, MODE, RMS><<<grid, 256, 0, st>>>(p);
  }
  return check_launch("row_quant");
}

#define VXK_DISPATCH_T(in_dtype, ...)                                          \
  switch (in_dtype) {                                                          \
    case VXK_F32: { using T = float; __VA_ARGS__; }                            \
    case VXK_F16: { using T = __half; __VA_ARGS__; }                           \
    case VXK_BF16: { using T = __nv_bfloat16; __VA_ARGS__; }                   \
    default: return fail(VXK_ERR_INVALID, "unsupported input dtype %d", in_dtype); \
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
  VXK_DISPATCH_T(in_dtype,
    if (q_dtype == VXK_Q_E4M3) return launch_row<T, VXK_Q_E4M3, MODE_TOKEN_DIV, false>(p, T_, v, stream);
    return launch_row<T, VXK_Q_E5M2, MODE_TOKEN_DIV, false>(p, T_, v, stream));
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
  VXK_DISPATCH_T(in_dtype,
    if (q_dtype == VXK_Q_E4M3) return launch_row<T, VXK_Q_E4M3, MODE_TOKEN_MUL, true>(p, T_, v, stream);
    return launch_row<T, VXK_Q_INT8, MODE_TOKEN_MUL, true>(p, T_, v, stream));
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
  VXK_DISPATCH_T(in_dtype,
    if (q_dtype == VXK_Q_E4M3) return launch_row<T, VXK_Q_E4M3, MODE_GROUP, true>(p, T_, v, stream);
    return launch_row<T, VXK_Q_INT8, MODE_GROUP, true>(p, T_, v, stream));
}

extern "C" int vxk_rms_norm_mxfp4_quant(void* out, const void* in, const void* weight, float* scales,
                                        double epsilon, void* residual, int64_t T_, int64_t H,
                                        int64_t group_size, int64_t ss0, int64_t ss1, int in_dtype,
                                        vxk_stream_t stream) {
  VXK_REQUIRE(T_ >= 0 && H >= 0, "negative shape");
  VXK_REQUIRE(group_size == 32, "MXFP4 requires group_size == 32, got %lld", (long long)group_size);
  VXK_REQUIRE(H % 32 == 0, "hidden_size must be divisible by group_size");
  RowParams p{};
  p.out = out; p.in = in; p.weight = weight; p.residual = residual; p.scales = scales;
  p.H = H; p.in_row_stride = H; p.nvec = static_cast<int>(H / kVec);
  p.group = 32; p.ss0 = ss0; p.ss1 = ss1; p.eps = static_cast<float>(epsilon);
  const bool v = vec_path_ok(p, in_dtype, /*packed4=*/true);
  VXK_DISPATCH_T(in_dtype,
    return launch_row<T, VXK_Q_E4M3, MODE_MXFP4, true>(p, T_, v, stream));
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
  VXK_DISPATCH_T(in_dtype,
    if (q_dtype == VXK_Q_E4M3) return launch_row<T, VXK_Q_E4M3, MODE_STATIC, true>(p, T_, v, stream);
    return launch_row<T, VXK_Q_E5M2, MODE_STATIC, true>(p, T_, v, stream));
}
