// Group-local / element-wise quantisation kernels: nothing is reduced across
// more than one 1 x G group, so the work is a flat stream of 8-element vectors.
//
//   a2  static_scaled_fp8_quant           (static_quant_kernel)
//   a3  dynamic_scaled_fp8_quant          (tensor_amax_kernel + static_quant_kernel)
//   a4  per_token_group_fp8_quant         (FMT_FP8_DIV / FMT_MX8)
//   a5  per_token_group_quant_mxfp4       (FMT_MXFP4)
//   a10 silu_and_mul_quant                (FMT_STATIC, PRO_SILU_T)
//   a11 silu_and_mul_per_block_quant      (FMT_FP8_DIV / int8, PRO_SILU_T)
//   a12 silu_and_mul_mxfp4_quant          (FMT_MXFP4, PRO_SILU_F32)
//
// Design: grid-stride over vector indices j in [0, T*H/8); a 1 x G group is
// G/8 consecutive lanes of one warp (G in {8..256}, power-of-two lane count),
// its amax is a segmented xor-shuffle.  Each thread issues all of its 128-bit
// loads (U vectors, x2 for gate/up) before the first use, single pass, 64-bit
// (fp8/int8) or 32-bit (fp4) coalesced streaming stores.
#include "vxk_common.cuh"

#include <cfloat>

namespace vxk {

std::atomic<int> g_silu_fast{0};

enum Pro { PRO_NONE = 0, PRO_SILU_T = 1, PRO_SILU_F32 = 2 };
enum Fmt { FMT_FP8_DIV = 0, FMT_MX8 = 1, FMT_MXFP4 = 2, FMT_STATIC = 3 };

struct FlatParams {
  void* out;
  const void* in;        // [T, H] or [T, 2H] (gate || up)
  float* scales;         // (t*ss0 + g*ss1)
  const float* scale_in; // scale_ub or static scale
  int64_t total;         // T * H / 8
  int64_t ss0, ss1;
  int nvec;              // H / 8
  int lanes;             // G / 8
  int ue8m0;
  float eps;
};

template <typename T, int PRO, bool FAST>
__device__ __forceinline__ void load_act(const FlatParams& p, int64_t t, int c, float (&v)[8]) {
  if constexpr (PRO == PRO_NONE) {
    load8<T, true>(static_cast<const T*>(p.in) + (t * p.nvec + c) * kVec, v);
  } else {
    const T* g = static_cast<const T*>(p.in) + (t * 2 * p.nvec + c) * kVec;
    float u[8];
    load8<T, true>(g, v);
    load8<T, true>(g + static_cast<int64_t>(p.nvec) * kVec, u);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if constexpr (PRO == PRO_SILU_T) {
        // F.silu(gate) * up evaluated in T: silu rounded to T, product rounded to T
        v[j] = round_to<T>(round_to<T>(silu_f32<FAST>(v[j])) * u[j]);
      } else {
        // gate_f32 * sigmoid(gate_f32) * up_f32, no intermediate rounding to T
        if constexpr (FAST) v[j] = silu_f32<true>(v[j]) * u[j];
        else v[j] = (v[j] * __fdiv_rn(1.0f, 1.0f + expf(-v[j]))) * u[j];
      }
    }
  }
}

template <typename T, int Q, int PRO, int FMT, bool FAST, int U>
__global__ void __launch_bounds__(256) flat_quant_kernel(const FlatParams p) {
  constexpr float kQMax = QTraits<Q>::kMax;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  const int64_t first = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int lanes = p.lanes;
  float r_static = 0.0f;
  if constexpr (FMT == FMT_STATIC) r_static = __fdiv_rn(1.0f, __ldg(p.scale_in));

  // warp-uniform trip count: lane 0's index decides, every lane runs the shuffles
  for (int64_t base = first - (threadIdx.x & 31); base < p.total; base += stride * U) {
    float v[U][8];
    int64_t jj[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      jj[u] = base + (threadIdx.x & 31) + u * stride;
      if (jj[u] < p.total) {
        load_act<T, PRO, FAST>(p, jj[u] / p.nvec, static_cast<int>(jj[u] % p.nvec), v[u]);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[u][j] = 0.0f;
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t j = jj[u];
      const bool valid = j < p.total;
      const int64_t t = j / p.nvec;
      const int c = static_cast<int>(j % p.nvec);
      float q[8];
      if constexpr (FMT == FMT_STATIC) {
#pragma unroll
        for (int k = 0; k < 8; ++k) q[k] = v[u][k] * r_static;
        if (valid) stg_stream(static_cast<uint8_t*>(p.out) + j * kVec, pack_q8<Q>(q));
      } else {
        float gmax = 0.0f;
#pragma unroll
        for (int k = 0; k < 8; ++k) gmax = fmaxf(gmax, fabsf(v[u][k]));
        gmax = seg_max(gmax, lanes);
        const bool leader = valid && (c % lanes == 0);
        float* sp = p.scales + t * p.ss0 + (c / lanes) * p.ss1;
        if constexpr (FMT == FMT_MXFP4) {
          const int e = rceil_e8m0(__fdiv_rn(gmax, 6.0f));
          const float d = e8m0_descale(e);
          if (leader) *sp = e8m0_scale(e);
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const float cl = fminf(fmaxf(v[u][k] * d, -6.0f), 6.0f);
            // bf16 inputs scaled by a power of two are already bf16 values
            q[k] = (PRO == PRO_NONE && std::is_same<T, __nv_bfloat16>::value)
                       ? cl : __bfloat162float(__float2bfloat16_rn(cl));
          }
          if (valid) stg_stream(static_cast<uint8_t*>(p.out) + j * 4, pack_e2m1x8(q));
        } else if constexpr (FMT == FMT_MX8) {
          const int e = rceil_e8m0(__fdiv_rn(gmax, kQMax));
          const float d = e8m0_descale(e);
          if (leader) *sp = e8m0_scale(e);
#pragma unroll
          for (int k = 0; k < 8; ++k) q[k] = v[u][k] * d;
          if (valid) stg_stream(static_cast<uint8_t*>(p.out) + j * kVec, pack_q8<Q>(q));
        } else {
          float s;
          if constexpr (PRO == PRO_NONE) {
            // a4: s = max(amax / qmax, eps)
            s = fmaxf(__fdiv_rn(gmax, kQMax), p.eps);
          } else if constexpr (Q == VXK_Q_INT8) {
            s = __fdiv_rn(fmaxf(gmax, 1e-10f), 127.0f);
          } else {
            s = __fdiv_rn(fmaxf(gmax, FLT_MIN), kQMax);
            if (p.scale_in != nullptr) s = fminf(s, __ldg(p.scale_in));
            if (p.ue8m0) s = exp2_ceil_log2(fmaxf(s, 1e-10f));
          }
          const float r = __fdiv_rn(1.0f, s);
          if (leader) *sp = s;
#pragma unroll
          for (int k = 0; k < 8; ++k) q[k] = div_by(v[u][k], s, r);
          if (valid) stg_stream(static_cast<uint8_t*>(p.out) + j * kVec, pack_q8<Q>(q));
        }
      }
    }
  }
}

// Generic group kernel (any G that divides H, any alignment): one warp per group.
template <typename T, int Q, int PRO, int FMT, bool FAST>
__global__ void __launch_bounds__(256) flat_quant_generic(const FlatParams p, int64_t H, int G,
                                                          int64_t num_groups) {
  constexpr float kQMax = QTraits<Q>::kMax;
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  const int gpr = static_cast<int>(H / G);
  auto value = [&](int64_t t, int h) -> float {
    if constexpr (PRO == PRO_NONE) {
      return to_f32<T>(static_cast<const T*>(p.in)[t * H + h]);
    } else {
      const T* row = static_cast<const T*>(p.in) + t * 2 * H;
      const float g = to_f32<T>(row[h]), u = to_f32<T>(row[H + h]);
      if constexpr (PRO == PRO_SILU_T) return round_to<T>(round_to<T>(silu_f32<FAST>(g)) * u);
      else if constexpr (FAST) return silu_f32<true>(g) * u;
      else return (g * __fdiv_rn(1.0f, 1.0f + expf(-g))) * u;
    }
  };
  for (int64_t gi = warp0; gi < num_groups; gi += nwarps) {
    const int64_t t = gi / gpr;
    const int g = static_cast<int>(gi % gpr);
    float gmax = 0.0f;
    if constexpr (FMT != FMT_STATIC) {
      for (int h = g * G + lane; h < (g + 1) * G; h += 32) gmax = fmaxf(gmax, fabsf(value(t, h)));
      gmax = warp_max(gmax);
    }
    if constexpr (FMT == FMT_STATIC) {
      const float r = __fdiv_rn(1.0f, __ldg(p.scale_in));
      uint8_t* o = static_cast<uint8_t*>(p.out) + t * H;
      for (int h = g * G + lane; h < (g + 1) * G; h += 32) o[h] = pack_q1<Q>(value(t, h) * r);
      continue;
    }
    float* sp = p.scales + t * p.ss0 + g * p.ss1;
    if constexpr (FMT == FMT_MXFP4) {
      const int e = rceil_e8m0(__fdiv_rn(gmax, 6.0f));
      const float d = e8m0_descale(e);
      if (lane == 0) *sp = e8m0_scale(e);
      uint8_t* o = static_cast<uint8_t*>(p.out) + t * (H / 2);
      for (int h2 = g * G / 2 + lane; h2 < (g + 1) * G / 2; h2 += 32) {
        float c0 = fminf(fmaxf(value(t, 2 * h2) * d, -6.0f), 6.0f);
        float c1 = fminf(fmaxf(value(t, 2 * h2 + 1) * d, -6.0f), 6.0f);
        c0 = __bfloat162float(__float2bfloat16_rn(c0));
        c1 = __bfloat162float(__float2bfloat16_rn(c1));
        o[h2] = static_cast<uint8_t>(cvt_e2m1x2(c1, c0));
      }
    } else {
      float s, r;
      bool mul = false;
      if constexpr (FMT == FMT_MX8) {
        const int e = rceil_e8m0(__fdiv_rn(gmax, kQMax));
        s = e8m0_scale(e);
        r = e8m0_descale(e);
        mul = true;
      } else {
        if constexpr (PRO == PRO_NONE) s = fmaxf(__fdiv_rn(gmax, kQMax), p.eps);
        else if constexpr (Q == VXK_Q_INT8) s = __fdiv_rn(fmaxf(gmax, 1e-10f), 127.0f);
        else {
          s = __fdiv_rn(fmaxf(gmax, FLT_MIN), kQMax);
          if (p.scale_in != nullptr) s = fminf(s, __ldg(p.scale_in));
          if (p.ue8m0) s = exp2_ceil_log2(fmaxf(s, 1e-10f));
        }
        r = __fdiv_rn(1.0f, s);
      }
      if (lane == 0) *sp = s;
      uint8_t* o = static_cast<uint8_t*>(p.out) + t * H;
      for (int h = g * G + lane; h < (g + 1) * G; h += 32) {
        const float x = value(t, h);
        o[h] = pack_q1<Q>(mul ? x * r : div_by(x, s, r));
      }
    }
  }
}

// ---- a2: static quant with an arbitrary (group_m, group_n) scale grid ----
struct StaticParams {
  void* out;
  const void* in;
  const float* scale;
  int64_t T, H, in_row_stride, out_row_stride, group_m, group_n, ss_i, ss_j;
};

// vector path: group_n % 8 == 0 so one scale per vector
template <typename T, int Q>
__global__ void __launch_bounds__(256) static_quant_vec_kernel(const StaticParams p) {
  const int nvec = static_cast<int>(p.H / kVec);
  const int64_t total = p.T * nvec;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t j = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; j < total;
       j += 2 * stride) {
    float v[2][8];
    const int64_t j1 = j + stride;
    const int64_t t0 = j / nvec, t1 = j1 / nvec;
    const int c0 = static_cast<int>(j % nvec), c1 = static_cast<int>(j1 % nvec);
    load8<T, true>(static_cast<const T*>(p.in) + t0 * p.in_row_stride + c0 * kVec, v[0]);
    if (j1 < total) load8<T, true>(static_cast<const T*>(p.in) + t1 * p.in_row_stride + c1 * kVec, v[1]);
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int64_t t = u ? t1 : t0;
      const int c = u ? c1 : c0;
      if (u == 1 && j1 >= total) break;
      const float s = __ldg(p.scale + (t / p.group_m) * p.ss_i + ((c * kVec) / p.group_n) * p.ss_j);
      const float r = __fdiv_rn(1.0f, s);
      float q[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) q[k] = v[u][k] * r;
      stg_stream(static_cast<uint8_t*>(p.out) + t * p.out_row_stride + c * kVec, pack_q8<Q>(q));
    }
  }
}

template <typename T, int Q>
__global__ void __launch_bounds__(256) static_quant_scalar_kernel(const StaticParams p) {
  const int64_t total = p.T * p.H;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int64_t t = i / p.H, h = i % p.H;
    const float s = __ldg(p.scale + (t / p.group_m) * p.ss_i + (h / p.group_n) * p.ss_j);
    const float x = to_f32<T>(static_cast<const T*>(p.in)[t * p.in_row_stride + h]);
    static_cast<uint8_t*>(p.out)[t * p.out_row_stride + h] = pack_q1<Q>(x * __fdiv_rn(1.0f, s));
  }
}

// ---- a3: whole-tensor amax -> scale = amax / qmax (atomic max on the fp32 bits) ----
template <typename T, int Q>
__global__ void __launch_bounds__(256) tensor_amax_kernel(const T* __restrict__ in, float* scale,
                                                          int64_t T_, int64_t H, int64_t in_row_stride,
                                                          bool vec) {
  __shared__ float scratch[33];
  float amax = 0.0f;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  const int64_t i0 = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (vec) {
    const int nvec = static_cast<int>(H / kVec);
    const int64_t total = T_ * nvec;
    for (int64_t j = i0; j < total; j += stride) {
      float v[8];
      load8<T, true>(in + (j / nvec) * in_row_stride + (j % nvec) * kVec, v);
#pragma unroll
      for (int k = 0; k < 8; ++k) amax = fmaxf(amax, fabsf(v[k]));
    }
  } else {
    const int64_t total = T_ * H;
    for (int64_t i = i0; i < total; i += stride)
      amax = fmaxf(amax, fabsf(to_f32<T>(in[(i / H) * in_row_stride + (i % H)])));
  }
  amax = block_reduce<true>(amax, scratch);
  if (threadIdx.x == 0) {
    const float s = __fdiv_rn(amax, QTraits<Q>::kMax);  // non-negative: int order == float order
    atomicMax(reinterpret_cast<int*>(scale), __float_as_int(s));
  }
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
static int flat_grid(int64_t work_items, int per_block) {
  const int64_t blocks = (work_items + per_block - 1) / per_block;
  const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
  return static_cast<int>(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}

template <typename T, int Q, int PRO, int FMT>
static int launch_flat(FlatParams p, int64_t T_, int64_t H, int G, bool vec_ok, cudaStream_t st) {
  if (T_ == 0 || H == 0) return VXK_OK;
  const bool fast = (PRO != PRO_NONE) && g_silu_fast.load(std::memory_order_relaxed) != 0;
  if (vec_ok) {
    p.nvec = static_cast<int>(H / kVec);
    p.total = T_ * p.nvec;
    p.lanes = G / kVec;
    constexpr int U = (PRO == PRO_NONE) ? 2 : 1;
    const int grid = flat_grid(p.total, 256 * U);
    if (fast) flat_quant_kernel<T, Q, PRO, FMT, (PRO != PRO_NONE), U><<<grid, 256, 0, st>>>(p);
    else flat_quant_kernel<T, Q, PRO, FMT, false, U><<<grid, 256, 0, st>>>(p);
  } else {
    const int64_t ngroups = T_ * (H / G);
    const int grid = flat_grid(ngroups, 8);
    if (fast) flat_quant_generic<T, Q, PRO, FMT, (PRO != PRO_NONE)><<<grid, 256, 0, st>>>(p, H, G, ngroups);
    else flat_quant_generic<T, Q, PRO, FMT, false><<<grid, 256, 0, st>>>(p, H, G, ngroups);
  }
  return check_launch("flat_quant");
}

#define VXK_DISPATCH_T(in_dtype, ...)                                          \
  switch (in_dtype) {                                                          \
    case VXK_F32: { using T = float; __VA_ARGS__; }                            \
    case VXK_F16: { using T = __half; __VA_ARGS__; }                           \
    case VXK_BF16: { using T = __nv_bfloat16; __VA_ARGS__; }                   \
    default: return fail(VXK_ERR_INVALID, "unsupported input dtype %d", in_dtype); \
  }

#define VXK_DISPATCH_T16(in_dtype, ...)                                        \
  switch (in_dtype) {                                                          \
    case VXK_F16: { using T = __half; __VA_ARGS__; }                           \
    case VXK_BF16: { using T = __nv_bfloat16; __VA_ARGS__; }                   \
    default: return fail(VXK_ERR_INVALID, "input must be f16 / bf16 (got dtype %d)", in_dtype); \
  }

static bool group_vec_ok(const void* in, const void* out, int64_t H, int64_t G, int in_dtype,
                         bool packed4) {
  const int lanes = static_cast<int>(G / kVec);
  if (G % kVec || lanes > 32 || (lanes & (lanes - 1))) return false;
  if (!aligned16(in)) return false;
  if (in_dtype != VXK_F32 && (H * 2) % 16) return false;
  if (reinterpret_cast<uintptr_t>(out) % (packed4 ? 4 : 8)) return false;
  return H % kVec == 0;
}

}  // namespace vxk

using namespace vxk;

extern "C" int vxk_set_silu_fast(int on) {
  const int prev = g_silu_fast.exchange(on ? 1 : 0);
  return prev;
}

extern "C" int vxk_per_token_group_fp8_quant(const void* in, void* out_q, float* out_s, int64_t T_,
                                             int64_t H, int64_t G, double eps, int scale_ue8m0,
                                             int64_t ss0, int64_t ss1, int in_dtype, int q_dtype,
                                             vxk_stream_t stream) {
  VXK_REQUIRE(T_ >= 0 && H >= 0, "negative shape");
  VXK_REQUIRE(G > 0 && H % G == 0, "the last dimension (%lld) must be divisible by group_size (%lld)",
              (long long)H, (long long)G);
  VXK_REQUIRE(q_dtype == VXK_Q_E4M3 || q_dtype == VXK_Q_E5M2, "output must be fp8 (e4m3 / e5m2)");
  FlatParams p{};
  p.out = out_q; p.in = in; p.scales = out_s; p.ss0 = ss0; p.ss1 = ss1;
  p.eps = static_cast<float>(eps);
  const bool v = group_vec_ok(in, out_q, H, G, in_dtype, false);
  VXK_DISPATCH_T(in_dtype,
    if (scale_ue8m0) {
      if (q_dtype == VXK_Q_E4M3) return launch_flat<T, VXK_Q_E4M3, PRO_NONE, FMT_MX8>(p, T_, H, (int)G, v, stream);
      return launch_flat<T, VXK_Q_E5M2, PRO_NONE, FMT_MX8>(p, T_, H, (int)G, v, stream);
    }
    if (q_dtype == VXK_Q_E4M3) return launch_flat<T, VXK_Q_E4M3, PRO_NONE, FMT_FP8_DIV>(p, T_, H, (int)G, v, stream);
    return launch_flat<T, VXK_Q_E5M2, PRO_NONE, FMT_FP8_DIV>(p, T_, H, (int)G, v, stream));
}

extern "C" int vxk_per_token_group_quant_mxfp4(const void* in, void* out_q, float* out_s, int64_t T_,
                                               int64_t H, int64_t G, double eps, int64_t ss0,
                                               int64_t ss1, int in_dtype, vxk_stream_t stream) {
  (void)eps;  // the torch reference (to_mxfp) has no epsilon; an all-zero group gets exponent byte 0
  VXK_REQUIRE(T_ >= 0 && H >= 0, "negative shape");
  VXK_REQUIRE(G == 32, "MXFP4 requires group_size == 32, got %lld", (long long)G);
  VXK_REQUIRE(H % 32 == 0, "the last dimension (%lld) must be divisible by 32", (long long)H);
  FlatParams p{};
  p.out = out_q; p.in = in; p.scales = out_s; p.ss0 = ss0; p.ss1 = ss1;
  const bool v = group_vec_ok(in, out_q, H, G, in_dtype, true);
  VXK_DISPATCH_T(in_dtype,
    return launch_flat<T, VXK_Q_E4M3, PRO_NONE, FMT_MXFP4>(p, T_, H, 32, v, stream));
}

extern "C" int vxk_silu_and_mul_quant(void* out, const void* in, const float* scale, int64_t T_,
                                      int64_t d, int in_dtype, int q_dtype, vxk_stream_t stream) {
  VXK_REQUIRE(T_ >= 0 && d >= 0, "negative shape");
  VXK_REQUIRE(q_dtype == VXK_Q_E4M3 || q_dtype == VXK_Q_E5M2, "output must be fp8 (e4m3 / e5m2)");
  VXK_REQUIRE(in_dtype == VXK_F16 || in_dtype == VXK_BF16, "silu_and_mul_quant supports f16 / bf16 input");
  FlatParams p{};
  p.out = out; p.in = in; p.scale_in = scale;
  if (T_ == 0 || d == 0) return VXK_OK;
  // element-wise: reuse the flat kernel with a "group" of one vector; d % 8 != 0 -> generic G = 1
  const bool v = (d % kVec == 0) && aligned16(in) && (reinterpret_cast<uintptr_t>(out) % 8 == 0);
  const int G = v ? kVec : static_cast<int>(d);  // generic path: one warp per row
  VXK_DISPATCH_T16(in_dtype,
    if (q_dtype == VXK_Q_E4M3) return launch_flat<T, VXK_Q_E4M3, PRO_SILU_T, FMT_STATIC>(p, T_, d, G, v, stream);
    return launch_flat<T, VXK_Q_E5M2, PRO_SILU_T, FMT_STATIC>(p, T_, d, G, v, stream));
}

extern "C" int vxk_silu_and_mul_per_block_quant(void* out, const void* in, float* scales, int64_t T_,
                                                int64_t H, int64_t G, const float* scale_ub,
                                                int64_t ss0, int64_t ss1, int scale_ue8m0, int in_dtype,
                                                int q_dtype, vxk_stream_t stream) {
  VXK_REQUIRE(T_ >= 0 && H >= 0, "negative shape");
  VXK_REQUIRE(G == 32 || G == 64 || G == 128, "group_size must be 32, 64 or 128, got %lld", (long long)G);
  VXK_REQUIRE(H % G == 0, "hidden_size (%lld) must be divisible by group_size (%lld)", (long long)H,
              (long long)G);
  VXK_REQUIRE(in_dtype == VXK_F16 || in_dtype == VXK_BF16, "input must be f16 / bf16");
  VXK_REQUIRE(!(scale_ub && q_dtype == VXK_Q_INT8), "scale_ub is only supported for FP8 output");
  FlatParams p{};
  p.out = out; p.in = in; p.scales = scales; p.scale_in = scale_ub; p.ss0 = ss0; p.ss1 = ss1;
  p.ue8m0 = scale_ue8m0;
  const bool v = group_vec_ok(in, out, H, G, in_dtype, false);
  VXK_DISPATCH_T16(in_dtype,
    switch (q_dtype) {
      case VXK_Q_E4M3: return launch_flat<T, VXK_Q_E4M3, PRO_SILU_T, FMT_FP8_DIV>(p, T_, H, (int)G, v, stream);
      case VXK_Q_E5M2: return launch_flat<T, VXK_Q_E5M2, PRO_SILU_T, FMT_FP8_DIV>(p, T_, H, (int)G, v, stream);
      case VXK_Q_INT8: return launch_flat<T, VXK_Q_INT8, PRO_SILU_T, FMT_FP8_DIV>(p, T_, H, (int)G, v, stream);
      default: return fail(VXK_ERR_INVALID, "unsupported quant dtype %d", q_dtype);
    });
}

extern "C" int vxk_silu_and_mul_mxfp4_quant(void* out, const void* in, float* scales, int64_t T_,
                                            int64_t H, int64_t G, double eps, int64_t ss0, int64_t ss1,
                                            int in_dtype, vxk_stream_t stream) {
  (void)eps;
  VXK_REQUIRE(T_ >= 0 && H >= 0, "negative shape");
  VXK_REQUIRE(G == 32, "MXFP4 requires group_size == 32, got %lld", (long long)G);
  VXK_REQUIRE(H % 32 == 0, "hidden_size (%lld) must be divisible by 32", (long long)H);
  VXK_REQUIRE(in_dtype == VXK_F16 || in_dtype == VXK_BF16, "input must be f16 / bf16");
  FlatParams p{};
  p.out = out; p.in = in; p.scales = scales; p.ss0 = ss0; p.ss1 = ss1;
  const bool v = group_vec_ok(in, out, H, G, in_dtype, true);
  VXK_DISPATCH_T16(in_dtype,
    return launch_flat<T, VXK_Q_E4M3, PRO_SILU_F32, FMT_MXFP4>(p, T_, H, 32, v, stream));
}

template <typename T, int Q>
static int launch_static(const StaticParams& p, bool vec, cudaStream_t st) {
  if (vec) {
    const int grid = flat_grid(p.T * (p.H / kVec), 512);
    static_quant_vec_kernel<T, Q><<<grid, 256, 0, st>>>(p);
  } else {
    const int grid = flat_grid(p.T * p.H, 1024);
    static_quant_scalar_kernel<T, Q><<<grid, 256, 0, st>>>(p);
  }
  return check_launch("static_quant");
}

static bool static_vec_ok(const StaticParams& p, int in_dtype) {
  const size_t es = in_dtype == VXK_F32 ? 4 : 2;
  return p.H % kVec == 0 && p.group_n % kVec == 0 && aligned16(p.in) &&
         (p.in_row_stride * es) % 16 == 0 && reinterpret_cast<uintptr_t>(p.out) % 8 == 0 &&
         p.out_row_stride % 8 == 0;
}

extern "C" int vxk_static_scaled_fp8_quant(void* out, const void* in, const float* scale, int64_t T_,
                                           int64_t H, int64_t in_row_stride, int64_t out_row_stride,
                                           int64_t group_m, int64_t group_n, int64_t ss_i, int64_t ss_j,
                                           int in_dtype, int q_dtype, vxk_stream_t stream) {
  VXK_REQUIRE(T_ >= 0 && H >= 0, "negative shape");
  VXK_REQUIRE(q_dtype == VXK_Q_E4M3 || q_dtype == VXK_Q_E5M2, "output must be fp8 (e4m3 / e5m2)");
  if (T_ == 0 || H == 0) return VXK_OK;
  VXK_REQUIRE(group_m > 0 && group_n > 0, "group shape must be positive");
  StaticParams p{out, in, scale, T_, H, in_row_stride, out_row_stride, group_m, group_n, ss_i, ss_j};
  const bool v = static_vec_ok(p, in_dtype);
  VXK_DISPATCH_T(in_dtype,
    if (q_dtype == VXK_Q_E4M3) return launch_static<T, VXK_Q_E4M3>(p, v, stream);
    return launch_static<T, VXK_Q_E5M2>(p, v, stream));
}

extern "C" int vxk_dynamic_scaled_fp8_quant(void* out, const void* in, float* scale, int64_t T_,
                                            int64_t H, int64_t in_row_stride, int64_t out_row_stride,
                                            int in_dtype, int q_dtype, vxk_stream_t stream) {
  VXK_REQUIRE(T_ >= 0 && H >= 0, "negative shape");
  VXK_REQUIRE(q_dtype == VXK_Q_E4M3 || q_dtype == VXK_Q_E5M2, "output must be fp8 (e4m3 / e5m2)");
  cudaError_t e = cudaMemsetAsync(scale, 0, sizeof(float), stream);
  if (e != cudaSuccess) return fail(VXK_ERR_CUDA, "memset: %s", cudaGetErrorString(e));
  if (T_ == 0 || H == 0) return VXK_OK;
  StaticParams p{out, in, scale, T_, H, in_row_stride, out_row_stride, T_, H, 0, 0};
  const bool v = static_vec_ok(p, in_dtype);
  const int grid = flat_grid(v ? T_ * (H / kVec) : T_ * H, v ? 1024 : 2048);
  VXK_DISPATCH_T(in_dtype, {
    int rc;
    if (q_dtype == VXK_Q_E4M3) {
      tensor_amax_kernel<T, VXK_Q_E4M3><<<grid, 256, 0, stream>>>(static_cast<const T*>(in), scale, T_, H, in_row_stride, v);
      if ((rc = check_launch("tensor_amax"))) return rc;
      return launch_static<T, VXK_Q_E4M3>(p, v, stream);
    }
    tensor_amax_kernel<T, VXK_Q_E5M2><<<grid, 256, 0, stream>>>(static_cast<const T*>(in), scale, T_, H, in_row_stride, v);
    if ((rc = check_launch("tensor_amax"))) return rc;
    return launch_static<T, VXK_Q_E5M2>(p, v, stream);
  });
}
