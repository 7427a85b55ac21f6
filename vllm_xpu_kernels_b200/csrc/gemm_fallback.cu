// Shape-agnostic scaled GEMM on CUDA cores: the path for operand layouts / sizes the
// tcgen05 kernels do not cover (tiny or unaligned K/N, N-major B for FP4, exotic scale
// grids).  32x32 output tile per CTA, fp32 accumulation, same scale semantics as the
// tensor-core path.  Correctness net, not a performance path: vxk_gemm_stats() reports
// how often it is taken so a benchmark can never silently run on it.
#include "gemm_common.cuh"

namespace vxk {

std::atomic<uint64_t> g_gemm_tc_calls{0}, g_gemm_fallback_calls{0};

__device__ __forceinline__ float fp8_to_f32(uint8_t v, bool e5m2) {
  uint16_t h;
  if (e5m2) {
    h = static_cast<uint16_t>(v) << 8;  // e5m2 is the top byte of an fp16
  } else {
    uint16_t in = v;
    uint32_t out;
    asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(out) : "h"(in));
    h = static_cast<uint16_t>(out & 0xffffu);
  }
  return __half2float(__ushort_as_half(h));
}

__device__ __forceinline__ float e2m1_to_f32(uint32_t nib) {
  // 0, .5, 1, 1.5, 2, 3, 4, 6 with sign bit 3
  const float mag[8] = {0.0f, 0.5f, 1.0f, 1.5f, 2.0f, 3.0f, 4.0f, 6.0f};
  const float m = mag[nib & 7u];
  return (nib & 8u) ? -m : m;
}

__device__ __forceinline__ float e8m0_to_f32(uint8_t e) {
  return __uint_as_float(e == 0 ? 0x00400000u : (static_cast<uint32_t>(e) << 23));
}

__device__ __forceinline__ float load_scale(const void* p, int64_t idx, int scale_dtype) {
  if (scale_dtype == VXK_S_E8M0) return e8m0_to_f32(static_cast<const uint8_t*>(p)[idx]);
  return static_cast<const float*>(p)[idx];
}

__device__ __forceinline__ float load_bias(const void* b, int64_t n, int dt) {
  if (dt == VXK_F32) return static_cast<const float*>(b)[n];
  if (dt == VXK_F16) return __half2float(static_cast<const __half*>(b)[n]);
  return __bfloat162float(static_cast<const __nv_bfloat16*>(b)[n]);
}

__device__ __forceinline__ void store_out(void* D, int64_t idx, float v, int out_dtype) {
  if (out_dtype == VXK_BF16) static_cast<__nv_bfloat16*>(D)[idx] = __float2bfloat16_rn(v);
  else static_cast<__half*>(D)[idx] = __float2half_rn(v);
}

constexpr int kFT = 32;  // fallback tile

// IS_FP4: A/B hold two E2M1 codes per byte along K (even k in the low nibble)
template <bool IS_FP4>
__global__ void __launch_bounds__(kFT * kFT) gemm_fallback_kernel(const GemmParams p) {
  __shared__ float sa[kFT][kFT + 1], sb[kFT][kFT + 1];
  const int tx = threadIdx.x % kFT, ty = threadIdx.x / kFT;
  const int64_t m0 = static_cast<int64_t>(blockIdx.y) * kFT, n0 = static_cast<int64_t>(blockIdx.x) * kFT;
  const int64_t m = m0 + ty, n = n0 + tx;
  const uint8_t* A = static_cast<const uint8_t*>(p.A);
  const uint8_t* B = static_cast<const uint8_t*>(p.B);
  const bool e5m2 = p.fp8_dtype == VXK_Q_E5M2;
  // scale granularity along K: block mode -> group_k, else the whole K (scales applied at the end)
  const int64_t gk = (p.a_scale_mode == 2) ? p.group_k : p.K;
  float total = 0.0f;
  for (int64_t kb = 0; kb < p.K; kb += gk) {
    const int64_t kend = (kb + gk < p.K) ? kb + gk : p.K;
    float acc = 0.0f;
    for (int64_t k0 = kb; k0 < kend; k0 += kFT) {
      // A tile: rows m0.., cols k0..; thread (ty, tx) loads A[m0+ty][k0+tx]
      {
        const int64_t mm = m0 + ty, kk = k0 + tx;
        float v = 0.0f;
        if (mm < p.M && kk < kend) {
          if constexpr (IS_FP4) {
            const uint8_t byte = A[mm * p.lda + (kk >> 1)];
            v = e2m1_to_f32((kk & 1) ? (byte >> 4) : (byte & 0xf));
          } else {
            v = fp8_to_f32(A[mm * p.lda + kk], e5m2);
          }
        }
        sa[ty][tx] = v;
      }
      {  // B tile: B(k, n) for k = k0+ty, n = n0+tx
        const int64_t kk = k0 + ty, nn = n0 + tx;
        float v = 0.0f;
        if (nn < p.N && kk < kend) {
          if constexpr (IS_FP4) {
            const uint8_t byte = B[nn * p.b_stride_n + (kk >> 1)];  // W[N, K/2] rows
            v = e2m1_to_f32((kk & 1) ? (byte >> 4) : (byte & 0xf));
          } else {
            v = fp8_to_f32(B[kk * p.b_stride_k + nn * p.b_stride_n], e5m2);
          }
        }
        sb[ty][tx] = v;
      }
      __syncthreads();
#pragma unroll
      for (int i = 0; i < kFT; ++i) acc = fmaf(sa[ty][i], sb[i][tx], acc);
      __syncthreads();
    }
    if (p.a_scale_mode == 2 && m < p.M && n < p.N) {
      const int64_t kbi = kb / gk;
      float s;
      if constexpr (IS_FP4) {
        s = load_scale(p.a_scale, m * p.sa0 + kbi * p.sa1, VXK_S_E8M0) *
            load_scale(p.b_scale, n * p.sb0 + kbi * p.sb1, VXK_S_E8M0);
      } else {
        s = load_scale(p.a_scale, m * p.sa0 + kbi * p.sa1, p.scale_dtype) *
            load_scale(p.b_scale, kbi * p.sb0 + (n / p.group_n) * p.sb1, p.scale_dtype);
      }
      total = fmaf(acc, s, total);
    } else {
      total += acc;
    }
  }
  if (m < p.M && n < p.N) {
    float v = total;
    if (p.a_scale_mode != 2) {
      const float sA = p.a_scale_mode == 1 ? static_cast<const float*>(p.a_scale)[m]
                                           : static_cast<const float*>(p.a_scale)[0];
      const float sB = p.b_scale_mode == 1 ? static_cast<const float*>(p.b_scale)[n]
                                           : static_cast<const float*>(p.b_scale)[0];
      v = v * sA * sB;
    }
    if (p.bias) v += load_bias(p.bias, n, p.bias_dtype);
    store_out(p.D, m * p.ldd + n, v, p.out_dtype);
  }
}

int launch_gemm_fallback(const GemmParams& p, bool is_fp4, cudaStream_t st) {
  dim3 grid(static_cast<unsigned>((p.N + kFT - 1) / kFT), static_cast<unsigned>((p.M + kFT - 1) / kFT));
  if (grid.y > 65535) return fail(VXK_ERR_UNSUPPORTED, "fallback GEMM: M too large (%lld)", (long long)p.M);
  if (is_fp4) gemm_fallback_kernel<true><<<grid, kFT * kFT, 0, st>>>(p);
  else gemm_fallback_kernel<false><<<grid, kFT * kFT, 0, st>>>(p);
  g_gemm_fallback_calls.fetch_add(1, std::memory_order_relaxed);
  return check_launch("gemm_fallback");
}

}  // namespace vxk

extern "C" void vxk_gemm_stats(uint64_t* tensor_core_calls, uint64_t* fallback_calls) {
  if (tensor_core_calls) *tensor_core_calls = vxk::g_gemm_tc_calls.load();
  if (fallback_calls) *fallback_calls = vxk::g_gemm_fallback_calls.load();
}
