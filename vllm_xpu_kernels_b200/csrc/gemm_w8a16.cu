// Host side of fp8_gemm_w8a16 (16-bit activations x FP8 weights): eligibility checks, TMA
// tensor maps, tile-width choice, launch of gemm_w8a16_kernel (gemm_w8a16.cuh) or of the
// CUDA-core kernel below for layouts TMA cannot express (N-major W, unaligned rows).
#include "gemm_w8a16_skinny.cuh"

#include <mutex>

namespace vxk {

int run_w8a16_dense_tmem(void* D, const void* A, const void* W, const float* scale, int scale_mode, int64_t group_k,
                         int64_t group_n, int64_t sb0, int64_t sb1, const void* bias, int64_t M, int64_t N, int64_t K,
                         int64_t lda, int64_t ldw, int64_t ldd, bool bf16, bool e5m2, cudaStream_t st);

// ---- CUDA-core path: any strides, any K / N; 32 x 32 outputs per CTA, fp32 accumulation ----
struct W16FallbackParams {
  void* D; const void* A; const uint8_t* B; const float* b_scale; const void* bias;
  int64_t M, N, K, lda, b_stride_k, b_stride_n, ldd;
  int b_scale_mode; int64_t group_k, group_n, sb0, sb1;
  int act_dtype, fp8_dtype, bias_dtype;
};

__device__ __forceinline__ float w16_fp8_to_f32(uint8_t v, bool e5m2) {
  uint16_t h;
  if (e5m2) {
    h = static_cast<uint16_t>(v) << 8;
  } else {
    uint32_t out;
    const uint16_t in = v;
    asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(out) : "h"(in));
    h = static_cast<uint16_t>(out & 0xffffu);
  }
  return __half2float(__ushort_as_half(h));
}

__global__ void __launch_bounds__(1024) gemm_w8a16_fallback_kernel(const W16FallbackParams p) {
  __shared__ float sa[32][33], sb[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t m0 = static_cast<int64_t>(blockIdx.y) * 32, n0 = static_cast<int64_t>(blockIdx.x) * 32;
  const int64_t m = m0 + ty, n = n0 + tx;
  const bool bf = p.act_dtype == VXK_BF16, e5m2 = p.fp8_dtype == VXK_Q_E5M2;
  float acc = 0.0f;
  for (int64_t k0 = 0; k0 < p.K; k0 += 32) {
    {
      const int64_t kk = k0 + tx;
      float v = 0.0f;
      if (m < p.M && kk < p.K)
        v = bf ? __bfloat162float(static_cast<const __nv_bfloat16*>(p.A)[m * p.lda + kk])
               : __half2float(static_cast<const __half*>(p.A)[m * p.lda + kk]);
      sa[ty][tx] = v;
    }
    {
      const int64_t kk = k0 + ty;
      float v = 0.0f;
      if (n < p.N && kk < p.K) {
        v = w16_fp8_to_f32(p.B[kk * p.b_stride_k + n * p.b_stride_n], e5m2);
        if (p.b_scale_mode == 2) {   // same single rounding to the activation type as the tensor-core path
          v *= p.b_scale[(kk / p.group_k) * p.sb0 + (n / p.group_n) * p.sb1];
          v = bf ? __bfloat162float(__float2bfloat16_rn(v)) : __half2float(__float2half_rn(v));
        }
      }
      sb[ty][tx] = v;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 32; ++i) acc = fmaf(sa[ty][i], sb[i][tx], acc);
    __syncthreads();
  }
  if (m < p.M && n < p.N) {
    if (p.b_scale_mode == 0) acc *= p.b_scale[0];
    else if (p.b_scale_mode == 1) acc *= p.b_scale[n];
    if (p.bias)
      acc += p.bias_dtype == VXK_F32   ? static_cast<const float*>(p.bias)[n]
             : p.bias_dtype == VXK_F16 ? __half2float(static_cast<const __half*>(p.bias)[n])
                                       : __bfloat162float(static_cast<const __nv_bfloat16*>(p.bias)[n]);
    if (bf) static_cast<__nv_bfloat16*>(p.D)[m * p.ldd + n] = __float2bfloat16_rn(acc);
    else static_cast<__half*>(p.D)[m * p.ldd + n] = __float2half_rn(acc);
  }
}

// ---- tensor-core path ----
template <int BLOCK_N, bool BF16, bool E5M2, bool SCALED>
static int launch_w16(const W16Params& wp, const CUtensorMap& ma, const CUtensorMap& mw, cudaStream_t st) {
  using L = W16Smem<BLOCK_N>;
  auto kernel = gemm_w8a16_kernel<BLOCK_N, BF16, E5M2, SCALED>;
  static PerDeviceOnce once;
  const cudaError_t attr_err = once.get([&](PerDeviceOnce::Entry& x) {
    x.err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotalBytes);
  }).err;
  if (attr_err != cudaSuccess)
    return fail(VXK_ERR_CUDA, "cudaFuncSetAttribute(smem=%d): %s", L::kTotalBytes, cudaGetErrorString(attr_err));
  const int tiles = wp.b.num_m_blocks * wp.b.num_n_blocks;
  const int grid = tiles < sm_count() ? tiles : sm_count();
  kernel<<<grid, kW16Threads, L::kTotalBytes, st>>>(wp, ma, mw);
  g_gemm_tc_calls.fetch_add(1, std::memory_order_relaxed);
  return check_launch("gemm_w8a16");
}

template <int BLOCK_N>
static int launch_w16_types(const W16Params& wp, const CUtensorMap& ma, const CUtensorMap& mw, bool bf16, bool e5m2,
                            bool scaled, cudaStream_t st) {
#define VXK_W16_CASE(B, E, S) \
  if (bf16 == B && e5m2 == E && scaled == S) return launch_w16<BLOCK_N, B, E, S>(wp, ma, mw, st);
  VXK_W16_CASE(true, false, false) VXK_W16_CASE(true, false, true)
  VXK_W16_CASE(true, true, false) VXK_W16_CASE(true, true, true)
  VXK_W16_CASE(false, false, false) VXK_W16_CASE(false, false, true)
  VXK_W16_CASE(false, true, false) VXK_W16_CASE(false, true, true)
#undef VXK_W16_CASE
  return fail(VXK_ERR_INVALID, "unreachable");
}

// Widest tile of {256, 128, 64, 32} that still gives every SM one: small-M calls are bound by
// streaming W from HBM, so all SMs must pull on it.
static int w16_bn(int64_t M, int64_t N) {
  const int64_t mt = (M + kBlockM - 1) / kBlockM, sms = sm_count();
  const int cand[4] = {256, 128, 64, 32};
  for (int i = 0; i < 4; ++i)
    if (N >= cand[i] && mt * ((N + cand[i] - 1) / cand[i]) >= sms) return cand[i];
  return 32;
}

static bool w16_force_fallback() {
  static const bool v = [] { const char* e = getenv("VXK_GEMM_FORCE_FALLBACK"); return e && e[0] == '1'; }();
  return v;
}

// ---- small-M path (M <= 128): operands swapped, K split over a cluster.  See gemm_w8a16_skinny.cuh ----
template <int MT, bool BF16, bool E5M2, bool SCALED, bool TA>
static int launch_w16_skinny_impl(const W16SkinnyParams& sp, const CUtensorMap& mw, const CUtensorMap& ma, int n_tiles,
                             cudaStream_t st) {
  using L = W16SkinnySmem<MT, TA>;
  auto kernel = gemm_w8a16_skinny_kernel<MT, BF16, E5M2, SCALED, TA>;
  static PerDeviceOnce once;
  const cudaError_t attr_err = once.get([&](PerDeviceOnce::Entry& x) {
    x.err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotalBytes);
  }).err;
  if (attr_err != cudaSuccess)
    return fail(VXK_ERR_CUDA, "cudaFuncSetAttribute(smem=%d): %s", L::kTotalBytes, cudaGetErrorString(attr_err));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(n_tiles, sp.splits, sp.m_tiles > 1 ? sp.m_tiles : 1);
  cfg.blockDim = dim3(kW16SkThreads);
  cfg.dynamicSmemBytes = L::kTotalBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1;
  attr[0].val.clusterDim.y = sp.splits;
  attr[0].val.clusterDim.z = 1;
  const int pdl = g_early_weight_loads.load(std::memory_order_relaxed);
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 2 : 1;
  const cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, sp, mw, ma);
  if (e != cudaSuccess) return fail(VXK_ERR_CUDA, "gemm_w8a16_skinny launch: %s", cudaGetErrorString(e));
  g_gemm_tc_calls.fetch_add(1, std::memory_order_relaxed);
  return check_launch("gemm_w8a16_skinny");
}

// widened operand in tensor memory (default) or in shared memory (VXK_W16_TMEM_A=0): see gemm_w8a16_skinny.cuh
template <int MT, bool BF16, bool E5M2, bool SCALED>
static int launch_w16_skinny(const W16SkinnyParams& sp, const CUtensorMap& mw, const CUtensorMap& ma, int n_tiles,
                             cudaStream_t st) {
  static const int ta = [] { const char* e = getenv("VXK_W16_TMEM_A"); return e ? atoi(e) : 1; }();
  return ta ? launch_w16_skinny_impl<MT, BF16, E5M2, SCALED, true>(sp, mw, ma, n_tiles, st)
            : launch_w16_skinny_impl<MT, BF16, E5M2, SCALED, false>(sp, mw, ma, n_tiles, st);
}

template <int MT>
static int launch_w16_skinny_types(const W16SkinnyParams& sp, const CUtensorMap& mw, const CUtensorMap& ma, int n_tiles,
                                   bool bf16, bool e5m2, cudaStream_t st) {
  const bool scaled = sp.b_scale_mode == 2;
#define VXK_W16S_CASE(B, E, S) \
  if (bf16 == B && e5m2 == E && scaled == S) return launch_w16_skinny<MT, B, E, S>(sp, mw, ma, n_tiles, st);
  VXK_W16S_CASE(true, false, false) VXK_W16S_CASE(true, false, true)
  VXK_W16S_CASE(true, true, false) VXK_W16S_CASE(true, true, true)
  VXK_W16S_CASE(false, false, false) VXK_W16S_CASE(false, false, true)
  VXK_W16S_CASE(false, true, false) VXK_W16S_CASE(false, true, true)
#undef VXK_W16S_CASE
  return fail(VXK_ERR_INVALID, "unreachable");
}

static int run_w16_skinny(W16SkinnyParams sp, const void* A, const void* W, int64_t M, int64_t N, int64_t K,
                          int64_t lda, int64_t ldw, bool bf16, bool e5m2, cudaStream_t st) {
  const int n_tiles = (int)((N + 127) / 128);
  sp.num_k_blocks = (int)(K / kW16SkBlockK);
  // K split: enough CTAs for one wave, every rank at least two raw stages, cluster sizes 1 / 2 / 4 / 8
  static const int env_splits = [] { const char* e = getenv("VXK_W16_SPLITS"); return e ? atoi(e) : 0; }();
  int splits = sm_count() / (n_tiles * (sp.m_tiles > 1 ? sp.m_tiles : 1));
  if (splits > 8) splits = 8;
  if (env_splits > 0) splits = env_splits > 8 ? 8 : env_splits;
  if (splits > sp.num_k_blocks / 2) splits = sp.num_k_blocks / 2;
  if (splits < 1) splits = 1;
  while (splits & (splits - 1)) --splits;
  const int per = (sp.num_k_blocks + splits - 1) / splits;
  splits = (sp.num_k_blocks + per - 1) / per;            // no empty rank
  sp.kb_per_split = per;
  sp.splits = splits;
  // L2 prefetch depth (raw stages ahead of the ring).  Measured 0 / 6 / 12 / 24: every depth > 0 is slower
  // (39.6 / 40.4 / 42.0 / 44.4 us at N = 28672) -- the kernel is not bound by HBM latency but by the SM-side
  // widening, and the extra requests only add contention.  Kept as a switch, off by default.
  static const int env_pf = [] { const char* e = getenv("VXK_W16_PREFETCH"); return e ? atoi(e) : 0; }();
  sp.prefetch = env_pf;
  const int mt = M <= 32 ? 32 : (M <= 64 ? 64 : 128);   // M > 128 only as sp.m_tiles tiles of 128
  CUtensorMap mw, ma;
  int rc;
  if ((rc = make_map_2d_u8(&mw, W, K, N, ldw, kW16SkBlockK, 128, true))) return rc;
  if ((rc = make_map_2d_u8(&ma, A, static_cast<uint64_t>(K) * 2, M, static_cast<uint64_t>(lda) * 2, 128, mt, true)))
    return rc;
  switch (mt) {
    case 32: return launch_w16_skinny_types<32>(sp, mw, ma, n_tiles, bf16, e5m2, st);
    case 64: return launch_w16_skinny_types<64>(sp, mw, ma, n_tiles, bf16, e5m2, st);
    default: return launch_w16_skinny_types<128>(sp, mw, ma, n_tiles, bf16, e5m2, st);
  }
}

}  // namespace vxk

using namespace vxk;

extern "C" int vxk_fp8_gemm_w8a16(void* D, const void* A, const void* B, const float* b_scale, const void* bias,
                                  int64_t M, int64_t N, int64_t K, int64_t lda, int64_t b_stride_k,
                                  int64_t b_stride_n, int64_t ldd, int b_scale_mode, int64_t group_k, int64_t group_n,
                                  int64_t sb0, int64_t sb1, int act_dtype, int fp8_dtype, int bias_dtype,
                                  int force_fallback, vxk_stream_t stream) {
  VXK_REQUIRE(M >= 0 && N >= 0 && K >= 0, "negative shape");
  VXK_REQUIRE(act_dtype == VXK_BF16 || act_dtype == VXK_F16, "activations must be bf16 or f16");
  VXK_REQUIRE(fp8_dtype == VXK_Q_E4M3 || fp8_dtype == VXK_Q_E5M2, "weight must be e4m3 or e5m2");
  VXK_REQUIRE(b_scale, "scale pointer must not be null (pass a one-element tensor of 1.0)");
  VXK_REQUIRE(b_scale_mode >= 0 && b_scale_mode <= 2, "b_scale_mode must be 0, 1 or 2");
  if (b_scale_mode == 2) VXK_REQUIRE(group_k > 0 && group_n > 0, "block scale groups must be positive");
  if (M == 0 || N == 0) return VXK_OK;
  cudaStream_t st = stream;

  const bool tc_ok = !force_fallback && !w16_force_fallback() && b_stride_k == 1 && aligned16(A) &&
                     (lda * 2) % 16 == 0 && lda > 0 && aligned16(B) && b_stride_n % 16 == 0 && b_stride_n > 0 &&
                     K >= 16 && M < (1LL << 31) && N < (1LL << 31) && K < (1LL << 30) &&
                     (b_scale_mode != 2 || group_k % kW16BlockK == 0);
  if (!tc_ok) {
    W16FallbackParams fp{D, A, static_cast<const uint8_t*>(B), b_scale, bias, M, N, K, lda, b_stride_k, b_stride_n,
                         ldd, b_scale_mode, group_k, group_n, sb0, sb1, act_dtype, fp8_dtype, bias_dtype};
    dim3 grid(static_cast<unsigned>((N + 31) / 32), static_cast<unsigned>((M + 31) / 32));
    if (grid.y > 65535) return fail(VXK_ERR_UNSUPPORTED, "fallback w8a16 GEMM: M too large (%lld)", (long long)M);
    gemm_w8a16_fallback_kernel<<<grid, 1024, 0, st>>>(fp);
    g_gemm_fallback_calls.fetch_add(1, std::memory_order_relaxed);
    return check_launch("gemm_w8a16_fallback");
  }

  // M <= 128: the small-M kernel.  129..256 tokens: two passes of it over 128-token halves -- the weight is
  // streamed twice, which still beats the general kernel's thin in-flight window at this M while the weight
  // tiles fit one wave (measured: 26 vs 37 us at N = K = 4096, but 138 vs 94 us at N = 28672).
  static const int skinny_max_m = [] { const char* e = getenv("VXK_W16_SKINNY_MAX_M"); return e ? atoi(e) : 256; }();
  static const int tiles_on = [] { const char* e = getenv("VXK_W16_MTILES"); return e ? atoi(e) : 1; }();
  const bool w16_tiles_fit = tiles_on && M <= 512 && ((N + 127) / 128) * ((M + 127) / 128) <= sm_count();
  if (((M <= skinny_max_m && M <= 256 && (M <= 128 || (N + 127) / 128 <= sm_count())) ||
       (M > 256 && M <= 2 * skinny_max_m && w16_tiles_fit)) &&
      K % kW16SkBlockK == 0 && K >= 512 && (b_scale_mode != 2 || group_k % kW16SkBlockK == 0)) {
    // token tiles side by side in one launch while they fit one wave (the weight comes from HBM once and from L2
    // for the other tiles, and no launch waits for the previous one); else the sequential passes below
    const int64_t tiles = (M + 127) / 128;
    const bool side_by_side = tiles > 1 && w16_tiles_fit;
    for (int64_t m0 = 0; m0 < M; m0 += side_by_side ? M : 128) {
      const int64_t mc = side_by_side ? M : (M - m0 < 128 ? M - m0 : 128);
      W16SkinnyParams sp{};
      sp.D = static_cast<uint16_t*>(D) + m0 * ldd; sp.b_scale = b_scale; sp.bias = bias; sp.M = (int)mc; sp.N = (int)N;
      sp.ldd = ldd; sp.b_scale_mode = b_scale_mode; sp.out_dtype = act_dtype; sp.bias_dtype = bias_dtype;
      sp.sb0 = sb0; sp.sb1 = sb1;
      sp.group_k = b_scale_mode == 2 ? (int)group_k : 1; sp.group_n = b_scale_mode == 2 ? (int)group_n : 1;
      sp.m_tiles = side_by_side ? static_cast<int>(tiles) : 1;
      const int rc = run_w16_skinny(sp, static_cast<const uint16_t*>(A) + m0 * lda, B, mc, N, K, lda, b_stride_n,
                                    act_dtype == VXK_BF16, fp8_dtype == VXK_Q_E5M2, st);
      if (rc) return rc;
    }
    return VXK_OK;
  }

  // Large M: the swap-AB kernel with the widened weight tile in TENSOR memory (gemm_grouped.cuh, GW_FP8T, one
  // "expert").  The kernel below keeps tokens on the 128-row side and the widened 256-column weight tile in shared
  // memory: per 64-deep step that is 128 KB of shared-memory traffic (TMA writes 32, converter reads 16 + writes
  // 32, MMA reads 48) in the 512 clk the MMAs take = 250 B/clk against the SM's 128 B/clk -- shared-memory bound
  // at 0.51 of the bf16 peak, which is where it measured (0.53-0.56).  With the weights on the 128-row side in TMEM
  // and 256 tokens as the N operand the same step moves 80 KB (156 B/clk).
  static const int dense_ta = [] { const char* e = getenv("VXK_W16_DENSE_TMEM_A"); return e ? atoi(e) : 1; }();
  const int want_bias_dtype = act_dtype == VXK_BF16 ? VXK_BF16 : VXK_F16;
  // (VXK_W16_DENSE_TMEM_A=2 keeps block-scaled weights on the shared-memory kernel, for A/B timing)
  const bool blk_ok = b_scale_mode != 2 || (group_k % 64 == 0 && dense_ta != 2);
  if (dense_ta && blk_ok && K % 64 == 0 && (!bias || bias_dtype == want_bias_dtype) && ldd % 2 == 0 && N % 2 == 0 && (reinterpret_cast<uintptr_t>(D) & 3) == 0) {
    const int rc = run_w8a16_dense_tmem(D, A, B, b_scale, b_scale_mode, group_k, group_n, sb0, sb1, bias, M, N, K, lda,
                                        b_stride_n, ldd, act_dtype == VXK_BF16, fp8_dtype == VXK_Q_E5M2, st);
    return rc;
  }

  W16Params wp{};
  TcParams& tp = wp.b;
  tp.D = D; tp.bias = bias; tp.M = (int)M; tp.N = (int)N; tp.K_bytes = (int)K; tp.ldd = ldd;
  tp.a_scale_mode = 0; tp.b_scale_mode = b_scale_mode; tp.out_dtype = act_dtype; tp.bias_dtype = bias_dtype;
  tp.fp8_fmt = fp8_dtype == VXK_Q_E5M2 ? 1 : 0;
  tp.b_scale = b_scale;
  tp.num_m_blocks = (int)((M + kBlockM - 1) / kBlockM);
  tp.num_k_blocks = (int)((K + kW16BlockK - 1) / kW16BlockK);
  tp.vec_store_ok = aligned16(D) && (ldd * 2) % 16 == 0;
  wp.blk_scale = b_scale; wp.sb0 = sb0; wp.sb1 = sb1;
  wp.group_k = b_scale_mode == 2 ? (int)group_k : 1; wp.group_n = b_scale_mode == 2 ? (int)group_n : 1;
  wp.K = (int)K;
  static const int env_bn = [] { const char* e = getenv("VXK_W16_BN"); return e ? atoi(e) : 0; }();
  const int bn = env_bn ? env_bn : w16_bn(M, N);
  tp.num_n_blocks = (int)((N + bn - 1) / bn);
  CUtensorMap ma, mw;
  int rc;
  if ((rc = make_map_2d_u8(&ma, A, static_cast<uint64_t>(K) * 2, M, static_cast<uint64_t>(lda) * 2, 128, kBlockM, true)))
    return rc;
  if ((rc = make_map_2d_u8(&mw, B, K, N, b_stride_n, kW16BlockK, bn, false))) return rc;
  const bool bf16 = act_dtype == VXK_BF16, e5m2 = fp8_dtype == VXK_Q_E5M2, scaled = b_scale_mode == 2;
  switch (bn) {
    case 256: return launch_w16_types<256>(wp, ma, mw, bf16, e5m2, scaled, st);
    case 128: return launch_w16_types<128>(wp, ma, mw, bf16, e5m2, scaled, st);
    case 64: return launch_w16_types<64>(wp, ma, mw, bf16, e5m2, scaled, st);
    default: return launch_w16_types<32>(wp, ma, mw, bf16, e5m2, scaled, st);
  }
}
