// Host side of the MoE grouped GEMM (vxk_moe_grouped_gemm): argument checks, TMA tensor maps (3-D over the
// expert-stacked weights), token-tile width from the average rows per expert, launch of gemm_grouped_kernel
// (gemm_grouped.cuh) or of the CUDA-core kernel below for shapes TMA cannot express.
#include "gemm_grouped.cuh"

namespace vxk {

static int gg_env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return e ? atoi(e) : dflt;
}

int make_map_3d_u8(CUtensorMap* map, const void* base, uint64_t d0_bytes, uint64_t d1, uint64_t d2, uint64_t pitch1,
                   uint64_t pitch2, uint32_t box0, uint32_t box1, int swizzle_bytes);

// ---- CUDA-core path: one thread per output, any N / K; also the cross-check of the tensor-core kernel ----
struct GgFallbackParams {
  void* D; const void* A; const void* B; const void* bias; const float* w_scale; const uint8_t* w_sf;
  const int32_t* rows;
  int64_t total_m, N, K, ldd;
  int E, wkind, bf16, e5m2;
};

__device__ __forceinline__ float gg_to_f32(const void* p, int64_t i, bool bf) {
  return bf ? __bfloat162float(static_cast<const __nv_bfloat16*>(p)[i]) : __half2float(static_cast<const __half*>(p)[i]);
}
__device__ __forceinline__ float gg_fp8_to_f32(uint8_t v, bool e5m2) {
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
__device__ __forceinline__ float gg_round16(float v, bool bf) {
  return bf ? __bfloat162float(__float2bfloat16_rn(v)) : __half2float(__float2half_rn(v));
}

__global__ void __launch_bounds__(256) gemm_grouped_fallback_kernel(const GgFallbackParams p) {
  const int64_t n = static_cast<int64_t>(blockIdx.x) * 64 + (threadIdx.x & 63);
  const int64_t m = static_cast<int64_t>(blockIdx.y) * 4 + (threadIdx.x >> 6);
  if (m >= p.total_m || n >= p.N) return;
  int e = -1;
  int64_t acc_rows = 0;
  for (int i = 0; i < p.E; ++i) {
    const int r = p.rows[i] < 0 ? 0 : p.rows[i];
    if (m < acc_rows + r) { e = i; break; }
    acc_rows += r;
  }
  if (e < 0) return;                       // rows past sum(rows_per_expert) are left untouched, like the reference
  const bool bf = p.bf16 != 0;
  float acc = 0.0f;
  const float kE2M1[8] = {0.0f, 0.5f, 1.0f, 1.5f, 2.0f, 3.0f, 4.0f, 6.0f};
  for (int64_t k = 0; k < p.K; ++k) {
    const float a = gg_to_f32(p.A, m * p.K + k, bf);
    float w;
    if (p.wkind == GW_16) {
      w = gg_to_f32(p.B, (static_cast<int64_t>(e) * p.K + k) * p.N + n, bf);
    } else if (p.wkind == GW_FP8) {
      const uint8_t c = static_cast<const uint8_t*>(p.B)[(static_cast<int64_t>(e) * p.K + k) * p.N + n];
      w = gg_fp8_to_f32(c, p.e5m2 != 0);
    } else {
      const uint8_t byte = static_cast<const uint8_t*>(p.B)[(static_cast<int64_t>(e) * p.N + n) * (p.K / 2) + k / 2];
      const uint32_t code = (k & 1) ? (byte >> 4) : (byte & 15);
      const float tv = (code & 8) ? -kE2M1[code & 7] : kE2M1[code & 7];
      const uint32_t s = p.w_sf[(static_cast<int64_t>(e) * p.N + n) * (p.K / 32) + k / 32];
      // the reference test's dequantisation: table value and 2^(s-127) both in the activation type, product rounded
      const float sc = gg_round16(__uint_as_float(s << 23), bf);
      w = gg_round16(tv * sc, bf);
    }
    acc = fmaf(a, w, acc);
  }
  if (p.wkind == GW_FP8) acc *= p.w_scale[e];
  if (p.bias) acc += gg_to_f32(p.bias, static_cast<int64_t>(e) * p.N + n, bf);
  if (bf) static_cast<__nv_bfloat16*>(p.D)[m * p.ldd + n] = __float2bfloat16_rn(acc);
  else static_cast<__half*>(p.D)[m * p.ldd + n] = __float2half_rn(acc);
}

// ---- tensor-core path ----
template <int MT, int WKIND, bool BF16, bool E5M2>
static int launch_gg(const GroupedParams& gp, const CUtensorMap& mw, const CUtensorMap& ma, int64_t max_tiles,
                     cudaStream_t st) {
  using L = GgSmem<MT, WKIND>;
  auto kernel = gemm_grouped_kernel<MT, WKIND, BF16, E5M2>;
  static PerDeviceOnce once;
  const cudaError_t attr_err = once.get([&](PerDeviceOnce::Entry& x) {
    x.err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotalBytes);
  }).err;
  if (attr_err != cudaSuccess)
    return fail(VXK_ERR_CUDA, "cudaFuncSetAttribute(smem=%d): %s", L::kTotalBytes, cudaGetErrorString(attr_err));
  const int grid = max_tiles < sm_count() ? static_cast<int>(max_tiles) : sm_count();
  kernel<<<grid, WKIND == GW_16 ? 256 : kGgConvThreads, L::kTotalBytes, st>>>(gp, mw, ma);
  g_gemm_tc_calls.fetch_add(1, std::memory_order_relaxed);
  return check_launch("gemm_grouped");
}

template <int MT>
static int launch_gg_types(const GroupedParams& gp, const CUtensorMap& mw, const CUtensorMap& ma, int64_t max_tiles,
                           int wkind, bool bf16, bool e5m2, cudaStream_t st) {
  if (wkind == GW_16)
    return bf16 ? launch_gg<MT, GW_16, true, false>(gp, mw, ma, max_tiles, st)
                : launch_gg<MT, GW_16, false, false>(gp, mw, ma, max_tiles, st);
  if (wkind == GW_MX4)
    return bf16 ? launch_gg<MT, GW_MX4, true, false>(gp, mw, ma, max_tiles, st)
                : launch_gg<MT, GW_MX4, false, false>(gp, mw, ma, max_tiles, st);
  if (wkind == GW_MX4T)
    return bf16 ? launch_gg<MT, GW_MX4T, true, false>(gp, mw, ma, max_tiles, st)
                : launch_gg<MT, GW_MX4T, false, false>(gp, mw, ma, max_tiles, st);
  if (wkind == GW_FP8T) {
    if (bf16)
      return e5m2 ? launch_gg<MT, GW_FP8T, true, true>(gp, mw, ma, max_tiles, st)
                  : launch_gg<MT, GW_FP8T, true, false>(gp, mw, ma, max_tiles, st);
    return e5m2 ? launch_gg<MT, GW_FP8T, false, true>(gp, mw, ma, max_tiles, st)
                : launch_gg<MT, GW_FP8T, false, false>(gp, mw, ma, max_tiles, st);
  }
  if (bf16)
    return e5m2 ? launch_gg<MT, GW_FP8, true, true>(gp, mw, ma, max_tiles, st)
                : launch_gg<MT, GW_FP8, true, false>(gp, mw, ma, max_tiles, st);
  return e5m2 ? launch_gg<MT, GW_FP8, false, true>(gp, mw, ma, max_tiles, st)
              : launch_gg<MT, GW_FP8, false, false>(gp, mw, ma, max_tiles, st);
}

// Dense fp8_gemm_w8a16 at large M through the grouped kernel (one expert owning every row, weights W[N, K]
// K-contiguous, widened operand in TMEM, 256-token tiles): see gemm_w8a16.cu for why.
int run_w8a16_dense_tmem(void* D, const void* A, const void* W, const float* scale, int scale_mode, int64_t group_k,
                         int64_t group_n, int64_t sb0, int64_t sb1, const void* bias, int64_t M, int64_t N, int64_t K,
                         int64_t lda, int64_t ldw, int64_t ldd, bool bf16, bool e5m2, cudaStream_t st) {
  const int mt = M <= 32 ? 32 : (M <= 64 ? 64 : (M <= 128 ? 128 : 256));
  const int block_k = mt <= 128 ? 128 : kGgBlockK;
  GroupedParams gp{};
  gp.D = D; gp.bias = bias; gp.rows = nullptr; gp.w_scale = scale; gp.scale_per_channel = scale_mode == 1;
  if (scale_mode == 2) {   // 2-D block scales: folded into the widening
    gp.blk_scale = scale; gp.w_scale = nullptr;
    gp.group_k = static_cast<int>(group_k); gp.group_n = static_cast<int>(group_n); gp.sb0 = sb0; gp.sb1 = sb1;
  }
  gp.total_m = static_cast<int>(M); gp.N = static_cast<int>(N); gp.K = static_cast<int>(K); gp.E = 1;
  gp.num_kb = static_cast<int>((K + block_k - 1) / block_k);
  gp.num_nt = static_cast<int>((N + 127) / 128);
  gp.ldd = ldd;
  const int64_t max_tiles = ((M + mt - 1) / mt) * gp.num_nt;
  CUtensorMap mw, ma;
  int rc;
  if ((rc = make_map_2d_u8(&ma, A, static_cast<uint64_t>(K) * 2, M, static_cast<uint64_t>(lda) * 2, 128, mt, true))) return rc;
  if ((rc = make_map_3d_u8(&mw, W, K, N, 1, ldw, N * ldw, block_k, 128, block_k))) return rc;
  switch (mt) {
    case 32: return launch_gg_types<32>(gp, mw, ma, max_tiles, GW_FP8T, bf16, e5m2, st);
    case 64: return launch_gg_types<64>(gp, mw, ma, max_tiles, GW_FP8T, bf16, e5m2, st);
    case 128: return launch_gg_types<128>(gp, mw, ma, max_tiles, GW_FP8T, bf16, e5m2, st);
    default: return launch_gg_types<256>(gp, mw, ma, max_tiles, GW_FP8T, bf16, e5m2, st);
  }
}

}  // namespace vxk

using namespace vxk;

extern "C" int vxk_moe_grouped_gemm(void* D, const void* A, const void* B, const void* scales, const void* bias,
                                    const int32_t* rows_per_expert, int64_t total_m, int64_t N, int64_t K,
                                    int64_t num_experts, int act_dtype, int weight_kind, int force_fallback,
                                    vxk_stream_t stream) {
  VXK_REQUIRE(total_m >= 0 && N >= 0 && K >= 0 && num_experts >= 0, "negative shape");
  VXK_REQUIRE(act_dtype == VXK_BF16 || act_dtype == VXK_F16, "activations must be bf16 or f16");
  VXK_REQUIRE(weight_kind >= VXK_MOE_W16 && weight_kind <= VXK_MOE_FP8_E5M2_NK, "unknown weight_kind");
  if (weight_kind == VXK_MOE_FP8_E4M3_NK || weight_kind == VXK_MOE_FP8_E5M2_NK) {
    // fp8 weights [E, N, K]: the TMEM-operand kernel (GW_FP8T)
    VXK_REQUIRE(N % 8 == 0, "N must be divisible by 8");
    VXK_REQUIRE(scales && rows_per_expert, "fp8 grouped gemm must have scales and rows_per_expert");
    if (total_m == 0 || N == 0 || num_experts == 0) return VXK_OK;
    VXK_REQUIRE(K > 0, "K must be positive");
    VXK_REQUIRE(!force_fallback && num_experts <= kGgMaxExperts && aligned16(A) && aligned16(B) && K % 16 == 0 &&
                    (reinterpret_cast<uintptr_t>(D) & 3) == 0 && total_m < (1LL << 30) && N < (1LL << 30) && K < (1LL << 30),
                "fp8 [E, N, K] grouped gemm needs K %% 16 == 0, 16-byte aligned operands and at most 512 experts");
    const int64_t avg = (total_m + num_experts - 1) / num_experts;
    const int mt = avg <= 32 ? 32 : (avg <= 64 ? 64 : (avg <= 128 ? 128 : 256));
    const int block_k = mt <= 128 ? 128 : kGgBlockK;
    GroupedParams gp{};
    gp.D = D; gp.bias = bias; gp.rows = rows_per_expert; gp.w_scale = static_cast<const float*>(scales);
    gp.total_m = static_cast<int>(total_m); gp.N = static_cast<int>(N); gp.K = static_cast<int>(K);
    gp.E = static_cast<int>(num_experts);
    gp.num_kb = static_cast<int>((K + block_k - 1) / block_k);
    gp.num_nt = static_cast<int>((N + 127) / 128);
    gp.ldd = N;
    const int64_t max_tiles = ((total_m + mt - 1) / mt + num_experts) * gp.num_nt;
    if (max_tiles >= (1LL << 31)) return fail(VXK_ERR_UNSUPPORTED, "grouped GEMM: too many tiles");
    CUtensorMap mw, ma;
    int rc;
    if ((rc = make_map_2d_u8(&ma, A, static_cast<uint64_t>(K) * 2, total_m, static_cast<uint64_t>(K) * 2, 128, mt, true)))
      return rc;
    if ((rc = make_map_3d_u8(&mw, B, K, N, num_experts, K, N * K, block_k, 128, block_k))) return rc;
    const bool bf = act_dtype == VXK_BF16, e5 = weight_kind == VXK_MOE_FP8_E5M2_NK;
    cudaStream_t s2 = stream;
    switch (mt) {
      case 32: return launch_gg_types<32>(gp, mw, ma, max_tiles, GW_FP8T, bf, e5, s2);
      case 64: return launch_gg_types<64>(gp, mw, ma, max_tiles, GW_FP8T, bf, e5, s2);
      case 128: return launch_gg_types<128>(gp, mw, ma, max_tiles, GW_FP8T, bf, e5, s2);
      default: return launch_gg_types<256>(gp, mw, ma, max_tiles, GW_FP8T, bf, e5, s2);
    }
  }
  VXK_REQUIRE(N % 8 == 0, "N must be divisible by 8");
  const int wkind = weight_kind == VXK_MOE_W16 ? GW_16 : (weight_kind == VXK_MOE_MXFP4 ? GW_MX4 : GW_FP8);
  if (wkind != GW_16) VXK_REQUIRE(scales, "fp8 / mxfp4 grouped gemm must have scales");
  if (wkind == GW_MX4) VXK_REQUIRE(K % 32 == 0, "mxfp4 grouped gemm: K must be a multiple of the group size 32");
  VXK_REQUIRE(rows_per_expert, "rows_per_expert must not be null");
  if (total_m == 0 || N == 0 || num_experts == 0) return VXK_OK;
  VXK_REQUIRE(K > 0, "K must be positive");
  cudaStream_t st = stream;
  const bool bf16 = act_dtype == VXK_BF16, e5m2 = weight_kind == VXK_MOE_FP8_E5M2;

  static const bool env_fb = gg_env_int("VXK_GEMM_FORCE_FALLBACK", 0) != 0;
  const int64_t wpitch = wkind == GW_16 ? N * 2 : (wkind == GW_FP8 ? N : K / 2);   // bytes per weight row
  const bool tc_ok = !force_fallback && !env_fb && num_experts <= kGgMaxExperts && aligned16(A) && aligned16(B) &&
                     (K * 2) % 16 == 0 && wpitch % 16 == 0 && (wkind != GW_MX4 || K % kGgBlockK == 0) &&
                     (reinterpret_cast<uintptr_t>(D) & 3) == 0 && total_m < (1LL << 30) && N < (1LL << 30) &&
                     K < (1LL << 30);
  if (!tc_ok) {
    GgFallbackParams fp{D, A, B, bias, static_cast<const float*>(scales), static_cast<const uint8_t*>(scales),
                        rows_per_expert, total_m, N, K, N, static_cast<int>(num_experts), wkind, bf16, e5m2};
    dim3 grid(static_cast<unsigned>((N + 63) / 64), static_cast<unsigned>((total_m + 3) / 4));
    if (grid.y > 65535) return fail(VXK_ERR_UNSUPPORTED, "fallback grouped GEMM: too many rows (%lld)", (long long)total_m);
    gemm_grouped_fallback_kernel<<<grid, 256, 0, st>>>(fp);
    g_gemm_fallback_calls.fetch_add(1, std::memory_order_relaxed);
    return check_launch("gemm_grouped_fallback");
  }

  // token-tile width from the average rows per expert (the reference picks its tile policy the same way,
  // grouped_gemm_xe2_interface.hpp:209,275-357): the per-expert counts are device data
  static const int env_mt = gg_env_int("VXK_GG_MT", 0);
  const int64_t avg = (total_m + num_experts - 1) / num_experts;
  int mt = avg <= 32 ? 32 : (avg <= 64 ? 64 : (avg <= 128 ? 128 : 256));
  if (env_mt == 32 || env_mt == 64 || env_mt == 128 || env_mt == 256) mt = env_mt;

  // mxfp4 with the widened operand in TMEM: 128-deep stages up to 128-token tiles, 64-deep at 256 (VXK_GG_TMEM_A=0:
  // the shared-memory operand everywhere; =1: TMEM only up to 128-token tiles)
  static const int env_ta = gg_env_int("VXK_GG_TMEM_A", 2);
  int launch_kind = wkind;
  if (wkind == GW_MX4 && env_ta && (mt <= 128 ? K % 128 == 0 : env_ta >= 2)) launch_kind = GW_MX4T;
  const int block_k = (launch_kind == GW_MX4T && mt <= 128) ? 128 : kGgBlockK;

  GroupedParams gp{};
  gp.D = D; gp.bias = bias; gp.rows = rows_per_expert;
  gp.w_scale = wkind == GW_FP8 ? static_cast<const float*>(scales) : nullptr;
  gp.w_sf = wkind == GW_MX4 ? static_cast<const uint8_t*>(scales) : nullptr;
  gp.total_m = static_cast<int>(total_m); gp.N = static_cast<int>(N); gp.K = static_cast<int>(K);
  gp.E = static_cast<int>(num_experts);
  gp.num_kb = static_cast<int>((K + block_k - 1) / block_k);
  gp.num_nt = static_cast<int>((N + 127) / 128);
  gp.ldd = N;
  gp.sf_vec = wkind == GW_MX4 && (K / 32) % 16 == 0 && aligned16(scales);
  static const int env_dbg = gg_env_int("VXK_GG_DEBUG", 0);
  gp.dbg = env_dbg;
  const int64_t max_tiles = ((total_m + mt - 1) / mt + num_experts) * gp.num_nt;
  if (max_tiles >= (1LL << 31)) return fail(VXK_ERR_UNSUPPORTED, "grouped GEMM: too many tiles");

  CUtensorMap mw, ma;
  int rc;
  if ((rc = make_map_2d_u8(&ma, A, static_cast<uint64_t>(K) * 2, total_m, static_cast<uint64_t>(K) * 2, 128, mt, true)))
    return rc;
  if (wkind == GW_MX4)
    rc = make_map_3d_u8(&mw, B, K / 2, N, num_experts, K / 2, N * (K / 2), block_k / 2, 128,
                        (launch_kind == GW_MX4T && mt <= 128) ? 64 : 0);
  else
    rc = make_map_3d_u8(&mw, B, wpitch, K, num_experts, wpitch, K * wpitch, 128, kGgBlockK, 128);
  if (rc) return rc;
  switch (mt) {
    case 32: return launch_gg_types<32>(gp, mw, ma, max_tiles, launch_kind, bf16, e5m2, st);
    case 64: return launch_gg_types<64>(gp, mw, ma, max_tiles, launch_kind, bf16, e5m2, st);
    case 128: return launch_gg_types<128>(gp, mw, ma, max_tiles, launch_kind, bf16, e5m2, st);
    default: return launch_gg_types<256>(gp, mw, ma, max_tiles, launch_kind, bf16, e5m2, st);
  }
}
