// Host side of the scaled GEMMs: eligibility checks, TMA tensor maps, scale-factor
// canonicalisation, launch of the tcgen05 kernels (gemm_tc.cuh) or of the CUDA-core
// fallback (gemm_fallback.cu), and the two C-ABI entry points.
#include "gemm_tc.cuh"

#include <mutex>

namespace vxk {

// ---- cuTensorMapEncodeTiled through the runtime (libcuda is not linked) ----
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return reinterpret_cast<EncodeTiledFn>(p);
  }();
  return fn;
}

// 2-D row-major byte tensor [rows, cols_bytes] with row pitch `pitch_bytes`, box [box_rows, box_cols]
static int make_map_2d_u8(CUtensorMap* map, const void* base, uint64_t cols_bytes, uint64_t rows,
                          uint64_t pitch_bytes, uint32_t box_cols, uint32_t box_rows, bool swizzle128) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return fail(VXK_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  const cuuint64_t dims[2] = {cols_bytes, rows};
  const cuuint64_t strides[1] = {pitch_bytes};
  const cuuint32_t box[2] = {box_cols, box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE,
                        swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(VXK_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) for [%llu x %llu] pitch %llu box [%u x %u]",
                static_cast<int>(r), (unsigned long long)rows, (unsigned long long)cols_bytes,
                (unsigned long long)pitch_bytes, box_rows, box_cols);
  return VXK_OK;
}

// ---- scale-factor canonicalisation: any e8m0 block grid -> bytes [rows, pitch] per 32 K-elements ----
__global__ void canon_sf_kernel(uint8_t* out, int64_t rows, int64_t groups32, int64_t pitch, const uint8_t* src,
                                int64_t row_div, int64_t row_stride, int64_t k_per_scale32, int64_t k_stride) {
  const int64_t total = rows * pitch;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t r = i / pitch, g = i % pitch;
    out[i] = (g < groups32) ? src[(r / row_div) * row_stride + (g / k_per_scale32) * k_stride] : 0;
  }
}

template <int KIND, int BLOCK_N>
static int launch_tc(const TcParams& tp, const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& msa,
                     const CUtensorMap& msb, cudaStream_t st) {
  using L = TcSmem<KIND, BLOCK_N>;
  auto kernel = gemm_tc_kernel<KIND, BLOCK_N>;
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [&] {
    attr_err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotalBytes);
  });
  if (attr_err != cudaSuccess)
    return fail(VXK_ERR_CUDA, "cudaFuncSetAttribute(smem=%d): %s", L::kTotalBytes, cudaGetErrorString(attr_err));
  static const int dbg = [] { const char* e = getenv("VXK_GEMM_DEBUG"); return e ? atoi(e) : 0; }();
  TcParams tpd = tp;
  tpd.debug = dbg;
  const int tiles = tp.num_m_blocks * tp.num_n_blocks;
  const int grid = tiles < sm_count() ? tiles : sm_count();
  kernel<<<grid, kTcThreads, L::kTotalBytes, st>>>(tpd, ma, mb, msa, msb);
  g_gemm_tc_calls.fetch_add(1, std::memory_order_relaxed);
  return check_launch("gemm_tc");
}

static bool tma_ok(const void* base, int64_t pitch_bytes) {
  return aligned16(base) && pitch_bytes % 16 == 0 && pitch_bytes > 0;
}

static int g_force_fallback = -1;
static bool force_fallback() {
  if (g_force_fallback < 0) {
    const char* e = getenv("VXK_GEMM_FORCE_FALLBACK");
    g_force_fallback = (e && e[0] == '1') ? 1 : 0;
  }
  return g_force_fallback == 1;
}

}  // namespace vxk

using namespace vxk;

extern "C" int vxk_set_gemm_force_fallback(int on) {
  const int prev = force_fallback() ? 1 : 0;
  g_force_fallback = on ? 1 : 0;
  return prev;
}

extern "C" int64_t vxk_fp8_gemm_workspace_bytes(int64_t M, int64_t N, int64_t K, int a_scale_mode, int scale_dtype) {
  if (a_scale_mode != 2 || scale_dtype != VXK_S_E8M0) return 0;
  const int64_t pitch = ((K + 31) / 32 + 15) / 16 * 16;
  return (M + N) * pitch + 256;
}

extern "C" int vxk_fp8_gemm(void* D, const void* A, const void* B, const void* a_scale, const void* b_scale,
                            const void* bias, int64_t M, int64_t N, int64_t K, int64_t lda, int64_t b_stride_k,
                            int64_t b_stride_n, int64_t ldd, int a_scale_mode, int b_scale_mode, int64_t group_k,
                            int64_t group_n, int64_t sa0, int64_t sa1, int64_t sb0, int64_t sb1, int scale_dtype,
                            int fp8_dtype, int out_dtype, int bias_dtype, void* workspace, int64_t workspace_bytes,
                            vxk_stream_t stream) {
  VXK_REQUIRE(M >= 0 && N >= 0 && K >= 0, "negative shape");
  VXK_REQUIRE(fp8_dtype == VXK_Q_E4M3 || fp8_dtype == VXK_Q_E5M2, "operands must be e4m3 or e5m2");
  VXK_REQUIRE(out_dtype == VXK_BF16 || out_dtype == VXK_F16, "output must be bf16 or f16");
  VXK_REQUIRE(a_scale && b_scale, "scale pointers must not be null (pass a one-element tensor of 1.0)");
  VXK_REQUIRE((a_scale_mode == 2) == (b_scale_mode == 2), "block scales must be given for both operands");
  if (a_scale_mode == 2) VXK_REQUIRE(group_k > 0 && group_n > 0, "block scale groups must be positive");
  if (M == 0 || N == 0) return VXK_OK;
  GemmParams p{D, A, B, a_scale, b_scale, bias, M, N, K, lda, b_stride_k, b_stride_n, ldd, a_scale_mode,
               b_scale_mode, group_k, group_n, sa0, sa1, sb0, sb1, scale_dtype, fp8_dtype, out_dtype, bias_dtype};
  cudaStream_t st = stream;

  const bool operands_ok = !force_fallback() && b_stride_k == 1 && tma_ok(A, lda) && tma_ok(B, b_stride_n) && K >= 32 &&
                           M < (1LL << 31) && N < (1LL << 31) && K < (1LL << 31);
  TcParams tp{};
  tp.D = D; tp.bias = bias; tp.M = (int)M; tp.N = (int)N; tp.K_bytes = (int)K; tp.ldd = ldd;
  tp.a_scale_mode = a_scale_mode; tp.b_scale_mode = b_scale_mode; tp.out_dtype = out_dtype;
  tp.bias_dtype = bias_dtype; tp.fp8_fmt = fp8_dtype == VXK_Q_E5M2 ? 1 : 0;
  tp.num_m_blocks = (int)((M + kBlockM - 1) / kBlockM);
  tp.num_k_blocks = (int)((K + kBlockKBytes - 1) / kBlockKBytes);
  tp.vec_store_ok = aligned16(D) && (ldd * 2) % 16 == 0;
  CUtensorMap ma, mb, msa{}, msb{};
  int rc;

  if (operands_ok && a_scale_mode != 2) {
    tp.a_scale = static_cast<const float*>(a_scale);
    tp.b_scale = static_cast<const float*>(b_scale);
    const bool wide = N >= 256 && (M * N) >= 148LL * 128 * 256;   // enough tiles to fill the SMs at 128x256
    const int bn = wide ? 256 : 128;
    tp.num_n_blocks = (int)((N + bn - 1) / bn);
    if ((rc = make_map_2d_u8(&ma, A, K, M, lda, kBlockKBytes, kBlockM, true))) return rc;
    if ((rc = make_map_2d_u8(&mb, B, K, N, b_stride_n, kBlockKBytes, bn, true))) return rc;
    return wide ? launch_tc<KIND_F8, 256>(tp, ma, mb, msa, msb, st) : launch_tc<KIND_F8, 128>(tp, ma, mb, msa, msb, st);
  }
  if (operands_ok && a_scale_mode == 2 && scale_dtype == VXK_S_E8M0 && group_k % 32 == 0 && workspace &&
      workspace_bytes >= vxk_fp8_gemm_workspace_bytes(M, N, K, a_scale_mode, scale_dtype)) {
    // native block-scaled MMA: expand the scale grids to one byte per (row, 32 K-elements)
    const int64_t groups32 = (K + 31) / 32, pitch = (groups32 + 15) / 16 * 16;
    uint8_t* ws = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
    uint8_t* sfa = ws;
    uint8_t* sfb = ws + M * pitch;
    sfb = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(sfb) + 15) & ~uintptr_t(15));
    const int blocks = 4 * sm_count();
    canon_sf_kernel<<<blocks, 256, 0, st>>>(sfa, M, groups32, pitch, static_cast<const uint8_t*>(a_scale), 1, sa0,
                                            group_k / 32, sa1);
    if ((rc = check_launch("canon_sfa"))) return rc;
    canon_sf_kernel<<<blocks, 256, 0, st>>>(sfb, N, groups32, pitch, static_cast<const uint8_t*>(b_scale), group_n, sb1,
                                            group_k / 32, sb0);
    if ((rc = check_launch("canon_sfb"))) return rc;
    const bool wide = N >= 192 && (M * N) >= 148LL * 128 * 192;
    const int bn = wide ? 192 : 128;
    tp.num_n_blocks = (int)((N + bn - 1) / bn);
    if ((rc = make_map_2d_u8(&ma, A, K, M, lda, kBlockKBytes, kBlockM, true))) return rc;
    if ((rc = make_map_2d_u8(&mb, B, K, N, b_stride_n, kBlockKBytes, bn, true))) return rc;
    if ((rc = make_map_2d_u8(&msa, sfa, pitch, M, pitch, 16, kBlockM, false))) return rc;
    if ((rc = make_map_2d_u8(&msb, sfb, pitch, N, pitch, 16, bn, false))) return rc;
    return wide ? launch_tc<KIND_MXF8, 192>(tp, ma, mb, msa, msb, st) : launch_tc<KIND_MXF8, 128>(tp, ma, mb, msa, msb, st);
  }
  return launch_gemm_fallback(p, /*is_fp4=*/false, st);
}

extern "C" int vxk_fp4_gemm(void* D, const void* A, const void* B, const void* a_scale, const void* b_scale,
                            const void* bias, int64_t M, int64_t N, int64_t K, int64_t lda_bytes, int64_t ldb_bytes,
                            int64_t ldd, int64_t sa0, int64_t sa1, int64_t sb0, int64_t sb1, int out_dtype,
                            int bias_dtype, vxk_stream_t stream) {
  VXK_REQUIRE(M >= 0 && N >= 0 && K >= 0, "negative shape");
  VXK_REQUIRE(K % 32 == 0, "K must be a multiple of the MX block (32)");
  VXK_REQUIRE(out_dtype == VXK_BF16 || out_dtype == VXK_F16, "output must be bf16 or f16");
  VXK_REQUIRE(a_scale && b_scale, "scale pointers must not be null");
  if (M == 0 || N == 0) return VXK_OK;
  GemmParams p{D, A, B, a_scale, b_scale, bias, M, N, K, lda_bytes, 1, ldb_bytes, ldd, 2, 2, 32, 1, sa0, sa1, sb0, sb1,
               VXK_S_E8M0, VXK_Q_E4M3, out_dtype, bias_dtype};
  cudaStream_t st = stream;
  const bool ok = !force_fallback() && tma_ok(A, lda_bytes) && tma_ok(B, ldb_bytes) && sa1 == 1 && sb1 == 1 &&
                  tma_ok(a_scale, sa0) && tma_ok(b_scale, sb0) && K >= 64 && M < (1LL << 31) && N < (1LL << 31);
  if (!ok) return launch_gemm_fallback(p, /*is_fp4=*/true, st);
  TcParams tp{};
  tp.D = D; tp.bias = bias; tp.M = (int)M; tp.N = (int)N; tp.K_bytes = (int)(K / 2); tp.ldd = ldd;
  tp.a_scale_mode = 2; tp.b_scale_mode = 2; tp.out_dtype = out_dtype; tp.bias_dtype = bias_dtype;
  tp.num_m_blocks = (int)((M + kBlockM - 1) / kBlockM);
  const bool wide = N >= 192 && (M * N) >= 148LL * 128 * 192;
  const int bn = wide ? 192 : 128;
  tp.num_n_blocks = (int)((N + bn - 1) / bn);
  tp.num_k_blocks = (int)((K / 2 + kBlockKBytes - 1) / kBlockKBytes);
  tp.vec_store_ok = aligned16(D) && (ldd * 2) % 16 == 0;
  CUtensorMap ma, mb, msa, msb;
  int rc;
  if ((rc = make_map_2d_u8(&ma, A, K / 2, M, lda_bytes, kBlockKBytes, kBlockM, true))) return rc;
  if ((rc = make_map_2d_u8(&mb, B, K / 2, N, ldb_bytes, kBlockKBytes, bn, true))) return rc;
  if ((rc = make_map_2d_u8(&msa, a_scale, K / 32, M, sa0, 16, kBlockM, false))) return rc;
  if ((rc = make_map_2d_u8(&msb, b_scale, K / 32, N, sb0, 16, bn, false))) return rc;
  return wide ? launch_tc<KIND_MXF4, 192>(tp, ma, mb, msa, msb, st) : launch_tc<KIND_MXF4, 128>(tp, ma, mb, msa, msb, st);
}
