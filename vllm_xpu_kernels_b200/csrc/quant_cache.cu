// Block-FP8 quantisation into / out of the paged indexer-K cache and the FP8 cache conversion
// (SURVEY.md 8f #3): the a4 arithmetic (1 x G groups, fp32 or UE8M0 scales, true division) with a
// paged scatter on the way in and a paged gather on the way out.
//
//   indexer_k_quant_and_cache        ref csrc/cache.cpp:530-612, launcher :1558-1590
//   cp_gather_indexer_k_quant_cache  ref csrc/cache.cpp:616-750, launcher :1625-1680
//   convert_fp8                      ref csrc/cache.cpp:1393-1540
//
// Cache layout (bytes), one block = block_size tokens:
//   [block_size * head_dim fp8 codes][block_size * head_dim / G fp32 scales]   (cache_stride = head_dim + 4 * head_dim / G)
// Arithmetic follows the reference TEST's torch function (tests/test_indexer_k_quant_and_cache.py:34-92):
//   s = max(amax, 1e-4) / 448 ; ue8m0: s = 2^ceil(log2 s) ; q = e4m3(clamp(x / s))     -- TRUE division
// (the SYCL kernel multiplies by nothing either: CopyWithScaleOp divides), with the exponent taken from
// the bit pattern like every other UE8M0 path here.
//
// HBM-bound byte shuffling: 16-byte loads of k, one 8-byte store of codes per thread, scales by the
// group's first lane; the gather moves 16 bytes per thread.  No tensor-core shaping (tier rule).
#include "vxk_common.cuh"

#include <cuda_fp8.h>

namespace vxk {

struct IndexerParams {
  const void* k;            // [T, head_dim]
  uint8_t* cache;           // [num_blocks, block_size, cache_stride] bytes
  const int64_t* slots;     // [T]
  int64_t T;
  int head_dim, G, block_size, ue8m0;
  int64_t k_row_stride;     // elements
  int64_t block_stride;     // bytes between cache blocks
  int lanes_log2;           // log2(G / 8) on the vector path
};

// vector path: a thread owns 8 consecutive elements, a group is G / 8 adjacent lanes
template <typename T>
__global__ void __launch_bounds__(256) indexer_k_quant_vec_kernel(const IndexerParams p) {
  const int nvec = p.head_dim / kVec;
  const int64_t total = p.T * nvec;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  // every lane of a warp runs the same number of iterations (shuffles below): round the bound up to a warp
  const int64_t total_w = (total + 31) / 32 * 32;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total_w; i += stride) {
    const bool live = i < total;
    const int64_t t = live ? i / nvec : 0;
    const int c = live ? static_cast<int>(i - t * nvec) : 0;
    float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (live) load8<T, true>(static_cast<const T*>(p.k) + t * p.k_row_stride + c * kVec, v);
    float gmax = 0.0f;
#pragma unroll
    for (int j = 0; j < 8; ++j) gmax = fmaxf(gmax, fabsf(v[j]));
    gmax = seg_max_pow2(gmax, p.lanes_log2);
    float s = div_const<448>(fmaxf(gmax, 1e-4f));
    if (p.ue8m0) s = exp2_ceil_log2(s);
    const float r = __frcp_rn(s);
    const int64_t slot = live ? __ldg(p.slots + t) : -1;
    if (slot < 0) continue;
    const int64_t blk = slot / p.block_size;
    const int off = static_cast<int>(slot - blk * p.block_size);
    uint8_t* base = p.cache + blk * p.block_stride;
    float q[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) q[j] = div_by<true>(v[j], s, r, 448.0f);
    *reinterpret_cast<uint2*>(base + static_cast<int64_t>(off) * p.head_dim + c * kVec) = pack_q8<VXK_Q_E4M3>(q);
    if ((c & ((1 << p.lanes_log2) - 1)) == 0) {
      const int g = (c * kVec) / p.G;
      const int groups = p.head_dim / p.G;
      *reinterpret_cast<float*>(base + static_cast<int64_t>(p.block_size) * p.head_dim +
                                (static_cast<int64_t>(off) * groups + g) * 4) = s;
    }
  }
}

// generic path (any G dividing head_dim, any alignment): one warp per group
template <typename T>
__global__ void __launch_bounds__(256) indexer_k_quant_generic_kernel(const IndexerParams p) {
  const int lane = threadIdx.x & 31;
  const int groups = p.head_dim / p.G;
  const int64_t total = p.T * groups;
  const int64_t warp0 = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  for (int64_t gi = warp0; gi < total; gi += nwarps) {
    const int64_t t = gi / groups;
    const int g = static_cast<int>(gi - t * groups);
    const T* row = static_cast<const T*>(p.k) + t * p.k_row_stride + g * p.G;
    float gmax = 0.0f;
    for (int h = lane; h < p.G; h += 32) gmax = fmaxf(gmax, fabsf(to_f32<T>(row[h])));
    gmax = warp_max(gmax);
    float s = __fdiv_rn(fmaxf(gmax, 1e-4f), 448.0f);
    if (p.ue8m0) s = exp2_ceil_log2(s);
    const int64_t slot = __ldg(p.slots + t);
    if (slot < 0) continue;
    const int64_t blk = slot / p.block_size;
    const int off = static_cast<int>(slot - blk * p.block_size);
    uint8_t* base = p.cache + blk * p.block_stride;
    uint8_t* o = base + static_cast<int64_t>(off) * p.head_dim + g * p.G;
    for (int h = lane; h < p.G; h += 32)
      o[h] = pack_q1<VXK_Q_E4M3>(fminf(fmaxf(__fdiv_rn(to_f32<T>(row[h]), s), -448.0f), 448.0f));
    if (lane == 0) {
      // the reference addresses the scale as a float index (byte offset / 4, cache.cpp:600-605)
      const int64_t byte_off = blk * p.block_stride + static_cast<int64_t>(p.block_size) * p.head_dim +
                               (static_cast<int64_t>(off) * p.head_dim + static_cast<int64_t>(g) * p.G) * 4 / p.G;
      reinterpret_cast<float*>(p.cache)[byte_off / 4] = s;
    }
  }
}

struct GatherParams {
  const uint8_t* cache;
  uint8_t* dst_k;
  uint8_t* dst_scale;
  const int32_t* block_table;   // [batch, max_blocks]
  const int32_t* cu_seq_lens;   // [batch + 1]
  int batch, max_blocks, num_tokens, head_dim, G, block_size;
  int64_t block_stride, dst_k_stride, dst_scale_stride;   // bytes
};

// one thread per 16-byte chunk of a token's codes; the chunk-0 thread also moves the token's scales
__global__ void __launch_bounds__(256) cp_gather_indexer_kernel(const GatherParams p, bool vec16) {
  const int chunks = (p.head_dim + 15) / 16;
  const int64_t total = static_cast<int64_t>(p.num_tokens) * chunks;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int tok = static_cast<int>(i / chunks);
    const int ch = static_cast<int>(i - static_cast<int64_t>(tok) * chunks);
    // batch of this token: the last b with cu_seq_lens[b] <= tok (binary search; ref: linear scan :676-687)
    int lo = 0, hi = p.batch;          // invariant: cu[lo] <= tok < cu[hi] when the token belongs to a sequence
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (__ldg(p.cu_seq_lens + mid) <= tok) lo = mid; else hi = mid;
    }
    const int seq_start = __ldg(p.cu_seq_lens + lo);
    if (tok < seq_start || tok >= __ldg(p.cu_seq_lens + lo + 1)) continue;   // not covered by any sequence
    const int pos = tok - seq_start;
    const int blk = __ldg(p.block_table + static_cast<int64_t>(lo) * p.max_blocks + pos / p.block_size);
    const int off = pos % p.block_size;
    const uint8_t* src = p.cache + static_cast<int64_t>(blk) * p.block_stride;
    const uint8_t* sk = src + static_cast<int64_t>(off) * p.head_dim + ch * 16;
    uint8_t* dk = p.dst_k + static_cast<int64_t>(tok) * p.dst_k_stride + ch * 16;
    const int n = min(16, p.head_dim - ch * 16);
    if (vec16 && n == 16) {
      *reinterpret_cast<uint4*>(dk) = *reinterpret_cast<const uint4*>(sk);
    } else {
      for (int b = 0; b < n; ++b) dk[b] = sk[b];
    }
    if (ch == 0) {
      const int groups = p.head_dim / p.G;
      const float* ss = reinterpret_cast<const float*>(src + static_cast<int64_t>(p.block_size) * p.head_dim) +
                        static_cast<int64_t>(off) * groups;
      float* ds = reinterpret_cast<float*>(p.dst_scale + static_cast<int64_t>(tok) * p.dst_scale_stride);
      for (int g = 0; g < groups; ++g) ds[g] = ss[g];
    }
  }
}

// ---- convert_fp8 ----
// MODE 0: "auto" = plain cast with a scale multiply (cache.cpp:1411-1414); MODE 1: fp8_e4m3,
// DEQ ? float(fp8) * scale : e4m3(clamp(float(x) / scale))  (:1420-1433)
template <typename Tin, typename Tout, int MODE>
__global__ void __launch_bounds__(256) convert_fp8_kernel(const Tin* __restrict__ src, Tout* __restrict__ dst,
                                                          float scale, int64_t n) {
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    if constexpr (MODE == 0) {
      float x;
      if constexpr (std::is_same<Tin, uint8_t>::value) x = static_cast<float>(src[i]);
      else x = to_f32<Tin>(src[i]);
      x = x * scale;
      if constexpr (std::is_same<Tout, uint8_t>::value) dst[i] = static_cast<uint8_t>(fminf(fmaxf(truncf(x), 0.0f), 255.0f));
      else dst[i] = from_f32<Tout>(x);
    } else if constexpr (std::is_same<Tin, uint8_t>::value) {
      const __half_raw h = __nv_cvt_fp8_to_halfraw(src[i], __NV_E4M3);
      dst[i] = from_f32<Tout>(__half2float(*reinterpret_cast<const __half*>(&h)) * scale);
    } else {
      const float x = fminf(fmaxf(__fdiv_rn(to_f32<Tin>(src[i]), scale), -448.0f), 448.0f);
      dst[i] = pack_q1<VXK_Q_E4M3>(x);
    }
  }
}

static int grid_for(int64_t items, int per_block) {
  const int64_t blocks = (items + per_block - 1) / per_block;
  const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
  return static_cast<int>(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}

}  // namespace vxk

using namespace vxk;

extern "C" int vxk_indexer_k_quant_and_cache(const void* k, void* kv_cache, const int64_t* slot_mapping,
                                             int64_t num_tokens, int64_t head_dim, int64_t k_row_stride,
                                             int64_t quant_block_size, int64_t cache_block_size,
                                             int64_t cache_stride, int64_t cache_block_stride_bytes,
                                             int use_ue8m0, int in_dtype, vxk_stream_t stream) {
  VXK_REQUIRE(num_tokens >= 0 && head_dim >= 0, "negative shape");
  VXK_REQUIRE(quant_block_size > 0 && head_dim % quant_block_size == 0,
              "head_dim must be divisible by quant_block_size");
  VXK_REQUIRE(cache_block_size > 0, "cache block size must be positive");
  const int64_t groups = head_dim / quant_block_size;
  VXK_REQUIRE(cache_stride >= head_dim + 4 * groups,
              "cache_stride (%lld) must hold head_dim fp8 codes + head_dim / quant_block_size fp32 scales",
              (long long)cache_stride);
  if (num_tokens == 0 || head_dim == 0) return VXK_OK;
  IndexerParams p{};
  p.k = k; p.cache = static_cast<uint8_t*>(kv_cache); p.slots = slot_mapping; p.T = num_tokens;
  p.head_dim = static_cast<int>(head_dim); p.G = static_cast<int>(quant_block_size);
  p.block_size = static_cast<int>(cache_block_size); p.ue8m0 = use_ue8m0;
  p.k_row_stride = k_row_stride; p.block_stride = cache_block_stride_bytes;
  const int64_t es = in_dtype == VXK_F32 ? 4 : 2;
  const int lanes = static_cast<int>(quant_block_size / kVec);
  const bool vec = quant_block_size % kVec == 0 && lanes <= 32 && (lanes & (lanes - 1)) == 0 && aligned16(k) &&
                   (k_row_stride * es) % 16 == 0 && reinterpret_cast<uintptr_t>(kv_cache) % 8 == 0 &&
                   cache_block_stride_bytes % 8 == 0 && head_dim % 8 == 0 && (cache_block_size * head_dim) % 4 == 0;
  cudaStream_t st = stream;
  if (vec) {
    p.lanes_log2 = 0;
    while ((1 << p.lanes_log2) < lanes) ++p.lanes_log2;
    const int grid = grid_for(num_tokens * (head_dim / kVec), 256);
    switch (in_dtype) {
      case VXK_F32: indexer_k_quant_vec_kernel<float><<<grid, 256, 0, st>>>(p); break;
      case VXK_F16: indexer_k_quant_vec_kernel<__half><<<grid, 256, 0, st>>>(p); break;
      case VXK_BF16: indexer_k_quant_vec_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(p); break;
      default: return fail(VXK_ERR_INVALID, "unsupported input dtype %d", in_dtype);
    }
  } else {
    const int grid = grid_for(num_tokens * groups, 8);
    switch (in_dtype) {
      case VXK_F32: indexer_k_quant_generic_kernel<float><<<grid, 256, 0, st>>>(p); break;
      case VXK_F16: indexer_k_quant_generic_kernel<__half><<<grid, 256, 0, st>>>(p); break;
      case VXK_BF16: indexer_k_quant_generic_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(p); break;
      default: return fail(VXK_ERR_INVALID, "unsupported input dtype %d", in_dtype);
    }
  }
  return check_launch("indexer_k_quant_and_cache");
}

extern "C" int vxk_cp_gather_indexer_k_quant_cache(const void* kv_cache, void* dst_k, void* dst_scale,
                                                   const int32_t* block_table, const int32_t* cu_seq_lens,
                                                   int64_t batch_size, int64_t max_blocks_per_seq,
                                                   int64_t num_tokens, int64_t head_dim,
                                                   int64_t quant_block_size, int64_t cache_block_size,
                                                   int64_t cache_block_stride_bytes, int64_t dst_k_stride_bytes,
                                                   int64_t dst_scale_stride_bytes, vxk_stream_t stream) {
  VXK_REQUIRE(num_tokens >= 0 && head_dim >= 0 && batch_size >= 0, "negative shape");
  VXK_REQUIRE(quant_block_size > 0 && head_dim % quant_block_size == 0,
              "head_dim must be divisible by quant_block_size");
  if (num_tokens == 0 || head_dim == 0 || batch_size == 0) return VXK_OK;
  GatherParams p{};
  p.cache = static_cast<const uint8_t*>(kv_cache); p.dst_k = static_cast<uint8_t*>(dst_k);
  p.dst_scale = static_cast<uint8_t*>(dst_scale); p.block_table = block_table; p.cu_seq_lens = cu_seq_lens;
  p.batch = static_cast<int>(batch_size); p.max_blocks = static_cast<int>(max_blocks_per_seq);
  p.num_tokens = static_cast<int>(num_tokens); p.head_dim = static_cast<int>(head_dim);
  p.G = static_cast<int>(quant_block_size); p.block_size = static_cast<int>(cache_block_size);
  p.block_stride = cache_block_stride_bytes; p.dst_k_stride = dst_k_stride_bytes;
  p.dst_scale_stride = dst_scale_stride_bytes;
  VXK_REQUIRE(reinterpret_cast<uintptr_t>(kv_cache) % 4 == 0 && cache_block_stride_bytes % 4 == 0 &&
                  (cache_block_size * head_dim) % 4 == 0 && reinterpret_cast<uintptr_t>(dst_scale) % 4 == 0 &&
                  dst_scale_stride_bytes % 4 == 0,
              "fp32 scales must be 4-byte aligned in the cache and in dst_scale");
  const bool vec16 = aligned16(kv_cache) && aligned16(dst_k) && cache_block_stride_bytes % 16 == 0 &&
                     head_dim % 16 == 0 && dst_k_stride_bytes % 16 == 0;
  const int64_t chunks = (head_dim + 15) / 16;
  cp_gather_indexer_kernel<<<grid_for(num_tokens * chunks, 256), 256, 0, stream>>>(p, vec16);
  return check_launch("cp_gather_indexer_k_quant_cache");
}

extern "C" int vxk_convert_fp8(void* dst, const void* src, double scale, int64_t n, int src_kind, int dst_kind,
                               int mode, vxk_stream_t stream) {
  // kinds: VXK_F32 / VXK_F16 / VXK_BF16, or 3 = raw byte (fp8 code / uint8)
  VXK_REQUIRE(n >= 0, "negative size");
  VXK_REQUIRE(mode == 0 || mode == 1, "mode must be 0 (auto) or 1 (fp8_e4m3)");
  VXK_REQUIRE((src_kind == 3) != (dst_kind == 3), "exactly one of src / dst must be the byte tensor");
  if (n == 0) return VXK_OK;
  const float sc = static_cast<float>(scale);
  const int grid = grid_for(n, 1024);
  cudaStream_t st = stream;
#define VXK_CVT(TIN, TOUT)                                                                                    \
  do {                                                                                                        \
    if (mode == 0) convert_fp8_kernel<TIN, TOUT, 0><<<grid, 256, 0, st>>>(static_cast<const TIN*>(src), static_cast<TOUT*>(dst), sc, n); \
    else convert_fp8_kernel<TIN, TOUT, 1><<<grid, 256, 0, st>>>(static_cast<const TIN*>(src), static_cast<TOUT*>(dst), sc, n);           \
    return check_launch("convert_fp8");                                                                       \
  } while (0)
  if (dst_kind == 3) {
    switch (src_kind) {
      case VXK_F32: VXK_CVT(float, uint8_t);
      case VXK_F16: VXK_CVT(__half, uint8_t);
      case VXK_BF16: VXK_CVT(__nv_bfloat16, uint8_t);
    }
  } else {
    switch (dst_kind) {
      case VXK_F32: VXK_CVT(uint8_t, float);
      case VXK_F16: VXK_CVT(uint8_t, __half);
      case VXK_BF16: VXK_CVT(uint8_t, __nv_bfloat16);
    }
  }
#undef VXK_CVT
  return fail(VXK_ERR_INVALID, "Unsupported data type combination for convert_fp8");
}
