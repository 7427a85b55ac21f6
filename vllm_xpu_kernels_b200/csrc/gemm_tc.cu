// Host side of the scaled GEMMs: eligibility checks, TMA tensor maps, scale-factor
// canonicalisation, launch of the tcgen05 kernels (gemm_tc.cuh) or of the CUDA-core
// fallback (gemm_fallback.cu), and the two C-ABI entry points.
#include "gemm_tc2_blk.cuh"
#include "gemm_skinny_mx4.cuh"
#include "gemm_skinny.cuh"


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
int make_map_2d_u8(CUtensorMap* map, const void* base, uint64_t cols_bytes, uint64_t rows,
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

// 3-D byte tensor [d2][d1][d0 bytes] (expert-stacked weights of the grouped GEMM), box [1][box1][box0]
int make_map_3d_u8(CUtensorMap* map, const void* base, uint64_t d0_bytes, uint64_t d1, uint64_t d2, uint64_t pitch1,
                   uint64_t pitch2, uint32_t box0, uint32_t box1, int swizzle_bytes) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return fail(VXK_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  const cuuint64_t dims[3] = {d0_bytes, d1, d2};
  const cuuint64_t strides[2] = {pitch1, pitch2};
  const cuuint32_t box[3] = {box0, box1, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE,
                        swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                        : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_NONE,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(VXK_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) for [%llu x %llu x %llu] box [%u x %u]",
                static_cast<int>(r), (unsigned long long)d2, (unsigned long long)d1, (unsigned long long)d0_bytes,
                box1, box0);
  return VXK_OK;
}

// packed scale-factor workspace: [chunks][128 words], box = `box_chunks` whole chunks
static int make_map_sf_chunks(CUtensorMap* map, const void* base, uint64_t chunks, uint32_t box_chunks) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return fail(VXK_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  const cuuint64_t dims[2] = {128, chunks};
  const cuuint64_t strides[1] = {512};
  const cuuint32_t box[2] = {128, box_chunks};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(VXK_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) for %llu scale-factor chunks", static_cast<int>(r),
                (unsigned long long)chunks);
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
  static PerDeviceOnce once;
  const cudaError_t attr_err = once.get([&](PerDeviceOnce::Entry& x) {
    x.err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotalBytes);
  }).err;
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

// ---- scale-factor packing for the CTA-pair kernels ----
// Any e8m0 scale grid -> chunks of 512 B in the layout tcgen05.cp.32x128b.warpx4 consumes:
//   chunk (kb, blk, w) = [32 lanes][4 words], word (lane, c) = the four scale bytes of row
//   blk*128 + c*32 + lane for the 32-element K groups kb*G + 4w .. 4w+3   (G groups per K slice)
// ordered [kb][blk][w] so that the chunks of consecutive 128-row blocks of one K slice are
// contiguous (one TMA box per operand per stage).  Rows / groups outside the tensor pack as 0.
__global__ void sf_pack_kernel(uint32_t* out, int64_t rows, int64_t groups32, int num_kb, int nblk, int sf_words,
                               const uint8_t* src, int64_t row_div, int64_t row_stride, int64_t k_per_scale32,
                               int64_t k_stride, int word_loads) {
  const int64_t total = static_cast<int64_t>(num_kb) * nblk * sf_words * 128;
  const int groups_per_kb = sf_words * 4;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i & 3), lane = static_cast<int>((i >> 2) & 31);
    const int64_t chunk = i >> 7;
    const int w = static_cast<int>(chunk % sf_words);
    const int64_t t = chunk / sf_words;
    const int blk = static_cast<int>(t % nblk);
    const int64_t kb = t / nblk;
    const int64_t row = static_cast<int64_t>(blk) * 128 + c * 32 + lane;
    const int64_t g0 = kb * groups_per_kb + w * 4;
    uint32_t v = 0;
    if (row < rows && g0 < groups32) {
      const uint8_t* rp = src + (row / row_div) * row_stride;
      if (word_loads && g0 + 4 <= groups32) {
        v = *reinterpret_cast<const uint32_t*>(rp + g0);
      } else {
#pragma unroll
        for (int b = 0; b < 4; ++b)
          if (g0 + b < groups32) v |= static_cast<uint32_t>(rp[((g0 + b) / k_per_scale32) * k_stride]) << (8 * b);
      }
    }
    out[i] = v;
  }
}

// The same packing for the common case -- row-major E8M0 grids, one byte per 32-element group, 16-byte aligned rows,
// K % 512 == 0 -- as a tiled transpose, BOTH operands in one launch.  A CTA owns one tile = 128 rows x 64 scale
// bytes (2048 K elements = 16 output chunks): it reads the tile as 512 coalesced 16-byte loads (four lanes per row
// segment, whole 32-byte sectors; sf_pack_kernel's 4-byte gathers use one eighth of every sector), parks it in shared
// memory (row pitch 17 words: conflict-free both ways) and writes each 512-byte chunk as one warp-wide 16-byte store.
// Blocks [0, a_tiles) pack operand A, the rest operand B.  ~6.4 us + 6.4 us + a launch gap -> one launch.
struct SfPack2Op {
  uint32_t* out;
  const uint8_t* src;
  int64_t rows, row_stride;
  int nblk, tiles;
};
__global__ void __launch_bounds__(256) sf_pack2_kernel(SfPack2Op a, SfPack2Op b, int64_t groups32, int num_kb,
                                                       int sf_words, int n_kq) {
  __shared__ uint32_t tile[128 * 17];
  const bool is_b = static_cast<int>(blockIdx.x) >= a.tiles;
  const SfPack2Op& o = is_b ? b : a;
  const int t = static_cast<int>(blockIdx.x) - (is_b ? a.tiles : 0);
  const int blk = t / n_kq, kq = t % n_kq;
  const int tid = threadIdx.x;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int r = (tid >> 2) + h * 64, seg = tid & 3;
    const int64_t row = static_cast<int64_t>(blk) * 128 + r;
    const int64_t byte0 = static_cast<int64_t>(kq) * 64 + seg * 16;      // = first group of the segment
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (row < o.rows && byte0 < groups32) v = __ldg(reinterpret_cast<const uint4*>(o.src + row * o.row_stride + byte0));
    uint32_t* d = tile + r * 17 + seg * 4;
    d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
  }
  __syncthreads();
  const int chunks = num_kb * sf_words;      // 512-byte chunks per 128-row block along K
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int lane = tid & 31, j = (tid >> 5) + h * 8;
    const int64_t jg = static_cast<int64_t>(kq) * 16 + j;                // chunk index along K = kb * sf_words + w
    if (jg < chunks) {
      const int64_t kb = jg / sf_words;
      const int w = static_cast<int>(jg % sf_words);
      const uint4 v = make_uint4(tile[lane * 17 + j], tile[(32 + lane) * 17 + j], tile[(64 + lane) * 17 + j],
                                 tile[(96 + lane) * 17 + j]);
      uint32_t* dst = o.out + (((kb * o.nblk + blk) * sf_words + w) << 7) + lane * 4;
      *reinterpret_cast<uint4*>(dst) = v;
    }
  }
}

static std::atomic<uint64_t> g_sf_pack_calls{0};   // scale re-packing launches (tests: the packed entry points issue none)

struct SfPack {
  int num_kb, sf_words, a_blocks, b_blocks;
  int64_t a_bytes, b_bytes;
};
static SfPack sf_pack_plan(int64_t M, int64_t N, int64_t K_elems, bool fp4) {
  SfPack s;
  const int64_t elems_per_kb = fp4 ? 2 * kBlockKBytes : kBlockKBytes;
  s.num_kb = static_cast<int>((K_elems + elems_per_kb - 1) / elems_per_kb);
  s.sf_words = fp4 ? 2 : 1;
  s.a_blocks = static_cast<int>((M + 127) / 128) + 1;     // +1: a tile's second block never leaves the grid
  s.b_blocks = static_cast<int>((N + 127) / 128) + 2;
  s.a_bytes = static_cast<int64_t>(s.num_kb) * s.a_blocks * s.sf_words * 512;
  s.b_bytes = static_cast<int64_t>(s.num_kb) * s.b_blocks * s.sf_words * 512;
  return s;
}

static int max_cluster_pairs(const void* kernel, int smem_bytes, int threads = kTc2Threads) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(2 * sm_count());
  cfg.blockDim = dim3(threads);
  cfg.dynamicSmemBytes = smem_bytes;
  cudaLaunchAttribute attr{};
  attr.id = cudaLaunchAttributeClusterDimension;
  attr.val.clusterDim.x = 2; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
  cfg.attrs = &attr;
  cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, kernel, &cfg) != cudaSuccess || n < 1) {
    cudaGetLastError();
    n = sm_count() / 2;
  }
  return n;
}

static int env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return e ? atoi(e) : dflt;
}

static int pair_count_hint() {   // SM pairs a cluster launch can keep resident (all TPCs on B200)
  return sm_count() / 2;
}

template <int KIND, int BLOCK_N, int NACC, int FUSE = 0>
static int launch_tc2(Tc2Params tp, const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& msa,
                      const CUtensorMap& msb, cudaStream_t st) {
  using L = Tc2Smem<KIND, BLOCK_N, NACC>;
  auto kernel = gemm_tc2_kernel<KIND, BLOCK_N, NACC, FUSE>;
  static PerDeviceOnce once;
  const PerDeviceOnce::Entry cfg_once = once.get([&](PerDeviceOnce::Entry& x) {
    x.err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotalBytes);
    if (x.err == cudaSuccess) x.aux = max_cluster_pairs(reinterpret_cast<const void*>(kernel), L::kTotalBytes);
  });
  const cudaError_t attr_err = cfg_once.err;
  const int pairs = cfg_once.aux;
  if (attr_err != cudaSuccess)
    return fail(VXK_ERR_CUDA, "cudaFuncSetAttribute(smem=%d): %s", L::kTotalBytes, cudaGetErrorString(attr_err));
  static const int dbg = env_int("VXK_GEMM_DEBUG", 0);
  // Row tiles per rasterisation group.  Measured at M = 8192 (profiles/r01e_groupm_sweep.txt): 16 is the best
  // overall; MXFP4 with a shallow K (<= 4096 operand bytes per row) gains 2-6 % from 32 (57344 columns:
  // 5.11 -> 5.41 PF/s; cause not profiled yet), while 32 costs 17-27 % on deep K and on FP8 with few
  // column tiles.
  static const int env_group_m = env_int("VXK_GEMM_GROUP_M", 0);
  const int group_m = env_group_m > 0 ? env_group_m : (KIND == KIND_MXF4 && tp.b.K_bytes <= 4096 ? 32 : 16);
  tp.b.debug = dbg;   // bring-up / timing switches: 1 skip tcgen05.cp, 8 skip scale-factor TMA
  tp.group_m = group_m;
  const int tiles = tp.num_m_pairs * tp.num_n_blocks;
  const int grid = 2 * (tiles < pairs ? tiles : pairs);
  kernel<<<grid, kTc2Threads, L::kTotalBytes, st>>>(tp, ma, mb, msa, msb);
  g_gemm_tc_calls.fetch_add(1, std::memory_order_relaxed);
  return check_launch("gemm_tc2");
}

template <int BLOCK_N, bool PIPE>
static int launch_tc2_blk(Tc2BlkParams tp, const CUtensorMap& ma, const CUtensorMap& mb, cudaStream_t st) {
  using L = Tc2Smem<KIND_F8, BLOCK_N, 1>;
  constexpr int kSmemBytes = L::kTotalBytes + 64;      // (room for a third pair of slice barriers)
  auto kernel = gemm_tc2_blk_kernel<BLOCK_N, PIPE>;
  static PerDeviceOnce once;
  const PerDeviceOnce::Entry cfg_once = once.get([&](PerDeviceOnce::Entry& x) {
    x.err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
    if (x.err == cudaSuccess) x.aux = max_cluster_pairs(reinterpret_cast<const void*>(kernel), kSmemBytes);
  });
  const cudaError_t attr_err = cfg_once.err;
  const int pairs = cfg_once.aux;
  if (attr_err != cudaSuccess)
    return fail(VXK_ERR_CUDA, "cudaFuncSetAttribute(smem=%d): %s", L::kTotalBytes, cudaGetErrorString(attr_err));
  static const int group_m = env_int("VXK_GEMM_GROUP_M", 16);
  static const int dbg = env_int("VXK_GEMM_DEBUG", 0);
  tp.b.debug = dbg;
  tp.group_m = group_m < 1 ? 1 : group_m;
  const int tiles = tp.num_m_pairs * tp.num_n_blocks;
  const int grid = 2 * (tiles < pairs ? tiles : pairs);
  kernel<<<grid, kTc2Threads, kSmemBytes, st>>>(tp, ma, mb);
  g_gemm_tc_calls.fetch_add(1, std::memory_order_relaxed);
  return check_launch("gemm_tc2_blk");
}

// Two W tiles per K loop (NACC = 2) cut the L2 bytes per flop by a quarter but quantise the
// work into tiles twice as large and expose part of the epilogue: take them when the last
// wave of super-tiles is not much emptier than the last wave of single tiles.
static bool prefer_two_acc(int64_t M, int64_t N, int64_t K_bytes, int bn) {
  static const int forced = env_int("VXK_GEMM_NACC", 0);
  if (forced == 1 || forced == 2) return forced == 2;
  const int64_t pairs = pair_count_hint();
  const int64_t mt = (M + 255) / 256;
  const int64_t t1 = mt * ((N + bn - 1) / bn), t2 = mt * ((N + 2 * bn - 1) / (2 * bn));
  const double c1 = static_cast<double>((t1 + pairs - 1) / pairs) * bn;
  const double c2 = static_cast<double>((t2 + pairs - 1) / pairs) * 2 * bn;
  return c2 <= c1 && K_bytes >= 12288;   // measured: pays off only when the K loop dwarfs the exposed epilogue
}

// Single-CTA tile width: the widest of {256 (fp8 only), 192 (block-scaled only), 128, 64, 32} that still
// gives every SM a tile.  Small-M GEMMs are bound by streaming W from HBM: what matters is that
// all SMs pull on it, not the shape the tensor core sees.
static int single_cta_bn(int64_t M, int64_t N, int widest) {
  const int64_t mt = (M + kBlockM - 1) / kBlockM, sms = sm_count();
  const int cand[5] = {widest, 128, 64, 32, 0};
  for (int i = 0; cand[i]; ++i) {
    if (cand[i] > widest) continue;
    if (N >= cand[i] && mt * ((N + cand[i] - 1) / cand[i]) >= sms) return cand[i];
  }
  return 32;
}

template <int KIND>
static int launch_tc_bn(int bn, const TcParams& tp, const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& msa,
                        const CUtensorMap& msb, cudaStream_t st) {
  switch (bn) {
    case 256: if constexpr (KIND == KIND_F8) return launch_tc<KIND, 256>(tp, ma, mb, msa, msb, st); else break;
    case 192: if constexpr (KIND != KIND_F8) return launch_tc<KIND, 192>(tp, ma, mb, msa, msb, st); else break;
    case 128: return launch_tc<KIND, 128>(tp, ma, mb, msa, msb, st);
    case 64: return launch_tc<KIND, 64>(tp, ma, mb, msa, msb, st);
    default: return launch_tc<KIND, 32>(tp, ma, mb, msa, msb, st);
  }
  return fail(VXK_ERR_INVALID, "no kernel for tile width %d", bn);
}

// Tile-shape choice by a wave count model: per-SM time of one tile ~ rows x cols handled by
// the SM, weighted by how well the shape feeds the tensor core (shared-memory / L2 bytes per
// MMA).  1-CTA 128 x bn1 tiles over all SMs against 256 x bn2 pair tiles over SM pairs.
static bool prefer_pair(int64_t M, int64_t N, int bn1, int bn2) {
  if (M <= 128) return false;
  const int64_t sms = sm_count();
  const int64_t t1 = ((M + 127) / 128) * ((N + bn1 - 1) / bn1);
  const int64_t t2 = ((M + 255) / 256) * ((N + bn2 - 1) / bn2);
  // the block-scaled single-CTA kernels (192-wide pair tiles as the alternative) lose more than this model's 1.35
  // against the pair kernel once K is deep: MXFP4 384 x 8192 x 28672 ran 87 us single against 68 us as a pair, 384 x
  // 8192 x 8192 29 vs 26 us, everything else measured within 3 % (profiles/r02c_pair_bias.txt) -> their single-CTA
  // cost is weighted 1.5 x.  VXK_GEMM_PAIR_BIAS overrides the weight for every kind.
  static const double env_bias = [] { const char* e = getenv("VXK_GEMM_PAIR_BIAS"); return e ? atof(e) : 0.0; }();
  const double bias = env_bias > 0.0 ? env_bias : (bn2 == 192 ? 1.5 : 1.0);
  const double c1 = static_cast<double>((t1 + sms - 1) / sms) * 128.0 * bn1 * (bn1 >= 256 ? 1.20 : 1.35) * bias;
  const double c2 = static_cast<double>((t2 + sms / 2 - 1) / (sms / 2)) * 128.0 * bn2;
  return c2 <= c1;
}

static bool tma_ok(const void* base, int64_t pitch_bytes) {
  return aligned16(base) && pitch_bytes % 16 == 0 && pitch_bytes > 0;
}

// testing / profiling aid: VXK_GEMM_SINGLE_CTA=1 keeps every GEMM on the single-CTA kernels
static bool force_single() {
  static const bool v = [] { const char* e = getenv("VXK_GEMM_SINGLE_CTA"); return e && e[0] == '1'; }();
  return v;
}

// fused all-gather destination of the NEXT vxk_fp8_gemm call of this thread (set by vxk_fp8_gemm_allgather)
struct PeerCfg { int n; void* ptr[8]; int multicast; };
static thread_local PeerCfg g_peer_cfg{0, {}, 0};

static int g_force_fallback = -1;
static bool force_fallback() {
  if (g_force_fallback < 0) {
    const char* e = getenv("VXK_GEMM_FORCE_FALLBACK");
    g_force_fallback = (e && e[0] == '1') ? 1 : 0;
  }
  return g_force_fallback == 1;
}

// Packs both scale grids into `workspace` and launches the CTA-pair block-scaled kernel.
// a_* / b_*: addressing of the source grids as in sf_pack_kernel (row_div, row_stride, K groups
// per scale, K stride).
template <int KIND>
static int run_pair_block_scaled(Tc2Params tp, const void* A, int64_t lda_bytes, const void* B, int64_t ldb_bytes,
                                 int64_t M, int64_t N, int64_t K_elems, const uint8_t* a_scale, int64_t a_row_stride,
                                 int64_t a_kps, int64_t a_k_stride, const uint8_t* b_scale, int64_t b_row_div,
                                 int64_t b_row_stride, int64_t b_kps, int64_t b_k_stride, void* workspace,
                                 cudaStream_t st, const void* a_packed = nullptr, const void* b_packed = nullptr) {
  constexpr bool fp4 = KIND == KIND_MXF4;
  constexpr int BN = 192;
  const SfPack sp = sf_pack_plan(M, N, K_elems, fp4);
  uint8_t* ws = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
  // a_packed / b_packed: the caller already holds that grid in tcgen05.cp order (vxk_mx_sf_pack, or written by
  // the quantisation kernel that produced the operand): no re-packing launch for it
  const uint32_t* sfa = a_packed ? static_cast<const uint32_t*>(a_packed) : reinterpret_cast<uint32_t*>(ws);
  const uint32_t* sfb = b_packed ? static_cast<const uint32_t*>(b_packed) : reinterpret_cast<uint32_t*>(ws + sp.a_bytes);
  const int64_t groups32 = (K_elems + 31) / 32;
  const int blocks = 8 * sm_count();
  const int a_words = a_kps == 1 && a_k_stride == 1 && a_row_stride % 4 == 0 && reinterpret_cast<uintptr_t>(a_scale) % 4 == 0;
  const int b_words = b_kps == 1 && b_k_stride == 1 && b_row_stride % 4 == 0 && reinterpret_cast<uintptr_t>(b_scale) % 4 == 0;
  int rc;
  // both grids in the plain row-major form: one tiled launch (sf_pack2_kernel)
  static const int env_pack2 = env_int("VXK_SF_PACK2", 1);
  auto plain = [&](const uint8_t* p, int words, int64_t row_div, int64_t row_stride) {
    return words && row_div == 1 && row_stride % 16 == 0 && reinterpret_cast<uintptr_t>(p) % 16 == 0 && groups32 % 16 == 0;
  };
  const bool a_fast = !a_packed && plain(a_scale, a_words, 1, a_row_stride);
  const bool b_fast = !b_packed && plain(b_scale, b_words, b_row_div, b_row_stride);
  bool a_done = a_packed != nullptr, b_done = b_packed != nullptr;
  if (env_pack2 && (a_fast || a_done) && (b_fast || b_done) && (a_fast || b_fast)) {
    const int n_kq = (sp.num_kb * sp.sf_words + 15) / 16;
    SfPack2Op oa{const_cast<uint32_t*>(sfa), a_scale, M, a_row_stride, sp.a_blocks, a_fast ? sp.a_blocks * n_kq : 0};
    SfPack2Op ob{const_cast<uint32_t*>(sfb), b_scale, N, b_row_stride, sp.b_blocks, b_fast ? sp.b_blocks * n_kq : 0};
    sf_pack2_kernel<<<oa.tiles + ob.tiles, 256, 0, st>>>(oa, ob, groups32, sp.num_kb, sp.sf_words, n_kq);
    if ((rc = check_launch("sf_pack2"))) return rc;
    g_sf_pack_calls.fetch_add(1, std::memory_order_relaxed);
    a_done = b_done = true;
  }
  if (!a_done) {
    sf_pack_kernel<<<blocks, 256, 0, st>>>(const_cast<uint32_t*>(sfa), M, groups32, sp.num_kb, sp.a_blocks, sp.sf_words,
                                           a_scale, 1, a_row_stride, a_kps, a_k_stride, a_words);
    if ((rc = check_launch("sf_pack_a"))) return rc;
    g_sf_pack_calls.fetch_add(1, std::memory_order_relaxed);
  }
  if (!b_done) {
    sf_pack_kernel<<<blocks, 256, 0, st>>>(const_cast<uint32_t*>(sfb), N, groups32, sp.num_kb, sp.b_blocks, sp.sf_words,
                                           b_scale, b_row_div, b_row_stride, b_kps, b_k_stride, b_words);
    if ((rc = check_launch("sf_pack_b"))) return rc;
    g_sf_pack_calls.fetch_add(1, std::memory_order_relaxed);
  }
  const bool two = prefer_two_acc(M, N, fp4 ? K_elems / 2 : K_elems, BN);
  const int tile_n = two ? 2 * BN : BN;
  tp.num_m_pairs = static_cast<int>((M + 255) / 256);
  tp.num_n_blocks = static_cast<int>((N + tile_n - 1) / tile_n);
  tp.sfa_blocks = sp.a_blocks;
  tp.sfb_blocks = sp.b_blocks;
  const uint64_t kbytes = fp4 ? K_elems / 2 : K_elems;
  CUtensorMap ma, mb, msa, msb;
  if ((rc = make_map_2d_u8(&ma, A, kbytes, M, lda_bytes, kBlockKBytes, kBlockM, true))) return rc;
  if ((rc = make_map_2d_u8(&mb, B, kbytes, N, ldb_bytes, kBlockKBytes, BN / 2, true))) return rc;
  if ((rc = make_map_sf_chunks(&msa, sfa, static_cast<uint64_t>(sp.num_kb) * sp.a_blocks * sp.sf_words, sp.sf_words)))
    return rc;
  if ((rc = make_map_sf_chunks(&msb, sfb, static_cast<uint64_t>(sp.num_kb) * sp.b_blocks * sp.sf_words,
                               (two ? Tc2Smem<KIND, BN, 2>::kSfbBlocks : 2) * sp.sf_words)))
    return rc;
  return two ? launch_tc2<KIND, BN, 2>(tp, ma, mb, msa, msb, st) : launch_tc2<KIND, BN, 1>(tp, ma, mb, msa, msb, st);
}

// Small-M FP8 GEMM (M <= 128): operands swapped, K split over a cluster.  See gemm_skinny.cuh.
template <int MT, bool BLK = false>
static int launch_skinny(SkinnyParams sp, const CUtensorMap& mw, const CUtensorMap& ma, int n_tiles, cudaStream_t st,
                         int m_tiles = 1) {
  using L = SkinnySmem<MT>;
  auto kernel = gemm_skinny_kernel<MT, BLK>;
  static PerDeviceOnce once;
  const cudaError_t attr_err = once.get([&](PerDeviceOnce::Entry& x) {
    x.err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotalBytes);
  }).err;
  if (attr_err != cudaSuccess)
    return fail(VXK_ERR_CUDA, "cudaFuncSetAttribute(smem=%d): %s", L::kTotalBytes, cudaGetErrorString(attr_err));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(n_tiles, sp.splits, m_tiles);
  cfg.blockDim = dim3(kSkinnyThreads);
  cfg.dynamicSmemBytes = L::kTotalBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1;
  attr[0].val.clusterDim.y = sp.splits;
  attr[0].val.clusterDim.z = 1;
  // programmatic dependent launch: prologue + weight prefetch overlap the previous kernel's tail
  const int pdl = g_early_weight_loads.load(std::memory_order_relaxed);
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 2 : 1;
  const cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, sp, mw, ma);
  if (e != cudaSuccess) return fail(VXK_ERR_CUDA, "gemm_skinny launch: %s", cudaGetErrorString(e));
  g_gemm_tc_calls.fetch_add(1, std::memory_order_relaxed);
  return check_launch("gemm_skinny");
}

// Token tiles of 128 for the small-M FP8 kernel at 256 < M <= 512 (1 = not eligible).  The CTA-pair kernel
// covers such a problem with M/256 x N/256 tiles -- 32 tiles on 64 SMs for 512 x 4096 -- and has no K split; the
// small-M kernel's 128 x 128 tiles with the cluster K split fill the machine (profiles/r02c_skinny_token_tiles.txt).
// VXK_SKINNY_MTILES=0 switches it off.
static int skinny_token_tiles(int64_t M, int n_tiles) {
  static const int on = env_int("VXK_SKINNY_MTILES", 1);
  // measured (graph replays, Llama-3-8B shapes): up to 256 tokens the single 256-token tile is as fast or faster
  // (13.5 vs 16.6 us at 129 x 6144 x 4096), and more than one wave of tiles loses to the CTA-pair kernel
  // (512 x 6144 x 4096: 31.9 vs 20.4 us)
  if (!on || M <= 256 || M > 512) return 1;
  const int t = static_cast<int>((M + 127) / 128);
  return n_tiles * t <= sm_count() ? t : 1;
}

// fp32-block small-M kernel: 64-token tiles side by side for 64 < M <= 256 (VXK_SKINNY_BLK_MAX_M, default 256)
static int skinny_blk_token_tiles(int64_t M, int n_tiles) {
  static const int max_m = env_int("VXK_SKINNY_BLK_MAX_M", 256);
  if (M <= 64 || M > max_m || M > 256) return 1;
  const int t = static_cast<int>((M + 63) / 64);
  // only while every tile can still split K over a cluster of two (measured, DeepSeek-V3 shapes: 65 x 1536 x 7168
  // 25.8 -> 10.9 us, 256 x 1536 x 7168 28.3 -> 18.2 us; without the split 128 x 7168 x 16384 64.2 -> 76.9 us)
  return n_tiles * t * 2 <= sm_count() ? t : 1;
}

struct SkinnyBlkScales { const float* sa; const float* sb; int64_t sa0, sa1, sb0, sb1; int group_k, group_n; };

static int run_skinny_f8(const TcParams& tp, const void* A, const void* W, int64_t M, int64_t N, int64_t K,
                         int64_t lda, int64_t ldw, cudaStream_t st, const SkinnyBlkScales* blk = nullptr) {
  SkinnyParams sp{};
  if (blk) {
    sp.blk_sa = blk->sa; sp.blk_sb = blk->sb; sp.sa0 = blk->sa0; sp.sa1 = blk->sa1; sp.sb0 = blk->sb0; sp.sb1 = blk->sb1;
    sp.group_k = blk->group_k; sp.group_n = blk->group_n;
  }
  sp.D = tp.D; sp.a_scale = tp.a_scale; sp.b_scale = tp.b_scale; sp.bias = tp.bias;
  sp.M = (int)M; sp.N = (int)N; sp.ldd = tp.ldd;
  sp.a_scale_mode = tp.a_scale_mode; sp.b_scale_mode = tp.b_scale_mode; sp.out_dtype = tp.out_dtype;
  sp.bias_dtype = tp.bias_dtype; sp.fp8_fmt = tp.fp8_fmt;
  sp.num_k_blocks = tp.num_k_blocks;
  const int n_tiles = (int)((N + 127) / 128);
  // K split: enough CTAs for one wave, every rank at least two ring stages, portable cluster size
  static const int env_splits = env_int("VXK_SKINNY_SPLITS", 0);
  // 256 < M <= 512: token tiles of 128 side by side (gridDim.z), see skinny_token_tiles(); one 256-token tile otherwise
  int m_tiles = (!blk && M > 128) ? skinny_token_tiles(M, n_tiles) : 1;
  int mt = M <= 32 ? 32 : (M <= 64 ? 64 : ((M <= 128 || m_tiles > 1) ? 128 : 256));
  if (blk && M > 64) { mt = 64; m_tiles = skinny_blk_token_tiles(M, n_tiles); }
  // two 256-token tiles instead of three or four 128-token ones while each can still split K over a cluster of two
  // (fewer operand bytes per flop: 512 x 4096 x 14336 38.4 -> 33.1 us; without the split fewer CTAs lose:
  // 384 x 6144 x 4096 17.3 -> 25.5 us).  VXK_SKINNY_MT256 = 0 / 1 forces the choice.
  static const int env_mt256 = env_int("VXK_SKINNY_MT256", -1);
  if (!blk && m_tiles > 1 && (env_mt256 >= 0 ? env_mt256 != 0 : n_tiles * 4 <= sm_count())) {
    mt = 256;
    m_tiles = static_cast<int>((M + 255) / 256);
  }
  const int ks = mt <= 128 ? 2 : 1;
  static const int env_pow2 = env_int("VXK_SKINNY_POW2", 1);
  int splits = sm_count() / (n_tiles * m_tiles);
  if (splits > 8) splits = 8;
  if (env_splits > 0) splits = env_splits > 8 ? 8 : env_splits;
  const int max_by_k = sp.num_k_blocks / (2 * ks);
  if (splits > max_by_k) splits = max_by_k;
  if (splits < 1) splits = 1;
  if (env_pow2) while (splits & (splits - 1)) --splits;   // cluster sizes 1 / 2 / 4 / 8
  int per = (sp.num_k_blocks + splits - 1) / splits;
  per = (per + ks - 1) / ks * ks;
  splits = (sp.num_k_blocks + per - 1) / per;          // no empty rank
  sp.kb_per_split = per;
  sp.splits = splits;
  CUtensorMap mw, ma;
  int rc;
  if ((rc = make_map_2d_u8(&mw, W, K, N, ldw, kBlockKBytes, 128, true))) return rc;
  if ((rc = make_map_2d_u8(&ma, A, K, M, lda, kBlockKBytes, mt, true))) return rc;
  if (blk) {
    switch (mt) {
      case 32: return launch_skinny<32, true>(sp, mw, ma, n_tiles, st);
      case 64: return launch_skinny<64, true>(sp, mw, ma, n_tiles, st, m_tiles);
      default: return launch_skinny<128, true>(sp, mw, ma, n_tiles, st);
    }
  }
  switch (mt) {
    case 32: return launch_skinny<32>(sp, mw, ma, n_tiles, st);
    case 64: return launch_skinny<64>(sp, mw, ma, n_tiles, st);
    case 128: return launch_skinny<128>(sp, mw, ma, n_tiles, st, m_tiles);
    default: return launch_skinny<256>(sp, mw, ma, n_tiles, st, m_tiles);
  }
}

// Small-M MXFP4 GEMM (M <= 128).  See gemm_skinny_mx4.cuh.
template <int MT>
static int launch_skinny_mx4(SkinnyMx4Params sp, const CUtensorMap& mw, const CUtensorMap& ma, const CUtensorMap& msw,
                             const CUtensorMap& msa, int n_tiles, cudaStream_t st, int m_tiles = 1) {
  using L = SkinnyMx4Smem<MT>;
  auto kernel = gemm_skinny_mx4_kernel<MT>;
  static PerDeviceOnce once;
  const cudaError_t attr_err = once.get([&](PerDeviceOnce::Entry& x) {
    x.err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotalBytes);
  }).err;
  if (attr_err != cudaSuccess)
    return fail(VXK_ERR_CUDA, "cudaFuncSetAttribute(smem=%d): %s", L::kTotalBytes, cudaGetErrorString(attr_err));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(n_tiles, sp.splits, m_tiles);
  cfg.blockDim = dim3(kSkinnyThreads);
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
  const cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, sp, mw, ma, msw, msa);
  if (e != cudaSuccess) return fail(VXK_ERR_CUDA, "gemm_skinny_mx4 launch: %s", cudaGetErrorString(e));
  g_gemm_tc_calls.fetch_add(1, std::memory_order_relaxed);
  return check_launch("gemm_skinny_mx4");
}

// A: [M, K/2] bytes, W: [N, K/2] bytes, scales natural [rows, K/32] bytes with 16-byte aligned rows
static int run_skinny_mx4(const TcParams& tp, const void* A, const void* W, const void* sa, const void* sw, int64_t M,
                          int64_t N, int64_t K, int64_t lda, int64_t ldw, int64_t sa0, int64_t sw0, cudaStream_t st) {
  SkinnyMx4Params sp{};
  sp.D = tp.D; sp.bias = tp.bias; sp.M = (int)M; sp.N = (int)N; sp.ldd = tp.ldd;
  sp.out_dtype = tp.out_dtype; sp.bias_dtype = tp.bias_dtype;
  sp.num_k_blocks = tp.num_k_blocks;
  const int n_tiles = (int)((N + 127) / 128);
  static const int env_splits = env_int("VXK_SKINNY_SPLITS", 0);
  const int mt = M <= 32 ? 32 : (M <= 64 ? 64 : 128);
  const int m_tiles = static_cast<int>((M + mt - 1) / mt);   // > 1 only for 128 < M <= 512 (skinny_mx4_token_tiles)
  constexpr int ks = 2;
  int splits = sm_count() / (n_tiles * m_tiles);
  if (splits > 8) splits = 8;
  if (env_splits > 0) splits = env_splits > 8 ? 8 : env_splits;
  const int max_by_k = sp.num_k_blocks / (2 * ks);
  if (splits > max_by_k) splits = max_by_k;
  if (splits < 1) splits = 1;
  while (splits & (splits - 1)) --splits;   // cluster sizes 1 / 2 / 4 / 8
  int per = (sp.num_k_blocks + splits - 1) / splits;
  per = (per + ks - 1) / ks * ks;
  splits = (sp.num_k_blocks + per - 1) / per;          // no empty rank
  sp.kb_per_split = per;
  sp.splits = splits;
  CUtensorMap mw, ma, msw, msa;
  int rc;
  if ((rc = make_map_2d_u8(&mw, W, K / 2, N, ldw, kBlockKBytes, 128, true))) return rc;
  if ((rc = make_map_2d_u8(&ma, A, K / 2, M, lda, kBlockKBytes, mt, true))) return rc;
  if ((rc = make_map_2d_u8(&msw, sw, K / 32, N, sw0, 16, 128, false))) return rc;
  if ((rc = make_map_2d_u8(&msa, sa, K / 32, M, sa0, 16, mt, false))) return rc;
  switch (mt) {
    case 32: return launch_skinny_mx4<32>(sp, mw, ma, msw, msa, n_tiles, st);
    case 64: return launch_skinny_mx4<64>(sp, mw, ma, msw, msa, n_tiles, st);
    default: return launch_skinny_mx4<128>(sp, mw, ma, msw, msa, n_tiles, st, m_tiles);
  }
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
  const int64_t single = (M + N) * pitch + 256;
  const SfPack sp = sf_pack_plan(M, N, K, false);
  const int64_t pair = sp.a_bytes + sp.b_bytes + 512;
  return single > pair ? single : pair;
}

// single-CTA kernel, scale rows that TMA cannot fetch directly (K/32 not a multiple of 16 bytes, e.g.
// K = 11008 or 2880): both grids are copied once to a 16-byte pitch
static int64_t fp4_canon_pitch(int64_t K) { return ((K + 31) / 32 + 15) / 16 * 16; }
static int64_t fp4_canon_bytes(int64_t M, int64_t N, int64_t K) { return (M + N) * fp4_canon_pitch(K) + 512; }

extern "C" int64_t vxk_fp4_gemm_workspace_bytes(int64_t M, int64_t N, int64_t K) {
  int64_t need = 0;
  if (M > 128) {             // CTA-pair kernel: scale factors re-packed in tcgen05.cp order
    const SfPack sp = sf_pack_plan(M, N, K, true);
    need = sp.a_bytes + sp.b_bytes + 512;
  }
  if ((K / 32) % 16 != 0) {  // (contiguous scale rows assumed; anything else is checked at call time)
    const int64_t c = fp4_canon_bytes(M, N, K);
    need = c > need ? c : need;
  }
  return need;               // 0: the single-CTA kernel reads the caller's scale tensors directly
}

// Fused gate_up projection: fp8_gemm -> silu_and_mul -> 1 x 128 block FP8 quantisation in one kernel (SURVEY 8f #2,
// second half).  See include/vxk_b200.h.  Returns VXK_ERR_UNSUPPORTED for shapes the fused kernel does not take (the
// torch shim then runs the two ops one after the other).
extern "C" int vxk_fp8_gemm_silu_mul_quant(void* out_q, float* out_s, const void* A, const void* B, const float* a_scale,
                                           const float* b_scale, const void* bias, int64_t M, int64_t inter, int64_t K,
                                           int64_t lda, int64_t ldb, int a_scale_mode, int b_scale_mode, int64_t ss0,
                                           int64_t ss1, int fp8_dtype, int mid_dtype, int bias_dtype, vxk_stream_t stream) {
  VXK_REQUIRE(M >= 0 && inter >= 0 && K >= 0, "negative shape");
  VXK_REQUIRE(fp8_dtype == VXK_Q_E4M3 || fp8_dtype == VXK_Q_E5M2, "operands must be e4m3 or e5m2");
  VXK_REQUIRE(mid_dtype == VXK_BF16 || mid_dtype == VXK_F16, "intermediate type must be bf16 or f16");
  VXK_REQUIRE(a_scale && b_scale, "scale pointers must not be null");
  VXK_REQUIRE(a_scale_mode >= 0 && a_scale_mode <= 1 && b_scale_mode >= 0 && b_scale_mode <= 1,
              "per-tensor / per-token / per-channel scales only");
  if (M == 0 || inter == 0) return VXK_OK;
  const int64_t N = 2 * inter;
  if (force_fallback() || inter % 128 != 0 || !tma_ok(A, lda) || !tma_ok(B, ldb) || K < 32 || !aligned16(out_q) ||
      M >= (1LL << 31) || N >= (1LL << 31) || K >= (1LL << 31))
    return fail(VXK_ERR_UNSUPPORTED, "fused gate_up kernel needs inter %% 128 == 0 and TMA-expressible operands");
  TcParams tp{};
  tp.D = nullptr; tp.bias = bias; tp.M = (int)M; tp.N = (int)N; tp.K_bytes = (int)K; tp.ldd = N;
  tp.a_scale_mode = a_scale_mode; tp.b_scale_mode = b_scale_mode; tp.out_dtype = mid_dtype;
  tp.bias_dtype = bias_dtype; tp.fp8_fmt = fp8_dtype == VXK_Q_E5M2 ? 1 : 0;
  tp.num_m_blocks = (int)((M + kBlockM - 1) / kBlockM);
  tp.num_k_blocks = (int)((K + kBlockKBytes - 1) / kBlockKBytes);
  tp.a_scale = a_scale; tp.b_scale = b_scale;
  tp.fuse_q = out_q; tp.fuse_s = out_s; tp.fuse_ss0 = ss0; tp.fuse_ss1 = ss1; tp.fuse_inter = (int)inter;
  Tc2Params t2{};
  t2.b = tp;
  t2.num_m_pairs = (int)((M + 255) / 256);
  t2.num_n_blocks = (int)(inter / 128);
  CUtensorMap ma, mb, msa{}, msb{};
  int rc;
  if ((rc = make_map_2d_u8(&ma, A, K, M, lda, kBlockKBytes, kBlockM, true))) return rc;
  if ((rc = make_map_2d_u8(&mb, B, K, N, ldb, kBlockKBytes, 128, true))) return rc;
  cudaStream_t st = stream;
  return mid_dtype == VXK_BF16 ? launch_tc2<KIND_F8, 256, 1, 1>(t2, ma, mb, msa, msb, st)
                               : launch_tc2<KIND_F8, 256, 1, 2>(t2, ma, mb, msa, msb, st);
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
  const PeerCfg peers = g_peer_cfg;
  g_peer_cfg.n = 0;
  if (peers.n > 0) {
    VXK_REQUIRE(operands_ok && !peers.multicast || (operands_ok && a_scale_mode != 2),
                "fused all-gather needs TMA-expressible operands (multicast: no block scales)");
    VXK_REQUIRE(a_scale_mode != 2 || (M > 128 && ((scale_dtype == VXK_S_F32 && group_k % 128 == 0 && group_n % 32 == 0 &&
                                                  K % group_k == 0) ||
                                                 (scale_dtype == VXK_S_E8M0 && group_k % 32 == 0 && workspace &&
                                                  workspace_bytes >= vxk_fp8_gemm_workspace_bytes(M, N, K, 2, scale_dtype)))),
                "fused all-gather with block scales needs M > 128 and a tensor-core-eligible scale grid");
    tp.n_peers = peers.n;
    tp.peer_multicast = peers.multicast;
    // a multicast mapping only takes the whole-row-segment stores of the CTA-pair kernel
    if (peers.multicast)
      VXK_REQUIRE(M > 256 && N % 256 == 0 && tp.vec_store_ok && !force_single(),
                  "multicast all-gather needs M > 256, N %% 256 == 0 and 16-byte aligned output rows");
    for (int i = 0; i < peers.n; ++i) {
      tp.peer_D[i] = peers.ptr[i];
      tp.vec_store_ok = tp.vec_store_ok && aligned16(peers.ptr[i]);
    }
  }
  CUtensorMap ma, mb, msa{}, msb{};
  int rc;

  if (operands_ok && a_scale_mode != 2) {
    tp.a_scale = static_cast<const float*>(a_scale);
    tp.b_scale = static_cast<const float*>(b_scale);
    static const int skinny_max_m = env_int("VXK_SKINNY_MAX_M", 512);
    // 129..256 tokens: only while the 128-row weight tiles fit one wave (else the general kernels win);
    // 257..512 tokens: only as token tiles of 128 (skinny_token_tiles)
    const int n_tiles128 = static_cast<int>((N + 127) / 128);
    if (peers.n == 0 && M <= skinny_max_m && K >= 512 &&
        (M <= 128 || (M <= 256 && n_tiles128 <= sm_count()) || (M <= 512 && skinny_token_tiles(M, n_tiles128) > 1)))
      return run_skinny_f8(tp, A, B, M, N, K, lda, b_stride_n, st);
    const int bn = single_cta_bn(M, N, 256);
    if (!force_single() && (prefer_pair(M, N, bn, 256) || peers.multicast)) {
      Tc2Params t2{};
      t2.b = tp;
      const bool two = prefer_two_acc(M, N, K, 256);
      t2.num_m_pairs = (int)((M + 255) / 256);
      t2.num_n_blocks = (int)((N + (two ? 511 : 255)) / (two ? 512 : 256));
      if ((rc = make_map_2d_u8(&ma, A, K, M, lda, kBlockKBytes, kBlockM, true))) return rc;
      if ((rc = make_map_2d_u8(&mb, B, K, N, b_stride_n, kBlockKBytes, 128, true))) return rc;
      return two ? launch_tc2<KIND_F8, 256, 2>(t2, ma, mb, msa, msb, st)
                 : launch_tc2<KIND_F8, 256, 1>(t2, ma, mb, msa, msb, st);
    }
    tp.num_n_blocks = (int)((N + bn - 1) / bn);
    if ((rc = make_map_2d_u8(&ma, A, K, M, lda, kBlockKBytes, kBlockM, true))) return rc;
    if ((rc = make_map_2d_u8(&mb, B, K, N, b_stride_n, kBlockKBytes, bn, true))) return rc;
    return launch_tc_bn<KIND_F8>(bn, tp, ma, mb, msa, msb, st);
  }
  // fp32 block scales at M <= 128 (config 5 in the decode regime): the small-M kernel with per-slice promotion
  static const int skinny_blk = env_int("VXK_SKINNY_BLK", 1);
  if (operands_ok && a_scale_mode == 2 && scale_dtype == VXK_S_F32 && group_k % 128 == 0 && K % group_k == 0 &&
      group_n > 0 && K >= 512 && peers.n == 0 && skinny_blk &&
      (M <= 64 || skinny_blk_token_tiles(M, static_cast<int>((N + 127) / 128)) > 1)) {
    // (a 128-token tile of this kernel measured slower than the general kernel; 65..256 tokens run as 64-token tiles
    //  side by side while they fit one wave)
    SkinnyBlkScales bs{static_cast<const float*>(a_scale), static_cast<const float*>(b_scale), sa0, sa1, sb0, sb1,
                       static_cast<int>(group_k), static_cast<int>(group_n)};
    return run_skinny_f8(tp, A, B, M, N, K, lda, b_stride_n, st, &bs);
  }
  if (operands_ok && a_scale_mode == 2 && scale_dtype == VXK_S_F32 && group_k % 128 == 0 && group_n % 32 == 0 &&
      K % group_k == 0) {
    // fp32 block scales: two-level accumulation (tensor core per K slice, CUDA cores across slices)
    static const int blk_pipe = env_int("VXK_BLK_PIPE", 1);
    // tile width: 192 (two TMEM buffers) or 160 (three); the fused all-gather pushes 384-byte row segments: 192 only
    static const int env_blk_bn = env_int("VXK_BLK_BN", 192);
    const int BN = (env_blk_bn == 160 && peers.n == 0) ? 160 : 192;
    Tc2BlkParams t2{};
    t2.b = tp;
    t2.sa = static_cast<const float*>(a_scale);
    t2.sb = static_cast<const float*>(b_scale);
    t2.sa0 = sa0; t2.sa1 = sa1; t2.sb0 = sb0; t2.sb1 = sb1;
    t2.group_k_blocks = (int)(group_k / 128);
    t2.group_n = (int)group_n;
    t2.num_m_pairs = (int)((M + 255) / 256);
    t2.num_n_blocks = (int)((N + BN - 1) / BN);
    if ((rc = make_map_2d_u8(&ma, A, K, M, lda, kBlockKBytes, kBlockM, true))) return rc;
    if ((rc = make_map_2d_u8(&mb, B, K, N, b_stride_n, kBlockKBytes, BN / 2, true))) return rc;
    if (BN == 160) return launch_tc2_blk<160, true>(t2, ma, mb, st);
    return blk_pipe ? launch_tc2_blk<192, true>(t2, ma, mb, st) : launch_tc2_blk<192, false>(t2, ma, mb, st);
  }
  if (operands_ok && a_scale_mode == 2 && scale_dtype == VXK_S_E8M0 && group_k % 32 == 0 && workspace &&
      workspace_bytes >= vxk_fp8_gemm_workspace_bytes(M, N, K, a_scale_mode, scale_dtype)) {
    const int bn = single_cta_bn(M, N, 192);
    if (!force_single() && (prefer_pair(M, N, bn, 192) || peers.n > 0)) {
      Tc2Params t2{};
      t2.b = tp;
      return run_pair_block_scaled<KIND_MXF8>(t2, A, lda, B, b_stride_n, M, N, K, static_cast<const uint8_t*>(a_scale),
                                              sa0, group_k / 32, sa1, static_cast<const uint8_t*>(b_scale), group_n,
                                              sb1, group_k / 32, sb0, workspace, st);
    }
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
    tp.num_n_blocks = (int)((N + bn - 1) / bn);
    if ((rc = make_map_2d_u8(&ma, A, K, M, lda, kBlockKBytes, kBlockM, true))) return rc;
    if ((rc = make_map_2d_u8(&mb, B, K, N, b_stride_n, kBlockKBytes, bn, true))) return rc;
    if ((rc = make_map_2d_u8(&msa, sfa, pitch, M, pitch, 16, kBlockM, false))) return rc;
    if ((rc = make_map_2d_u8(&msb, sfb, pitch, N, pitch, 16, bn, false))) return rc;
    return launch_tc_bn<KIND_MXF8>(bn, tp, ma, mb, msa, msb, st);
  }
  return launch_gemm_fallback(p, /*is_fp4=*/false, st);
}

extern "C" int vxk_fp8_gemm_allgather(void* const* peer_out, int n_peers, int self_rank, void* multicast_out,
                                      int64_t col_offset, int64_t out_ld, const void* A, const void* B,
                                      const void* a_scale, const void* b_scale, const void* bias, int64_t M, int64_t N,
                                      int64_t K, int64_t lda, int64_t b_stride_k, int64_t b_stride_n, int a_scale_mode,
                                      int b_scale_mode, int64_t group_k, int64_t group_n, int64_t sa0, int64_t sa1,
                                      int64_t sb0, int64_t sb1, int scale_dtype, int fp8_dtype, int out_dtype,
                                      int bias_dtype, void* workspace, int64_t workspace_bytes, vxk_stream_t stream) {
  VXK_REQUIRE(peer_out && n_peers >= 1 && n_peers <= 8, "1..8 peer outputs");
  VXK_REQUIRE(col_offset >= 0 && col_offset + N <= out_ld, "this rank's columns must lie inside a row of the output");
  VXK_REQUIRE(self_rank >= 0 && self_rank < n_peers, "self_rank must index peer_out");
  int others = 0;
  for (int i = 0; i < n_peers; ++i) {
    VXK_REQUIRE(peer_out[i], "null peer output");
    if (i != self_rank) g_peer_cfg.ptr[others++] = static_cast<uint16_t*>(peer_out[i]) + col_offset;
  }
  g_peer_cfg.n = others;
  g_peer_cfg.multicast = 0;
  if (multicast_out && others > 0) {     // one NVLS store instead of (ranks - 1) unicast stores
    g_peer_cfg.ptr[0] = static_cast<uint16_t*>(multicast_out) + col_offset;
    g_peer_cfg.n = 1;
    g_peer_cfg.multicast = 1;
  }
  void* own = static_cast<uint16_t*>(peer_out[self_rank]) + col_offset;
  const int rc = vxk_fp8_gemm(own, A, B, a_scale, b_scale, bias, M, N, K, lda, b_stride_k, b_stride_n, out_ld,
                              a_scale_mode, b_scale_mode, group_k, group_n, sa0, sa1, sb0, sb1, scale_dtype, fp8_dtype,
                              out_dtype, bias_dtype, workspace, workspace_bytes, stream);
  g_peer_cfg.n = 0;
  return rc;
}

extern "C" int64_t vxk_mx_sf_packed_bytes(int64_t rows, int64_t K, int kind, int operand) {
  if (rows < 0 || K < 0 || (kind != 0 && kind != 1) || (operand != 0 && operand != 1)) return -1;
  const SfPack sp = operand == 0 ? sf_pack_plan(rows, 0, K, kind == 1) : sf_pack_plan(0, rows, K, kind == 1);
  return operand == 0 ? sp.a_bytes : sp.b_bytes;
}

extern "C" int vxk_mx_sf_pack(void* dst, const void* src, int64_t rows, int64_t K, int64_t s_row, int64_t s_k,
                              int kind, int operand, vxk_stream_t stream) {
  VXK_REQUIRE(dst && src, "null pointer");
  VXK_REQUIRE(rows >= 0 && K >= 0 && K % 32 == 0, "rows >= 0 and K a multiple of 32");
  VXK_REQUIRE((kind == 0 || kind == 1) && (operand == 0 || operand == 1), "kind / operand must be 0 or 1");
  VXK_REQUIRE(reinterpret_cast<uintptr_t>(dst) % 16 == 0, "packed scale buffer must be 16-byte aligned");
  const SfPack sp = operand == 0 ? sf_pack_plan(rows, 0, K, kind == 1) : sf_pack_plan(0, rows, K, kind == 1);
  const int nblk = operand == 0 ? sp.a_blocks : sp.b_blocks;
  const int words = s_k == 1 && s_row % 4 == 0 && reinterpret_cast<uintptr_t>(src) % 4 == 0;
  sf_pack_kernel<<<8 * sm_count(), 256, 0, stream>>>(static_cast<uint32_t*>(dst), rows, K / 32, sp.num_kb, nblk,
                                                     sp.sf_words, static_cast<const uint8_t*>(src), 1, s_row, 1, s_k, words);
  return check_launch("mx_sf_pack");
}

extern "C" uint64_t vxk_sf_pack_launches(void) { return g_sf_pack_calls.load(std::memory_order_relaxed); }

static int fp4_gemm_impl(void* D, const void* A, const void* B, const void* a_scale, const void* b_scale,
                         const void* bias, int64_t M, int64_t N, int64_t K, int64_t lda_bytes, int64_t ldb_bytes,
                         int64_t ldd, int64_t sa0, int64_t sa1, int64_t sb0, int64_t sb1, int out_dtype,
                         int bias_dtype, void* workspace, int64_t workspace_bytes, vxk_stream_t stream,
                         const void* a_packed, const void* b_packed) {
  VXK_REQUIRE(M >= 0 && N >= 0 && K >= 0, "negative shape");
  VXK_REQUIRE(reinterpret_cast<uintptr_t>(a_packed) % 16 == 0 && reinterpret_cast<uintptr_t>(b_packed) % 16 == 0,
              "packed scale buffers must be 16-byte aligned");
  VXK_REQUIRE(K % 32 == 0, "K must be a multiple of the MX block (32)");
  VXK_REQUIRE(out_dtype == VXK_BF16 || out_dtype == VXK_F16, "output must be bf16 or f16");
  VXK_REQUIRE(a_scale && b_scale, "scale pointers must not be null");
  if (M == 0 || N == 0) return VXK_OK;
  GemmParams p{D, A, B, a_scale, b_scale, bias, M, N, K, lda_bytes, 1, ldb_bytes, ldd, 2, 2, 32, 1, sa0, sa1, sb0, sb1,
               VXK_S_E8M0, VXK_Q_E4M3, out_dtype, bias_dtype};
  cudaStream_t st = stream;
  // the operands must be TMA-expressible; the scale grids only when the single-CTA kernel fetches them directly
  const bool ok = !force_fallback() && tma_ok(A, lda_bytes) && tma_ok(B, ldb_bytes) && K >= 64 && M < (1LL << 31) &&
                  N < (1LL << 31);
  const bool scales_direct = sa1 == 1 && sb1 == 1 && tma_ok(a_scale, sa0) && tma_ok(b_scale, sb0);
  const PeerCfg peers = g_peer_cfg;
  g_peer_cfg.n = 0;
  if (peers.n > 0)
    VXK_REQUIRE(ok && M > 128 && !force_single() && !peers.multicast && workspace &&
                    workspace_bytes >= vxk_fp4_gemm_workspace_bytes(M, N, K),
                "fused all-gather: needs the CTA-pair MXFP4 path (M > 128, TMA-expressible operands, workspace)");
  if (!ok) return launch_gemm_fallback(p, /*is_fp4=*/true, st);
  TcParams tp{};
  tp.D = D; tp.bias = bias; tp.M = (int)M; tp.N = (int)N; tp.K_bytes = (int)(K / 2); tp.ldd = ldd;
  tp.n_peers = peers.n;
  for (int i = 0; i < peers.n; ++i) tp.peer_D[i] = peers.ptr[i];
  tp.a_scale_mode = 2; tp.b_scale_mode = 2; tp.out_dtype = out_dtype; tp.bias_dtype = bias_dtype;
  tp.num_m_blocks = (int)((M + kBlockM - 1) / kBlockM);
  const int bn = single_cta_bn(M, N, 192);
  tp.num_n_blocks = (int)((N + bn - 1) / bn);
  tp.num_k_blocks = (int)((K / 2 + kBlockKBytes - 1) / kBlockKBytes);
  tp.vec_store_ok = aligned16(D) && (ldd * 2) % 16 == 0;
  for (int i = 0; i < peers.n; ++i) tp.vec_store_ok = tp.vec_store_ok && aligned16(peers.ptr[i]);
  // decode regime: swap-AB split-K kernel straight from the caller's scale tensors
  static const int skinny_mx4 = env_int("VXK_SKINNY_MX4", 1);
  // (65..128 tokens: only while the 128-row weight tiles fit one wave -- measured, profiles/r02_mxfp4_smallm.txt)
  // 129..512 tokens: two to four 128-token tiles side by side (gridDim.z) while they fit one wave
  static const int skinny_mx4_max_m = env_int("VXK_SKINNY_MX4_MAX_M", 512);
  const int64_t n_tiles128 = (N + 127) / 128;
  if (skinny_mx4 && peers.n == 0 && K >= 1024 && scales_direct &&
      (M <= 64 || (M <= 128 && n_tiles128 <= sm_count()) ||
       (M <= 512 && M <= skinny_mx4_max_m && n_tiles128 * ((M + 127) / 128) <= sm_count())))
    return run_skinny_mx4(tp, A, B, a_scale, b_scale, M, N, K, lda_bytes, ldb_bytes, sa0, sb0, st);
  const bool ws_ok = (a_packed && b_packed) || (workspace && workspace_bytes >= vxk_fp4_gemm_workspace_bytes(M, N, K));
  if (!force_single() && (prefer_pair(M, N, bn, 192) || peers.n > 0) && ws_ok) {
    Tc2Params t2{};
    t2.b = tp;
    return run_pair_block_scaled<KIND_MXF4>(t2, A, lda_bytes, B, ldb_bytes, M, N, K,
                                            static_cast<const uint8_t*>(a_scale), sa0, 1, sa1,
                                            static_cast<const uint8_t*>(b_scale), 1, sb0, 1, sb1, workspace, st,
                                            a_packed, b_packed);
  }
  CUtensorMap ma, mb, msa, msb;
  int rc;
  const void* sfa = a_scale;
  const void* sfb = b_scale;
  int64_t pa = sa0, pb = sb0;
  if (!scales_direct) {
    if (!workspace || workspace_bytes < fp4_canon_bytes(M, N, K)) return launch_gemm_fallback(p, /*is_fp4=*/true, st);
    const int64_t groups32 = K / 32, pitch = fp4_canon_pitch(K);
    uint8_t* ws = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
    uint8_t* wa = ws;
    uint8_t* wb = ws + M * pitch;
    const int blocks = 4 * sm_count();
    canon_sf_kernel<<<blocks, 256, 0, st>>>(wa, M, groups32, pitch, static_cast<const uint8_t*>(a_scale), 1, sa0, 1, sa1);
    if ((rc = check_launch("canon_sfa"))) return rc;
    canon_sf_kernel<<<blocks, 256, 0, st>>>(wb, N, groups32, pitch, static_cast<const uint8_t*>(b_scale), 1, sb0, 1, sb1);
    if ((rc = check_launch("canon_sfb"))) return rc;
    sfa = wa; sfb = wb; pa = pb = pitch;
  }
  if ((rc = make_map_2d_u8(&ma, A, K / 2, M, lda_bytes, kBlockKBytes, kBlockM, true))) return rc;
  if ((rc = make_map_2d_u8(&mb, B, K / 2, N, ldb_bytes, kBlockKBytes, bn, true))) return rc;
  if ((rc = make_map_2d_u8(&msa, sfa, scales_direct ? K / 32 : pa, M, pa, 16, kBlockM, false))) return rc;
  if ((rc = make_map_2d_u8(&msb, sfb, scales_direct ? K / 32 : pb, N, pb, 16, bn, false))) return rc;
  return launch_tc_bn<KIND_MXF4>(bn, tp, ma, mb, msa, msb, st);
}

extern "C" int vxk_fp4_gemm(void* D, const void* A, const void* B, const void* a_scale, const void* b_scale,
                            const void* bias, int64_t M, int64_t N, int64_t K, int64_t lda_bytes, int64_t ldb_bytes,
                            int64_t ldd, int64_t sa0, int64_t sa1, int64_t sb0, int64_t sb1, int out_dtype,
                            int bias_dtype, void* workspace, int64_t workspace_bytes, vxk_stream_t stream) {
  return fp4_gemm_impl(D, A, B, a_scale, b_scale, bias, M, N, K, lda_bytes, ldb_bytes, ldd, sa0, sa1, sb0, sb1,
                       out_dtype, bias_dtype, workspace, workspace_bytes, stream, nullptr, nullptr);
}

// ---- vxk_fp4_gemm_static_scratch: vxk_fp4_gemm for callers that may not allocate (include/vxk_b200.h) ----
// The scale-factor re-packing scratch lives in a file-scope device array instead of a caller buffer.  Every device
// that loads the library holds these 4 MiB until the process exits.  Nothing in it has to start at zero:
// sf_pack_kernel / canon_sf_kernel rewrite every word of the region a call uses (zeros outside the tensor) before
// the GEMM reads it, so the previous call's contents are never seen.  All calls on a device share the array: two
// calls that overlap on the GPU (different streams) race.
#define VXK_FP4_STATIC_SCRATCH_BYTES (4 << 20)
static __device__ __align__(256) unsigned char g_fp4_static_scratch[VXK_FP4_STATIC_SCRATCH_BYTES];

// Device address of the array on the current device.  In host code `g_fp4_static_scratch` is only the host-side
// stand-in of the symbol; cudaGetSymbolAddress (once per device, with that device current) gives the address a
// kernel can use.
static int fp4_static_scratch_ptr(void** out) {
  static std::mutex mu;
  static void* ptr[64] = {nullptr};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(VXK_ERR_CUDA, "cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= 64) return fail(VXK_ERR_UNSUPPORTED, "static scratch: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> g(mu);
  if (!ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_fp4_static_scratch);
    if (e != cudaSuccess || !p)
      return fail(VXK_ERR_CUDA, "cudaGetSymbolAddress(static scratch) on device %d: %s", dev, cudaGetErrorString(e));
    ptr[dev] = p;
  }
  *out = ptr[dev];
  return VXK_OK;
}

extern "C" int64_t vxk_fp4_gemm_static_scratch_bytes(void) { return VXK_FP4_STATIC_SCRATCH_BYTES; }

extern "C" int vxk_fp4_gemm_static_scratch(void* D, const void* A, const void* B, const void* a_scale,
                                           const void* b_scale, const void* bias, int64_t M, int64_t N, int64_t K,
                                           int64_t lda_bytes, int64_t ldb_bytes, int64_t ldd, int64_t sa0, int64_t sa1,
                                           int64_t sb0, int64_t sb1, int out_dtype, int bias_dtype,
                                           vxk_stream_t stream) {
  VXK_REQUIRE(M >= 0 && N >= 0 && K >= 0, "negative shape");
  const int64_t need = vxk_fp4_gemm_workspace_bytes(M, N, K);
  if (need + 256 > VXK_FP4_STATIC_SCRATCH_BYTES)
    return fail(VXK_ERR_INVALID,
                "vxk_fp4_gemm_static_scratch: M = %lld, N = %lld, K = %lld needs %lld bytes of scale-factor scratch, the "
                "static array holds %lld; call vxk_fp4_gemm with a caller-provided workspace",
                (long long)M, (long long)N, (long long)K, (long long)(need + 256),
                (long long)VXK_FP4_STATIC_SCRATCH_BYTES);
  void* ws = nullptr;
  const int rc = fp4_static_scratch_ptr(&ws);
  if (rc) return rc;
  return fp4_gemm_impl(D, A, B, a_scale, b_scale, bias, M, N, K, lda_bytes, ldb_bytes, ldd, sa0, sa1, sb0, sb1,
                       out_dtype, bias_dtype, ws, VXK_FP4_STATIC_SCRATCH_BYTES, stream, nullptr, nullptr);
}

extern "C" int vxk_fp4_gemm_sf(void* D, const void* A, const void* B, const void* a_scale, const void* b_scale,
                               const void* a_scale_packed, const void* b_scale_packed, const void* bias, int64_t M,
                               int64_t N, int64_t K, int64_t lda_bytes, int64_t ldb_bytes, int64_t ldd, int64_t sa0,
                               int64_t sa1, int64_t sb0, int64_t sb1, int out_dtype, int bias_dtype, void* workspace,
                               int64_t workspace_bytes, vxk_stream_t stream) {
  return fp4_gemm_impl(D, A, B, a_scale, b_scale, bias, M, N, K, lda_bytes, ldb_bytes, ldd, sa0, sa1, sb0, sb1,
                       out_dtype, bias_dtype, workspace, workspace_bytes, stream, a_scale_packed, b_scale_packed);
}

extern "C" int vxk_fp4_gemm_allgather(void* const* peer_out, int n_peers, int self_rank, int64_t col_offset,
                                      int64_t out_ld, const void* A, const void* B, const void* a_scale,
                                      const void* b_scale, const void* bias, int64_t M, int64_t N, int64_t K,
                                      int64_t lda_bytes, int64_t ldb_bytes, int64_t sa0, int64_t sa1, int64_t sb0,
                                      int64_t sb1, int out_dtype, int bias_dtype, void* workspace,
                                      int64_t workspace_bytes, vxk_stream_t stream) {
  VXK_REQUIRE(peer_out && n_peers >= 1 && n_peers <= 8, "1..8 peer outputs");
  VXK_REQUIRE(self_rank >= 0 && self_rank < n_peers, "self_rank must index peer_out");
  VXK_REQUIRE(col_offset >= 0 && col_offset + N <= out_ld, "this rank's columns must lie inside a row of the output");
  int others = 0;
  for (int i = 0; i < n_peers; ++i) {
    VXK_REQUIRE(peer_out[i], "null peer output");
    if (i != self_rank) g_peer_cfg.ptr[others++] = static_cast<uint16_t*>(peer_out[i]) + col_offset;
  }
  g_peer_cfg.n = others;
  g_peer_cfg.multicast = 0;
  void* own = static_cast<uint16_t*>(peer_out[self_rank]) + col_offset;
  const int rc = vxk_fp4_gemm(own, A, B, a_scale, b_scale, bias, M, N, K, lda_bytes, ldb_bytes, out_ld, sa0, sa1, sb0,
                              sb1, out_dtype, bias_dtype, workspace, workspace_bytes, stream);
  g_peer_cfg.n = 0;
  return rc;
}
