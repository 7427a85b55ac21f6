// MoE grouped GEMM on tcgen05 (device side): for every expert e
//
//   D[rows of e, :] = A[rows of e, :] * dequant(B[e]) (+ bias[e])          A, D: bf16 / f16, fp32 accumulate
//
//   ref: cutlass_grouped_gemm_interface  csrc/xpu/grouped_gemm/grouped_gemm_interface.cpp:12-59
//        -> cutlass_grouped_gemm_xe2_impl  csrc/xpu/grouped_gemm/xe_2/grouped_gemm_xe2_interface.hpp:160-357
//        recipes (vllm_xpu_kernels/fused_moe_interface.py:32-45): "bf16" (16-bit weights [E, K, N]),
//        "fp8" (e4m3 / e5m2 weights [E, K, N] + one fp32 scale per expert), "mxfp4" (packed E2M1 weights
//        [E, N, K/2] + E8M0 scale bytes [E, N, K/32]); activations stay 16-bit in all of them.
//
// rows_per_expert lives on the device (no host sync), so the tile list is built inside the kernel: every CTA
// scans rows_per_expert once (warp 3), and all roles walk the same persistent tile sequence
//   tile = (expert e, token tile t of MT rows, weight tile of 128 output columns)
// ordered so that consecutive CTAs share the weight tile (t fastest).
//
// Operands are swapped like the small-M kernels (UMMA M = 128 weight rows, UMMA N = MT tokens, the accumulator is
// D^T): a token tile may start at any row (TMA box at an arbitrary row coordinate; rows of the next expert that
// ride along are computed and never stored), MT is chosen from the average rows per expert like the reference's
// policies (grouped_gemm_xe2_interface.hpp:275-357), and a warp's D^T rows are consecutive output columns, so the
// epilogue's stores are contiguous.
//
//   warp 0      TMA producer: weight tile + MT x 64 activation box of one 64-deep K slice per ring stage
//   warp 1      MMA issuer: tcgen05.mma.kind::f16 128 x MT x 16, four per stage, two TMEM accumulators
//   warp 2      TMEM allocator          warp 3  tile-list scan
//   warps 4-7   epilogue: tcgen05.ld, per-expert scale, bias, 16-bit, lane pairs swap so a lane stores 2 columns
//   warps 8-23  converters (fp8 / mxfp4 only): raw tile -> 16-bit operand tile (exact widening)
//
// Weight operand layouts:
//   [E, K, N] weights (16-bit, fp8) are N-contiguous = "MN-major" for the tensor core: the operand tile is
//   [2 blocks of 64 columns][64 k rows][128 B], SWIZZLE_128B; for 16-bit weights TMA writes it directly
//   (two boxes), for fp8 the converters write it from a raw 64 x 128 B box.
//   [E, N, K/2] mxfp4 weights are K-contiguous: raw box 128 rows x 32 B, widened into the usual K-major tile; the
//   E8M0 scale is folded into the widening (value * 2^(s-127) is exact in bf16; for f16 the product is rounded
//   the way the reference test dequantises, tests/fused_moe/test_grouped_gemm.py:341-377).
#pragma once

#include "gemm_w8a16.cuh"

namespace vxk {

constexpr int GW_16 = 0, GW_FP8 = 1, GW_MX4 = 2;
constexpr int GW_MX4T = 3;   // mxfp4 with the widened operand in TMEM
constexpr int GW_FP8T = 4;   // fp8 weights [E, N, K] (K-contiguous: the dense fp8_gemm_w8a16 layout), operand in TMEM
constexpr int kGgBlockK = 64;
#ifndef VXK_GG_STAGE_CAP
#define VXK_GG_STAGE_CAP 10
#endif
#ifndef VXK_GG_CONV_SLOTS
#define VXK_GG_CONV_SLOTS 3
#endif
constexpr int kGgMaxExperts = 512;
constexpr int kGgConvWarps = 16;          // 4 per scheduler: the widening chains are latency bound with fewer
constexpr int kGgConvThreads = 256 + 32 * kGgConvWarps;

struct GroupedParams {
  void* D;
  const void* bias;        // [E, N] activation dtype, or null
  const float* w_scale;    // GW_FP8: [E]
  const uint8_t* w_sf;     // GW_MX4: [E, N, K / 32]
  const int32_t* rows;     // [E]; null: one expert owning all total_m rows (dense call)
  int scale_per_channel;   // GW_FP8T: w_scale is [E, N] (per output channel) instead of [E]
  const float* blk_scale;  // GW_FP8T, dense call only: 2-D block scales, folded in while widening (w_scale unused);
  int group_k, group_n;    //   scale of (k, n) = blk_scale[(k / group_k) * sb0 + (n / group_n) * sb1], group_k % 64 == 0
  int64_t sb0, sb1;
  int total_m, N, K, E;
  int num_kb, num_nt;
  int64_t ldd;
  int sf_vec;              // GW_MX4: scale rows are 16-byte aligned and K / 32 is a multiple of 16
  int dbg;                 // VXK_GG_DEBUG (profiling only, shared-memory operand paths; results are garbage): 1 = no widening
                           // arithmetic, 2 = no proxy fence, 4 = converters only pass the barriers on
};

__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1,
                                            int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}

// D[tmem] (+)= A[tmem] * B[smem]: the 128-row operand is read from tensor memory (lane = row, one 32-bit column =
// two consecutive k), which takes it off the shared-memory read path that bounds the small token tiles
__device__ __forceinline__ void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// registers -> TMEM: this warp's 32 lanes x 16 consecutive columns
__device__ __forceinline__ void tmem_st_32x32b_x16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// MN-major operand tile, SWIZZLE_128B: 64-element column blocks `lbo` bytes apart, groups of 8 k rows 1024 B apart
__device__ __forceinline__ uint64_t make_mnmajor_sw128_desc(uint32_t smem_addr, uint32_t lbo) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr >> 4) & 0x3fffu);
  d |= static_cast<uint64_t>((lbo >> 4) & 0x3fffu) << 16;
  d |= static_cast<uint64_t>(1024u >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}

// 8 E2M1 codes (one 32-bit word, even element in the low nibble) -> four packed 16-bit pairs, times the group
// scale `sc2` (packed pair of the activation type)
template <bool BF16>
__device__ __forceinline__ void e2m1x8_widen(uint32_t word, uint32_t sc2, uint32_t (&w)[4]) {
  uint32_t h[4];
  asm("{\n\t.reg .b8 b0, b1, b2, b3;\n\tmov.b32 {b0, b1, b2, b3}, %4;\n\t"
      "cvt.rn.f16x2.e2m1x2 %0, b0;\n\tcvt.rn.f16x2.e2m1x2 %1, b1;\n\t"
      "cvt.rn.f16x2.e2m1x2 %2, b2;\n\tcvt.rn.f16x2.e2m1x2 %3, b3;\n\t}"
      : "=r"(h[0]), "=r"(h[1]), "=r"(h[2]), "=r"(h[3])
      : "r"(word));
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if constexpr (BF16) {
      // f16 -> f32 (two conversions on the FMA pipe) -> packed bf16 (one pack), all exact.  Moving the f16 fields
      // into bf16 positions with shift / mask / merge + a re-bias multiply is one instruction shorter but all of it
      // sits on the half-rate ALU pipe, which is what bounds the converters (69 % busy, profiles/r02_grouped_mx4t.csv):
      // measured 190 -> 173 us at M = 256.  sc2 = bf16 bits of 2^(s-127), the reference's own factor.
      const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&h[i]));
      const __nv_bfloat162 b = __floats2bfloat162_rn(f.x, f.y);
      asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(w[i]) : "r"(*reinterpret_cast<const uint32_t*>(&b)), "r"(sc2));
    } else {
      asm("mul.rn.f16x2 %0, %1, %2;" : "=r"(w[i]) : "r"(h[i]), "r"(sc2));
    }
  }
}

template <int MT, int WKIND>
struct GgSmem {
  static constexpr bool kTmemA = WKIND == GW_MX4T || WKIND == GW_FP8T;
  static constexpr bool kConv = WKIND != GW_16;
  // K elements per ring stage: 128 with the operand in TMEM, 64 otherwise and for 256-token tiles (whose
  // activation boxes are 32 KB per 64 k)
  static constexpr int kBlockK = (kTmemA && MT <= 128) ? 128 : kGgBlockK;
  static constexpr int kOpBytes = 128 * 128;            // 128 weight rows x 64 elements of 16 bits
  static constexpr int kRawW = WKIND == GW_16 ? kOpBytes : WKIND == GW_FP8 ? 64 * 128
                               : WKIND == GW_FP8T ? 128 * kBlockK : WKIND == GW_MX4T ? 128 * (kBlockK / 2) : 128 * 32;
  static constexpr int kABox = MT * 128;                // one 64-deep activation slice
  static constexpr int kABoxes = kBlockK / 64;
  static constexpr int kStageBytes = (kRawW + kABoxes * kABox + 1023) / 1024 * 1024;
  static constexpr int kTxBytes = kRawW + kABoxes * kABox;
  // Ring depth / operand slots were swept (10 / 15 / 24 stages, 2 / 3 / 6 / 8 slots, 8 / 16 converter warps): no
  // effect.  What bounds the small-MT shapes is the tensor core's read of the 128-row operand from shared memory:
  // 4 KB per tcgen05.mma at ~32 B / clk = ~128 clk per 16-deep step whatever MT is, i.e. 16 weights / clk / SM
  // (measured: 550-600 clk per 64-deep stage with the converters doing nothing, for 16-bit, fp8 and mxfp4 alike;
  // profiles/r02_grouped_gemm.txt).  16-bit weights therefore sit at the HBM roofline (32 B / clk / SM), fp8 is
  // capped at ~4.5 TB/s and mxfp4 at ~2.4 TB/s of weight stream; the way past it is the widened operand in TMEM
  // (GW_MX4T: the converters write tensor memory, four 64-column slots next to the accumulators).
  static constexpr int kConvSlots = !kConv ? 0 : kTmemA ? 4 : VXK_GG_CONV_SLOTS;
  static constexpr int kConvSmemBytes = (kConv && !kTmemA) ? kConvSlots * kOpBytes : 0;
  static constexpr int kSchedBytes = 2 * (kGgMaxExperts + 1) * 4;
  static constexpr int kStagesRaw = (220 * 1024 - kConvSmemBytes - kSchedBytes) / kStageBytes;
  static constexpr int kStages = kStagesRaw > VXK_GG_STAGE_CAP ? VXK_GG_STAGE_CAP : kStagesRaw;
  static constexpr int kBarrierBytes = (2 * kStages + 2 * 8 + 4) * 8 + 16;
  static constexpr int kTotalBytes = kStages * kStageBytes + kConvSmemBytes + kBarrierBytes + kSchedBytes + 1024;
  static_assert(kStages >= 3, "ring too shallow");
  // TMEM-A: operand slots in columns [256, 512), accumulators in [0, 256): two buffers up to MT = 128, one at 256
  static constexpr int kAccBufs = (kTmemA && MT > 128) ? 1 : 2;
  static constexpr int kSlotCols = kBlockK / 2;         // one TMEM column = two consecutive k
};

struct GgTile { int e, row0, rows, n0; };

template <int MT>
__device__ __forceinline__ GgTile gg_tile(const int* tile_start, const int* row_start, int E, int num_nt, int idx) {
  int lo = 0, hi = E - 1;                     // smallest e whose linear range ends after idx
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (tile_start[mid + 1] * num_nt > idx) hi = mid; else lo = mid + 1;
  }
  const int e = lo;
  const int te = tile_start[e + 1] - tile_start[e];
  const int local = idx - tile_start[e] * num_nt;
  const int nt = local / te, t = local - nt * te;
  const int rows_e = row_start[e + 1] - row_start[e];
  GgTile r;
  r.e = e;
  r.row0 = row_start[e] + t * MT;
  r.rows = min(MT, rows_e - t * MT);
  r.n0 = nt * 128;
  return r;
}

template <int MT, int WKIND, bool BF16, bool E5M2>
__global__ void __launch_bounds__(WKIND == GW_16 ? 256 : kGgConvThreads, 1)
gemm_grouped_kernel(const GroupedParams p, const __grid_constant__ CUtensorMap map_w,
                    const __grid_constant__ CUtensorMap map_a) {
  using L = GgSmem<MT, WKIND>;
  constexpr int S = L::kStages;
  constexpr int C = L::kConv ? L::kConvSlots : 1;
  constexpr bool kMnMajor = WKIND == GW_16 || WKIND == GW_FP8;
  constexpr bool kTmemA = L::kTmemA;
  constexpr uint32_t kASlotCol = 256;           // TMEM-A: operand slots live in columns [256, 512)
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  auto stage_w = [&](int s) { return smem + s * L::kStageBytes; };
  auto stage_a = [&](int s) { return smem + s * L::kStageBytes + L::kRawW; };
  uint8_t* conv_base = smem + S * L::kStageBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(conv_base + L::kConvSmemBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + S;
  uint64_t* conv_full = bars + 2 * S;
  uint64_t* conv_empty = bars + 2 * S + 8;
  uint64_t* tmem_full_bar = bars + 2 * S + 16;
  uint64_t* tmem_empty_bar = bars + 2 * S + 18;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 2 * S + 20);
  int* tile_start = reinterpret_cast<int*>(reinterpret_cast<uint8_t*>(bars) + L::kBarrierBytes);
  int* row_start = tile_start + (kGgMaxExperts + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_w);
    tma_prefetch_desc(&map_a);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < S; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 8; ++i) {
      mbar_init(&conv_full[i], kGgConvWarps);
      mbar_init(&conv_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full_bar[i], 1);
      mbar_init(&tmem_empty_bar[i], 128);
    }
    fence_barrier_init();
  }
  constexpr uint32_t kCols = kTmemA ? 512 : (2 * MT < 32 ? 32 : 2 * MT);
  if (warp == 2) tmem_alloc<kCols>(tmem_ptr_smem);
  if (warp == 3) {
    // exclusive prefix sums of rows and of token tiles per expert
    int carry_r = 0, carry_t = 0;
    for (int base = 0; base < p.E; base += 32) {
      const int e = base + lane;
      int r = e < p.E ? (p.rows ? __ldg(p.rows + e) : p.total_m) : 0;
      r = r < 0 ? 0 : r;
      int sr = r, stl = (r + MT - 1) / MT;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int a = __shfl_up_sync(0xffffffffu, sr, o), b = __shfl_up_sync(0xffffffffu, stl, o);
        if (lane >= o) { sr += a; stl += b; }
      }
      if (e < p.E) {
        row_start[e + 1] = carry_r + sr;
        tile_start[e + 1] = carry_t + stl;
      }
      carry_r += __shfl_sync(0xffffffffu, sr, 31);
      carry_t += __shfl_sync(0xffffffffu, stl, 31);
    }
    if (lane == 0) { row_start[0] = 0; tile_start[0] = 0; }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  const int num_tiles = tile_start[p.E] * p.num_nt;

  if (warp == 0) {
    // ===================== TMA producer =====================
    int stage = 0;
    uint32_t phase = 0;
    for (int idx = blockIdx.x; idx < num_tiles; idx += gridDim.x) {
      const GgTile t = gg_tile<MT>(tile_start, row_start, p.E, p.num_nt, idx);
      for (int kb = 0; kb < p.num_kb; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        if (elect_one()) {
          mbar_arrive_expect_tx(&full_bar[stage], L::kTxBytes);
          if constexpr (WKIND == GW_16) {        // byte coordinates: (column bytes, k, expert)
            tma_load_3d(stage_w(stage), &map_w, &full_bar[stage], t.n0 * 2, kb * kGgBlockK, t.e);
            tma_load_3d(stage_w(stage) + 8192, &map_w, &full_bar[stage], t.n0 * 2 + 128, kb * kGgBlockK, t.e);
          } else if constexpr (WKIND == GW_FP8) {
            tma_load_3d(stage_w(stage), &map_w, &full_bar[stage], t.n0, kb * kGgBlockK, t.e);
          } else if constexpr (WKIND == GW_FP8T) {   // (k bytes, weight row, expert)
            tma_load_3d(stage_w(stage), &map_w, &full_bar[stage], kb * L::kBlockK, t.n0, t.e);
          } else {                                // (k bytes, weight row, expert), two codes per byte
            tma_load_3d(stage_w(stage), &map_w, &full_bar[stage], kb * (L::kBlockK / 2), t.n0, t.e);
          }
#pragma unroll
          for (int h = 0; h < L::kABoxes; ++h)
            tma_load_2d(stage_a(stage) + h * L::kABox, &map_a, &full_bar[stage], (kb * L::kABoxes + h) * 128, t.row0);
        }
        __syncwarp();
        if (++stage == S) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer: D^T[128 x MT] += W16[128 x 16] * A[MT x 16]^T =====================
    // (a_format != b_format -- f16 weights against bf16 activations, which would make the fp8 widening one cvt per
    // two codes -- is rejected by the hardware: illegal instruction, measured)
    int stage = 0, cs = 0, iter = 0;
    uint32_t phase = 0, cphase = 0;
    for (int idx = blockIdx.x; idx < num_tiles; idx += gridDim.x, ++iter) {
      // UMMA N = the token rows this tile really has, rounded up to 16: the last token tile of an expert is partial
      // (half a tile wasted per expert on average), and the MMA time scales with N
      const GgTile t = gg_tile<MT>(tile_start, row_start, p.E, p.num_nt, idx);
      const int umma_n = min(MT, max(32, (t.rows + 15) & ~15));   // (N = 16 measured slower than 32 on the fp8 path)
      const uint32_t idesc = make_idesc_f16(BF16 ? 1 : 0, 128, umma_n) | (kMnMajor ? (1u << 15) : 0u);
      const int as = iter % L::kAccBufs;
      const uint32_t aphase = (iter / L::kAccBufs) & 1;
      mbar_wait(&tmem_empty_bar[as], aphase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + as * MT;
      for (int kb = 0; kb < p.num_kb; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        if constexpr (L::kConv) mbar_wait(&conv_full[cs], cphase);
        tc_fence_after();
        if constexpr (kTmemA) {
          if (elect_one()) {
#pragma unroll
            for (int h = 0; h < L::kABoxes; ++h) {
              const uint64_t da = make_kmajor_sw128_desc(smem_u32(stage_a(stage) + h * L::kABox));
#pragma unroll
              for (int k = 0; k < 4; ++k)         // 16 k = 8 TMEM columns of the operand slot per MMA
                umma_f16_ts(tmem_d, tmem_base + kASlotCol + cs * L::kSlotCols + (h * 4 + k) * 8,
                            da + static_cast<uint64_t>(2 * k), idesc, (kb > 0 || h > 0 || k > 0) ? 1u : 0u);
            }
            umma_commit(&conv_empty[cs]);
            umma_commit(&empty_bar[stage]);
            if (kb == p.num_kb - 1) umma_commit(&tmem_full_bar[as]);
          }
        } else if (elect_one()) {
          const uint32_t w_addr = L::kConv ? smem_u32(conv_base + cs * L::kOpBytes) : smem_u32(stage_w(stage));
          const uint64_t dw = kMnMajor ? make_mnmajor_sw128_desc(w_addr, 8192) : make_kmajor_sw128_desc(w_addr);
          const uint64_t da = make_kmajor_sw128_desc(smem_u32(stage_a(stage)));
#pragma unroll
          for (int k = 0; k < 4; ++k)             // 16 k per MMA: 2 groups of 8 k rows (MN-major) / 32 B (K-major)
            umma_f16(tmem_d, dw + static_cast<uint64_t>(kMnMajor ? 128 * k : 2 * k), da + static_cast<uint64_t>(2 * k),
                     idesc, (kb > 0 || k > 0) ? 1u : 0u);
          if constexpr (L::kConv) umma_commit(&conv_empty[cs]);
          umma_commit(&empty_bar[stage]);
          if (kb == p.num_kb - 1) umma_commit(&tmem_full_bar[as]);
        }
        __syncwarp();
        if constexpr (L::kConv) { if (++cs == C) { cs = 0; cphase ^= 1; } }
        if (++stage == S) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp >= 8) {
    // ===================== converters =====================
    if constexpr (WKIND == GW_FP8T) {
      // fp8 [N, K] weights, operand in TMEM: thread (quadrant q = warp % 4, lane) owns weight row 32 q + lane, the
      // four warps of a quadrant split the stage's K.  raw: 128 rows x kBlockK bytes, swizzled with a span of one row
      // (SWIZZLE_128B / SWIZZLE_64B): consecutive rows reading the same logical chunk hit different bank groups.
      constexpr int kBK = L::kBlockK;
      constexpr int kChunks = kBK / 64;            // 16-byte chunks per thread and stage (2 or 1)
      const int q = warp & 3, j = (warp - 8) >> 2, row = q * 32 + lane;
      const int sw = kBK == 128 ? (row & 7) : ((row >> 1) & 3);
      int stage = 0, cs = 0;
      uint32_t phase = 0, cphase = 0;
      for (int idx = blockIdx.x; idx < num_tiles; idx += gridDim.x) {
        const float* blk = nullptr;                // block-scaled dense call: this row's scale column
        if (p.blk_scale) {
          const GgTile t = gg_tile<MT>(tile_start, row_start, p.E, p.num_nt, idx);
          blk = p.blk_scale + static_cast<int64_t>(min(t.n0 + row, p.N - 1) / p.group_n) * p.sb1;
        }
        for (int kb = 0; kb < p.num_kb; ++kb) {
          // the thread's 16 * kChunks codes lie inside one scale group (group_k % 64 == 0); fetched before the wait
          float bs = 1.0f;
          if (blk) bs = __ldg(blk + static_cast<int64_t>(min(kb * kBK + kChunks * j * 16, p.K - 1) / p.group_k) * p.sb0);
          mbar_wait(&full_bar[stage], phase);
          const uint32_t raw = smem_u32(stage_w(stage)) + row * kBK;
          uint32_t w[kChunks * 8];
#pragma unroll
          for (int c = 0; c < kChunks; ++c) {
            const uint4 v = lds128(raw + (((kChunks * j + c) ^ sw) << 4));
            uint32_t t8[8];
            if (blk) widen16<BF16, E5M2, true>(v, bs, t8);   // one rounding to the activation type, like gemm_w8a16.cuh
            else widen16<BF16, E5M2, false>(v, 1.0f, t8);
#pragma unroll
            for (int i = 0; i < 8; ++i) w[c * 8 + i] = t8[i];
          }
          mbar_wait(&conv_empty[cs], cphase ^ 1);
          tc_fence_after();
          const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + kASlotCol + cs * L::kSlotCols +
                                 j * (kChunks * 8);
          if constexpr (kChunks == 2) {
            tmem_st_32x32b_x16(taddr, w);
          } else {
            asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr),
                         "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7])
                         : "memory");
          }
          tmem_st_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&conv_full[cs]);
          if (++cs == C) { cs = 0; cphase ^= 1; }
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
    } else if constexpr (kTmemA) {
      // mxfp4, operand in TMEM: a warp can only write the 32 TMEM lanes of its quadrant (warp % 4), so thread
      // (quadrant q, lane l) owns weight row 32 q + l; the four warps of a quadrant split the stage's K.
      //   128-deep stages (MT <= 128): raw 128 rows x 64 B, SWIZZLE_64B (16-byte chunk ^= (row >> 1) & 3: eight
      //     consecutive rows reading the same logical chunk hit eight bank groups); a thread takes one 32-code group
      //   64-deep stages (MT = 256): raw 128 rows x 32 B, linear; a thread takes 16 codes (half a scale group)
      constexpr int kBK = L::kBlockK;
      constexpr int kWords = kBK / 8;              // 16-bit pairs (= TMEM columns) per thread and stage: 16 or 8
      constexpr int kGrp = kBK / 32;               // scale groups per stage: 4 or 2
      const int q = warp & 3, j = (warp - 8) >> 2, row = q * 32 + lane;
      const int grp = kBK == 128 ? j : (j >> 1);   // this thread's scale group within the stage
      int stage = 0, cs = 0;
      uint32_t phase = 0, cphase = 0;
      for (int idx = blockIdx.x; idx < num_tiles; idx += gridDim.x) {
        const GgTile t = gg_tile<MT>(tile_start, row_start, p.E, p.num_nt, idx);
        const int n = min(t.n0 + row, p.N - 1);
        const uint8_t* sfp = p.w_sf + (static_cast<int64_t>(t.e) * p.N + n) * (p.K / 32);
        // scale bytes: 16 per fetch = 16 / kGrp stages, this thread's group of each; fetched one batch ahead
        constexpr int kBatch = 16 / kGrp;          // stages per fetch: 4 or 8
        auto sf_load = [&](int kb) {
          if (p.sf_vec) return __ldg(reinterpret_cast<const uint4*>(sfp + kb * kGrp));
          uint32_t wv[4] = {0, 0, 0, 0};
          for (int i = 0; i < 16; ++i)
            if (kb * kGrp + i < p.K / 32) wv[i >> 2] |= static_cast<uint32_t>(__ldg(sfp + kb * kGrp + i)) << (8 * (i & 3));
          return make_uint4(wv[0], wv[1], wv[2], wv[3]);
        };
        uint4 sf_nxt = sf_load(0);
        uint32_t sf_lo = 0, sf_hi = 0;             // this thread's next scale bytes, the next one lowest
        for (int kb = 0; kb < p.num_kb; ++kb) {
          if ((kb & (kBatch - 1)) == 0) {
            if constexpr (kGrp == 4) {
              const uint32_t sel = 0x4000u | (static_cast<uint32_t>(4 + grp) << 4) | static_cast<uint32_t>(grp);
              const uint32_t lo = __byte_perm(sf_nxt.x, sf_nxt.y, sel), hi = __byte_perm(sf_nxt.z, sf_nxt.w, sel);
              sf_lo = __byte_perm(lo, hi, 0x5410u);
            } else {
              const uint32_t sel = grp ? 0x7531u : 0x6420u;
              sf_lo = __byte_perm(sf_nxt.x, sf_nxt.y, sel);
              sf_hi = __byte_perm(sf_nxt.z, sf_nxt.w, sel);
            }
            if (kb + kBatch < p.num_kb) sf_nxt = sf_load(kb + kBatch);
          }
          const uint32_t sb = sf_lo & 0xffu;
          sf_lo = __funnelshift_r(sf_lo, sf_hi, 8);
          sf_hi >>= 8;
          uint32_t h;
          if constexpr (BF16) {
            h = sb << 7;                          // bf16 bits of 2^(sb-127): the reference's (scale << 7).view(bf16)
          } else {
            const __half x = __float2half_rn(__uint_as_float(sb << 23));
            h = *reinterpret_cast<const uint16_t*>(&x);
          }
          const uint32_t sc2 = h | (h << 16);
          mbar_wait(&full_bar[stage], phase);
          uint32_t vv[kWords / 4];
          if constexpr (kBK == 128) {
            const uint4 v = lds128(smem_u32(stage_w(stage)) + row * 64 + ((j ^ ((row >> 1) & 3)) << 4));
            vv[0] = v.x; vv[1] = v.y; vv[2] = v.z; vv[3] = v.w;
          } else {
            asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(vv[0]), "=r"(vv[1])
                         : "r"(smem_u32(stage_w(stage)) + row * 32 + j * 8));
          }
          uint32_t w[kWords];
#pragma unroll
          for (int i = 0; i < kWords / 4; ++i) {
            uint32_t t4[4];
            e2m1x8_widen<BF16>(vv[i], sc2, t4);
            w[4 * i] = t4[0]; w[4 * i + 1] = t4[1]; w[4 * i + 2] = t4[2]; w[4 * i + 3] = t4[3];
          }
          mbar_wait(&conv_empty[cs], cphase ^ 1);
          tc_fence_after();
          const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + kASlotCol + cs * L::kSlotCols + j * kWords;
          if constexpr (kWords == 16) {
            tmem_st_32x32b_x16(taddr, w);
          } else {
            asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr),
                         "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7])
                         : "memory");
          }
          tmem_st_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&conv_full[cs]);
          if (++cs == C) { cs = 0; cphase ^= 1; }
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
    } else if constexpr (L::kConv) {
      const int ct = threadIdx.x - 256;          // 0..511
      int stage = 0, cs = 0;
      uint32_t phase = 0, cphase = 0;
      for (int idx = blockIdx.x; idx < num_tiles; idx += gridDim.x) {
        const GgTile t = gg_tile<MT>(tile_start, row_start, p.E, p.num_nt, idx);
        // mxfp4: this thread's scale bytes, 8 K slices (16 bytes of the row, every other one is this half's) per
        // fetch, fetched one group ahead so the load latency never sits in the per-stage chain
        const uint8_t* sfp = nullptr;
        uint4 sf_nxt = make_uint4(0, 0, 0, 0);
        uint32_t sf_lo = 0, sf_hi = 0;            // this thread's 8 scale bytes of the current group, next one lowest
        auto sf_load = [&](int kb) {
          if (p.sf_vec) return __ldg(reinterpret_cast<const uint4*>(sfp + kb * 2));
          uint32_t wv[4] = {0, 0, 0, 0};
          for (int i = 0; i < 16; ++i)
            if (kb * 2 + i < p.K / 32) wv[i >> 2] |= static_cast<uint32_t>(__ldg(sfp + kb * 2 + i)) << (8 * (i & 3));
          return make_uint4(wv[0], wv[1], wv[2], wv[3]);
        };
        if constexpr (WKIND == GW_MX4) {
          const int n = min(t.n0 + (ct >> 2), p.N - 1);
          sfp = p.w_sf + (static_cast<int64_t>(t.e) * p.N + n) * (p.K / 32);
          sf_nxt = sf_load(0);
        }
        for (int kb = 0; kb < p.num_kb; ++kb) {
          uint32_t sc2 = 0;
          if constexpr (WKIND == GW_MX4) {
            if ((kb & 7) == 0) {
              const uint32_t sel = (ct & 2) ? 0x7531u : 0x6420u;          // odd / even bytes: this half's groups
              sf_lo = __byte_perm(sf_nxt.x, sf_nxt.y, sel);
              sf_hi = __byte_perm(sf_nxt.z, sf_nxt.w, sel);
              if (kb + 8 < p.num_kb) sf_nxt = sf_load(kb + 8);
            }
            const uint32_t s = sf_lo & 0xffu;
            sf_lo = __funnelshift_r(sf_lo, sf_hi, 8);
            sf_hi >>= 8;
            uint32_t h;
            if constexpr (BF16) {
              h = s << 7;                        // bf16 bits of 2^(s-127): the reference's (scale << 7).view(bf16)
            } else {
              const __half x = __float2half_rn(__uint_as_float(s << 23));
              h = *reinterpret_cast<const uint16_t*>(&x);
            }
            sc2 = h | (h << 16);
          }
          mbar_wait(&full_bar[stage], phase);
          const uint32_t raw = smem_u32(stage_w(stage));
          if (p.dbg & 4) {
            mbar_wait(&conv_empty[cs], cphase ^ 1);
          } else if constexpr (WKIND == GW_FP8) {
            // raw: 64 k rows x 128 column bytes (SWIZZLE_128B); this thread: 16 columns of k row k0
            const int j = ct & 7, k0 = ct >> 3, sw = k0 & 7;
            const uint4 v0 = lds128(raw + k0 * 128 + ((j ^ sw) << 4));
            uint32_t w0[8];
            if (p.dbg & 1) {
#pragma unroll
              for (int i = 0; i < 8; ++i) w0[i] = v0.x;
            } else {
              widen16<BF16, E5M2, false>(v0, 1.0f, w0);
            }
            mbar_wait(&conv_empty[cs], cphase ^ 1);
            // a quarter warp (one k row, j = 0..7) writes both 64-column blocks; j >= 4 stores its odd chunk first
            // so that the eight 16-byte stores of one instruction fall on eight different bank groups
            const uint32_t dst = smem_u32(conv_base + cs * L::kOpBytes) + (j >> 2) * 8192 + k0 * 128;
            const uint32_t f = j >> 2;
            const uint32_t c0 = ((((j & 3) * 2 + f) ^ sw) << 4), c1 = ((((j & 3) * 2 + (f ^ 1)) ^ sw) << 4);
            if (f) {
              sts128(dst + c0, w0[4], w0[5], w0[6], w0[7]);
              sts128(dst + c1, w0[0], w0[1], w0[2], w0[3]);
            } else {
              sts128(dst + c0, w0[0], w0[1], w0[2], w0[3]);
              sts128(dst + c1, w0[4], w0[5], w0[6], w0[7]);
            }
          } else {
            // raw: 128 weight rows x 32 bytes (64 codes), linear; this thread: 16 codes (half a scale group)
            const int row = ct >> 2, qtr = ct & 3, sw = row & 7;
            uint32_t v0, v1;
            asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v0), "=r"(v1) : "r"(raw + row * 32 + qtr * 8));
            uint32_t w[2][4];
            e2m1x8_widen<BF16>(v0, sc2, w[0]);
            e2m1x8_widen<BF16>(v1, sc2, w[1]);
            mbar_wait(&conv_empty[cs], cphase ^ 1);
            const uint32_t dst = smem_u32(conv_base + cs * L::kOpBytes) + row * 128;
#pragma unroll
            for (int q = 0; q < 2; ++q)
              sts128(dst + (((qtr * 2 + q) ^ sw) << 4), w[q][0], w[q][1], w[q][2], w[q][3]);
          }
          if (!(p.dbg & 2)) fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(&conv_full[cs]);
          if (++cs == C) { cs = 0; cphase ^= 1; }
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    const int q = warp & 3;
    const int col = q * 32 + lane;               // D^T row = output column within the tile
    int iter = 0;
    for (int idx = blockIdx.x; idx < num_tiles; idx += gridDim.x, ++iter) {
      const GgTile t = gg_tile<MT>(tile_start, row_start, p.E, p.num_nt, idx);
      const int as = iter % L::kAccBufs;
      const uint32_t aphase = (iter / L::kAccBufs) & 1;
      const int n = t.n0 + col;
      const bool n_ok = n < p.N;
      float sc = 1.0f, bv = 0.0f;
      if constexpr (WKIND == GW_FP8) sc = __ldg(p.w_scale + t.e);
      if constexpr (WKIND == GW_FP8T)
        sc = p.blk_scale ? 1.0f
             : p.scale_per_channel ? (n_ok ? __ldg(p.w_scale + static_cast<int64_t>(t.e) * p.N + n) : 0.0f)
                                   : __ldg(p.w_scale + t.e);
      if (p.bias && n_ok) {
        const int64_t bi = static_cast<int64_t>(t.e) * p.N + n;
        bv = BF16 ? __bfloat162float(static_cast<const __nv_bfloat16*>(p.bias)[bi])
                  : __half2float(static_cast<const __half*>(p.bias)[bi]);
      }
      // lane pairs (2i, 2i+1) swap so that the even lane stores columns (n, n+1) of even token rows and the odd
      // lane the same two columns of odd token rows: 4-byte stores, 64 contiguous bytes per row per warp
      const int n2 = n & ~1;
      const bool pair_ok = n2 + 1 < p.N;
      uint16_t* dbase = static_cast<uint16_t*>(p.D) + n2;
      mbar_wait_backoff(&tmem_full_bar[as], aphase);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < MT / 32; ++c) {
        if (c * 32 >= t.rows) break;
        uint32_t r[32];
        tmem_ld_32x32b_x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * MT + c * 32, r);
        tmem_ld_wait();
        uint32_t hv[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const float v = fmaf(__uint_as_float(r[j]), sc, bv);
          if constexpr (BF16) {
            const __nv_bfloat16 x = __float2bfloat16_rn(v);
            hv[j] = *reinterpret_cast<const uint16_t*>(&x);
          } else {
            const __half x = __float2half_rn(v);
            hv[j] = *reinterpret_cast<const uint16_t*>(&x);
          }
        }
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          const uint32_t give = (lane & 1) ? hv[j] : hv[j + 1];
          const uint32_t got = __shfl_xor_sync(0xffffffffu, give, 1);
          const int tr = c * 32 + j + (lane & 1);            // token row within the tile
          const uint32_t packed = (lane & 1) ? (got | (hv[j + 1] << 16)) : (hv[j] | (got << 16));
          if (tr < t.rows && n_ok && t.row0 + tr < p.total_m) {      // (sum(rows_per_expert) may exceed the rows of A / D)
            uint16_t* d = dbase + static_cast<int64_t>(t.row0 + tr) * p.ldd;
            if (pair_ok) *reinterpret_cast<uint32_t*>(d) = packed;
            else d[0] = static_cast<uint16_t>(packed);
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&tmem_empty_bar[as]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<kCols>(tmem_base);
  }
}

}  // namespace vxk
