// Small-M fp8_gemm_w8a16 for sm_100a: M <= 128 tokens of bf16 / f16 activations against a large FP8
// weight -- the decode regime this op exists for, bound by streaming W[N, K] from HBM once at
// ONE byte per element (ref: fp8_gemm_w8a16.h:13-145).
//
// Same plan as gemm_skinny.cuh (operands swapped: UMMA M = 128 weight rows, UMMA N = MT tokens, the
// accumulator is D^T; K split over a thread-block cluster with a deterministic DSMEM reduction;
// programmatic dependent launch with the weight boxes requested before griddepcontrol.wait), plus
// the widening stage in between:
//
//   RAW ring   S stages of {128 W rows x 128 FP8 bytes (SWIZZLE_128B), two MT x 128 B activation boxes}:
//              this is what HBM has to deliver, and all of it is payload -- 16 KB of weight bytes per
//              stage, 64 KB in flight per SM at MT = 32 (the general kernel has 2 KB per stage at M = 1
//              because its stage is dominated by a 128-row activation box of padding)
//   CONV ring  2 slots of two K-major SWIZZLE_128B operand tiles (128 rows x 64 elements of bf16 / f16 each)
//   warps 8-23 converters (16 warps): raw stage -> conv slot (exact widening, widen16 of gemm_w8a16.cuh;
//              128x128 block scales folded in); a thread widens 32 codes of one weight row per stage in
//              registers BEFORE it waits for the slot, then stores, ONE generic->async proxy fence and
//              one arrive per warp per stage (the fence, not the arithmetic, is what a converter warp
//              waits on: six half-stage slots with a fence each measured slower than two full ones)
//   warp 1     tcgen05.mma.kind::f16 128 x MT x 16, eight per stage; its commits free the conv slot and
//              the raw stage (activation boxes)
#pragma once

#include "gemm_skinny.cuh"
#include "gemm_w8a16.cuh"
#include "gemm_grouped.cuh"   // umma_f16_ts, tmem_st_32x32b_x16

namespace vxk {

constexpr int kW16SkThreads = 768;
constexpr int kW16SkConvWarps = 16;   // warps 8-23
constexpr int kW16SkWarps = kW16SkThreads / 32;
constexpr int kW16SkBlockK = 128;   // K elements per raw stage

struct W16SkinnyParams {
  void* D;
  const float* b_scale;   // scalar (mode 0), per channel [N] (mode 1), block grid (mode 2)
  const void* bias;
  int M, N;
  int64_t ldd;
  int b_scale_mode, out_dtype, bias_dtype;
  int64_t sb0, sb1;
  int group_k, group_n;   // block mode: group_k % 128 == 0
  int num_k_blocks;       // 128-element K blocks in total
  int kb_per_split;
  int splits;
  int prefetch;           // raw stages requested into L2 ahead of the ring (cp.async.bulk.prefetch.tensor)
  int m_tiles;            // token tiles of MT side by side in gridDim.z (0 / 1: one tile)
};

// TMA prefetch of one box into L2: no shared memory, no barrier -- it only starts the HBM read early
__device__ __forceinline__ void tma_prefetch_2d(const CUtensorMap* m, int c0, int c1) {
  asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];" ::"l"(reinterpret_cast<uint64_t>(m)),
               "r"(c0), "r"(c1)
               : "memory");
}

// TA = the widened operand lives in TENSOR MEMORY instead of shared memory: a tcgen05.mma that reads its 128-row
// operand from shared memory takes ~130-150 clk per 16-deep step whatever the token count (measured on the grouped
// GEMM, profiles/r02_grouped_gemm.txt) = 16 weights / clk / SM = 4.5 TB/s of FP8 weight stream at best; from TMEM it
// is ~80 clk.  The converters write their 32 lanes with tcgen05.st (no proxy fence, no shared-memory round trip),
// four 64-column operand slots sit next to the accumulator, and the shared memory the slots used goes to the ring.
template <int MT, bool TA = false>
struct W16SkinnySmem {
  static constexpr int kRawW = 128 * kW16SkBlockK;          // 16 KB
  static constexpr int kABox = MT * 128;                    // one 64-element activation slice
  static constexpr int kRawStage = kRawW + 2 * kABox;
  static constexpr int kConvSlot = 2 * 128 * 128;           // two operand tiles (128 rows x 64 elements each), 32 KB
  static constexpr int kConvSlots = TA ? 4 : 2;
  static constexpr int kConvSmem = TA ? 0 : kConvSlots * kConvSlot;
  static constexpr int kStagesRaw = (196 * 1024 - kConvSmem) / kRawStage;
  static constexpr int kStages = kStagesRaw > (TA ? 10 : 8) ? (TA ? 10 : 8) : kStagesRaw;
  static constexpr int kBarrierBytes = (2 * kStages + 2 * kConvSlots + 1) * 8 + 16;
  static constexpr int kTotalBytes = kStages * kRawStage + kConvSmem + kBarrierBytes + 1024;
  static_assert(kRawStage % 1024 == 0, "SWIZZLE_128B tiles need 1024-byte alignment");
  static_assert(kStages >= 2, "ring too shallow");
  static_assert(kStages * kRawStage + kConvSmem >= MT * 128 * 4, "the partial tile is parked in the rings");
};

template <int MT, bool BF16, bool E5M2, bool SCALED, bool TA = false>
__global__ void __launch_bounds__(kW16SkThreads, 1)
gemm_w8a16_skinny_kernel(const W16SkinnyParams p_in, const __grid_constant__ CUtensorMap map_w,
                         const __grid_constant__ CUtensorMap map_a) {
  // token tiles (gridDim.z > 1, 128 < M <= 512): tile z owns tokens [z * MT, z * MT + MT), as in gemm_skinny_kernel
  W16SkinnyParams p = p_in;
  const int m0 = static_cast<int>(blockIdx.z) * MT;
  p.D = static_cast<uint16_t*>(p_in.D) + static_cast<int64_t>(m0) * p_in.ldd;
  p.M = min(MT, p_in.M - m0);
  using L = W16SkinnySmem<MT, TA>;
  constexpr uint32_t kASlotCol = 256;        // TA: operand slots in TMEM columns [256, 512)
  constexpr int S = L::kStages;
  constexpr int C = L::kConvSlots;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* conv_base = smem + S * L::kRawStage;
  uint64_t* bars = reinterpret_cast<uint64_t*>(conv_base + L::kConvSmem);
  uint64_t* full_bar = bars;                 // raw stage landed (TMA)
  uint64_t* empty_bar = bars + S;            // raw stage free (MMA commit)
  uint64_t* conv_full = bars + 2 * S;        // conv slot written (converter warps)
  uint64_t* conv_empty = bars + 2 * S + C;   // conv slot free (MMA commit)
  uint64_t* acc_bar = bars + 2 * S + 2 * C;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 2 * S + 2 * C + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * 128;
  const int rank = blockIdx.y;
  const int kb0 = rank * p.kb_per_split;
  const int kb1 = min(p.num_k_blocks, kb0 + p.kb_per_split);
  const int nk = kb1 - kb0;                  // raw stages this CTA consumes (>= 1)

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_w);
    tma_prefetch_desc(&map_a);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < S; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < C; ++i) {
      mbar_init(&conv_full[i], kW16SkConvWarps);
      mbar_init(&conv_empty[i], 1);
    }
    mbar_init(acc_bar, 1);
    fence_barrier_init();
  }
  constexpr uint32_t kCols = TA ? 512 : (MT < 32 ? 32 : MT);
  if (warp == 2) tmem_alloc<kCols>(tmem_ptr_smem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    // ===================== TMA producer =====================
    const int pre = min(nk, S);
    auto load_w = [&](int it, int stg) {
      uint8_t* st = smem + stg * L::kRawStage;
      mbar_arrive_expect_tx(&full_bar[stg], L::kRawStage);
      tma_load_2d(st, &map_w, &full_bar[stg], (kb0 + it) * kW16SkBlockK, n0);
    };
    auto load_a = [&](int it, int stg) {
      uint8_t* st = smem + stg * L::kRawStage + L::kRawW;
      tma_load_2d(st, &map_a, &full_bar[stg], (kb0 + it) * 256, m0);                // byte coordinates
      tma_load_2d(st + L::kABox, &map_a, &full_bar[stg], (kb0 + it) * 256 + 128, m0);
    };
    // Optional (p.prefetch > 0, off by default -- measured slower at every depth): request the weight stream
    // into L2 `prefetch` stages ahead of the ring, so that the ring's own loads hit L2.
    const int pf = p.prefetch;
    if (elect_one()) {
      for (int it = 0; it < pre; ++it) load_w(it, it);      // the weight does not depend on the previous kernel
      for (int it = pre; it < min(nk, pre + pf); ++it) tma_prefetch_2d(&map_w, (kb0 + it) * kW16SkBlockK, n0);
    }
    __syncwarp();
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (elect_one()) {
      for (int it = 0; it < pre; ++it) load_a(it, it);
    }
    __syncwarp();
    int stage = pre == S ? 0 : pre;
    uint32_t phase = pre == S ? 1 : 0;
    for (int it = pre; it < nk; ++it) {
      mbar_wait(&empty_bar[stage], phase ^ 1);
      if (elect_one()) {
        load_w(it, stage);
        load_a(it, stage);
        if (it + pf < nk) tma_prefetch_2d(&map_w, (kb0 + it + pf) * kW16SkBlockK, n0);
      }
      __syncwarp();
      if (++stage == S) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer: D^T[128 x MT] += W16[128 x 16] * A[MT x 16]^T =====================
    const uint32_t idesc = make_idesc_f16(BF16 ? 1 : 0, 128, MT);
    int stage = 0, cs = 0;
    uint32_t phase = 0, cphase = 0;
    for (int it = 0; it < nk; ++it) {
      mbar_wait(&full_bar[stage], phase);      // activation boxes
      mbar_wait(&conv_full[cs], cphase);       // widened weight tiles
      tc_fence_after();
      if constexpr (TA) {
        if (elect_one()) {
          const uint32_t a_addr = smem_u32(smem + stage * L::kRawStage + L::kRawW);
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const uint64_t da = make_kmajor_sw128_desc(a_addr + h * L::kABox);
#pragma unroll
            for (int k = 0; k < 4; ++k)          // 16 k = 8 TMEM columns of the operand slot per MMA
              umma_f16_ts(tmem_base, tmem_base + kASlotCol + cs * 64 + (h * 4 + k) * 8, da + static_cast<uint64_t>(2 * k),
                          idesc, (it > 0 || h > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&conv_empty[cs]);
          umma_commit(&empty_bar[stage]);
          if (it == nk - 1) umma_commit(acc_bar);
        }
      } else if (elect_one()) {
        const uint32_t w_addr = smem_u32(conv_base + cs * L::kConvSlot);
        const uint32_t a_addr = smem_u32(smem + stage * L::kRawStage + L::kRawW);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const uint64_t dw = make_kmajor_sw128_desc(w_addr + h * 128 * 128);
          const uint64_t da = make_kmajor_sw128_desc(a_addr + h * L::kABox);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_f16(tmem_base, dw + static_cast<uint64_t>(2 * k), da + static_cast<uint64_t>(2 * k), idesc,
                     (it > 0 || h > 0 || k > 0) ? 1u : 0u);
        }
        umma_commit(&conv_empty[cs]);
        umma_commit(&empty_bar[stage]);
        if (it == nk - 1) umma_commit(acc_bar);
      }
      __syncwarp();
      if (++cs == C) { cs = 0; cphase ^= 1; }
      if (++stage == S) { stage = 0; phase ^= 1; }
    }
  } else if (warp >= 8) {
    // ===================== converters =====================
    const int ct = threadIdx.x - 256;          // 0..511: weight row ct / 4, 16-byte chunks jq and 4 + jq of its 128 bytes
    constexpr bool scaled = SCALED;
    // TA: a warp writes the 32 TMEM lanes of its quadrant (warp % 4): row = 32 q + lane, the four warps of a quadrant
    // take the four 32-byte quarters of the 128-byte raw row (chunks 2 jq, 2 jq + 1)
    const int row = TA ? (warp & 3) * 32 + lane : ct >> 2, jq = TA ? (warp - 8) >> 2 : ct & 3, sw = row & 7;
    int stage = 0, cs = 0;
    uint32_t phase = 0, cphase = 0;
    for (int it = 0; it < nk; ++it) {
      mbar_wait(&full_bar[stage], phase);
      const uint32_t raw = smem_u32(smem + stage * L::kRawStage) + row * 128;   // SWIZZLE_128B as TMA wrote it
      uint4 v[2];
#pragma unroll
      for (int h = 0; h < 2; ++h) v[h] = lds128(raw + (((TA ? 2 * jq + h : h * 4 + jq) ^ sw) << 4));
      float sc = 1.0f;
      if constexpr (scaled) {
        const int n = min(n0 + row, p.N - 1);
        sc = __ldg(p.b_scale + static_cast<int64_t>(((kb0 + it) * kW16SkBlockK) / p.group_k) * p.sb0 +
                   static_cast<int64_t>(n / p.group_n) * p.sb1);
      }
      uint32_t w[2][8];
#pragma unroll
      for (int h = 0; h < 2; ++h) widen16<BF16, E5M2, SCALED>(v[h], sc, w[h]);
      mbar_wait(&conv_empty[cs], cphase ^ 1);
      if constexpr (TA) {
        tc_fence_after();
        uint32_t wv[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) wv[i] = w[i >> 3][i & 7];
        tmem_st_32x32b_x16(tmem_base + (static_cast<uint32_t>((warp & 3) * 32) << 16) + kASlotCol + cs * 64 + jq * 16, wv);
        tmem_st_wait();
        tc_fence_before();
      } else {
        const uint32_t dst = smem_u32(conv_base + cs * L::kConvSlot) + row * 128;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          sts128(dst + h * (128 * 128) + (((2 * jq) ^ sw) << 4), w[h][0], w[h][1], w[h][2], w[h][3]);
          sts128(dst + h * (128 * 128) + (((2 * jq + 1) ^ sw) << 4), w[h][4], w[h][5], w[h][6], w[h][7]);
        }
        fence_proxy_async_smem();
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&conv_full[cs]);
      if (++cs == C) { cs = 0; cphase ^= 1; }
      if (++stage == S) { stage = 0; phase ^= 1; }
    }
  }

  // ===================== epilogue =====================
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const bool bf = p.out_dtype == VXK_BF16;
  float* park = reinterpret_cast<float*>(smem);
  if (warp >= 4 && warp < 8) {
    const int q = warp & 3;
    const int row = q * 32 + lane;
    mbar_wait_backoff(acc_bar, 0);
    tc_fence_after();
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#pragma unroll 1
    for (int c = 0; c < MT / 32; ++c) {
      if (c * 32 >= p.M) break;
      uint32_t r[32];
      tmem_ld_32x32b_x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c * 32, r);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) park[(c * 32 + j) * 128 + row] = __uint_as_float(r[j]);
    }
    tc_fence_before();
  }

  {
    cluster_sync_all();          // every CTA's partial tile is parked
    const uint32_t park_addr = smem_u32(park);
    const int nc = n0 + lane * 4;
    float swv[4], bv[4];
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const bool ok = nc + t < p.N;
      swv[t] = p.b_scale_mode == 1 ? (ok ? __ldg(p.b_scale + nc + t) : 0.0f)
               : p.b_scale_mode == 0 ? __ldg(p.b_scale) : 1.0f;
      bv[t] = 0.0f;
      if (p.bias && ok)
        bv[t] = (p.bias_dtype == VXK_F32)   ? __ldg(static_cast<const float*>(p.bias) + nc + t)
                : (p.bias_dtype == VXK_F16) ? __half2float(static_cast<const __half*>(p.bias)[nc + t])
                                            : __bfloat162float(static_cast<const __nv_bfloat16*>(p.bias)[nc + t]);
    }
    uint32_t peer[8];
#pragma unroll
    for (int s = 0; s < 8; ++s) peer[s] = mapa_u32(park_addr + lane * 16, s < p.splits ? s : 0);
    for (int m = rank + p.splits * warp; m < p.M; m += p.splits * kW16SkWarps) {
      float4 part[8];
#pragma unroll
      for (int s = 0; s < 8; ++s)
        if (s < p.splits)
          asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];"
                       : "=f"(part[s].x), "=f"(part[s].y), "=f"(part[s].z), "=f"(part[s].w)
                       : "r"(peer[s] + m * 512));
      float acc[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
      for (int s = 0; s < 8; ++s)
        if (s < p.splits) {
          acc[0] += part[s].x; acc[1] += part[s].y; acc[2] += part[s].z; acc[3] += part[s].w;
        }
      uint16_t h[4];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        float v = acc[t] * swv[t];
        if (p.bias) v += bv[t];
        if (bf) {
          const __nv_bfloat16 x = __float2bfloat16_rn(v);
          h[t] = *reinterpret_cast<const uint16_t*>(&x);
        } else {
          const __half x = __float2half_rn(v);
          h[t] = *reinterpret_cast<const uint16_t*>(&x);
        }
      }
      uint16_t* d = static_cast<uint16_t*>(p.D) + static_cast<int64_t>(m) * p.ldd + nc;
      if (nc + 3 < p.N && (reinterpret_cast<uintptr_t>(d) & 7) == 0) {
        *reinterpret_cast<uint2*>(d) = make_uint2(h[0] | (uint32_t(h[1]) << 16), h[2] | (uint32_t(h[3]) << 16));
      } else {
#pragma unroll
        for (int t = 0; t < 4; ++t)
          if (nc + t < p.N) d[t] = h[t];
      }
    }
    cluster_sync_all();          // nobody leaves while its shared memory is still being read
  }
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<kCols>(tmem_base);
  }
}

}  // namespace vxk
