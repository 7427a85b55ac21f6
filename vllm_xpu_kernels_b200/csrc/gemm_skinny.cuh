// Small-M ("skinny") FP8 GEMM for sm_100a: M <= 128 tokens against a large weight -- the decode
// regime, bound by streaming W[N, K] from HBM once (bytes = K*N + ...).
//
//   D[m, n] = sA(m) * sW(n) * sum_k A[m, k] * W[n, k]  (+ bias[n])       ref: fp8_gemm_w8a8.h:12-184
//
// The general kernels (gemm_tc.cuh / gemm_tc2.cuh) put the tokens on the 128-row UMMA M side: at
// M = 1..64 a 128-row activation box is almost all padding, the weight box shrinks to 32 rows x
// 128 B so that every SM gets a tile, and with 32 KB of weight bytes in flight per SM the stream
// is latency-bound at 0.9 TB/s (benchmarks/bench_gemm_sweep.py, round 1d).  This kernel swaps the
// operands and splits K:
//   * UMMA M = 128 rows of W, UMMA N = MT tokens (32 / 64 / 128): the accumulator is D^T.
//     Every shared-memory byte of the ring is weight data that HBM has to deliver anyway.
//   * a ring stage holds KS = 2 consecutive 128-byte K slices of the 128 W rows (two TMA boxes
//     on the same DRAM pages back to back) -- 32 KB per stage, up to 192 KB in flight per SM.
//   * N / 128 row tiles alone cannot fill 148 SMs (N = 4096 -> 32 tiles): K is split over a
//     thread-block CLUSTER (1 x splits), every CTA accumulates its K range in TMEM, parks the
//     fp32 partial tile in its own (drained) ring memory, and after one cluster barrier CTA r
//     reduces token columns r, r + splits, ... over distributed shared memory in a fixed order
//     (deterministic; no atomics, no workspace, no second kernel).
//   * a warp finishes one token row at a time: 128 columns as 32 float4 per CTA of the cluster,
//     scales / bias / conversion, one 256-byte contiguous store.
#pragma once

#include "gemm_tc2.cuh"   // cluster helpers

namespace vxk {

constexpr int kSkinnyThreads = 256;

struct SkinnyParams {
  void* D;                // [M, N] bf16 / f16, row stride ldd
  const float* a_scale;   // per token [M] (mode 1) or scalar (mode 0)
  const float* b_scale;   // per channel [N] (mode 1) or scalar (mode 0)
  const void* bias;       // [N] or null
  int M, N;
  int64_t ldd;
  int a_scale_mode, b_scale_mode, out_dtype, bias_dtype, fp8_fmt;
  int num_k_blocks;       // 128-byte K slices in total
  int kb_per_split;       // slices per cluster rank (a multiple of the slices per stage)
  int splits;             // cluster size along y
  // BLK (fp32 block scales, DeepSeek-V3 1 x 128 / 128 x 128; fp8_gemm_w8a8.h:28-41,86-134):
  const float* blk_sa;    // sA[m, k / group_k] at blk_sa[m * sa0 + kg * sa1]
  const float* blk_sb;    // sW[k / group_k, n / group_n] at blk_sb[kg * sb0 + ng * sb1]
  int64_t sa0, sa1, sb0, sb1;
  int group_k, group_n;   // group_k % 128 == 0
};

// slices per ring stage: two for MT <= 128, one for the 256-token tile (its A slice alone is 32 KB)
template <int MT>
struct SkinnySmem {
  static constexpr int kKS = MT <= 128 ? 2 : 1;
  static constexpr int kWSlice = 128 * kBlockKBytes;                 // 16 KB
  static constexpr int kASlice = MT * kBlockKBytes;
  static constexpr int kStageBytes = kKS * (kWSlice + kASlice);
  static constexpr int kStages = (196 * 1024) / kStageBytes > 8 ? 8 : (196 * 1024) / kStageBytes;
  static constexpr int kBarrierBytes = (2 * kStages + 5) * 8 + 16;
  static constexpr int kTotalBytes = kStages * kStageBytes + kBarrierBytes + 1024;
  static_assert(kStageBytes % 1024 == 0, "SWIZZLE_128B tiles need 1024-byte alignment");
  static_assert(kStages * kStageBytes >= MT * 128 * 4, "the partial tile is parked in the ring");
};

__device__ __forceinline__ float ld_dsmem_f32(uint32_t cluster_addr) {
  float v;
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(cluster_addr));
  return v;
}
__device__ __forceinline__ uint32_t mapa_u32(uint32_t cta_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(cta_addr), "r"(rank));
  return r;
}

// BLK: fp32 block scales (BASELINE config 5 in the decode regime).  Non-power-of-two scales have no tcgen05 mode, so
// every 128-deep K slice is a fresh accumulation into one of two TMEM buffers; the four epilogue warps pull each
// partial product into registers (thread = output column, MT token columns), fold it into fp32 accumulators with
// sA[m, kb] * sW[kb, n / group_n] (lane j fetches token j's scale one slice ahead, the warp shuffles it out) and hand the
// buffer back, while the MMA warp is on the next slice.  The register accumulators are then parked and reduced over
// the cluster exactly like the plain kernel's TMEM tile.
template <int MT, bool BLK = false>
__global__ void __launch_bounds__(kSkinnyThreads, 1)
gemm_skinny_kernel(const SkinnyParams p_in, const __grid_constant__ CUtensorMap map_w,
                   const __grid_constant__ CUtensorMap map_a) {
  static_assert(!BLK || MT <= 128, "block-scaled small-M kernel: two MT-column TMEM buffers, MT register accumulators");
  // Token tiles (gridDim.z > 1, 256 < M <= 512): tile z owns tokens [z * MT, z * MT + MT); everything below sees its
  // own tile as the whole problem (output, per-token scales and the block-scale grid shifted, M clipped), and the
  // tiles of one weight tile run side by side, so the weight comes from HBM once and from L2 for the other tiles.
  SkinnyParams p = p_in;
  const int m0 = static_cast<int>(blockIdx.z) * MT;
  if (m0 > 0) {
    p.D = static_cast<uint16_t*>(p_in.D) + static_cast<int64_t>(m0) * p_in.ldd;
    if (p_in.a_scale_mode == 1) p.a_scale = p_in.a_scale + m0;
    if (BLK) p.blk_sa = p_in.blk_sa + static_cast<int64_t>(m0) * p_in.sa0;
  }
  p.M = min(MT, p_in.M - m0);
  using L = SkinnySmem<MT>;
  constexpr int S = L::kStages;
  constexpr int kSkinnyKS = L::kKS;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S * L::kStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + S;
  uint64_t* acc_bar = bars + 2 * S;
  uint64_t* pfull = bars + 2 * S + 1;       // BLK: partial product of a K slice landed in TMEM buffer i
  uint64_t* pempty = bars + 2 * S + 3;      // BLK: buffer i drained into registers
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 2 * S + 5);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * 128;
  const int rank = blockIdx.y;                       // cluster = (1, splits, 1): rank within the cluster
  const int kb0 = rank * p.kb_per_split;
  const int kb1 = min(p.num_k_blocks, kb0 + p.kb_per_split);
  const int num_stages_k = (kb1 - kb0 + kSkinnyKS - 1) / kSkinnyKS;   // ring stages this CTA consumes (>= 1)

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_w);
    tma_prefetch_desc(&map_a);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < S; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_init(acc_bar, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&pfull[i], 1);
      mbar_init(&pempty[i], 128);
    }
    fence_barrier_init();
  }
  constexpr uint32_t kCols = BLK ? (2 * MT < 32 ? 32 : 2 * MT) : (MT < 32 ? 32 : MT);   // TMEM allocation: power of two >= 32
  if (warp == 2) tmem_alloc<kCols>(tmem_ptr_smem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    // ===================== TMA producer =====================
    // Programmatic dependent launch: the weight is constant, so the W boxes of the first ring pass are
    // requested BEFORE griddepcontrol.wait -- while the previous kernel in the stream (the one that
    // produces the activations) is still draining; only the activation boxes wait for it.
    int stage = 0;
    uint32_t phase = 0;
    const int pre = min(num_stages_k, S);
    auto load_w = [&](int it, int stg) {
      const int kb = kb0 + it * kSkinnyKS;
      const int ns = min(kSkinnyKS, kb1 - kb);
      uint8_t* st = smem + stg * L::kStageBytes;
      mbar_arrive_expect_tx(&full_bar[stg], ns * (L::kWSlice + L::kASlice));
      for (int s = 0; s < ns; ++s) tma_load_2d(st + s * L::kWSlice, &map_w, &full_bar[stg], (kb + s) * kBlockKBytes, n0);
    };
    auto load_a = [&](int it, int stg) {
      const int kb = kb0 + it * kSkinnyKS;
      const int ns = min(kSkinnyKS, kb1 - kb);
      uint8_t* st = smem + stg * L::kStageBytes;
      for (int s = 0; s < ns; ++s)
        tma_load_2d(st + kSkinnyKS * L::kWSlice + s * L::kASlice, &map_a, &full_bar[stg], (kb + s) * kBlockKBytes, m0);
    };
    if (elect_one()) {
      for (int it = 0; it < pre; ++it) load_w(it, it);
    }
    __syncwarp();
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (elect_one()) {
      for (int it = 0; it < pre; ++it) load_a(it, it);
    }
    __syncwarp();
    stage = pre == S ? 0 : pre;
    phase = pre == S ? 1 : 0;
    for (int it = pre; it < num_stages_k; ++it) {
      mbar_wait(&empty_bar[stage], phase ^ 1);
      if (elect_one()) {
        load_w(it, stage);
        load_a(it, stage);
      }
      __syncwarp();
      if (++stage == S) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer: D^T[128 x MT] += W_tile[128 x 32B] * A_tile[MT x 32B]^T =====================
    const uint32_t idesc = make_idesc_f8(p.fp8_fmt, 128, MT);
    int stage = 0;
    uint32_t phase = 0;
    for (int it = 0; it < num_stages_k; ++it) {
      const int ns = min(kSkinnyKS, kb1 - (kb0 + it * kSkinnyKS));
      mbar_wait(&full_bar[stage], phase);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t st = smem_u32(smem + stage * L::kStageBytes);
        for (int s = 0; s < ns; ++s) {
          const uint64_t dw = make_kmajor_sw128_desc(st + s * L::kWSlice);
          const uint64_t da = make_kmajor_sw128_desc(st + kSkinnyKS * L::kWSlice + s * L::kASlice);
          if constexpr (BLK) {
            const int sl = it * kSkinnyKS + s, buf = sl & 1;         // slice counter of this CTA
            mbar_wait(&pempty[buf], ((sl >> 1) & 1) ^ 1);
            tc_fence_after();
#pragma unroll
            for (int k = 0; k < kBlockKBytes / kUmmaKBytes; ++k)
              umma_f8f6f4(tmem_base + buf * MT, dw + static_cast<uint64_t>(k * (kUmmaKBytes >> 4)),
                          da + static_cast<uint64_t>(k * (kUmmaKBytes >> 4)), idesc, k > 0 ? 1u : 0u);
            umma_commit(&pfull[buf]);
          } else {
#pragma unroll
            for (int k = 0; k < kBlockKBytes / kUmmaKBytes; ++k)
              umma_f8f6f4(tmem_base, dw + static_cast<uint64_t>(k * (kUmmaKBytes >> 4)),
                          da + static_cast<uint64_t>(k * (kUmmaKBytes >> 4)), idesc, (it > 0 || s > 0 || k > 0) ? 1u : 0u);
          }
        }
        umma_commit(&empty_bar[stage]);
        if (!BLK && it == num_stages_k - 1) umma_commit(acc_bar);
      }
      __syncwarp();
      if (++stage == S) { stage = 0; phase ^= 1; }
    }
  }

  // ===================== epilogue =====================
  // warps 4-7: thread (q, lane) owns accumulator lane row = q * 32 + lane (output column n0 + row);
  // the fp32 tile is parked in the drained ring as [token][128 columns]
  asm volatile("griddepcontrol.wait;" ::: "memory");   // scales / bias / D may depend on the previous kernel
  const bool bf = p.out_dtype == VXK_BF16;
  const float sa_all = (BLK || p.a_scale_mode == 1) ? 1.0f : __ldg(p.a_scale);
  float* park = reinterpret_cast<float*>(smem);
  if (BLK && warp >= 4) {
    if constexpr (BLK) {
      const int q = warp & 3;
      const int row = q * 32 + lane;
      const int ng = min(n0 + row, p.N - 1) / p.group_n;
      float acc[MT];
#pragma unroll
      for (int j = 0; j < MT; ++j) acc[j] = 0.0f;
      const int nsl = kb1 - kb0;
      auto load_scales = [&](int sl, float& sw, float (&sa)[MT / 32]) {
        const int kg = ((kb0 + sl) * kBlockKBytes) / p.group_k;
        sw = __ldg(p.blk_sb + static_cast<int64_t>(kg) * p.sb0 + static_cast<int64_t>(ng) * p.sb1);
#pragma unroll
        for (int c = 0; c < MT / 32; ++c) {
          const int m = c * 32 + lane;
          sa[c] = m < p.M ? __ldg(p.blk_sa + static_cast<int64_t>(m) * p.sa0 + static_cast<int64_t>(kg) * p.sa1) : 0.0f;
        }
      };
      float sw_n, sa_n[MT / 32];
      load_scales(0, sw_n, sa_n);
      for (int sl = 0; sl < nsl; ++sl) {
        const int buf = sl & 1;
        const float sw = sw_n;
        float sa[MT / 32];
#pragma unroll
        for (int c = 0; c < MT / 32; ++c) sa[c] = sa_n[c] * sw;          // sA[m, kb] * sW[kb, n / group_n], per lane's token
        if (sl + 1 < nsl) load_scales(sl + 1, sw_n, sa_n);
        mbar_wait(&pfull[buf], (sl >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < MT / 32; ++c) {
          if (c * 32 < p.M) {
            uint32_t r[32];
            tmem_ld_32x32b_x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * MT + c * 32, r);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j)
              acc[c * 32 + j] = fmaf(__uint_as_float(r[j]), __shfl_sync(0xffffffffu, sa[c], j), acc[c * 32 + j]);
          }
        }
        tc_fence_before();
        mbar_arrive(&pempty[buf]);
      }
      asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
      // every slice of this CTA has been consumed: the ring is drained, park the register tile in it
#pragma unroll
      for (int j = 0; j < MT; ++j)
        if (j < p.M) park[j * 128 + row] = acc[j];
    }
  } else if (warp >= 4) {
    const int q = warp & 3;
    const int row = q * 32 + lane;
    mbar_wait(acc_bar, 0);
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
    // (a launch without the cluster attribute is a cluster of one: the barrier degenerates to a CTA barrier)
    cluster_sync_all();          // every CTA's partial tile is parked (release / acquire at cluster scope)
    // CTA `rank` finishes token rows rank, rank + splits, ...; all 8 warps take rows round-robin.  A warp
    // reads one parked row (128 columns) as 32 float4 from every CTA of the cluster -- all loads of a
    // row are issued before the first add (DSMEM latency is paid once per row, not once per load) --
    // and sums them in rank order (deterministic).
    const uint32_t park_addr = smem_u32(park);
    const int nc = n0 + lane * 4;                       // this lane's 4 output columns
    float swv[4], bv[4];
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const bool ok = nc + t < p.N;
      swv[t] = BLK ? 1.0f : ((p.b_scale_mode == 1) ? (ok ? __ldg(p.b_scale + nc + t) : 0.0f) : __ldg(p.b_scale));
      bv[t] = 0.0f;
      if (p.bias && ok)
        bv[t] = (p.bias_dtype == VXK_F32)   ? __ldg(static_cast<const float*>(p.bias) + nc + t)
                : (p.bias_dtype == VXK_F16) ? __half2float(static_cast<const __half*>(p.bias)[nc + t])
                                            : __bfloat162float(static_cast<const __nv_bfloat16*>(p.bias)[nc + t]);
    }
    uint32_t peer[8];
#pragma unroll
    for (int s = 0; s < 8; ++s) peer[s] = mapa_u32(park_addr + lane * 16, s < p.splits ? s : 0);
    for (int m = rank + p.splits * warp; m < p.M; m += p.splits * 8) {
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
      const float sa = BLK ? 1.0f : ((p.a_scale_mode == 1) ? __ldg(p.a_scale + m) : sa_all);
      uint16_t h[4];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        float v = (acc[t] * sa) * swv[t];
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
