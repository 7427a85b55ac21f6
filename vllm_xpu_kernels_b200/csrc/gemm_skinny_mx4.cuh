// Small-M ("skinny") MXFP4 GEMM for sm_100a: M <= 128 tokens against a large E2M1 weight with one E8M0 scale
// per 32 elements -- the decode regime of BASELINE config 4, bound by streaming W[N, K/2] (+ N * K / 32 scale
// bytes) from HBM once.
//
//   D[m, n] = sum_k (sA[m, k/32] * A[m, k]) * (sW[n, k/32] * W[n, k])  (+ bias[n])     ref: fp4_gemm_w4a4.h:12-167
//
// Same structure as gemm_skinny.cuh (operands swapped: UMMA M = 128 weight rows, UMMA N = MT tokens; two
// 128-byte K slices = 512 elements per ring stage; K split over a thread-block cluster, partial tiles parked
// in the drained ring and reduced over distributed shared memory in rank order), with the block scales on top:
//   * a stage covers 16 scale groups, i.e. exactly one 16-byte TMA box row per weight row / token row of the
//     caller's NATURAL [rows, K/32] scale tensors -- no workspace, no packing launch;
//   * warp 3 re-tiles the two raw boxes into the 512-byte chunks tcgen05.cp.32x128b.warpx4 takes (word c of
//     lane l = the four scale bytes of row c * 32 + l), four chunks per operand and stage;
//   * the MMA thread copies them into TMEM (tcgen05.cp and tcgen05.mma execute in issue order, so the scale
//     columns of stage i are only overwritten after the MMAs of stage i - 1 have read them) and issues
//     kind::mxf4.block_scale.scale_vec::2X, 128 x MT x 64 elements, eight per stage.
// The general kernel (gemm_tc.cuh) at M <= 128 measured 12-30 % of the HBM roofline on the Llama-70B shapes
// (profiles/r02_mxfp4_smallm.txt): 128-token activation boxes that are mostly padding and 32 KB in flight.
#pragma once

#include "gemm_skinny.cuh"

namespace vxk {

struct SkinnyMx4Params {
  void* D;                // [M, N] bf16 / f16, row stride ldd
  const void* bias;       // [N] or null
  int M, N;
  int64_t ldd;
  int out_dtype, bias_dtype;
  int num_k_blocks;       // 128-byte (256-element) K slices in total
  int kb_per_split;       // slices per cluster rank (a multiple of 2)
  int splits;             // cluster size along y
};

template <int MT>
struct SkinnyMx4Smem {
  static constexpr int kKS = 2;                                      // K slices per stage (= 16 scale groups)
  static constexpr int kWSlice = 128 * kBlockKBytes;                 // 16 KB
  static constexpr int kASlice = MT * kBlockKBytes;
  static constexpr int kRawSfw = 128 * 16;                           // one 16-byte box row per weight row
  static constexpr int kRawSfa = MT * 16;
  static constexpr int kSfT = 4 * 512;                               // four tcgen05.cp chunks per operand
  static constexpr int kOffA = kKS * kWSlice;
  static constexpr int kOffRawW = kOffA + kKS * kASlice;
  static constexpr int kOffRawA = kOffRawW + kRawSfw;
  static constexpr int kOffSfwT = kOffRawA + kRawSfa;
  static constexpr int kOffSfaT = kOffSfwT + kSfT;
  static constexpr int kStageBytesRaw = kOffSfaT + kSfT;
  static constexpr int kStageBytes = (kStageBytesRaw + 1023) / 1024 * 1024;
  static constexpr int kTxBytes = kKS * (kWSlice + kASlice) + kRawSfw + kRawSfa;
  static constexpr int kStages = (196 * 1024) / kStageBytes > 8 ? 8 : (196 * 1024) / kStageBytes;
  static constexpr int kBarrierBytes = (3 * kStages + 1) * 8 + 16;
  static constexpr int kTotalBytes = kStages * kStageBytes + kBarrierBytes + 1024;
  static_assert(kStages * kStageBytes >= MT * 128 * 4, "the partial tile is parked in the ring");
};

template <int MT>
__global__ void __launch_bounds__(kSkinnyThreads, 1)
gemm_skinny_mx4_kernel(const SkinnyMx4Params p_in, const __grid_constant__ CUtensorMap map_w,
                       const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_sfw,
                       const __grid_constant__ CUtensorMap map_sfa) {
  // token tiles (gridDim.z > 1, 128 < M <= 512): tile z owns tokens [z * MT, z * MT + MT), as in gemm_skinny_kernel
  SkinnyMx4Params p = p_in;
  const int m0 = static_cast<int>(blockIdx.z) * MT;
  p.D = static_cast<uint16_t*>(p_in.D) + static_cast<int64_t>(m0) * p_in.ldd;
  p.M = min(MT, p_in.M - m0);
  using L = SkinnyMx4Smem<MT>;
  constexpr int S = L::kStages;
  constexpr int KS = L::kKS;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S * L::kStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + S;
  uint64_t* sf_bar = bars + 2 * S;          // scale chunks of the stage are re-tiled
  uint64_t* acc_bar = bars + 3 * S;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 3 * S + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * 128;
  const int rank = blockIdx.y;
  const int kb0 = rank * p.kb_per_split;
  const int kb1 = min(p.num_k_blocks, kb0 + p.kb_per_split);
  const int num_stages_k = (kb1 - kb0 + KS - 1) / KS;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_w);
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_sfw);
    tma_prefetch_desc(&map_sfa);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < S; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
      mbar_init(&sf_bar[i], 1);
    }
    mbar_init(acc_bar, 1);
    fence_barrier_init();
  }
  // TMEM: accumulator MT columns, then 16 columns of weight scales and 16 of token scales (4 words x 4 columns)
  constexpr uint32_t kCols = MT <= 32 ? 64 : (MT <= 64 ? 128 : 256);
  constexpr uint32_t kSfwCol = MT;
  constexpr uint32_t kSfaCol = MT + 16;
  static_assert(kSfaCol + 16 <= kCols, "TMEM budget");
  if (warp == 2) tmem_alloc<kCols>(tmem_ptr_smem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    // ===================== TMA producer =====================
    // Programmatic dependent launch: weight data and weight scales are constants of the model, requested
    // before griddepcontrol.wait; the token boxes wait for the kernel that produced them.
    int stage = 0;
    uint32_t phase = 0;
    const int pre = min(num_stages_k, S);
    auto load_w = [&](int it, int stg) {
      const int kb = kb0 + it * KS;
      const int ns = min(KS, kb1 - kb);
      uint8_t* st = smem + stg * L::kStageBytes;
      mbar_arrive_expect_tx(&full_bar[stg], ns * (L::kWSlice + L::kASlice) + L::kRawSfw + L::kRawSfa);
      for (int s = 0; s < ns; ++s) tma_load_2d(st + s * L::kWSlice, &map_w, &full_bar[stg], (kb + s) * kBlockKBytes, n0);
      tma_load_2d(st + L::kOffRawW, &map_sfw, &full_bar[stg], kb * 8, n0);
    };
    auto load_a = [&](int it, int stg) {
      const int kb = kb0 + it * KS;
      const int ns = min(KS, kb1 - kb);
      uint8_t* st = smem + stg * L::kStageBytes;
      for (int s = 0; s < ns; ++s)
        tma_load_2d(st + L::kOffA + s * L::kASlice, &map_a, &full_bar[stg], (kb + s) * kBlockKBytes, m0);
      tma_load_2d(st + L::kOffRawA, &map_sfa, &full_bar[stg], kb * 8, m0);
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
    // ===================== MMA issuer: D^T[128 x MT] += (sW W_tile)[128 x 64] * (sA A_tile)[MT x 64]^T =====================
    const uint32_t idesc = make_idesc_block_scaled(1, 128, MT);
    int stage = 0;
    uint32_t phase = 0;
    for (int it = 0; it < num_stages_k; ++it) {
      const int ns = min(KS, kb1 - (kb0 + it * KS));
      mbar_wait(&sf_bar[stage], phase);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t st = smem_u32(smem + stage * L::kStageBytes);
#pragma unroll
        for (int w = 0; w < 4; ++w) {
          utccp_32x128b_warpx4(tmem_base + kSfwCol + w * 4, make_sf_desc(st + L::kOffSfwT + w * 512));
          utccp_32x128b_warpx4(tmem_base + kSfaCol + w * 4, make_sf_desc(st + L::kOffSfaT + w * 512));
        }
        for (int s = 0; s < ns; ++s) {
          const uint64_t dw = make_kmajor_sw128_desc(st + s * L::kWSlice);
          const uint64_t da = make_kmajor_sw128_desc(st + L::kOffA + s * L::kASlice);
#pragma unroll
          for (int k = 0; k < kBlockKBytes / kUmmaKBytes; ++k) {
            // MMA k of slice s covers scale groups s * 8 + 2k, + 1: bytes {0,1} / {2,3} of word s * 2 + k / 2
            const uint32_t w = s * 2 + (k >> 1), id = (k & 1) * 2;
            umma_mxf4(tmem_base, dw + static_cast<uint64_t>(k * (kUmmaKBytes >> 4)),
                      da + static_cast<uint64_t>(k * (kUmmaKBytes >> 4)), idesc_with_sf_ids(idesc, id, id),
                      (it > 0 || s > 0 || k > 0) ? 1u : 0u, tmem_base + kSfwCol + w * 4, tmem_base + kSfaCol + w * 4);
          }
        }
        umma_commit(&empty_bar[stage]);
        if (it == num_stages_k - 1) umma_commit(acc_bar);
      }
      __syncwarp();
      if (++stage == S) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 3) {
    // ===================== scale-factor re-tiler =====================
    // raw box: [row][16 bytes] = words 0..3 of the row (groups 4 w .. 4 w + 3 of this stage)
    // chunk w: [lane][4 words], word c = raw word w of row c * 32 + lane (token rows >= MT: zero)
    int stage = 0;
    uint32_t phase = 0;
    for (int it = 0; it < num_stages_k; ++it) {
      mbar_wait(&full_bar[stage], phase);
      uint8_t* st = smem + stage * L::kStageBytes;
      const uint4* raw_w = reinterpret_cast<const uint4*>(st + L::kOffRawW);
      const uint4* raw_a = reinterpret_cast<const uint4*>(st + L::kOffRawA);
      uint4* out_w = reinterpret_cast<uint4*>(st + L::kOffSfwT);
      uint4* out_a = reinterpret_cast<uint4*>(st + L::kOffSfaT);
      uint4 rw[4], ra[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        rw[c] = raw_w[c * 32 + lane];
        ra[c] = (c * 32 < MT) ? raw_a[c * 32 + lane] : make_uint4(0, 0, 0, 0);
      }
      out_w[0 * 32 + lane] = make_uint4(rw[0].x, rw[1].x, rw[2].x, rw[3].x);
      out_w[1 * 32 + lane] = make_uint4(rw[0].y, rw[1].y, rw[2].y, rw[3].y);
      out_w[2 * 32 + lane] = make_uint4(rw[0].z, rw[1].z, rw[2].z, rw[3].z);
      out_w[3 * 32 + lane] = make_uint4(rw[0].w, rw[1].w, rw[2].w, rw[3].w);
      out_a[0 * 32 + lane] = make_uint4(ra[0].x, ra[1].x, ra[2].x, ra[3].x);
      out_a[1 * 32 + lane] = make_uint4(ra[0].y, ra[1].y, ra[2].y, ra[3].y);
      out_a[2 * 32 + lane] = make_uint4(ra[0].z, ra[1].z, ra[2].z, ra[3].z);
      out_a[3 * 32 + lane] = make_uint4(ra[0].w, ra[1].w, ra[2].w, ra[3].w);
      fence_proxy_async_smem();   // generic-proxy writes -> visible to tcgen05.cp (async proxy)
      __syncwarp();
      if (lane == 0) mbar_arrive(&sf_bar[stage]);
      if (++stage == S) { stage = 0; phase ^= 1; }
    }
  }

  // ===================== epilogue (see gemm_skinny.cuh; the scales were applied inside the MMA) =====================
  asm volatile("griddepcontrol.wait;" ::: "memory");   // bias / D may depend on the previous kernel
  const bool bf = p.out_dtype == VXK_BF16;
  float* park = reinterpret_cast<float*>(smem);
  if (warp >= 4) {
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
    cluster_sync_all();          // every CTA's partial tile is parked
    const uint32_t park_addr = smem_u32(park);
    const int nc = n0 + lane * 4;                       // this lane's 4 output columns
    float bv[4];
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const bool ok = nc + t < p.N;
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
      uint16_t h[4];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        float v = acc[t];
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
