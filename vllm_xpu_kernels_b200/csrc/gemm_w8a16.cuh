// fp8_gemm_w8a16 on tcgen05 (device side): bf16 / f16 activations x FP8 weights.
//
//   D[M,N] = epilogue( sum_k A[m,k] * dequant(W[n,k]) )     A, W K-major
//   ref: fp8_gemm_w8a16  csrc/xpu/onednn/onednn_matmul.cpp:81-99 -> dnnl_matmul_w8a16_fp8
//        csrc/xpu/onednn/fp8_gemm_w8a16.h:13-145
//
// The tensor core has no mixed 16-bit x 8-bit mode, so the FP8 weight tile is widened to the
// activation type INSIDE the main loop -- the weight stream from HBM stays 1 byte / element,
// which is what bounds the small-M (decode) shapes this op exists for:
//
//   warp 0       TMA producer: A tile (128 rows x 64 elements, SWIZZLE_128B) and the RAW FP8 W tile
//                (BLOCK_N rows x 64 bytes, linear) of one K slice into a ring stage
//   warps 8-15   converters: raw FP8 -> bf16 / f16 (exact; e4m3 and e5m2 are subsets of both) written
//                as the K-major SWIZZLE_128B operand tile of the same stage, generic->async proxy
//                fence, arrive on the stage's conv barrier.  128x128 block scales are folded in
//                here (fp32 multiply, one rounding to the activation type -- the reference test's
//                own dequantisation, test_fp8_gemm_onednn.py:367-410); per-tensor / per-channel
//                scales are applied to the fp32 accumulator in the epilogue instead.
//   warp 1       MMA issuer: tcgen05.mma.kind::f16 128 x BLOCK_N x 16, four per stage, two TMEM
//                accumulators; tcgen05.commit frees the stage
//   warp 2       TMEM allocator
//   warps 4-7    epilogue (shared with the FP8 kernels: epilogue_store_32)
#pragma once

#include "gemm_tc.cuh"

#ifndef VXK_W16_BF16_VIA_F32
#define VXK_W16_BF16_VIA_F32 1
#endif

namespace vxk {

constexpr int kW16Threads = 512;
constexpr int kW16ConvWarps = 8;
constexpr int kW16BlockK = 64;   // elements per K slice: 128 B of A, 64 B of raw W

struct W16Params {
  TcParams b;             // D, b_scale (fp32: scalar / [N]), bias, M, N, ldd, b_scale_mode, out_dtype (= act dtype), ...
  const float* blk_scale; // block mode: sB[kb * sb0 + nb * sb1]
  int64_t sb0, sb1;
  int group_k, group_n;   // block mode; group_k % 64 == 0
  int K;                  // elements
};

__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// kind::f16 instruction descriptor: c_format f32, a/b_format 0 = f16, 1 = bf16, both K-major
__host__ __device__ constexpr uint32_t make_idesc_f16(int is_bf16, int umma_m, int umma_n) {
  return (1u << 4) | (static_cast<uint32_t>(is_bf16) << 7) | (static_cast<uint32_t>(is_bf16) << 10) |
         (static_cast<uint32_t>(umma_n >> 3) << 17) | (static_cast<uint32_t>(umma_m >> 4) << 24);
}

// two FP8 codes (low 16 bits of `pair`) -> packed f16x2 (exact)
template <bool E5M2>
__device__ __forceinline__ uint32_t fp8x2_to_f16x2(uint32_t pair) {
  uint32_t out;
  const uint16_t in = static_cast<uint16_t>(pair);
  if constexpr (E5M2) asm("cvt.rn.f16x2.e5m2x2 %0, %1;" : "=r"(out) : "h"(in));
  else asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(out) : "h"(in));
  return out;
}
// four FP8 codes (one 32-bit word) -> two packed bf16x2 words (exact) without going through fp32:
// sign / exponent / mantissa bits are put where bf16 keeps them (the exponent still carrying the FP8
// bias), then one packed multiply by 2^(127 - bias) fixes the bias and normalises FP8 subnormals in
// the same step.  The 4-bit (e4m3) / 3-bit (e5m2) shift is done on all four codes at once, split into
// the byte that becomes the high half-byte (sign + upper magnitude bits) and the byte that becomes the
// low one; two byte permutes interleave them: 7 ALU-pipe instructions + 2 multiplies per 4 codes (the
// widening is bound by the half-rate ALU pipe, ncu profiles/r01e_w8a16_*.csv).
template <bool E5M2>
__device__ __forceinline__ void fp8x4_to_bf16x4(uint32_t word, uint32_t& lo, uint32_t& hi) {
  constexpr int sh = E5M2 ? 3 : 4;
  constexpr uint32_t hmask = E5M2 ? 0x0f0f0f0fu : 0x07070707u;
  constexpr uint32_t lmask = E5M2 ? 0xe0e0e0e0u : 0xf0f0f0f0u;
  const uint32_t x = ((word >> sh) & hmask) | (word & 0x80808080u);   // high bytes: s000 0eee (e4m3)
  const uint32_t y = (word << (8 - sh)) & lmask;                      // low bytes:  emmm 0000
  const uint32_t t0 = __byte_perm(y, x, 0x5140u);                     // codes 0, 1
  const uint32_t t1 = __byte_perm(y, x, 0x7362u);                     // codes 2, 3
  constexpr uint32_t fix = E5M2 ? 0x77807780u : 0x7b807b80u;          // bf16 2^112 / 2^120, packed
  asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(lo) : "r"(t0), "r"(fix));
  asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(hi) : "r"(t1), "r"(fix));
}

// shared-space accesses with 32-bit addresses (a generic pointer costs 64-bit address arithmetic and the
// generic load / store path in the converters' inner loop)
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
// wait of a warp that has nothing to do for a long time (epilogue warps during the main loop): poll with
// back-off so that the polling does not take issue slots from the converter warps on the same scheduler
__device__ __forceinline__ void mbar_wait_backoff(uint64_t* bar, uint32_t phase) {
  while (!mbar_try_wait(bar, phase)) __nanosleep(256);
}

// 16 FP8 codes -> 16 activation-type values (8 packed words), optional fp32 scale folded in.
template <bool BF16, bool E5M2, bool SCALED>
__device__ __forceinline__ void widen16(const uint4& raw, float s, uint32_t (&w)[8]) {
  const uint32_t in[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if constexpr (SCALED) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const uint32_t f16 = fp8x2_to_f16x2<E5M2>(in[i] >> (16 * h));
        float2 f = __half22float2(*reinterpret_cast<const __half2*>(&f16));
        f.x *= s; f.y *= s;
        if constexpr (BF16) {
          const __nv_bfloat162 o = __floats2bfloat162_rn(f.x, f.y);
          w[2 * i + h] = *reinterpret_cast<const uint32_t*>(&o);
        } else {
          const __half2 o = __floats2half2_rn(f.x, f.y);
          w[2 * i + h] = *reinterpret_cast<const uint32_t*>(&o);
        }
      }
    } else if constexpr (BF16) {
#if VXK_W16_BF16_VIA_F32
      // fp8 -> f16 (one cvt per pair) -> f32 (FMA pipe) -> packed bf16: 2 ALU-pipe + 2 FMA-pipe instructions per pair
      // against 3.5 + 1 for the bit-placement form; the converters are bound by the half-rate ALU pipe
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const uint32_t f16 = fp8x2_to_f16x2<E5M2>(in[i] >> (16 * h));
        const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&f16));
        const __nv_bfloat162 o = __floats2bfloat162_rn(f.x, f.y);
        w[2 * i + h] = *reinterpret_cast<const uint32_t*>(&o);
      }
#else
      fp8x4_to_bf16x4<E5M2>(in[i], w[2 * i], w[2 * i + 1]);
#endif
    } else {
      w[2 * i] = fp8x2_to_f16x2<E5M2>(in[i]);
      w[2 * i + 1] = fp8x2_to_f16x2<E5M2>(in[i] >> 16);
    }
  }
}

template <int BLOCK_N>
struct W16Smem {
  static constexpr int kABytes = kBlockM * 128;
  static constexpr int kRawBytes = BLOCK_N * kW16BlockK;
  static constexpr int kConvBytes = BLOCK_N * 128;
  static constexpr int kStageBytesRaw = kABytes + kConvBytes + kRawBytes;
  static constexpr int kStageBytes = (kStageBytesRaw + 1023) / 1024 * 1024;
  static constexpr int kTxBytes = kABytes + kRawBytes;
  static constexpr int kStages = (200 * 1024) / kStageBytes > 8 ? 8 : (200 * 1024) / kStageBytes;
  static constexpr int kBarrierBytes = (3 * kStages + 4) * 8 + 16;
  static constexpr int kTotalBytes = kStages * kStageBytes + kBarrierBytes + 1024;
};

template <int BLOCK_N, bool BF16, bool E5M2, bool SCALED>
__global__ void __launch_bounds__(kW16Threads, 1)
gemm_w8a16_kernel(const W16Params wp, const __grid_constant__ CUtensorMap map_a,
                  const __grid_constant__ CUtensorMap map_w) {
  using L = W16Smem<BLOCK_N>;
  constexpr int S = L::kStages;
  const TcParams& p = wp.b;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  auto stage_a = [&](int s) { return smem + s * L::kStageBytes; };
  auto stage_w = [&](int s) { return stage_a(s) + L::kABytes; };            // widened operand tile
  auto stage_raw = [&](int s) { return stage_w(s) + L::kConvBytes; };       // FP8 bytes as TMA delivers them
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S * L::kStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + S;
  uint64_t* conv_bar = bars + 2 * S;
  uint64_t* tmem_full_bar = bars + 3 * S;
  uint64_t* tmem_empty_bar = bars + 3 * S + 2;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 3 * S + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_w);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < S; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
      mbar_init(&conv_bar[i], kW16ConvWarps);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full_bar[i], 1);
      mbar_init(&tmem_empty_bar[i], 128);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<kTmemCols>(tmem_ptr_smem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  constexpr uint32_t kAccCols = BLOCK_N;
  const int num_tiles = p.num_m_blocks * p.num_n_blocks;

  if (warp == 0) {
    // ===================== TMA producer =====================
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m0 = (tile % p.num_m_blocks) * kBlockM;
      const int n0 = (tile / p.num_m_blocks) * BLOCK_N;
      for (int kb = 0; kb < p.num_k_blocks; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        if (elect_one()) {
          mbar_arrive_expect_tx(&full_bar[stage], L::kTxBytes);
          tma_load_2d(stage_a(stage), &map_a, &full_bar[stage], kb * 128, m0);          // byte coordinates
          tma_load_2d(stage_raw(stage), &map_w, &full_bar[stage], kb * kW16BlockK, n0);
        }
        __syncwarp();
        if (++stage == S) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    const uint32_t idesc = make_idesc_f16(BF16 ? 1 : 0, kBlockM, BLOCK_N);
    int stage = 0;
    uint32_t phase = 0;
    int iter = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++iter) {
      const int as = iter & 1;
      const uint32_t aphase = (iter >> 1) & 1;
      mbar_wait(&tmem_empty_bar[as], aphase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + as * kAccCols;
      for (int kb = 0; kb < p.num_k_blocks; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        mbar_wait(&conv_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t da = make_kmajor_sw128_desc(smem_u32(stage_a(stage)));
          const uint64_t db = make_kmajor_sw128_desc(smem_u32(stage_w(stage)));
#pragma unroll
          for (int k = 0; k < 4; ++k)                                        // 16 elements = 32 B per MMA
            umma_f16(tmem_d, da + static_cast<uint64_t>(2 * k), db + static_cast<uint64_t>(2 * k), idesc,
                     (kb > 0 || k > 0) ? 1u : 0u);
          umma_commit(&empty_bar[stage]);
          if (kb == p.num_k_blocks - 1) umma_commit(&tmem_full_bar[as]);
        }
        __syncwarp();
        if (++stage == S) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp >= 8) {
    // ===================== converters =====================
    const int ct = threadIdx.x - 256;                      // 0..255
    constexpr int kChunks = BLOCK_N * 4;                   // 16-byte chunks of the raw tile
    constexpr int kIter = (kChunks + 255) / 256;
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int n0 = (tile / p.num_m_blocks) * BLOCK_N;
      for (int kb = 0; kb < p.num_k_blocks; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        const uint32_t raw = smem_u32(stage_raw(stage));
        const uint32_t dst = smem_u32(stage_w(stage));
        uint4 v[kIter];
#pragma unroll
        for (int i = 0; i < kIter; ++i) {
          const int c = i * 256 + ct;
          if (kChunks % 256 == 0 || c < kChunks) v[i] = lds128(raw + c * 16);
        }
#pragma unroll
        for (int i = 0; i < kIter; ++i) {
          const int c = i * 256 + ct;
          if (kChunks % 256 == 0 || c < kChunks) {
            const int row = c >> 2, j = c & 3;
            float s = 1.0f;
            if constexpr (SCALED) {
              const int n = n0 + row;
              const int nn = n < p.N ? n : p.N - 1;
              s = __ldg(wp.blk_scale + static_cast<int64_t>((kb * kW16BlockK) / wp.group_k) * wp.sb0 +
                        static_cast<int64_t>(nn / wp.group_n) * wp.sb1);
            }
            uint32_t w[8];
            widen16<BF16, E5M2, SCALED>(v[i], s, w);
            const uint32_t rowp = dst + row * 128;
            const int sw = row & 7;
            sts128(rowp + (((2 * j) ^ sw) << 4), w[0], w[1], w[2], w[3]);
            sts128(rowp + (((2 * j + 1) ^ sw) << 4), w[4], w[5], w[6], w[7]);
          }
        }
        fence_proxy_async_smem();          // generic-proxy writes -> visible to tcgen05.mma (async proxy)
        __syncwarp();
        if (lane == 0) mbar_arrive(&conv_bar[stage]);
        if (++stage == S) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    const int q = warp & 3;
    const int row_in_tile = q * 32 + lane;
    int iter = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++iter) {
      const int as = iter & 1;
      const uint32_t aphase = (iter >> 1) & 1;
      const int m0 = (tile % p.num_m_blocks) * kBlockM;
      const int n0 = (tile / p.num_m_blocks) * BLOCK_N;
      const int m = m0 + row_in_tile;
      mbar_wait_backoff(&tmem_full_bar[as], aphase);
      tc_fence_after();
      const float row_scale = (!SCALED && p.b_scale_mode == 0) ? __ldg(p.b_scale) : 1.0f;
      const bool per_col = !SCALED && p.b_scale_mode == 1;
#pragma unroll 1
      for (int c = 0; c < BLOCK_N / 32; ++c) {
        uint32_t r[32];
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * kAccCols + c * 32;
        tmem_ld_32x32b_x32(taddr, r);
        tmem_ld_wait();
        const int nbase = n0 + c * 32;
        if (m < p.M && nbase < p.N) epilogue_store_32(p, r, m, nbase, row_scale, per_col);
      }
      tc_fence_before();
      mbar_arrive(&tmem_empty_bar[as]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<kTmemCols>(tmem_base);
  }
}

}  // namespace vxk
