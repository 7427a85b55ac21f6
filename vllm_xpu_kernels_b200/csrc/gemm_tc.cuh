// tcgen05 scaled-GEMM kernel for sm_100a (device side).
//
//   D[M,N] = epilogue( sum_k A[m,k] * W[n,k] )      A, W K-major (row = m / n, K contiguous)
//
// One persistent CTA per SM, 256 threads, warp-specialised:
//   warp 0      TMA producer: A / W tiles (128 x 128 B, SWIZZLE_128B) and, for the
//               block-scaled kinds, the raw UE8M0 scale bytes, into a STAGES-deep ring
//   warp 1      MMA issuer: one elected thread issues tcgen05.mma (UMMA 128 x BLOCK_N x 32 B)
//               into one of two TMEM accumulator buffers, tcgen05.commit frees ring slots
//   warp 2      TMEM allocator; for block-scaled kinds re-tiles the scale bytes in shared
//               memory into the 32-lane x 128-bit layout tcgen05.cp wants
//   warps 4-7   epilogue: tcgen05.ld the 128 x BLOCK_N fp32 accumulator (one TMEM lane =
//               one output row per thread), apply per-token / per-channel scales and bias,
//               convert to bf16 / f16, 128-bit stores.  Runs concurrently with the next
//               tile's main loop (double-buffered accumulator).
//
// KIND_F8   : kind::f8f6f4, scales applied in the epilogue (per tensor / token / channel)
// KIND_MXF8 : kind::mxf8f6f4.block_scale, UE8M0 scale per 32 K-elements of each row
// KIND_MXF4 : kind::mxf4.block_scale.scale_vec::2X, packed E2M1, UE8M0 scale per 32
#pragma once

#include "gemm_common.cuh"

namespace vxk {

enum GemmKind { KIND_F8 = 0, KIND_MXF8 = 1, KIND_MXF4 = 2 };

constexpr int kBlockM = 128;
constexpr int kBlockKBytes = 128;          // one 128 B swizzle atom of K per stage
constexpr int kUmmaKBytes = 32;            // K bytes consumed by one tcgen05.mma
constexpr int kTcThreads = 256;
constexpr uint32_t kTmemCols = 512;

struct TcParams {
  void* D;
  const float* a_scale;   // KIND_F8: [M] or scalar
  const float* b_scale;   // KIND_F8: [N] or scalar
  const void* bias;
  int M, N, K_bytes;      // K extent in BYTES of operand storage (fp8: K, fp4: K/2)
  int64_t ldd;
  int a_scale_mode, b_scale_mode, out_dtype, bias_dtype;
  int fp8_fmt;            // 0 = E4M3, 1 = E5M2 (fp8 kinds)
  int num_m_blocks, num_n_blocks, num_k_blocks;
  int vec_store_ok;       // D rows 16-byte aligned -> 128-bit stores
  int n_peers;            // > 0: fused all-gather -- the tile also goes to peer_D[0..n_peers), the OTHER ranks'
  void* peer_D[8];        //      peer-mapped [M, N_total] outputs (offset to this rank's columns, row stride ldd like D)
  int peer_multicast;     // 1: peer_D[0] is an NVLS multicast mapping of all ranks' outputs (one multimem.st reaches all)
  int debug;              // bring-up switches (env VXK_GEMM_DEBUG): 1 skip tcgen05.cp, 2 plain MMA, 4 skip re-tiler
  // fused gate_up GEMM -> silu_and_mul -> 1x128 block FP8 quantisation (gemm_tc2.cuh, FUSE): W is [2 * inter, K] (gate rows,
  // then up rows), N = 2 * inter; D is not written, the e4m3 codes [M, inter] and fp32 scales go to fuse_q / fuse_s
  void* fuse_q;
  float* fuse_s;          // scale of (row m, group g) at fuse_s[m * fuse_ss0 + g * fuse_ss1]
  int64_t fuse_ss0, fuse_ss1;
  int fuse_inter;
};

// One thread's 32 consecutive accumulator columns of output row m: scales, bias, conversion
// to bf16 / f16 and the stores (128-bit when the row is 16-byte aligned and the chunk is whole).
// Every array index is a compile-time constant (no local memory).
// (the arithmetic half: 32 accumulator columns -> 16 packed words of the output type)
__device__ __forceinline__ void epilogue_pack_32(const TcParams& p, const uint32_t (&r)[32], int nbase, float row_scale,
                                                 bool per_col, uint32_t (&w)[16]) {
  float v[32];
  const bool whole = nbase + 32 <= p.N;
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * row_scale;
  if (per_col) {
    if (whole && (reinterpret_cast<uintptr_t>(p.b_scale) & 15) == 0) {   // nbase % 32 == 0: 128-bit broadcast loads
      const float4* s4 = reinterpret_cast<const float4*>(p.b_scale + nbase);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 s = __ldg(s4 + j);
        v[4 * j] *= s.x; v[4 * j + 1] *= s.y; v[4 * j + 2] *= s.z; v[4 * j + 3] *= s.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (nbase + j < p.N) v[j] *= __ldg(p.b_scale + nbase + j);
    }
  }
  if (p.bias) {
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const int n = nbase + j;
      if (n < p.N)
        v[j] += (p.bias_dtype == VXK_F32)   ? __ldg(static_cast<const float*>(p.bias) + n)
                : (p.bias_dtype == VXK_F16) ? __half2float(static_cast<const __half*>(p.bias)[n])
                                            : __bfloat162float(static_cast<const __nv_bfloat16*>(p.bias)[n]);
    }
  }
  const bool bf = p.out_dtype == VXK_BF16;
#pragma unroll
  for (int t = 0; t < 16; ++t) {
    if (bf) {
      __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * t], v[2 * t + 1]);
      w[t] = *reinterpret_cast<uint32_t*>(&h);
    } else {
      __half2 h = __floats2half2_rn(v[2 * t], v[2 * t + 1]);
      w[t] = *reinterpret_cast<uint32_t*>(&h);
    }
  }
}

__device__ __forceinline__ void epilogue_store_32(const TcParams& p, const uint32_t (&r)[32], int m, int nbase,
                                                  float row_scale, bool per_col, bool direct_peers = true) {
  uint32_t w[16];
  epilogue_pack_32(p, r, nbase, row_scale, per_col, w);
  const bool whole = nbase + 32 <= p.N;
  const int64_t off = static_cast<int64_t>(m) * p.ldd + nbase;
  if (p.n_peers > 0 && direct_peers) {
    // fused all-gather, simple form: the 32-column piece is stored straight into every other rank's output
    // through its NVLink peer mapping.  One thread = one row, so a warp's store touches 32 rows: correct for
    // any shape / alignment but NVLink-inefficient; the CTA-pair kernel copies whole tile rows instead
    // (gemm_tc2.cuh) and only falls back to this for unaligned outputs.
#pragma unroll 1
    for (int pe = 0; pe < p.n_peers; ++pe) {
      uint16_t* drow = static_cast<uint16_t*>(p.peer_D[pe]) + off;
      if (p.vec_store_ok && whole) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          *reinterpret_cast<uint4*>(drow + 8 * j) = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (nbase + j < p.N) drow[j] = static_cast<uint16_t>(j & 1 ? w[j >> 1] >> 16 : w[j >> 1] & 0xffffu);
      }
    }
  }
  uint16_t* drow = static_cast<uint16_t*>(p.D) + off;
  if (p.vec_store_ok && whole) {
#pragma unroll
    for (int j = 0; j < 4; ++j)
      *reinterpret_cast<uint4*>(drow + 8 * j) = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
  } else {
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (nbase + j < p.N) drow[j] = static_cast<uint16_t>(j & 1 ? w[j >> 1] >> 16 : w[j >> 1] & 0xffffu);
  }
}

template <int KIND, int BLOCK_N>
struct TcSmem {
  static constexpr int kScaled = KIND != KIND_F8;
  static constexpr int kSfGroups = KIND == KIND_MXF4 ? 8 : 4;        // 32-element groups per stage
  static constexpr int kSfWords = kSfGroups / 4;                     // 32-bit SF words per row per stage
  static constexpr int kABytes = kBlockM * kBlockKBytes;
  static constexpr int kBBytes = BLOCK_N * kBlockKBytes;
  static constexpr int kRawSfaBytes = kScaled ? kBlockM * 16 : 0;    // TMA box: 16 scale bytes per row
  static constexpr int kNB = (BLOCK_N + 127) / 128;                  // 128-row scale blocks of the B tile
  static constexpr int kRawSfbBytes = kScaled ? kNB * 128 * 16 : 0;  // sized for whole 128-row blocks
  static constexpr int kSfaTBytes = kScaled ? kSfWords * (kBlockM / 128) * 512 : 0;
  static constexpr int kSfbTBytes = kScaled ? kSfWords * kNB * 512 : 0;
  static constexpr int kStageBytesRaw = kABytes + kBBytes + kRawSfaBytes + kRawSfbBytes + kSfaTBytes + kSfbTBytes;
  static constexpr int kStageBytes = (kStageBytesRaw + 1023) / 1024 * 1024;
  static constexpr int kTxBytes = kABytes + kBBytes + kRawSfaBytes + (kScaled ? BLOCK_N * 16 : 0);
  static constexpr int kStages = (200 * 1024) / kStageBytes > 8 ? 8 : (200 * 1024) / kStageBytes;
  static constexpr int kBarrierBytes = (3 * kStages + 4) * 8 + 16;
  static constexpr int kTotalBytes = kStages * kStageBytes + kBarrierBytes + 1024;  // + alignment slack
  static_assert(kStageBytes % 1024 == 0, "stage must keep 1024 B alignment for SWIZZLE_128B");
};

template <int KIND, int BLOCK_N>
__global__ void __launch_bounds__(kTcThreads, 1)
gemm_tc_kernel(const TcParams p, const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
               const __grid_constant__ CUtensorMap map_sfa, const __grid_constant__ CUtensorMap map_sfb) {
  using L = TcSmem<KIND, BLOCK_N>;
  constexpr int S = L::kStages;
  constexpr bool kScaled = L::kScaled;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  auto stage_a = [&](int s) { return smem + s * L::kStageBytes; };
  auto stage_b = [&](int s) { return stage_a(s) + L::kABytes; };
  auto stage_raw_sfa = [&](int s) { return stage_b(s) + L::kBBytes; };
  auto stage_raw_sfb = [&](int s) { return stage_raw_sfa(s) + L::kRawSfaBytes; };
  auto stage_sfa_t = [&](int s) { return stage_raw_sfb(s) + L::kRawSfbBytes; };
  auto stage_sfb_t = [&](int s) { return stage_sfa_t(s) + L::kSfaTBytes; };
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S * L::kStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + S;
  uint64_t* sf_bar = bars + 2 * S;
  uint64_t* tmem_full_bar = bars + 3 * S;
  uint64_t* tmem_empty_bar = bars + 3 * S + 2;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 3 * S + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
    if constexpr (kScaled) {
      tma_prefetch_desc(&map_sfa);
      tma_prefetch_desc(&map_sfb);
    }
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < S; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
      mbar_init(&sf_bar[i], 1);
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
  constexpr uint32_t kAccCols = BLOCK_N;                       // fp32 accumulator columns per buffer
  constexpr uint32_t kSfaCol = 2 * kAccCols;                   // SFA: kSfWords * 4 columns
  constexpr uint32_t kSfbCol = kSfaCol + 8;                    // SFB: kSfWords * 4 * (BLOCK_N/128) columns
  constexpr int NB = L::kNB;
  static_assert(!kScaled || kSfbCol + 8 * NB <= kTmemCols, "TMEM budget");

  const int num_tiles = p.num_m_blocks * p.num_n_blocks;

  if (warp == 0) {
    // ===================== TMA producer (whole warp converged, one elected lane issues) =====================
    {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m0 = (tile % p.num_m_blocks) * kBlockM;
        const int n0 = (tile / p.num_m_blocks) * BLOCK_N;
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if (elect_one()) {
            mbar_arrive_expect_tx(&full_bar[stage], L::kTxBytes);
            tma_load_2d(stage_a(stage), &map_a, &full_bar[stage], kb * kBlockKBytes, m0);
            tma_load_2d(stage_b(stage), &map_b, &full_bar[stage], kb * kBlockKBytes, n0);
            if constexpr (kScaled) {
              tma_load_2d(stage_raw_sfa(stage), &map_sfa, &full_bar[stage], (kb * L::kSfGroups) & ~15, m0);
              tma_load_2d(stage_raw_sfb(stage), &map_sfb, &full_bar[stage], (kb * L::kSfGroups) & ~15, n0);
            }
          }
          __syncwarp();
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (whole warp converged, one elected lane issues) =====================
    {
      const uint32_t idesc = make_idesc_block_scaled(KIND == KIND_MXF4 ? 1 : p.fp8_fmt, kBlockM, BLOCK_N);
      const uint32_t idesc_f8 = make_idesc_f8(p.fp8_fmt, kBlockM, BLOCK_N);
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
          mbar_wait(kScaled ? &sf_bar[stage] : &full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint64_t da = make_kmajor_sw128_desc(smem_u32(stage_a(stage)));
            const uint64_t db = make_kmajor_sw128_desc(smem_u32(stage_b(stage)));
            if (kScaled && !(p.debug & 1)) {
#pragma unroll
              for (int w = 0; w < L::kSfWords; ++w) {
                utccp_32x128b_warpx4(tmem_base + kSfaCol + w * 4, make_sf_desc(smem_u32(stage_sfa_t(stage) + w * 512)));
#pragma unroll
                for (int j = 0; j < NB; ++j)
                  utccp_32x128b_warpx4(tmem_base + kSfbCol + (w * NB + j) * 4,
                                       make_sf_desc(smem_u32(stage_sfb_t(stage) + (w * NB + j) * 512)));
              }
            }
#pragma unroll
            for (int k = 0; k < kBlockKBytes / kUmmaKBytes; ++k) {
              const uint64_t dak = da + static_cast<uint64_t>(k * (kUmmaKBytes >> 4));
              const uint64_t dbk = db + static_cast<uint64_t>(k * (kUmmaKBytes >> 4));
              const uint32_t acc = (kb > 0 || k > 0) ? 1u : 0u;
              if (KIND == KIND_F8 || (p.debug & 2)) {
                umma_f8f6f4(tmem_d, dak, dbk, idesc_f8, acc);
              } else if constexpr (KIND == KIND_MXF8) {
                // one SF byte per row per MMA: byte k of the row's 32-bit word
                umma_mxf8f6f4(tmem_d, dak, dbk, idesc_with_sf_ids(idesc, k, k), acc, tmem_base + kSfaCol,
                              tmem_base + kSfbCol);
              } else {
                // two SF bytes per row per MMA: bytes {0,1} or {2,3} of word k/2
                const uint32_t w = k >> 1, id = (k & 1) * 2;
                umma_mxf4(tmem_d, dak, dbk, idesc_with_sf_ids(idesc, id, id), acc, tmem_base + kSfaCol + w * 4,
                          tmem_base + kSfbCol + w * NB * 4);
              }
            }
            umma_commit(&empty_bar[stage]);                 // frees the ring slot when the MMAs retire
            if (kb == p.num_k_blocks - 1) umma_commit(&tmem_full_bar[as]);
          }
          __syncwarp();
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 2) {
    // ===================== scale-factor re-tiler =====================
    if constexpr (kScaled) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          // raw: [row][16 B] fetched at a 16-byte aligned group index (TMA faults on unaligned
          // inner coordinates of a 16-byte box); this stage's groups start at byte (kb*G) % 16
          const int wofs = ((kb * L::kSfGroups) & 15) >> 2;
          // out (per word w, per 128-row block j): [lane][4 words] with word c = row j*128 + c*32 + lane
          const uint32_t* raw_a = reinterpret_cast<const uint32_t*>(stage_raw_sfa(stage));
          const uint32_t* raw_b = reinterpret_cast<const uint32_t*>(stage_raw_sfb(stage));
          uint4* out_a = reinterpret_cast<uint4*>(stage_sfa_t(stage));
          uint4* out_b = reinterpret_cast<uint4*>(stage_sfb_t(stage));
#pragma unroll
          for (int w = 0; w < L::kSfWords && !(p.debug & 4); ++w) {
            uint4 v;
            v.x = raw_a[(0 * 32 + lane) * 4 + wofs + w];
            v.y = raw_a[(1 * 32 + lane) * 4 + wofs + w];
            v.z = raw_a[(2 * 32 + lane) * 4 + wofs + w];
            v.w = raw_a[(3 * 32 + lane) * 4 + wofs + w];
            out_a[w * 32 + lane] = v;
#pragma unroll
            for (int j = 0; j < NB; ++j) {
              uint4 u;
              u.x = raw_b[(j * 128 + 0 * 32 + lane) * 4 + wofs + w];
              u.y = raw_b[(j * 128 + 1 * 32 + lane) * 4 + wofs + w];
              u.z = raw_b[(j * 128 + 2 * 32 + lane) * 4 + wofs + w];
              u.w = raw_b[(j * 128 + 3 * 32 + lane) * 4 + wofs + w];
              out_b[(w * NB + j) * 32 + lane] = u;
            }
          }
          fence_proxy_async_smem();   // generic-proxy writes -> visible to tcgen05.cp (async proxy)
          __syncwarp();
          if (lane == 0) mbar_arrive(&sf_bar[stage]);
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    const int q = warp & 3;                 // TMEM lane quarter this warp may access
    const int row_in_tile = q * 32 + lane;
    int iter = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++iter) {
      const int as = iter & 1;
      const uint32_t aphase = (iter >> 1) & 1;
      const int m0 = (tile % p.num_m_blocks) * kBlockM;
      const int n0 = (tile / p.num_m_blocks) * BLOCK_N;
      const int m = m0 + row_in_tile;
      mbar_wait(&tmem_full_bar[as], aphase);
      tc_fence_after();
      float row_scale = 1.0f;
      if constexpr (KIND == KIND_F8) {
        row_scale = (p.a_scale_mode == 1) ? (m < p.M ? __ldg(p.a_scale + m) : 0.0f) : __ldg(p.a_scale);
        if (p.b_scale_mode == 0) row_scale *= __ldg(p.b_scale);
      }
      const bool per_col = (KIND == KIND_F8) && (p.b_scale_mode == 1);
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
