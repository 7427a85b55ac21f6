// FP8 GEMM with fp32 block scales (DeepSeek-V3 recipe: activations scaled per 1 x group_k,
// weights per group_k x group_n, any fp32 values -- not powers of two, so the block-scaled MMA
// kinds do not apply).  CTA-pair tcgen05 kernel with two-level accumulation:
//
//   for every K slice of 128:   P   = A[:, slice] * W[:, slice]^T        tensor core -> TMEM
//                               acc += sA[m, slice] * sW[slice, n] * P    CUDA cores, fp32 registers
//
// Same producer / ring / CTA-pair structure as gemm_tc2.cuh (256 x 192 tile per pair, plain
// kind::f8f6f4 MMAs), but the two TMEM buffers alternate per K SLICE: while the tensor core
// fills one, the eight promotion warps of each CTA pull the other into registers
// (tcgen05.ld, 32-column chunks through two rotating register buffers so that the FMAs of one
// chunk overlap the load of the next), hand the buffer back as soon as its last chunk has landed,
// and fold it into their accumulators (96 columns per thread, packed fp32x2 FMAs).  The promotion
// warps take 224 registers (setmaxnreg), the role warps give theirs up.
// What bounds it (round 2, profiles/r02_blk_analysis.txt).  Per K slice and CTA pair at 8192 x 36864 x 7168:
//   m  = 344 ns  MMA side alone (promotion warps hand every buffer straight back, VXK_GEMM_DEBUG=16): the L2 -> SM
//                operand stream of a 256 x 192 tile, not the tensor pipe (278 ns at the measured FP8 peak)
//   d  = 255 ns  draining 96 KB per CTA through tcgen05.ld: 223 B/clk/SM measured WITH the tensor pipe busy
//                (253 idle, benchmarks/tmem_contention.cu) -- the full-rate FP8 MMA needs 256 B/clk/SM, so this
//                algorithm can never pass ~85 % of the tensor peak on sm_100
//   h  = 335 ns  the two hand-overs (commit -> mbarrier -> promotion warps, and sixteen arrivals across the
//                pair -> MMA warp -> issue -> first MMA)
// With two TMEM buffers every buffer goes round the loop MMA -> hand-over -> drain -> hand-over, so the period
// per slice is (m + d + h) / 2 = 467 ns -- exactly what the kernel measures (1995 TFLOP/s, 0.60 of the FP8 peak).
// More buffers would take h off the critical path, but the TMEM has 512 columns: four 128-column buffers
// (256 x 128 tiles) were built and measured at 1200-1420 TFLOP/s (eight or sixteen promotion warps, register
// double buffer) -- the 128-wide tile's MMA side alone reaches only 1515 TFLOP/s; 256-wide tiles leave no
// registers for the staging buffers (128 accumulators per thread: 700 B of spills, 1860 TFLOP/s with 16-column
// chunks).  Three 160-column buffers (256 x 160 tiles, columns split 96 + 64 over the two promotion warps of a lane
// quarter; VXK_BLK_BN=160) take h off the critical path as intended but lose more on the MMA side (fewer flops per
// operand byte) and on the 96 / 64 imbalance: 1427 vs 1756 TFLOP/s at 8192 x 36864 x 7168, slower on all six
// DeepSeek-V3 shapes (profiles/r02_blk_analysis.txt).  192 x 2 is the measured optimum of this design space.
// ref: fp8_gemm_w8a8.h:28-41,86-134 (block scale shapes), tests/test_fp8_gemm_onednn.py:270-358.
#pragma once

#include "gemm_tc2.cuh"

namespace vxk {

struct Tc2BlkParams {
  TcParams b;
  const float* sa;        // sA[m * sa0 + kg * sa1]
  const float* sb;        // sW[kg * sb0 + ng * sb1]
  int64_t sa0, sa1, sb0, sb1;
  int group_k_blocks;     // group_k / 128
  int group_n;
  int num_m_pairs, num_n_blocks, group_m;
};

template <int BLOCK_N, bool PIPE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kTc2Threads, 1)
gemm_tc2_blk_kernel(const Tc2BlkParams pp, const __grid_constant__ CUtensorMap map_a,
                    const __grid_constant__ CUtensorMap map_b) {
  using L = Tc2Smem<KIND_F8, BLOCK_N, 1>;
  constexpr int S = L::kStages;
  const TcParams& p = pp.b;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr int kOffB = L::kABytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S * L::kStageBytes);
  uint64_t* full_bar = bars;                 // leader CTA
  uint64_t* empty_bar = bars + S;            // per CTA
  // TMEM buffers that rotate per K slice: two 192-column ones, or THREE 160-column ones (a 256 x 160 tile): with three
  // the hand-over latency h of the analysis above leaves the critical path (period max(m, d, (m + d + h) / 3))
  constexpr int kNB = BLOCK_N <= 160 ? 3 : 2;
  uint64_t* p_full_bar = bars + 2 * S;             // per CTA: partial product of one K slice is in TMEM buffer i
  uint64_t* p_empty_bar = bars + 2 * S + kNB;      // leader CTA: buffer i has been pulled into registers by all 16 warps
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 2 * S + 2 * kNB);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < S; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < kNB; ++i) {
      mbar_init(&p_full_bar[i], 1);
      mbar_init(&p_empty_bar[i], 16);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc_pair<kTmemCols>(tmem_ptr_smem);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  constexpr uint32_t kAccCols = BLOCK_N;
  static_assert(kNB * kAccCols <= kTmemCols, "TMEM budget");

  const int num_tiles = pp.num_m_pairs * pp.num_n_blocks;
  const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
  const int KB = p.num_k_blocks;

  if (warp < 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
    if (warp == 0) {
      // ===================== TMA producer (both CTAs) =====================
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = pair; tile < num_tiles; tile += num_pairs) {
        int mt, nt;
        tile_coords(tile, pp.num_m_pairs, pp.num_n_blocks, pp.group_m, mt, nt);
        const int m0 = mt * (2 * kBlockM) + static_cast<int>(rank) * kBlockM;
        const int nb = nt * BLOCK_N + static_cast<int>(rank) * (BLOCK_N / 2);
        for (int kb = 0; kb < KB; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if (elect_one()) {
            uint8_t* st = smem + stage * L::kStageBytes;
            if (leader) mbar_arrive_expect_tx(&full_bar[stage], 2 * (L::kABytes + L::kBBytes));
            tma_load_2d_pair(st, &map_a, &full_bar[stage], kb * kBlockKBytes, m0);
            tma_load_2d_pair(st + kOffB, &map_b, &full_bar[stage], kb * kBlockKBytes, nb);
          }
          __syncwarp();
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
    } else if (warp == 1 && leader) {
      // ===================== MMA issuer (leader CTA) =====================
      const uint32_t idesc = make_idesc_f8(p.fp8_fmt, 2 * kBlockM, BLOCK_N);
      const uint32_t lo_base = desc_lo(smem);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t g = 0;   // K slices issued so far (all tiles)
      for (int tile = pair; tile < num_tiles; tile += num_pairs) {
        for (int kb = 0; kb < KB; ++kb, ++g) {
          const uint32_t buf = g % kNB;
          mbar_wait(&p_empty_bar[buf], ((g / kNB) & 1) ^ 1);
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t lo_a = lo_base + stage * (L::kStageBytes >> 4);
            const uint32_t lo_b = lo_a + (kOffB >> 4);
            const uint32_t tmem_d = tmem_base + buf * kAccCols;
#pragma unroll
            for (int k = 0; k < kBlockKBytes / kUmmaKBytes; ++k)
              umma2_issue<KIND_F8>(tmem_d, lo_a + k * 2, lo_b + k * 2, idesc, k > 0 ? 1u : 0u, 0, 0);
            umma_commit_pair(&empty_bar[stage]);
            umma_commit_pair(&p_full_bar[buf]);
          }
          __syncwarp();
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 224;");
    // ===================== promotion + epilogue warps (both CTAs) =====================
    const int e = warp - 4;
    const int q = e & 3;                    // TMEM lane quarter
    const int h = e >> 2;                   // column part of the tile: halves, or 96 + 64 of a 160-wide tile
    auto promote = [&](auto nch_tag) {
    constexpr int kCols = decltype(nch_tag)::value * 32;      // columns per thread
    const int col_off = BLOCK_N == 160 ? h * 96 : h * kCols;
    constexpr int CW = 32;                  // tcgen05.ld chunk (columns)
    constexpr int kChunks = kCols / CW;
    constexpr int kSc = kCols / 32;         // 32-column scale blocks per thread
    static_assert(kCols % 32 == 0, "column halves must be whole 32-column chunks");
    const bool dbg_idle = p.debug & 16;     // timing experiment: hand every buffer straight back, no promotion
    const int row_in_tile = static_cast<int>(rank) * kBlockM + q * 32 + lane;
    // fused all-gather: needs 16-byte aligned output rows and N % 8 == 0 (else direct peer stores)
    const bool row_copy = p.n_peers > 0 && p.vec_store_ok && p.N % 8 == 0;
    uint32_t g = 0;
    for (int tile = pair; tile < num_tiles; tile += num_pairs) {
      int mt, nt;
      tile_coords(tile, pp.num_m_pairs, pp.num_n_blocks, pp.group_m, mt, nt);
      const int m = mt * (2 * kBlockM) + row_in_tile;
      const int n_first = nt * BLOCK_N + col_off;
      const bool row_ok = m < p.M;
      const float* sa_row = pp.sa + static_cast<int64_t>(row_ok ? m : 0) * pp.sa0;
      // scale column groups of this thread's chunks (a 32-column chunk never straddles a group: group_n % 32 == 0)
      int ng[kSc];
#pragma unroll
      for (int c = 0; c < kSc; ++c) {
        const int n = n_first + c * 32;
        ng[c] = (n < p.N ? n : 0) / pp.group_n;
      }
      float acc[kCols];
#pragma unroll
      for (int j = 0; j < kCols; ++j) acc[j] = 0.0f;
      // Scales are fetched one slice ahead and kept RAW: the warp issues in order, so forming the
      // product sA * sW right after the loads would stall every slice on an L2 round trip (~600 clk,
      // measured with clock64 stamps); the product is formed when the slice is consumed.
      struct RawScales { float a; float b[kSc]; };
      auto load_scales = [&](int kb, RawScales& s) {
        const int kg = pp.group_k_blocks == 1 ? kb : kb / pp.group_k_blocks;
        s.a = row_ok ? __ldg(sa_row + static_cast<int64_t>(kg) * pp.sa1) : 0.0f;
#pragma unroll
        for (int c = 0; c < kSc; ++c) s.b[c] = __ldg(pp.sb + kg * pp.sb0 + ng[c] * pp.sb1);
      };
      RawScales s_next;
      load_scales(0, s_next);
      const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + col_off;
      if (dbg_idle) {
        for (int kb = 0; kb < KB; ++kb, ++g) {
          const uint32_t buf = g % kNB;
          mbar_wait(&p_full_bar[buf], (g / kNB) & 1);
          tc_fence_after();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(&p_empty_bar[buf], 0);
        }
      } else if constexpr (!PIPE) {
        // all chunks of a slice at once: shortest hand-back latency, no overlap of ld and FMAs inside the warp
        for (int kb = 0; kb < KB; ++kb, ++g) {
          const uint32_t buf = g % kNB;
          float s[kChunks];
          asm volatile("" : "+f"(s_next.a));
#pragma unroll
          for (int c = 0; c < kChunks; ++c) s[c] = s_next.a * s_next.b[c];
          mbar_wait(&p_full_bar[buf], (g / kNB) & 1);
          tc_fence_after();
          uint32_t r[kChunks][32];
#pragma unroll
          for (int c = 0; c < kChunks; ++c) tmem_ld_32x32b_x32(t_lane + buf * kAccCols + c * 32, r[c]);
          tmem_ld_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(&p_empty_bar[buf], 0);   // buffer is in registers: hand it back first
          if (kb + 1 < KB) load_scales(kb + 1, s_next);
#pragma unroll
          for (int c = 0; c < kChunks; ++c)
#pragma unroll
            for (int j = 0; j < 32; j += 2)
              fma2(acc[c * 32 + j], acc[c * 32 + j + 1], __uint_as_float(r[c][j]), __uint_as_float(r[c][j + 1]), s[c], s[c]);
        }
      } else {
        // Software pipeline over (K slice, 32-column chunk) units: while the FMAs of one chunk run, the
        // tcgen05.ld of the next chunk -- of the next slice after the last one -- is already in flight
        // (two rotating 32-register buffers; tcgen05.wait::ld has no finer granularity than "all").
        // A slice's TMEM buffer goes back to the MMA warp as soon as its last chunk has landed.
        uint32_t r[2][CW];
        mbar_wait(&p_full_bar[g % kNB], (g / kNB) & 1);
        tc_fence_after();
        tmem_ld_32x32b<CW>(t_lane + (g % kNB) * kAccCols, r[0]);
        auto slice = [&](auto kp_tag, int kb) {
          constexpr int KP = decltype(kp_tag)::value;      // parity of the slice index within the tile
          const uint32_t buf = g % kNB;
          float s[kSc];
          asm volatile("" : "+f"(s_next.a));
#pragma unroll
          for (int c = 0; c < kSc; ++c) s[c] = s_next.a * s_next.b[c];
          const bool more = kb + 1 < KB;
#pragma unroll
          for (int c = 0; c < kChunks; ++c) {
            constexpr int kOdd = kChunks & 1;              // chunks per slice odd: buffer roles flip every slice
            const int cur = (KP * kOdd + c) & 1;
            const float sc = s[c * CW / 32];
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < CW; ++j) asm volatile("" : "+r"(r[cur][j]));   // values are defined from here on
            bool issued = false;
            if (c + 1 < kChunks) {
              tmem_ld_32x32b<CW>(t_lane + buf * kAccCols + (c + 1) * CW, r[cur ^ 1]);
              issued = true;
            } else {
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive_cluster(&p_empty_bar[buf], 0);   // whole slice is in registers: hand it back
              if (more) {
                load_scales(kb + 1, s_next);
                const uint32_t gn = g + 1;
                if (__all_sync(0xffffffffu, mbar_try_wait(&p_full_bar[gn % kNB], (gn / kNB) & 1))) {
                  tc_fence_after();
                  tmem_ld_32x32b<CW>(t_lane + (gn % kNB) * kAccCols, r[cur ^ 1]);
                  issued = true;
                }
              }
            }
#pragma unroll
            for (int j = 0; j < CW; j += 2)
              fma2(acc[c * CW + j], acc[c * CW + j + 1], __uint_as_float(r[cur][j]), __uint_as_float(r[cur][j + 1]), sc, sc);
            if (c + 1 == kChunks && more && !issued) {     // next slice was not ready: wait for it now
              const uint32_t gn = g + 1;
              mbar_wait(&p_full_bar[gn % kNB], (gn / kNB) & 1);
              tc_fence_after();
              tmem_ld_32x32b<CW>(t_lane + (gn % kNB) * kAccCols, r[cur ^ 1]);
            }
          }
          ++g;
        };
        int kb = 0;
        for (; kb + 1 < KB; kb += 2) {
          slice(std::integral_constant<int, 0>{}, kb);
          slice(std::integral_constant<int, 1>{}, kb + 1);
        }
        if (kb < KB) slice(std::integral_constant<int, 0>{}, kb);
      }
      if (row_ok) {
#pragma unroll
        for (int c = 0; c < kSc; ++c) {
          const int nbase = n_first + c * 32;
          if (nbase < p.N) {
            uint32_t bits[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) bits[j] = __float_as_uint(acc[c * 32 + j]);
            epilogue_store_32(p, bits, m, nbase, 1.0f, false, !row_copy);
          }
        }
      }
      if (row_copy) {
        // fused all-gather (see gemm_tc2.cuh): push this CTA's 128 x BLOCK_N part of the finished tile to the
        // other ranks as whole 384-byte row segments re-read from the local output (L2)
        asm volatile("bar.sync 1, 256;" ::: "memory");
        const int mrow0 = mt * (2 * kBlockM) + static_cast<int>(rank) * kBlockM;
        const int cw = lane * 8;
        const int nc = nt * BLOCK_N + cw;
        if (cw < BLOCK_N && nc + 8 <= p.N) {
#pragma unroll 1
          for (int rr = e; rr < kBlockM; rr += 16) {
            const int ma = mrow0 + rr, mb = ma + 8;
            uint4 va = make_uint4(0, 0, 0, 0), vb = va;
            if (ma < p.M) va = ld_global_v4(static_cast<const uint16_t*>(p.D) + static_cast<int64_t>(ma) * p.ldd + nc);
            if (mb < p.M) vb = ld_global_v4(static_cast<const uint16_t*>(p.D) + static_cast<int64_t>(mb) * p.ldd + nc);
#pragma unroll 1
            for (int pe = 0; pe < p.n_peers; ++pe) {
              uint16_t* pd = static_cast<uint16_t*>(p.peer_D[pe]) + nc;
              if (ma < p.M) *reinterpret_cast<uint4*>(pd + static_cast<int64_t>(ma) * p.ldd) = va;
              if (mb < p.M) *reinterpret_cast<uint4*>(pd + static_cast<int64_t>(mb) * p.ldd) = vb;
            }
          }
        }
      }
    }
    };   // promote
    if constexpr (BLOCK_N == 160) {
      if (h == 0) promote(std::integral_constant<int, 3>{});
      else promote(std::integral_constant<int, 2>{});
    } else {
      promote(std::integral_constant<int, BLOCK_N / 64>{});
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_pair<kTmemCols>(tmem_base);
  }
}

}  // namespace vxk
