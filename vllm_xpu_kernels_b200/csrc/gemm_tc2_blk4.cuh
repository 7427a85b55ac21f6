// FP8 GEMM with fp32 block scales, second generation: four 128-column TMEM buffers per CTA pair and a
// register double buffer in the promotion warps.
//
//   for every K slice of 128:   P   = A[:, slice] * W[:, slice]^T        tensor core -> TMEM buffer g % 4
//                               acc += sA[m, slice] * sW[slice, n] * P    CUDA cores, fp32 registers
//
// What bounds this algorithm on sm_100 is the TMEM -> register path: every K slice the whole 128 x N fp32
// partial product of a CTA goes through tcgen05.ld (4 B per 256 flop, i.e. 256 B/clk/SM at the full FP8
// rate, against a drain rate of ~340 B/clk/SM measured with an idle tensor pipe, benchmarks/tmem_bw.cu).
// gemm_tc2_blk.cuh (two 192-column buffers) keeps one tcgen05.ld per warp in flight and reaches ~130 B/clk;
// a 16-warp variant with one 32-column block per warp measured worse still (every warp handles every slice,
// so the per-warp ld -> wait -> arrive -> FMA chain, not the warp count, sets the slice rate:
// profiles/r02_blk_gen_sweep.txt).  Here
//   * the TMEM holds FOUR partial products (4 x 128 columns = all 512): the MMA warp runs up to three
//     slices ahead of the drain, so a hand-over round trip never idles the tensor pipe;
//   * each of the eight promotion warps owns 32 lanes x 64 columns and TWO 64-register staging buffers:
//     the tcgen05.ld of slice g + 1 is issued before the FMAs of slice g, so a whole slice (64 KB per CTA)
//     is always in flight on the TMEM read path.
// Tile per pair: 256 x 128 (UMMA 256 x 128 x 32, 256 clk of tensor work per K slice).
// ref: fp8_gemm_w8a8.h:28-41,86-134 (block scale shapes), tests/test_fp8_gemm_onednn.py:270-358.
#pragma once

#include "gemm_tc2_blk.cuh"

namespace vxk {

constexpr int kBlk4Threads = kTc2Threads;   // 4 role warps + 8 promotion warps
constexpr int kBlk4N = 128;
constexpr int kBlk4Bufs = 4;

struct Blk4Smem {
  static constexpr int kABytes = kBlockM * kBlockKBytes;             // this CTA's 128 rows of A
  static constexpr int kBBytes = (kBlk4N / 2) * kBlockKBytes;        // this CTA's half of the W tile
  static constexpr int kStageBytes = kABytes + kBBytes;              // 24 KB, a multiple of 1024
  static constexpr int kStages = 8;
  static constexpr int kBarriers = 2 * kStages + 2 * kBlk4Bufs;
  static constexpr int kTotalBytes = kStages * kStageBytes + kBarriers * 8 + 16 + 1024;
  static_assert(kStageBytes % 1024 == 0, "stage must keep 1024 B alignment for SWIZZLE_128B");
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kBlk4Threads, 1)
gemm_tc2_blk4_kernel(const Tc2BlkParams pp, const __grid_constant__ CUtensorMap map_a,
                     const __grid_constant__ CUtensorMap map_b) {
  using L = Blk4Smem;
  constexpr int S = L::kStages;
  constexpr int NB = kBlk4Bufs;
  constexpr int BLOCK_N = kBlk4N;
  const TcParams& p = pp.b;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr int kOffB = L::kABytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S * L::kStageBytes);
  uint64_t* full_bar = bars;                   // leader CTA
  uint64_t* empty_bar = bars + S;              // per CTA
  uint64_t* p_full_bar = bars + 2 * S;         // per CTA: partial product of one K slice is in TMEM buffer i
  uint64_t* p_empty_bar = bars + 2 * S + NB;   // leader CTA: buffer i has been pulled into registers by all 16 warps
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 2 * S + 2 * NB);

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
    for (int i = 0; i < NB; ++i) {
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
  static_assert(NB * BLOCK_N <= kTmemCols, "TMEM budget");

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
            if (leader) mbar_arrive_expect_tx(&full_bar[stage], 2 * L::kStageBytes);
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
          const uint32_t buf = g % NB;
          mbar_wait(&p_empty_bar[buf], ((g / NB) & 1) ^ 1);
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t lo_a = lo_base + stage * (L::kStageBytes >> 4);
            const uint32_t lo_b = lo_a + (kOffB >> 4);
            const uint32_t tmem_d = tmem_base + buf * BLOCK_N;
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
    const int q = e & 3;                    // TMEM lane quarter (warp % 4)
    const int h = e >> 2;                   // column half of the tile
    constexpr int kCols = BLOCK_N / 2;      // 64 columns per thread
    const int row_in_tile = static_cast<int>(rank) * kBlockM + q * 32 + lane;
    // fused all-gather: needs 16-byte aligned output rows and N % 8 == 0 (else direct peer stores)
    const bool row_copy = p.n_peers > 0 && p.vec_store_ok && p.N % 8 == 0;
    const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + h * kCols;
    uint32_t g = 0;
    // the first slice of this CTA's first tile: in flight before any scale arrives
    uint32_t ra[kCols], rb[kCols];
    if (pair < num_tiles) {
      mbar_wait(&p_full_bar[0], 0);
      tc_fence_after();
      tmem_ld_32x32b_x32(t_lane, *reinterpret_cast<uint32_t(*)[32]>(&ra[0]));
      tmem_ld_32x32b_x32(t_lane + 32, *reinterpret_cast<uint32_t(*)[32]>(&ra[32]));
    }
    for (int tile = pair; tile < num_tiles; tile += num_pairs) {
      int mt, nt;
      tile_coords(tile, pp.num_m_pairs, pp.num_n_blocks, pp.group_m, mt, nt);
      const int m = mt * (2 * kBlockM) + row_in_tile;
      const int n_first = nt * BLOCK_N + h * kCols;
      const bool row_ok = m < p.M;
      const bool last_tile = tile + num_pairs >= num_tiles;
      // a 32-column chunk never straddles a scale group (group_n % 32 == 0)
      const int ng0 = (n_first < p.N ? n_first : 0) / pp.group_n;
      const int ng1 = (n_first + 32 < p.N ? n_first + 32 : 0) / pp.group_n;
      const float* sa_row = pp.sa + static_cast<int64_t>(row_ok ? m : 0) * pp.sa0;
      const float* sb_c0 = pp.sb + static_cast<int64_t>(ng0) * pp.sb1;
      const float* sb_c1 = pp.sb + static_cast<int64_t>(ng1) * pp.sb1;
      float acc[kCols];
#pragma unroll
      for (int j = 0; j < kCols; ++j) acc[j] = 0.0f;
      // raw scales two slices ahead (an L2 round trip is longer than one slice of tensor work); the product
      // sA * sW is formed when the slice is consumed so that the loads never stall the in-order issue
      struct Raw { float a, b0, b1; };
      auto load_scales = [&](int kb) {
        const int64_t kg = pp.group_k_blocks == 1 ? kb : kb / pp.group_k_blocks;
        Raw r;
        r.a = __ldg(sa_row + kg * pp.sa1);
        r.b0 = __ldg(sb_c0 + kg * pp.sb0);
        r.b1 = __ldg(sb_c1 + kg * pp.sb0);
        return r;
      };
      Raw s0 = load_scales(0), s1 = load_scales(KB > 1 ? 1 : 0);
      // one K slice: `cur` holds (or is about to hold) this slice, `nxt` receives the next one while the FMAs run
      auto slice = [&](uint32_t (&cur)[kCols], uint32_t (&nxt)[kCols], int kb) {
        const uint32_t buf = g % NB;
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < kCols; ++j) asm volatile("" : "+r"(cur[j]));   // values are defined from here on
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(&p_empty_bar[buf], 0);   // slice is in registers: hand the buffer back
        const bool more = kb + 1 < KB || !last_tile;                 // the next slice may belong to the next tile
        bool issued = false;
        const uint32_t gn = g + 1;
        if (more && __all_sync(0xffffffffu, mbar_try_wait(&p_full_bar[gn % NB], (gn / NB) & 1))) {
          tc_fence_after();
          tmem_ld_32x32b_x32(t_lane + (gn % NB) * BLOCK_N, *reinterpret_cast<uint32_t(*)[32]>(&nxt[0]));
          tmem_ld_32x32b_x32(t_lane + (gn % NB) * BLOCK_N + 32, *reinterpret_cast<uint32_t(*)[32]>(&nxt[32]));
          issued = true;
        }
        const float c0 = s0.a * s0.b0, c1 = s0.a * s0.b1;
        s0 = s1;
        s1 = load_scales(kb + 2 < KB ? kb + 2 : KB - 1);
#pragma unroll
        for (int j = 0; j < 32; j += 2)
          fma2(acc[j], acc[j + 1], __uint_as_float(cur[j]), __uint_as_float(cur[j + 1]), c0, c0);
#pragma unroll
        for (int j = 32; j < 64; j += 2)
          fma2(acc[j], acc[j + 1], __uint_as_float(cur[j]), __uint_as_float(cur[j + 1]), c1, c1);
        if (more && !issued) {     // next slice was not ready: wait for it now
          mbar_wait(&p_full_bar[gn % NB], (gn / NB) & 1);
          tc_fence_after();
          tmem_ld_32x32b_x32(t_lane + (gn % NB) * BLOCK_N, *reinterpret_cast<uint32_t(*)[32]>(&nxt[0]));
          tmem_ld_32x32b_x32(t_lane + (gn % NB) * BLOCK_N + 32, *reinterpret_cast<uint32_t(*)[32]>(&nxt[32]));
        }
        ++g;
      };
      // register buffer roles alternate per slice; g parity (not kb parity) says which one holds the current slice
      for (int kb = 0; kb < KB; ++kb) {
        if ((g & 1) == 0) slice(ra, rb, kb);
        else slice(rb, ra, kb);
      }
      if (row_ok) {
#pragma unroll
        for (int c = 0; c < 2; ++c) {
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
        // fused all-gather (see gemm_tc2.cuh): push this CTA's 128 x 128 part of the finished tile to the other
        // ranks as whole 256-byte row segments re-read from the local output (L2); half a warp per row
        asm volatile("bar.sync 1, 256;" ::: "memory");
        const int mrow0 = mt * (2 * kBlockM) + static_cast<int>(rank) * kBlockM;
        const int nc = nt * BLOCK_N + (lane & 15) * 8;
        if (nc + 8 <= p.N) {
#pragma unroll 1
          for (int rr = e * 2 + (lane >> 4); rr < kBlockM; rr += 16) {
            const int ma = mrow0 + rr;
            if (ma < p.M) {
              const uint4 va = ld_global_v4(static_cast<const uint16_t*>(p.D) + static_cast<int64_t>(ma) * p.ldd + nc);
#pragma unroll 1
              for (int pe = 0; pe < p.n_peers; ++pe)
                *reinterpret_cast<uint4*>(static_cast<uint16_t*>(p.peer_D[pe]) + static_cast<int64_t>(ma) * p.ldd + nc) = va;
            }
          }
        }
      }
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
