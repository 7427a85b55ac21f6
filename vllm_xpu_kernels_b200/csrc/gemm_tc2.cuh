// tcgen05 scaled-GEMM kernel for sm_100a, CTA-pair version (cta_group::2).
//
//   D[M,N] = epilogue( sum_k A[m,k] * W[n,k] )      A, W K-major (row = m / n, K contiguous)
//
// Two CTAs on the two SMs of a TPC (cluster 2x1x1) work on one 256 x BLOCK_N output tile:
// CTA r holds A rows [m0 + 128 r, +128) and W rows [n0 + r BLOCK_N/2, + BLOCK_N/2) of every K
// slice in its shared memory; the pair's tensor cores read both halves of W, so each operand
// byte is fetched from L2 once per 256 x BLOCK_N tile (1.5x less L2->SM traffic than the 128 x
// 256 single-CTA tile) and each SM's shared memory feeds half of the W reads.  CTA r keeps the
// fp32 accumulator of its 128 rows x BLOCK_N columns in its own TMEM (two buffers, so the
// epilogue of tile i overlaps the main loop of tile i+1).
//
// Per CTA, 256 threads, warp-specialised:
//   warp 0      TMA producer (both CTAs): own A / W halves (SWIZZLE_128B boxes) and, for the
//               block-scaled kinds, the scale-factor chunks (already in tcgen05.cp layout, see
//               sf_pack_kernel) into a STAGES-deep ring; every transaction signals the LEADER
//               CTA's full barrier (cta_group::2 TMA, peer bit cleared)
//   warp 1      MMA issuer (leader CTA only): one thread issues tcgen05.cp (scale factors ->
//               TMEM of both CTAs) and tcgen05.mma.cta_group::2 (UMMA 256 x BLOCK_N x 32 B);
//               tcgen05.commit multicasts the "slot free" / "accumulator full" arrivals to both
//   warp 2      TMEM allocator
//   warps 4-11  epilogue (both CTAs): tcgen05.ld own accumulator rows (two warps per TMEM lane
//               quarter, each half of the columns), scales / bias, bf16 / f16 stores; then one
//               arrival per warp on the leader's "accumulator empty" barrier
// The producer and MMA warps run their loops with all 32 lanes converged and elect one lane per
// instruction group: in a lane-0-only branch the compiler cannot prove the tcgen05 / TMA operands
// warp-uniform and wraps every instruction in an elect / broadcast loop -- the single issuing
// thread then takes longer per K slice than the tensor core does.
#pragma once

#include "gemm_tc.cuh"
#include "quant_silu.cuh"

namespace vxk {

// NACC = 2: the pair works on a 256 x (2 BLOCK_N) super-tile, both TMEM buffers are accumulators
// of the same K loop (A is fetched once for two W tiles: another 25 % off the L2->SM bytes per
// flop, which is what bounds these kernels -- B200's L2 delivers ~6.3 KB/clk chip-wide); the
// epilogue then overlaps only the first K slices of the next tile.  NACC = 1: one 256 x BLOCK_N
// tile per K loop, the two TMEM buffers alternate between tiles (full epilogue overlap).
template <int KIND, int BLOCK_N, int NACC>
struct Tc2Smem {
  static constexpr bool kScaled = KIND != KIND_F8;
  static constexpr int kSfGroups = KIND == KIND_MXF4 ? 8 : 4;        // 32-element groups per stage
  static constexpr int kSfWords = kSfGroups / 4;                     // 32-bit SF words per row per stage
  static constexpr int kABytes = kBlockM * kBlockKBytes;             // this CTA's 128 rows of A
  static constexpr int kBHalfBytes = (BLOCK_N / 2) * kBlockKBytes;   // this CTA's half of one W tile
  static constexpr int kBBytes = NACC * kBHalfBytes;
  static constexpr int kSfbBlocks = NACC == 2 ? (2 * BLOCK_N + 127) / 128 : 2;   // 128-row blocks: any 192 / 256 window needs 2
  static constexpr int kSfaBytes = kScaled ? kSfWords * 512 : 0;     // one 128-row block
  static constexpr int kSfbBytes = kScaled ? kSfbBlocks * kSfWords * 512 : 0;
  static constexpr int kStageBytesRaw = kABytes + kBBytes + kSfaBytes + kSfbBytes;
  static constexpr int kStageBytes = (kStageBytesRaw + 1023) / 1024 * 1024;
  static constexpr int kTxBytes = kStageBytesRaw;                    // per CTA
  static constexpr int kStages = (220 * 1024) / kStageBytes > 8 ? 8 : (220 * 1024) / kStageBytes;
  static constexpr int kBarrierBytes = (2 * kStages + 4) * 8 + 16;
  static constexpr int kTotalBytes = kStages * kStageBytes + kBarrierBytes + 1024;  // + alignment slack
  static_assert(kStageBytes % 1024 == 0, "stage must keep 1024 B alignment for SWIZZLE_128B");
  static_assert(BLOCK_N % 32 == 0 && BLOCK_N <= 256, "UMMA N for M = 256");
};

// ---- cluster / peer helpers ----
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// plain (coherent, L1-allocating) 128-bit global load the compiler may not turn into a read-only-path load
__device__ __forceinline__ uint4 ld_global_v4(const void* ptr) {
  uint4 v;
  asm volatile("ld.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(ptr) : "memory");
  return v;
}
// 128-bit store through an NVLS multicast mapping: the NVSwitch replicates it into every rank's buffer
__device__ __forceinline__ void multimem_st_v4(void* mc_ptr, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_ptr), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;   // shared::cluster address -> same offset in the pair's even CTA

__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1)
      : "memory");
}
// arrive on the barrier at the same shared-memory offset in CTA `rank` of the cluster.  Default
// (CTA-scope) semantics on purpose: the arrival only says "this warp's tcgen05.ld of the buffer has
// completed" (ordered by tcgen05.wait::ld + tcgen05.fence::before_thread_sync); cluster scope
// compiles to MEMBAR.ALL.GPU and holds the signal back until the warp's global accesses drain.
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t* bar, uint32_t rank) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\tmapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t}" ::"r"(smem_u32(bar)),
      "r"(rank)
      : "memory");
}
template <uint32_t COLS>
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* dst_smem) {  // one whole warp in EACH CTA of the pair
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "n"(COLS)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <uint32_t COLS>
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS) : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {   // arrives on `bar` in both CTAs
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"(static_cast<uint16_t>(3))
      : "memory");
}
__device__ __forceinline__ void umma2_f8f6f4(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma2_mxf8f6f4(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                               uint32_t accumulate, uint32_t tmem_sfa, uint32_t tmem_sfb) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::mxf8f6f4.block_scale [%0], %1, %2, %3, [%5], [%6], p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(tmem_sfa), "r"(tmem_sfb)
      : "memory");
}
__device__ __forceinline__ void umma2_mxf4(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                           uint32_t accumulate, uint32_t tmem_sfa, uint32_t tmem_sfb) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::mxf4.block_scale.scale_vec::2X [%0], %1, %2, %3, [%5], [%6], p;\n\t}" ::"r"(
          tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(tmem_sfa), "r"(tmem_sfb)
      : "memory");
}
__device__ __forceinline__ void utccp2_32x128b_warpx4(uint32_t tmem_dst, uint64_t smem_desc) {
  asm volatile("tcgen05.cp.cta_group::2.32x128b.warpx4 [%0], %1;" ::"r"(tmem_dst), "l"(smem_desc) : "memory");
}

// ---- descriptors as (lo, hi) halves: hi is a compile-time constant, lo = (smem address >> 4) ----
constexpr uint32_t kDescHiSw128 = (1024u >> 4) | (1u << 14) | (2u << 29);  // SBO 1024 B, version 1, SWIZZLE_128B
constexpr uint32_t kDescHiSf = (128u >> 4) | (1u << 14);                   // SBO 128 B, version 1, no swizzle
__device__ __forceinline__ uint32_t desc_lo(const void* smem_ptr) { return (smem_u32(smem_ptr) >> 4) & 0x3fffu; }

template <int KIND>
__device__ __forceinline__ void umma2_issue(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t idesc,
                                            uint32_t accumulate, uint32_t tmem_sfa, uint32_t tmem_sfb) {
  if constexpr (KIND == KIND_F8) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "mov.b64 da, {%1, %5};\n\tmov.b64 db, {%2, %5};\n\t"
        "tcgen05.mma.cta_group::2.kind::f8f6f4 [%0], da, db, %3, p;\n\t}" ::"r"(tmem_d),
        "r"(a_lo), "r"(b_lo), "r"(idesc), "r"(accumulate), "n"(kDescHiSw128)
        : "memory");
  } else if constexpr (KIND == KIND_MXF8) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "mov.b64 da, {%1, %7};\n\tmov.b64 db, {%2, %7};\n\t"
        "tcgen05.mma.cta_group::2.kind::mxf8f6f4.block_scale [%0], da, db, %3, [%5], [%6], p;\n\t}" ::"r"(tmem_d),
        "r"(a_lo), "r"(b_lo), "r"(idesc), "r"(accumulate), "r"(tmem_sfa), "r"(tmem_sfb), "n"(kDescHiSw128)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "mov.b64 da, {%1, %7};\n\tmov.b64 db, {%2, %7};\n\t"
        "tcgen05.mma.cta_group::2.kind::mxf4.block_scale.scale_vec::2X [%0], da, db, %3, [%5], [%6], p;\n\t}" ::"r"(
            tmem_d),
        "r"(a_lo), "r"(b_lo), "r"(idesc), "r"(accumulate), "r"(tmem_sfa), "r"(tmem_sfb), "n"(kDescHiSw128)
        : "memory");
  }
}
__device__ __forceinline__ void utccp2_lo(uint32_t tmem_dst, uint32_t sf_lo) {
  asm volatile(
      "{\n\t.reg .b64 d;\n\tmov.b64 d, {%1, %2};\n\t"
      "tcgen05.cp.cta_group::2.32x128b.warpx4 [%0], d;\n\t}" ::"r"(tmem_dst),
      "r"(sf_lo), "n"(kDescHiSf)
      : "memory");
}

struct Tc2Params {
  TcParams b;             // shared fields (D, scales, bias, M, N, K_bytes, ldd, modes, dtypes, num_k_blocks)
  int num_m_pairs;        // ceil(M / 256)
  int num_n_blocks;       // ceil(N / (NACC * BLOCK_N))
  int sfa_blocks;         // 128-row blocks per K slice in the packed SFA / SFB workspaces
  int sfb_blocks;
  int group_m;            // tile rasterisation: GROUP_M row-tiles share their column tiles back to back
};

constexpr int kTc2Threads = 384;   // 4 role warps + 8 epilogue warps

// tile index -> (m pair-tile, n tile): groups of `group_m` row tiles sweep all column tiles, so
// the tiles in flight at any time (one per SM pair) share a few A panels and W panels in L2
__device__ __forceinline__ void tile_coords(int tile, int num_m, int num_n, int group_m, int& mt, int& nt) {
  const int per_group = group_m * num_n;
  const int g = tile / per_group;
  const int first_m = g * group_m;
  const int rows = min(group_m, num_m - first_m);
  const int r = tile - g * per_group;
  mt = first_m + r % rows;
  nt = r / rows;
}

// FUSE (KIND_F8, 256-wide tiles, NACC = 1 only): 0 = plain GEMM; 1 / 2 = the gate_up projection with silu_and_mul and the
// 1 x 128 block FP8 quantisation in the epilogue (bf16 / f16 intermediate type).  The pair's tile is 128 gate columns
// (CTA 0's half of W: rows [g0, g0 + 128)) next to the 128 up columns they multiply (CTA 1's half: rows
// [inter + g0, + 128)), so every epilogue thread holds whole (gate, up) pairs of its output row: it rounds both to the
// intermediate type exactly as the plain epilogue would have stored them, runs the SAME prologue / division helpers
// as silu_and_mul_per_block_quant (quant_silu.cuh) and stores 64 e4m3 codes; the two warps that share a row exchange
// their 64-column maxima through shared memory.  Bit-identical to fp8_gemm followed by silu_and_mul_per_block_quant,
// without the [M, 2 inter] 16-bit round trip through HBM.
template <int KIND, int BLOCK_N, int NACC, int FUSE = 0>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kTc2Threads, 1)
gemm_tc2_kernel(const Tc2Params pp, const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                const __grid_constant__ CUtensorMap map_sfa, const __grid_constant__ CUtensorMap map_sfb) {
  using L = Tc2Smem<KIND, BLOCK_N, NACC>;
  constexpr int S = L::kStages;
  constexpr bool kScaled = L::kScaled;
  constexpr int kTileN = NACC * BLOCK_N;
  const TcParams& p = pp.b;
  extern __shared__ uint8_t smem_raw[];
  // both CTAs place every buffer at the same shared-memory offset (the pair's MMA and the
  // peer-bit barrier addressing rely on it): the dynamic segment starts at the same offset in both
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr int kOffB = L::kABytes, kOffSfa = kOffB + L::kBBytes, kOffSfb = kOffSfa + L::kSfaBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S * L::kStageBytes);
  uint64_t* full_bar = bars;                    // used in the leader CTA only
  uint64_t* empty_bar = bars + S;               // per CTA
  uint64_t* tmem_full_bar = bars + 2 * S;       // per CTA ([1] unused when NACC == 2)
  uint64_t* tmem_empty_bar = bars + 2 * S + 2;  // used in the leader CTA only
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 2 * S + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;

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
      mbar_init(&full_bar[i], 1);     // the leader producer's arrive.expect_tx (bytes of both CTAs)
      mbar_init(&empty_bar[i], 1);    // multicast tcgen05.commit
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full_bar[i], 1);                    // multicast tcgen05.commit
      mbar_init(&tmem_empty_bar[i], NACC == 2 ? 8 : 16);  // epilogue warps of both CTAs that drain buffer i
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc_pair<kTmemCols>(tmem_ptr_smem);
  tc_fence_before();
  cluster_sync_all();   // barriers initialised and TMEM allocated in both CTAs before any cross-CTA signal
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  constexpr uint32_t kAccCols = BLOCK_N;                       // fp32 accumulator columns per buffer
  constexpr uint32_t kSfaCol = 2 * kAccCols;                   // SFA: kSfWords * 4 columns (8 reserved)
  constexpr uint32_t kSfbCol = kSfaCol + 8;                    // SFB: per word kSfbBlocks * 4 columns
  static_assert(!kScaled || kSfbCol + L::kSfWords * L::kSfbBlocks * 4 <= kTmemCols, "TMEM budget");
  static_assert(2 * kAccCols <= kTmemCols, "TMEM budget");

  const int num_tiles = pp.num_m_pairs * pp.num_n_blocks;
  const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs, whole warp converged) =====================
    int stage = 0;
    uint32_t phase = 0;
    const bool sf_loads = kScaled && !(p.debug & 8);
    const uint32_t tx_bytes = 2 * (sf_loads ? L::kTxBytes : L::kABytes + L::kBBytes);
    for (int tile = pair; tile < num_tiles; tile += num_pairs) {
      int mt, nt;
      tile_coords(tile, pp.num_m_pairs, pp.num_n_blocks, pp.group_m, mt, nt);
      const int m0 = mt * (2 * kBlockM) + static_cast<int>(rank) * kBlockM;
      const int n0 = nt * kTileN;
      const int nb = FUSE ? nt * (BLOCK_N / 2) + static_cast<int>(rank) * p.fuse_inter
                          : n0 + static_cast<int>(rank) * (BLOCK_N / 2);
      for (int kb = 0; kb < p.num_k_blocks; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        if (elect_one()) {
          uint8_t* st = smem + stage * L::kStageBytes;
          if (leader) mbar_arrive_expect_tx(&full_bar[stage], tx_bytes);
          tma_load_2d_pair(st, &map_a, &full_bar[stage], kb * kBlockKBytes, m0);
#pragma unroll
          for (int a = 0; a < NACC; ++a)
            tma_load_2d_pair(st + kOffB + a * L::kBHalfBytes, &map_b, &full_bar[stage], kb * kBlockKBytes,
                             nb + a * BLOCK_N);
          if (sf_loads) {
            tma_load_2d_pair(st + kOffSfa, &map_sfa, &full_bar[stage], 0,
                             (kb * pp.sfa_blocks + (m0 >> 7)) * L::kSfWords);
            tma_load_2d_pair(st + kOffSfb, &map_sfb, &full_bar[stage], 0,
                             (kb * pp.sfb_blocks + (n0 >> 7)) * L::kSfWords);
          }
        }
        __syncwarp();
        if (++stage == S) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA, whole warp converged) =====================
    if (leader) {
      const uint32_t idesc0 = KIND == KIND_F8 ? make_idesc_f8(p.fp8_fmt, 2 * kBlockM, BLOCK_N)
                                              : make_idesc_block_scaled(KIND == KIND_MXF4 ? 1 : p.fp8_fmt, 2 * kBlockM, BLOCK_N);
      const uint32_t lo_base = desc_lo(smem);
      const bool do_cp = kScaled && !(p.debug & 1);
      int stage = 0;
      uint32_t phase = 0;
      int iter = 0;
      for (int tile = pair; tile < num_tiles; tile += num_pairs, ++iter) {
        int mt, nt;
        tile_coords(tile, pp.num_m_pairs, pp.num_n_blocks, pp.group_m, mt, nt);
        // NACC == 1: tile `iter` accumulates into buffer iter & 1; NACC == 2: into both
        const int as = NACC == 2 ? 0 : (iter & 1);
        const uint32_t aphase = NACC == 2 ? (iter & 1) : ((iter >> 1) & 1);
        const uint32_t sfb_shift = static_cast<uint32_t>((nt * kTileN) & 127) >> 5;   // tile starts inside a 128-row SF block
        mbar_wait(&tmem_empty_bar[as], aphase ^ 1);
        tc_fence_after();
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t lo_a = lo_base + stage * (L::kStageBytes >> 4);
          const uint32_t accf = kb > 0 ? 1u : 0u;
#pragma unroll
          for (int a = 0; a < NACC; ++a) {
            if (NACC == 2 && a == 1 && kb == 0) {     // second accumulator: drained a little later
              mbar_wait(&tmem_empty_bar[1], aphase ^ 1);
              tc_fence_after();
            }
            if (elect_one()) {
              if (a == 0 && do_cp) {
#pragma unroll
                for (int w = 0; w < L::kSfWords; ++w) {
                  utccp2_lo(tmem_base + kSfaCol + w * 4, lo_a + ((kOffSfa + w * 512) >> 4));
#pragma unroll
                  for (int j = 0; j < L::kSfbBlocks; ++j)
                    utccp2_lo(tmem_base + kSfbCol + (w * L::kSfbBlocks + j) * 4,
                              lo_a + ((kOffSfb + (j * L::kSfWords + w) * 512) >> 4));
                }
              }
              const uint32_t tmem_d = tmem_base + (as + a) * kAccCols;
              const uint32_t lo_b = lo_a + ((kOffB + a * L::kBHalfBytes) >> 4);
              // SFB columns of this accumulator's W tile: 32 rows per column
              const uint32_t sfb_a = sfb_shift + a * (BLOCK_N / 32);
#pragma unroll
              for (int k = 0; k < kBlockKBytes / kUmmaKBytes; ++k) {
                const uint32_t acc = (k > 0) ? 1u : accf;
                if constexpr (KIND == KIND_F8) {
                  umma2_issue<KIND>(tmem_d, lo_a + k * 2, lo_b + k * 2, idesc0, acc, 0, 0);
                } else if constexpr (KIND == KIND_MXF8) {
                  umma2_issue<KIND>(tmem_d, lo_a + k * 2, lo_b + k * 2, idesc_with_sf_ids(idesc0, k, k), acc,
                                    tmem_base + kSfaCol, tmem_base + kSfbCol + sfb_a);
                } else {
                  const uint32_t w = k >> 1, id = (k & 1) * 2;
                  umma2_issue<KIND>(tmem_d, lo_a + k * 2, lo_b + k * 2, idesc_with_sf_ids(idesc0, id, id), acc,
                                    tmem_base + kSfaCol + w * 4, tmem_base + kSfbCol + w * L::kSfbBlocks * 4 + sfb_a);
                }
              }
              if (a == NACC - 1) {
                umma_commit_pair(&empty_bar[stage]);          // frees the ring slot in both CTAs when the MMAs retire
                if (kb == p.num_k_blocks - 1) umma_commit_pair(&tmem_full_bar[as]);
              }
            }
            __syncwarp();
          }
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue (both CTAs) =====================
    const int e = warp - 4;
    const int q = e & 3;                    // TMEM lane quarter this warp may access (warp % 4)
    const int h = e >> 2;                   // NACC == 2: accumulator; NACC == 1: column half
    const int row_in_tile = static_cast<int>(rank) * kBlockM + q * 32 + lane;
    constexpr int kChunks = NACC == 2 ? BLOCK_N / 32 : BLOCK_N / 64;
    const bool row_copy = p.n_peers > 0 && p.vec_store_ok;   // else (unaligned output rows): direct peer stores
    const bool mcast = p.peer_multicast != 0;
    int iter = 0;
    for (int tile = pair; tile < num_tiles; tile += num_pairs, ++iter) {
      int mt, nt;
      tile_coords(tile, pp.num_m_pairs, pp.num_n_blocks, pp.group_m, mt, nt);
      const int as = NACC == 2 ? 0 : (iter & 1);
      const uint32_t aphase = NACC == 2 ? (iter & 1) : ((iter >> 1) & 1);
      const int m = mt * (2 * kBlockM) + row_in_tile;
      const int n0 = nt * kTileN;
      mbar_wait(&tmem_full_bar[as], aphase);
      tc_fence_after();
      float row_scale = 1.0f;
      if constexpr (KIND == KIND_F8) {
        row_scale = (p.a_scale_mode == 1) ? (m < p.M ? __ldg(p.a_scale + m) : 0.0f) : __ldg(p.a_scale);
        if (p.b_scale_mode == 0) row_scale *= __ldg(p.b_scale);
      }
      const bool per_col = (KIND == KIND_F8) && (p.b_scale_mode == 1);
      if constexpr (FUSE != 0) {
        static_assert(FUSE == 0 || (KIND == KIND_F8 && BLOCK_N == 256 && NACC == 1), "fused epilogue: FP8, 256-wide, NACC 1");
        using T = std::conditional_t<FUSE == 1, __nv_bfloat16, __half>;
        constexpr int kFast = FUSE == 1 ? 1 : 0;      // the silu auto mode of quant_flat.cu: bf16 approximate (guarded), f16 exact
        __shared__ float xchg[2][2][kBlockM];          // [tile parity][column half][row of this CTA]
        const int g0 = nt * 128;                       // first of this tile's 128 output columns
        const int rl = q * 32 + lane;                  // row within this CTA
        // this warp: output columns [g0 + 64 h, + 64): gate accumulator columns 64 h .., up columns 128 + 64 h ..
        uint32_t gw[2][16], uw[2][16];
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t r[32];
          const uint32_t tbase = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * kAccCols + h * 64 + c * 32;
          tmem_ld_32x32b_x32(tbase, r);
          tmem_ld_wait();
          epilogue_pack_32(p, r, g0 + h * 64 + c * 32, row_scale, per_col, gw[c]);
          tmem_ld_32x32b_x32(tbase + 128, r);
          tmem_ld_wait();
          epilogue_pack_32(p, r, p.fuse_inter + g0 + h * 64 + c * 32, row_scale, per_col, uw[c]);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(&tmem_empty_bar[as], 0);     // accumulator handed back before the arithmetic
        uint32_t yw[8][4];
        float gmax = 0.0f;
#pragma unroll
        for (int vv = 0; vv < 8; ++vv) {
          const int c = vv >> 2, o = (vv & 3) * 4;
          RawV<T> gr, ur;
          gr.a = make_uint4(gw[c][o], gw[c][o + 1], gw[c][o + 2], gw[c][o + 3]);
          ur.a = make_uint4(uw[c][o], uw[c][o + 1], uw[c][o + 2], uw[c][o + 3]);
          float v[8];
          prologue<T, PRO_SILU_T, kFast>(gr, ur, v, yw[vv]);
#pragma unroll
          for (int k = 0; k < 8; ++k) gmax = fmaxf(gmax, fabsf(v[k]));
        }
        xchg[iter & 1][h][rl] = gmax;
        asm volatile("bar.sync 1, 256;" ::: "memory");
        gmax = fmaxf(gmax, xchg[iter & 1][h ^ 1][rl]);
        // silu_and_mul_per_block_quant (quant_flat.cu, PRO_SILU_T / FMT_FP8_DIV, e4m3): s = max(gmax, FLT_MIN) / 448
        const float sc = div_qmax<VXK_Q_E4M3>(fmaxf(gmax, 1.17549435e-38f));
        const float rc = __frcp_rn(sc);
        if (m < p.M) {
          if (h == 0) p.fuse_s[static_cast<int64_t>(m) * p.fuse_ss0 + static_cast<int64_t>(nt) * p.fuse_ss1] = sc;
          uint8_t* orow = static_cast<uint8_t*>(p.fuse_q) + static_cast<int64_t>(m) * p.fuse_inter + g0 + h * 64;
#pragma unroll
          for (int vv = 0; vv < 8; vv += 2) {
            uint2 pk[2];
#pragma unroll
            for (int t2 = 0; t2 < 2; ++t2) {
              float v[8];
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const float2 f = Pair<T>::unpack(yw[vv + t2][k]);
                v[2 * k] = f.x;
                v[2 * k + 1] = f.y;
              }
              if (__builtin_expect(!(rc <= 3.402823466e+38f), 0)) {
                float qv[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) qv[k] = slow_div(v[k], sc);
                pk[t2] = pack_q8<VXK_Q_E4M3>(qv);
              } else {
                pk[t2] = div_pack_fp8_signed<VXK_Q_E4M3>(v, yw[vv + t2], sc, rc);
              }
            }
            *reinterpret_cast<uint4*>(orow + vv * 8) = make_uint4(pk[0].x, pk[0].y, pk[1].x, pk[1].y);
          }
        }
        continue;
      }
      // columns [col0, col0 + 32 kChunks) of the tile; TMEM column = same offset from buffer `as`
      const int col0 = NACC == 2 ? h * BLOCK_N : h * (BLOCK_N / 2);
#pragma unroll 1
      for (int c = 0; c < kChunks; ++c) {
        uint32_t r[32];
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * kAccCols + col0 + c * 32;
        tmem_ld_32x32b_x32(taddr, r);
        tmem_ld_wait();
        const int nbase = n0 + col0 + c * 32;
        if (m < p.M && nbase < p.N) epilogue_store_32(p, r, m, nbase, row_scale, per_col, !row_copy);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&tmem_empty_bar[NACC == 2 ? h : as], 0);
      if (row_copy) {
        // Fused all-gather: this CTA's 128 x kTileN part of the tile now sits in the local output
        // (still in L2).  The eight epilogue warps push it to the other ranks as WHOLE ROW SEGMENTS -- a warp
        // reads up to 512 contiguous bytes of one row (16 B per lane) and stores them to every peer mapping, so the
        // NVLink writes are full 384 / 512-byte runs instead of the 16-byte pieces a row-per-thread store gives.
        // The accumulator was handed back above: the transfer overlaps the next tile's main loop.
        asm volatile("bar.sync 1, 256;" ::: "memory");     // all epilogue warps' local stores of this tile are visible
        const int mrow0 = mt * (2 * kBlockM) + static_cast<int>(rank) * kBlockM;
#pragma unroll 1
        for (int seg = 0; seg < (kTileN + 255) / 256; ++seg) {
          const int cw = seg * 256 + lane * 8;            // column inside the tile (192-wide tiles: 24 lanes)
          const int nc = n0 + cw;
          if (cw >= kTileN || nc >= p.N) continue;
          const bool whole8 = nc + 8 <= p.N;
#pragma unroll 1
          for (int rr = e; rr < kBlockM; rr += 16) {
            const int ma = mrow0 + rr, mb = ma + 8;
            const uint16_t* la = static_cast<const uint16_t*>(p.D) + static_cast<int64_t>(ma) * p.ldd + nc;
            const uint16_t* lb = static_cast<const uint16_t*>(p.D) + static_cast<int64_t>(mb) * p.ldd + nc;
            uint4 va = make_uint4(0, 0, 0, 0), vb = va;
            if (whole8) {
              if (ma < p.M) va = ld_global_v4(la);
              if (mb < p.M) vb = ld_global_v4(lb);
#pragma unroll 1
              for (int pe = 0; pe < p.n_peers; ++pe) {
                uint16_t* pa = static_cast<uint16_t*>(p.peer_D[pe]) + static_cast<int64_t>(ma) * p.ldd + nc;
                uint16_t* pb = static_cast<uint16_t*>(p.peer_D[pe]) + static_cast<int64_t>(mb) * p.ldd + nc;
                if (mcast) {      // one store, replicated by the switch: egress = the slab once, not (ranks - 1) times
                  if (ma < p.M) multimem_st_v4(pa, va);
                  if (mb < p.M) multimem_st_v4(pb, vb);
                } else {
                  if (ma < p.M) *reinterpret_cast<uint4*>(pa) = va;
                  if (mb < p.M) *reinterpret_cast<uint4*>(pb) = vb;
                }
              }
            } else {
              for (int j = 0; nc + j < p.N; ++j)
                for (int pe = 0; pe < p.n_peers; ++pe) {
                  uint16_t* pd = static_cast<uint16_t*>(p.peer_D[pe]);
                  if (ma < p.M) pd[static_cast<int64_t>(ma) * p.ldd + nc + j] = *(volatile const uint16_t*)(la + j);
                  if (mb < p.M) pd[static_cast<int64_t>(mb) * p.ldd + nc + j] = *(volatile const uint16_t*)(lb + j);
                }
            }
          }
        }
      }
    }
  }

  tc_fence_before();
  cluster_sync_all();    // both CTAs are done with the pair's TMEM / shared memory
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_pair<kTmemCols>(tmem_base);
  }
}

}  // namespace vxk
