// Group-local / element-wise quantisation kernels: nothing is reduced across
// more than one 1 x G group, so the work is a flat stream of 8-element vectors.
//
//   a2  static_scaled_fp8_quant           (static_quant_kernel)
//   a3  dynamic_scaled_fp8_quant          (tensor_amax_kernel + static_quant_kernel)
//   a4  per_token_group_fp8_quant         (FMT_FP8_DIV / FMT_MX8)
//   a5  per_token_group_quant_mxfp4       (FMT_MXFP4)
//   a10 silu_and_mul_quant                (FMT_STATIC, PRO_SILU_T)
//   a11 silu_and_mul_per_block_quant      (FMT_FP8_DIV / int8, PRO_SILU_T)
//   a12 silu_and_mul_mxfp4_quant          (FMT_MXFP4, PRO_SILU_F32)
//
// Design: grid-stride over vector indices j in [0, T*H/8); a 1 x G group is
// G/8 consecutive lanes of one warp (G in {8..256}, power-of-two lane count),
// its amax is a segmented xor-shuffle.  Each thread issues all of its 128-bit
// loads (U vectors, x2 for gate/up) before the first use, single pass, 64-bit
// (fp8/int8) or 32-bit (fp4) coalesced streaming stores.
#include "quant_silu.cuh"

#include <cfloat>

namespace vxk {

// silu evaluation (see DESIGN.md "silu"): -1 = auto (default), 0 = expf + IEEE divide everywhere,
// 1 = ex2.approx / rcp.approx everywhere.  Auto picks, per op and input type, the cheapest form that is
// PROVEN bit-identical to the reference on every input: a10 / a11 round silu(gate) to the 16-bit input
// type right away, so the function has 65536 inputs and tests/test_silu_exhaustive_gpu.py checks all of
// them against torch -- the approximate form passes for bf16 (8-bit significand) and not for f16
// (11-bit), which therefore takes the accurate form.  a12 keeps the fp32 product (no rounding to hide
// the last bits; torch's own fp32 sigmoid differs between hosts there) and uses the approximate form.
std::atomic<int> g_silu_fast{-1};

// (enum Pro and the gate / up prologue + exact-division helpers live in quant_silu.cuh)
enum Fmt { FMT_FP8_DIV = 0, FMT_MX8 = 1, FMT_MXFP4 = 2, FMT_STATIC = 3 };

struct FlatParams {
  void* out;
  const void* in;        // [T, H] or [T, 2H] (gate || up)
  float* scales;         // (t*ss0 + g*ss1)
  const float* scale_in; // scale_ub or static scale
  int64_t T;
  int64_t ss0, ss1;
  int nvec;              // H / 8
  int lanes_log2;        // log2(G / 8)
  int ue8m0;
  float eps;
  // thread -> (row-in-block, column) mapping: tx column threads, rpb rows per block
  int tx, rpb;
  int ncb;               // column blocks per row = ceil(nvec / (tx * U)) = gridDim.y
  uint8_t* sf_packed;    // MXFP4 only, nullable: E8M0 bytes also go here in the GEMM's tcgen05.cp order
  int sf_nblk;           // 128-row blocks of that buffer (sf_pack_blocks(T, 0))
};

// Work decomposition: blockIdx.y picks a fixed block of U*tx column vectors, the CTA grid-strides
// over row blocks of rpb rows (blockIdx.x).  Every per-thread column quantity (validity, scale
// offset, leader flag) is loop-invariant and the three row pointers advance by a constant, so the
// loop body is loads + arithmetic + stores only.  All 128-bit loads of an iteration (U vectors, x2
// for gate/up) are issued before the first use.  These kernels are instruction-issue / XU-pipe
// bound on B200 rather than HBM bound unless the body is kept minimal: packed 16-bit multiply for
// silu*up, constant-divisor and reciprocal-based exact divisions, compile-time segmented shuffles.
// LL = log2(lanes per group) or -1 (runtime p.lanes_log2).
// Register bound: 5 CTAs of 256 threads per SM (48 registers).  Left to itself ptxas gives the MXFP4 / f32
// instantiations 79-86 registers (2-3 CTAs); 6 CTAs (40 registers) spills and measured 2-4 % slower than 5, 7 CTAs
// 10-18 % slower (profiles/r02_flat_occupancy.txt).
#ifndef VXK_FLAT_MIN_CTAS
#define VXK_FLAT_MIN_CTAS 5
#endif
template <typename T, int Q, int PRO, int FMT, int FAST, int U, int LL>
__global__ void __launch_bounds__(256, VXK_FLAT_MIN_CTAS) flat_quant_kernel(const FlatParams p) {
  constexpr float kQMax = QTraits<Q>::kMax;
  constexpr bool kGated = PRO != PRO_NONE;
  constexpr int kOutVecBytes = (FMT == FMT_MXFP4) ? 4 : kVec;
  constexpr bool kSignPatch = FMT == FMT_FP8_DIV && Q != VXK_Q_INT8 && sizeof(T) == 2;
  const int tx = p.tx;
  const int ry = threadIdx.x / tx;            // once per thread
  const int cx = threadIdx.x - ry * tx;
  const int ll = LL >= 0 ? LL : p.lanes_log2;
  const int nvec = p.nvec;
  const int c0 = blockIdx.y * (U * tx) + cx;
  const bool sat = p.scale_in != nullptr;
  // programmatic dependent launch (see launch_pdl): nothing above touches global memory
  pdl_launch_dependents();
  pdl_wait();
  float r_static = 0.0f, ub = 0.0f;
  if constexpr (FMT == FMT_STATIC) r_static = __frcp_rn(__ldg(p.scale_in));
  else if (sat) ub = __ldg(p.scale_in);
  const int ue8m0 = p.ue8m0;
  const float eps = p.eps;

  const int64_t in_row = static_cast<int64_t>(nvec) * kVec * (kGated ? 2 : 1);
  int64_t t = static_cast<int64_t>(blockIdx.x) * p.rpb + ry;
  const int64_t t_step = static_cast<int64_t>(gridDim.x) * p.rpb;
  const T* __restrict__ gp = static_cast<const T*>(p.in) + t * in_row + static_cast<int64_t>(c0) * kVec;
  uint8_t* op = static_cast<uint8_t*>(p.out) + (t * nvec + c0) * kOutVecBytes;
  float* sp = p.scales + t * p.ss0;
  const int64_t g_step = t_step * in_row, o_step = t_step * nvec * kOutVecBytes, s_step = t_step * p.ss0;
  const int u_in = tx * kVec, u_out = tx * kOutVecBytes, up_ofs = nvec * kVec;

  bool ok[U], leader[U];
  int64_t s_ofs[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int c = c0 + u * tx;
    ok[u] = c < nvec;
    leader[u] = ok[u] && ((c & ((1 << ll) - 1)) == 0);
    s_ofs[u] = static_cast<int64_t>(c >> ll) * p.ss1;
  }

  for (; t < p.T; t += t_step, gp += g_step, op += o_step, sp += s_step) {   // warp-uniform (tx % 32 == 0)
    RawV<T> g[U], up[kGated ? U : 1];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (ok[u]) {
        g[u] = ldg_raw<T>(gp + u * u_in);
        if constexpr (kGated) up[u] = ldg_raw<T>(gp + u * u_in + up_ofs);
      } else {
        g[u] = RawV<T>{};
        if constexpr (kGated) up[u] = RawV<T>{};
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float v[8], q[8];
      uint32_t yw[4] = {0, 0, 0, 0};
      prologue<T, PRO, FAST>(g[u], up[kGated ? u : 0], v, yw);
      uint8_t* o = op + u * u_out;
      if constexpr (FMT == FMT_STATIC) {
#pragma unroll
        for (int k = 0; k < 8; k += 2) {
          q[k] = v[k];
          q[k + 1] = v[k + 1];
          mul2(q[k], q[k + 1], r_static, r_static);
        }
        if (ok[u]) stg_stream(o, pack_q8<Q>(q));
      } else {
        float gmax = 0.0f;
#pragma unroll
        for (int k = 0; k < 8; ++k) gmax = fmaxf(gmax, fabsf(v[k]));
        // every lane takes part (invalid ones carry 0)
        if constexpr (LL >= 0) gmax = seg_max_c<(1 << (LL >= 0 ? LL : 0))>(gmax);
        else gmax = seg_max(gmax, 1 << ll);
        float* sq = sp + s_ofs[u];
        if constexpr (FMT == FMT_MXFP4) {
          const int e = rceil_e8m0(div_const<6>(gmax));
          const float d = e8m0_descale(e);
          if (leader[u]) {
            *sq = e8m0_scale(e);
            if (p.sf_packed) p.sf_packed[sf_packed_offset<2>(t, (c0 + u * tx) >> ll, p.sf_nblk)] = static_cast<uint8_t>(e);
          }
          // (the reference's clamp to +-6 is what the saturating conversion does anyway)
#pragma unroll
          for (int k = 0; k < 8; k += 2) {
            q[k] = v[k];
            q[k + 1] = v[k + 1];
            mul2(q[k], q[k + 1], d, d);
          }
          // bf16 inputs scaled by a power of two are already bf16 values; everything else
          // takes the reference's intermediate rounding to bf16 (mx_utils.py:346)
          if constexpr (!(PRO == PRO_NONE && std::is_same<T, __nv_bfloat16>::value)) {
#pragma unroll
            for (int k = 0; k < 8; k += 2) {
              const __nv_bfloat162 b2 = __floats2bfloat162_rn(q[k], q[k + 1]);
              const uint32_t w = *reinterpret_cast<const uint32_t*>(&b2);
              q[k] = bf16_lo(w);
              q[k + 1] = bf16_hi(w);
            }
          }
          if (ok[u]) stg_stream(o, pack_e2m1x8(q));
        } else if constexpr (FMT == FMT_MX8) {
          const int e = rceil_e8m0(div_qmax<Q>(gmax));
          const float d = e8m0_descale(e);
          if (leader[u]) *sq = e8m0_scale(e);
#pragma unroll
          for (int k = 0; k < 8; k += 2) {
            q[k] = v[k];
            q[k + 1] = v[k + 1];
            mul2(q[k], q[k + 1], d, d);
          }
          if (ok[u]) stg_stream(o, pack_q8<Q>(q));
        } else {
          float s;
          if constexpr (PRO == PRO_NONE) {
            s = fmaxf(div_qmax<Q>(gmax), eps);            // a4: s = max(amax / qmax, eps)
          } else if constexpr (Q == VXK_Q_INT8) {
            s = div_qmax<Q>(fmaxf(gmax, 1e-10f));
          } else {
            s = div_qmax<Q>(fmaxf(gmax, FLT_MIN));
            if (sat) s = fminf(s, ub);
            if (ue8m0) s = exp2_ceil_log2(fmaxf(s, 1e-10f));
          }
          const float r = __frcp_rn(s);
          if (leader[u]) *sq = s;
          uint2 packed;
          if (__builtin_expect(!(r <= FLT_MAX), 0)) {
            // s is subnormal or zero (an all-zero / vanishing group): 1/s overflows, take the
            // IEEE divide (0 / s = 0, matching the reference)
#pragma unroll
            for (int k = 0; k < 8; ++k) q[k] = slow_div(v[k], s);
            packed = pack_q8<Q>(q);
          } else if (Q != VXK_Q_INT8 && sat) {
#pragma unroll
            for (int k = 0; k < 8; ++k) q[k] = div_by<true>(v[k], s, r, kQMax);
            packed = pack_q8<Q>(q);
          } else if constexpr (kSignPatch) {
            packed = div_pack_fp8_signed<Q>(v, yw, s, r);
          } else {
#pragma unroll
            for (int k = 0; k < 8; ++k) q[k] = div_by<false>(v[k], s, r);
            packed = pack_q8<Q>(q);
          }
          if (ok[u]) stg_stream(o, packed);
        }
      }
    }
  }
}

// Generic group kernel (any G that divides H, any alignment): one warp per group.
template <typename T, int Q, int PRO, int FMT, int FAST>
__global__ void __launch_bounds__(256) flat_quant_generic(const FlatParams p, int64_t H, int G,
                                                          int64_t num_groups) {
  constexpr float kQMax = QTraits<Q>::kMax;
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  const int gpr = static_cast<int>(H / G);
  auto value = [&](int64_t t, int h) -> float {
    if constexpr (PRO == PRO_NONE) {
      return to_f32<T>(static_cast<const T*>(p.in)[t * H + h]);
    } else {
      const T* row = static_cast<const T*>(p.in) + t * 2 * H;
      const float g = to_f32<T>(row[h]), u = to_f32<T>(row[H + h]);
      if constexpr (PRO == PRO_SILU_T) return round_to<T>(round_to<T>(silu_f32<FAST>(g)) * u);
      else if constexpr (FAST) return silu_f32<1>(g) * u;
      else return (g * __fdiv_rn(1.0f, 1.0f + expf(-g))) * u;
    }
  };
  for (int64_t gi = warp0; gi < num_groups; gi += nwarps) {
    const int64_t t = gi / gpr;
    const int g = static_cast<int>(gi % gpr);
    float gmax = 0.0f;
    if constexpr (FMT != FMT_STATIC) {
      for (int h = g * G + lane; h < (g + 1) * G; h += 32) gmax = fmaxf(gmax, fabsf(value(t, h)));
      gmax = warp_max(gmax);
    }
    if constexpr (FMT == FMT_STATIC) {
      const float r = __fdiv_rn(1.0f, __ldg(p.scale_in));
      uint8_t* o = static_cast<uint8_t*>(p.out) + t * H;
      for (int h = g * G + lane; h < (g + 1) * G; h += 32) o[h] = pack_q1<Q>(value(t, h) * r);
      continue;
    }
    float* sp = p.scales + t * p.ss0 + g * p.ss1;
    if constexpr (FMT == FMT_MXFP4) {
      const int e = rceil_e8m0(__fdiv_rn(gmax, 6.0f));
      const float d = e8m0_descale(e);
      if (lane == 0) {
        *sp = e8m0_scale(e);
        if (p.sf_packed) p.sf_packed[sf_packed_offset<2>(t, g, p.sf_nblk)] = static_cast<uint8_t>(e);
      }
      uint8_t* o = static_cast<uint8_t*>(p.out) + t * (H / 2);
      for (int h2 = g * G / 2 + lane; h2 < (g + 1) * G / 2; h2 += 32) {
        float c0 = fminf(fmaxf(value(t, 2 * h2) * d, -6.0f), 6.0f);
        float c1 = fminf(fmaxf(value(t, 2 * h2 + 1) * d, -6.0f), 6.0f);
        c0 = __bfloat162float(__float2bfloat16_rn(c0));
        c1 = __bfloat162float(__float2bfloat16_rn(c1));
        o[h2] = static_cast<uint8_t>(cvt_e2m1x2(c1, c0));
      }
    } else {
      float s, r;
      bool mul = false;
      if constexpr (FMT == FMT_MX8) {
        const int e = rceil_e8m0(__fdiv_rn(gmax, kQMax));
        s = e8m0_scale(e);
        r = e8m0_descale(e);
        mul = true;
      } else {
        if constexpr (PRO == PRO_NONE) s = fmaxf(__fdiv_rn(gmax, kQMax), p.eps);
        else if constexpr (Q == VXK_Q_INT8) s = __fdiv_rn(fmaxf(gmax, 1e-10f), 127.0f);
        else {
          s = __fdiv_rn(fmaxf(gmax, FLT_MIN), kQMax);
          if (p.scale_in != nullptr) s = fminf(s, __ldg(p.scale_in));
          if (p.ue8m0) s = exp2_ceil_log2(fmaxf(s, 1e-10f));
        }
        r = __fdiv_rn(1.0f, s);
      }
      if (lane == 0) *sp = s;
      uint8_t* o = static_cast<uint8_t*>(p.out) + t * H;
      for (int h = g * G + lane; h < (g + 1) * G; h += 32) {
        const float x = value(t, h);
        o[h] = pack_q1<Q>(mul ? x * r : (!(r <= FLT_MAX) ? __fdiv_rn(x, s) : (Q == VXK_Q_INT8 ? div_by<false>(x, s, r) : div_by<true>(x, s, r, kQMax))));
      }
    }
  }
}

// ---- a2: static quant with an arbitrary (group_m, group_n) scale grid ----
struct StaticParams {
  void* out;
  const void* in;
  const float* scale;
  int64_t T, H, in_row_stride, out_row_stride, group_m, group_n, ss_i, ss_j;
};

// vector path: group_n % 8 == 0 so one scale per vector; persistent CTAs over rows,
// 32-bit index arithmetic only (host guarantees T < 2^32, H < 2^31)
template <typename T, int Q>
__global__ void __launch_bounds__(256) static_quant_vec_kernel(const StaticParams p) {
  const int nvec = static_cast<int>(p.H / kVec);
  const unsigned gm = static_cast<unsigned>(p.group_m), gnv = static_cast<unsigned>(p.group_n / kVec);
  pdl_launch_dependents();
  pdl_wait();   // the scale may come from the kernel right before this one (dynamic_scaled_fp8_quant)
  for (int64_t t = blockIdx.x; t < p.T; t += gridDim.x) {
    const T* __restrict__ xrow = static_cast<const T*>(p.in) + t * p.in_row_stride;
    uint8_t* orow = static_cast<uint8_t*>(p.out) + t * p.out_row_stride;
    const float* srow = p.scale + static_cast<int64_t>(static_cast<unsigned>(t) / gm) * p.ss_i;
    for (int c0 = threadIdx.x; c0 < nvec; c0 += 4 * blockDim.x) {
      float v[4][8];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int c = c0 + u * blockDim.x;
        if (c < nvec) load8<T, true>(xrow + c * kVec, v[u]);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int c = c0 + u * blockDim.x;
        if (c < nvec) {
          const float s = __ldg(srow + static_cast<int64_t>(static_cast<unsigned>(c) / gnv) * p.ss_j);
          const float r = __fdiv_rn(1.0f, s);
          float q[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) q[k] = v[u][k] * r;
          stg_stream(orow + c * kVec, pack_q8<Q>(q));
        }
      }
    }
  }
}

template <typename T, int Q>
__global__ void __launch_bounds__(256) static_quant_scalar_kernel(const StaticParams p) {
  const int64_t total = p.T * p.H;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int64_t t = i / p.H, h = i % p.H;
    const float s = __ldg(p.scale + (t / p.group_m) * p.ss_i + (h / p.group_n) * p.ss_j);
    const float x = to_f32<T>(static_cast<const T*>(p.in)[t * p.in_row_stride + h]);
    static_cast<uint8_t*>(p.out)[t * p.out_row_stride + h] = pack_q1<Q>(x * __fdiv_rn(1.0f, s));
  }
}

// ---- a3: whole-tensor amax -> scale = amax / qmax (atomic max on the fp32 bits) ----
template <typename T, int Q>
__global__ void __launch_bounds__(256) tensor_amax_kernel(const T* __restrict__ in, float* scale,
                                                          int64_t T_, int64_t H, int64_t in_row_stride,
                                                          bool vec) {
  __shared__ float scratch[33];
  float amax = 0.0f;
  for (int64_t t = blockIdx.x; t < T_; t += gridDim.x) {
    const T* __restrict__ xrow = in + t * in_row_stride;
    if (vec) {
      const int nvec = static_cast<int>(H / kVec);
      for (int c0 = threadIdx.x; c0 < nvec; c0 += 4 * blockDim.x) {
        float v[4][8];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int c = c0 + u * blockDim.x;
          if (c < nvec) load8<T, true>(xrow + c * kVec, v[u]);
          else {
#pragma unroll
            for (int k = 0; k < 8; ++k) v[u][k] = 0.0f;
          }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
          for (int k = 0; k < 8; ++k) amax = fmaxf(amax, fabsf(v[u][k]));
      }
    } else {
      for (int h = threadIdx.x; h < H; h += blockDim.x) amax = fmaxf(amax, fabsf(to_f32<T>(xrow[h])));
    }
  }
  amax = block_reduce<true>(amax, scratch);
  if (threadIdx.x == 0) {
    const float s = __fdiv_rn(amax, QTraits<Q>::kMax);  // non-negative: int order == float order
    atomicMax(reinterpret_cast<int*>(scale), __float_as_int(s));
  }
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
static int flat_grid(int64_t work_items, int per_block) {
  const int64_t blocks = (work_items + per_block - 1) / per_block;
  const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
  return static_cast<int>(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}

// resident CTAs per SM of one kernel instantiation (queried once per instantiation)
template <typename K>
static int resident_ctas(K kernel, int threads) {
  int n = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, threads, 0) != cudaSuccess || n < 1) n = 1;
  return n;
}

template <typename T, int Q, int PRO, int FMT, int FAST, int U, int LL>
static void launch_flat_vec(const FlatParams& p, int64_t row_blocks, int ncb, int threads, cudaStream_t st) {
  auto kernel = flat_quant_kernel<T, Q, PRO, FMT, FAST, U, LL>;
  static thread_local int occ[9] = {0};       // by threads / 32
  int& o = occ[threads / 32];
  if (o == 0) o = resident_ctas(kernel, threads);
  // one wave of resident CTAs: a CTA that only starts when another retires would run its whole
  // share of the rows at a fraction of the machine
  int64_t gx = static_cast<int64_t>(sm_count()) * o / ncb;
  if (gx < 1) gx = 1;
  if (gx > row_blocks) gx = row_blocks;
  launch_pdl(kernel, dim3(static_cast<unsigned>(gx), static_cast<unsigned>(ncb)), dim3(threads), 0, st, p);
}

template <typename T, int Q, int PRO, int FMT, int FAST, int U>
static void launch_flat_ll(const FlatParams& p, int64_t row_blocks, int ncb, int threads, cudaStream_t st) {
  switch (p.lanes_log2) {   // the group sizes of the reference's kernels get unrolled shuffles
    case 0: return launch_flat_vec<T, Q, PRO, FMT, FAST, U, 0>(p, row_blocks, ncb, threads, st);
    case 2: return launch_flat_vec<T, Q, PRO, FMT, FAST, U, 2>(p, row_blocks, ncb, threads, st);
    case 3: return launch_flat_vec<T, Q, PRO, FMT, FAST, U, 3>(p, row_blocks, ncb, threads, st);
    case 4: return launch_flat_vec<T, Q, PRO, FMT, FAST, U, 4>(p, row_blocks, ncb, threads, st);
    default: return launch_flat_vec<T, Q, PRO, FMT, FAST, U, -1>(p, row_blocks, ncb, threads, st);
  }
}

template <typename T, int Q, int PRO, int FMT>
static int launch_flat(FlatParams p, int64_t T_, int64_t H, int G, bool vec_ok, cudaStream_t st) {
  if (T_ == 0 || H == 0) return VXK_OK;
  int mode = (PRO != PRO_NONE) ? g_silu_fast.load(std::memory_order_relaxed) : 0;
  if (mode < 0) mode = (PRO == PRO_SILU_F32 || std::is_same<T, __nv_bfloat16>::value) ? 1 : 0;   // auto
  p.T = T_;
  if (vec_ok) {
    constexpr int U = (PRO == PRO_NONE) ? 4 : 2;   // 4 x 16 B loads in flight per thread either way
    p.nvec = static_cast<int>(H / kVec);
    const int lanes = G / kVec;
    p.lanes_log2 = 0;
    while ((1 << p.lanes_log2) < lanes) ++p.lanes_log2;
    // column threads per row: the multiple of 32 (hence of `lanes`) that leaves the fewest idle
    // lanes in the last column block; short rows share a CTA (rpb rows per block)
    // Among equally tight choices a divisor of 256 wins: 8-warp CTAs.  (Registers are handed out to CTAs in units
    // of four warps, so a 7-warp CTA -- tx = 224, which H = 14336 would otherwise pick -- occupies the register file
    // like an 8-warp one: 35 resident warps instead of 40, measured 4-5 % slower on the bench's dominant kernel,
    // profiles/r02_flat_occupancy.txt.)
    int best_tx = 256;
    double best_idle = 2.0;
    static const int kTxOrder[8] = {256, 128, 64, 32, 224, 192, 160, 96};
    for (int i = 0; i < 8; ++i) {
      const int tx = kTxOrder[i];
      const int64_t cols = (static_cast<int64_t>(p.nvec) + tx * U - 1) / (tx * U) * (tx * U);
      const double idle = static_cast<double>(cols - p.nvec) / static_cast<double>(cols);
      if (idle < best_idle - 1e-9) { best_idle = idle; best_tx = tx; }
    }
    static const int env_tx = [] { const char* e = getenv("VXK_FLAT_TX"); return e ? atoi(e) : 0; }();   // (A/B timing)
    if (env_tx >= 32 && env_tx <= 256 && env_tx % 32 == 0) best_tx = env_tx;
    p.tx = best_tx;
    p.rpb = 256 / p.tx;
    p.ncb = (p.nvec + p.tx * U - 1) / (p.tx * U);
    const int64_t row_blocks = (T_ + p.rpb - 1) / p.rpb;
    if (row_blocks > 0x7fffffffLL || p.ncb > 65535)
      return fail(VXK_ERR_INVALID, "problem too large (%lld row blocks x %d column blocks)", (long long)row_blocks, p.ncb);
    const int threads = p.tx * p.rpb;
    if constexpr (PRO == PRO_NONE) {
      launch_flat_ll<T, Q, PRO, FMT, 0, U>(p, row_blocks, p.ncb, threads, st);
    } else {
      if (mode == 0) launch_flat_ll<T, Q, PRO, FMT, 0, U>(p, row_blocks, p.ncb, threads, st);
      else launch_flat_ll<T, Q, PRO, FMT, 1, U>(p, row_blocks, p.ncb, threads, st);
    }
  } else {
    const int64_t ngroups = T_ * (H / G);
    const int grid = flat_grid(ngroups, 8);
    if (mode != 0) flat_quant_generic<T, Q, PRO, FMT, (PRO != PRO_NONE) ? 1 : 0><<<grid, 256, 0, st>>>(p, H, G, ngroups);
    else flat_quant_generic<T, Q, PRO, FMT, 0><<<grid, 256, 0, st>>>(p, H, G, ngroups);
  }
  return check_launch("flat_quant");
}

#define VXK_DISPATCH_T(in_dtype, ...)                                          \
  switch (in_dtype) {                                                          \
    case VXK_F32: { using T = float; __VA_ARGS__; }                            \
    case VXK_F16: { using T = __half; __VA_ARGS__; }                           \
    case VXK_BF16: { using T = __nv_bfloat16; __VA_ARGS__; }                   \
    default: return fail(VXK_ERR_INVALID, "unsupported input dtype %d", in_dtype); \
  }

#define VXK_DISPATCH_T16(in_dtype, ...)                                        \
  switch (in_dtype) {                                                          \
    case VXK_F16: { using T = __half; __VA_ARGS__; }                           \
    case VXK_BF16: { using T = __nv_bfloat16; __VA_ARGS__; }                   \
    default: return fail(VXK_ERR_INVALID, "input must be f16 / bf16 (got dtype %d)", in_dtype); \
  }

static bool group_vec_ok(const void* in, const void* out, int64_t H, int64_t G, int in_dtype,
                         bool packed4) {
  const int lanes = static_cast<int>(G / kVec);
  if (G % kVec || lanes > 32 || (lanes & (lanes - 1))) return false;
  if (!aligned16(in)) return false;
  if (in_dtype != VXK_F32 && (H * 2) % 16) return false;
  if (reinterpret_cast<uintptr_t>(out) % (packed4 ? 4 : 8)) return false;
  return H % kVec == 0;
}

}  // namespace vxk

using namespace vxk;

extern "C" int vxk_set_silu_fast(int on) {
  return g_silu_fast.exchange(on < 0 ? -1 : (on ? 1 : 0));
}

extern "C" int vxk_per_token_group_fp8_quant(const void* in, void* out_q, float* out_s, int64_t T_,
                                             int64_t H, int64_t G, double eps, int scale_ue8m0,
                                             int64_t ss0, int64_t ss1, int in_dtype, int q_dtype,
                                             vxk_stream_t stream) {
  VXK_REQUIRE(T_ >= 0 && H >= 0, "negative shape");
  VXK_REQUIRE(G > 0 && H % G == 0, "the last dimension (%lld) must be divisible by group_size (%lld)",
              (long long)H, (long long)G);
  VXK_REQUIRE(q_dtype == VXK_Q_E4M3 || q_dtype == VXK_Q_E5M2, "output must be fp8 (e4m3 / e5m2)");
  FlatParams p{};
  p.out = out_q; p.in = in; p.scales = out_s; p.ss0 = ss0; p.ss1 = ss1;
  p.eps = static_cast<float>(eps);
  const bool v = group_vec_ok(in, out_q, H, G, in_dtype, false);
  VXK_DISPATCH_T(in_dtype,
    if (scale_ue8m0) {
      if (q_dtype == VXK_Q_E4M3) return launch_flat<T, VXK_Q_E4M3, PRO_NONE, FMT_MX8>(p, T_, H, (int)G, v, stream);
      return launch_flat<T, VXK_Q_E5M2, PRO_NONE, FMT_MX8>(p, T_, H, (int)G, v, stream);
    }
    if (q_dtype == VXK_Q_E4M3) return launch_flat<T, VXK_Q_E4M3, PRO_NONE, FMT_FP8_DIV>(p, T_, H, (int)G, v, stream);
    return launch_flat<T, VXK_Q_E5M2, PRO_NONE, FMT_FP8_DIV>(p, T_, H, (int)G, v, stream));
}

extern "C" int vxk_per_token_group_quant_mxfp4_sf(const void* in, void* out_q, float* out_s, void* sf_packed,
                                                  int64_t T_, int64_t H, int64_t G, double eps, int64_t ss0,
                                                  int64_t ss1, int in_dtype, vxk_stream_t stream) {
  (void)eps;  // the torch reference (to_mxfp) has no epsilon; an all-zero group gets exponent byte 0
  VXK_REQUIRE(T_ >= 0 && H >= 0, "negative shape");
  VXK_REQUIRE(G == 32, "MXFP4 requires group_size == 32, got %lld", (long long)G);
  VXK_REQUIRE(H % 32 == 0, "the last dimension (%lld) must be divisible by 32", (long long)H);
  FlatParams p{};
  p.out = out_q; p.in = in; p.scales = out_s; p.ss0 = ss0; p.ss1 = ss1;
  p.sf_packed = static_cast<uint8_t*>(sf_packed); p.sf_nblk = sf_pack_blocks(T_, 0);
  const bool v = group_vec_ok(in, out_q, H, G, in_dtype, true);
  VXK_DISPATCH_T(in_dtype,
    return launch_flat<T, VXK_Q_E4M3, PRO_NONE, FMT_MXFP4>(p, T_, H, 32, v, stream));
}
extern "C" int vxk_per_token_group_quant_mxfp4(const void* in, void* out_q, float* out_s, int64_t T_,
                                               int64_t H, int64_t G, double eps, int64_t ss0,
                                               int64_t ss1, int in_dtype, vxk_stream_t stream) {
  return vxk_per_token_group_quant_mxfp4_sf(in, out_q, out_s, nullptr, T_, H, G, eps, ss0, ss1, in_dtype, stream);
}

extern "C" int vxk_silu_and_mul_quant(void* out, const void* in, const float* scale, int64_t T_,
                                      int64_t d, int in_dtype, int q_dtype, vxk_stream_t stream) {
  VXK_REQUIRE(T_ >= 0 && d >= 0, "negative shape");
  VXK_REQUIRE(q_dtype == VXK_Q_E4M3 || q_dtype == VXK_Q_E5M2, "output must be fp8 (e4m3 / e5m2)");
  VXK_REQUIRE(in_dtype == VXK_F16 || in_dtype == VXK_BF16, "silu_and_mul_quant supports f16 / bf16 input");
  FlatParams p{};
  p.out = out; p.in = in; p.scale_in = scale;
  if (T_ == 0 || d == 0) return VXK_OK;
  // element-wise: reuse the flat kernel with a "group" of one vector; d % 8 != 0 -> generic G = 1
  const bool v = (d % kVec == 0) && aligned16(in) && (reinterpret_cast<uintptr_t>(out) % 8 == 0);
  const int G = v ? kVec : static_cast<int>(d);  // generic path: one warp per row
  VXK_DISPATCH_T16(in_dtype,
    if (q_dtype == VXK_Q_E4M3) return launch_flat<T, VXK_Q_E4M3, PRO_SILU_T, FMT_STATIC>(p, T_, d, G, v, stream);
    return launch_flat<T, VXK_Q_E5M2, PRO_SILU_T, FMT_STATIC>(p, T_, d, G, v, stream));
}

extern "C" int vxk_silu_and_mul_per_block_quant(void* out, const void* in, float* scales, int64_t T_,
                                                int64_t H, int64_t G, const float* scale_ub,
                                                int64_t ss0, int64_t ss1, int scale_ue8m0, int in_dtype,
                                                int q_dtype, vxk_stream_t stream) {
  VXK_REQUIRE(T_ >= 0 && H >= 0, "negative shape");
  VXK_REQUIRE(G == 32 || G == 64 || G == 128, "group_size must be 32, 64 or 128, got %lld", (long long)G);
  VXK_REQUIRE(H % G == 0, "hidden_size (%lld) must be divisible by group_size (%lld)", (long long)H,
              (long long)G);
  VXK_REQUIRE(in_dtype == VXK_F16 || in_dtype == VXK_BF16, "input must be f16 / bf16");
  VXK_REQUIRE(!(scale_ub && q_dtype == VXK_Q_INT8), "scale_ub is only supported for FP8 output");
  FlatParams p{};
  p.out = out; p.in = in; p.scales = scales; p.scale_in = scale_ub; p.ss0 = ss0; p.ss1 = ss1;
  p.ue8m0 = scale_ue8m0;
  const bool v = group_vec_ok(in, out, H, G, in_dtype, false);
  VXK_DISPATCH_T16(in_dtype,
    switch (q_dtype) {
      case VXK_Q_E4M3: return launch_flat<T, VXK_Q_E4M3, PRO_SILU_T, FMT_FP8_DIV>(p, T_, H, (int)G, v, stream);
      case VXK_Q_E5M2: return launch_flat<T, VXK_Q_E5M2, PRO_SILU_T, FMT_FP8_DIV>(p, T_, H, (int)G, v, stream);
      case VXK_Q_INT8: return launch_flat<T, VXK_Q_INT8, PRO_SILU_T, FMT_FP8_DIV>(p, T_, H, (int)G, v, stream);
      default: return fail(VXK_ERR_INVALID, "unsupported quant dtype %d", q_dtype);
    });
}

extern "C" int vxk_silu_and_mul_mxfp4_quant(void* out, const void* in, float* scales, int64_t T_,
                                            int64_t H, int64_t G, double eps, int64_t ss0, int64_t ss1,
                                            int in_dtype, vxk_stream_t stream) {
  return vxk_silu_and_mul_mxfp4_quant_sf(out, in, scales, nullptr, T_, H, G, eps, ss0, ss1, in_dtype, stream);
}
extern "C" int vxk_silu_and_mul_mxfp4_quant_sf(void* out, const void* in, float* scales, void* sf_packed, int64_t T_,
                                               int64_t H, int64_t G, double eps, int64_t ss0, int64_t ss1,
                                               int in_dtype, vxk_stream_t stream) {
  (void)eps;
  VXK_REQUIRE(T_ >= 0 && H >= 0, "negative shape");
  VXK_REQUIRE(G == 32, "MXFP4 requires group_size == 32, got %lld", (long long)G);
  VXK_REQUIRE(H % 32 == 0, "hidden_size (%lld) must be divisible by 32", (long long)H);
  VXK_REQUIRE(in_dtype == VXK_F16 || in_dtype == VXK_BF16, "input must be f16 / bf16");
  FlatParams p{};
  p.out = out; p.in = in; p.scales = scales; p.ss0 = ss0; p.ss1 = ss1;
  p.sf_packed = static_cast<uint8_t*>(sf_packed); p.sf_nblk = sf_pack_blocks(T_, 0);
  const bool v = group_vec_ok(in, out, H, G, in_dtype, true);
  VXK_DISPATCH_T16(in_dtype,
    return launch_flat<T, VXK_Q_E4M3, PRO_SILU_F32, FMT_MXFP4>(p, T_, H, 32, v, stream));
}

template <typename T, int Q>
static int launch_static(const StaticParams& p, bool vec, cudaStream_t st) {
  if (vec) {
    const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
    const int nvec = static_cast<int>(p.H / kVec);
    const int threads = nvec >= 1024 ? 256 : (nvec >= 512 ? 128 : 64);
    launch_pdl(static_quant_vec_kernel<T, Q>, dim3(static_cast<unsigned>(p.T < cap ? p.T : cap)), dim3(threads), 0, st, p);
  } else {
    const int grid = flat_grid(p.T * p.H, 1024);
    static_quant_scalar_kernel<T, Q><<<grid, 256, 0, st>>>(p);
  }
  return check_launch("static_quant");
}

static bool static_vec_ok(const StaticParams& p, int in_dtype) {
  const size_t es = in_dtype == VXK_F32 ? 4 : 2;
  return p.H % kVec == 0 && p.group_n % kVec == 0 && p.T < 0xffffffffLL && p.H < 0x7fffffffLL && aligned16(p.in) &&
         (p.in_row_stride * es) % 16 == 0 && reinterpret_cast<uintptr_t>(p.out) % 8 == 0 &&
         p.out_row_stride % 8 == 0;
}

extern "C" int vxk_static_scaled_fp8_quant(void* out, const void* in, const float* scale, int64_t T_,
                                           int64_t H, int64_t in_row_stride, int64_t out_row_stride,
                                           int64_t group_m, int64_t group_n, int64_t ss_i, int64_t ss_j,
                                           int in_dtype, int q_dtype, vxk_stream_t stream) {
  VXK_REQUIRE(T_ >= 0 && H >= 0, "negative shape");
  VXK_REQUIRE(q_dtype == VXK_Q_E4M3 || q_dtype == VXK_Q_E5M2, "output must be fp8 (e4m3 / e5m2)");
  if (T_ == 0 || H == 0) return VXK_OK;
  VXK_REQUIRE(group_m > 0 && group_n > 0, "group shape must be positive");
  StaticParams p{out, in, scale, T_, H, in_row_stride, out_row_stride, group_m, group_n, ss_i, ss_j};
  const bool v = static_vec_ok(p, in_dtype);
  if (v && ss_i == 0 && ss_j == 0 && in_row_stride == H && out_row_stride == H) {
    // per-tensor scale on contiguous rows: the flat streaming kernel
    FlatParams f{};
    f.out = out; f.in = in; f.scale_in = scale;
    VXK_DISPATCH_T(in_dtype,
      if (q_dtype == VXK_Q_E4M3) return launch_flat<T, VXK_Q_E4M3, PRO_NONE, FMT_STATIC>(f, T_, H, kVec, true, stream);
      return launch_flat<T, VXK_Q_E5M2, PRO_NONE, FMT_STATIC>(f, T_, H, kVec, true, stream));
  }
  VXK_DISPATCH_T(in_dtype,
    if (q_dtype == VXK_Q_E4M3) return launch_static<T, VXK_Q_E4M3>(p, v, stream);
    return launch_static<T, VXK_Q_E5M2>(p, v, stream));
}

extern "C" int vxk_dynamic_scaled_fp8_quant(void* out, const void* in, float* scale, int64_t T_,
                                            int64_t H, int64_t in_row_stride, int64_t out_row_stride,
                                            int in_dtype, int q_dtype, vxk_stream_t stream) {
  VXK_REQUIRE(T_ >= 0 && H >= 0, "negative shape");
  VXK_REQUIRE(q_dtype == VXK_Q_E4M3 || q_dtype == VXK_Q_E5M2, "output must be fp8 (e4m3 / e5m2)");
  cudaError_t e = cudaMemsetAsync(scale, 0, sizeof(float), stream);
  if (e != cudaSuccess) return fail(VXK_ERR_CUDA, "memset: %s", cudaGetErrorString(e));
  if (T_ == 0 || H == 0) return VXK_OK;
  StaticParams p{out, in, scale, T_, H, in_row_stride, out_row_stride, T_, H, 0, 0};
  const bool v = static_vec_ok(p, in_dtype);
  const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
  const unsigned grid = static_cast<unsigned>(T_ < cap ? T_ : cap);
  VXK_DISPATCH_T(in_dtype, {
    int rc;
    if (q_dtype == VXK_Q_E4M3) {
      tensor_amax_kernel<T, VXK_Q_E4M3><<<grid, 256, 0, stream>>>(static_cast<const T*>(in), scale, T_, H, in_row_stride, v);
      if ((rc = check_launch("tensor_amax"))) return rc;
      return launch_static<T, VXK_Q_E4M3>(p, v, stream);
    }
    tensor_amax_kernel<T, VXK_Q_E5M2><<<grid, 256, 0, stream>>>(static_cast<const T*>(in), scale, T_, H, in_row_stride, v);
    if ((rc = check_launch("tensor_amax"))) return rc;
    return launch_static<T, VXK_Q_E5M2>(p, v, stream);
  });
}

// ---- testing aid: the silu evaluation of the fused kernels, element by element --------------
// out_t[i] = T(silu(in[i])) as a10 / a11 compute it (mode 0 / 1 as vxk_set_silu_fast), out_f32[i] =
// the fp32 value before the rounding to T (what a12 multiplies by `up`).  mode 2: the approximate
// form with caller-supplied constants, silu = (x * rcp(1 + ex2(x * c))) * beta -- used once to search
// for constants whose rounding to bf16 agrees with torch on every input (tests/test_silu_exhaustive_gpu.py).
namespace vxk {
template <typename T>
__global__ void silu_probe_kernel(const T* __restrict__ in, T* out_t, float* out_f32, int64_t n, int mode, float c,
                                  float beta) {
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const float x = to_f32<T>(in[i]);
    float s;
    if (mode == 0) s = silu_f32<0>(x);
    else if (mode == 1) {
      // the approximate form as the fused kernels use it: bf16 vectors holding a gate <= -87.25 take the
      // accurate form (prologue<PRO_SILU_T, FAST>)
      if (std::is_same<T, __nv_bfloat16>::value && !(x > -87.25f)) s = silu_f32<0>(x);
      else s = silu_f32<1>(x);
    }
    else s = (x * rcp_approx(1.0f + ex2_approx(x * c))) * beta;
    if (out_t) out_t[i] = from_f32<T>(s);
    if (out_f32) out_f32[i] = s;
  }
}
}  // namespace vxk

extern "C" int vxk_debug_silu(void* out_t, float* out_f32, const void* in, int64_t n, int in_dtype, int mode,
                              float c, float beta, vxk_stream_t stream) {
  VXK_REQUIRE(n >= 0 && mode >= 0 && mode <= 2, "bad arguments");
  if (n == 0) return VXK_OK;
  const int grid = flat_grid(n, 256);
  VXK_DISPATCH_T16(in_dtype,
    silu_probe_kernel<T><<<grid, 256, 0, stream>>>(static_cast<const T*>(in), static_cast<T*>(out_t), out_f32, n, mode, c, beta);
    return check_launch("silu_probe"));
}
