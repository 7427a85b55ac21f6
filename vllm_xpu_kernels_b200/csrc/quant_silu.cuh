// Device helpers shared by the fused silu_and_mul + quant kernels (quant_flat.cu) and the GEMM epilogue that
// fuses the same arithmetic behind the gate_up projection (gemm_tc2.cuh, FUSE): the gate / up prologue
// (y = T(T(silu(gate)) * up) and friends) and the exact division + fp8 packing of the 1 x G block quantisation.
// Every floating-point step is an explicit instruction (packed asm, intrinsics), so the results do not depend on the
// including file's -fmad setting.
#pragma once

#include <type_traits>

#include "vxk_common.cuh"

namespace vxk {

enum Pro { PRO_NONE = 0, PRO_SILU_T = 1, PRO_SILU_F32 = 2 };

// ---- packed 16-bit helpers (bf16x2 / half2 held in a 32-bit register) ----
template <typename T> struct Pair;
template <> struct Pair<__nv_bfloat16> {
  using V = __nv_bfloat162;
  static __device__ __forceinline__ float2 unpack(uint32_t w) { return make_float2(bf16_lo(w), bf16_hi(w)); }
  static __device__ __forceinline__ V pack(float lo, float hi) { return __floats2bfloat162_rn(lo, hi); }
};
template <> struct Pair<__half> {
  using V = __half2;
  static __device__ __forceinline__ float2 unpack(uint32_t w) {
    return __half22float2(*reinterpret_cast<const __half2*>(&w));
  }
  static __device__ __forceinline__ V pack(float lo, float hi) { return __floats2half2_rn(lo, hi); }
};

template <typename T> struct RawV { uint4 a; };
template <> struct RawV<float> { uint4 a, b; };
template <typename T>
__device__ __forceinline__ RawV<T> ldg_raw(const T* p) {
  RawV<T> r;
  r.a = ldg_stream(p);
  if constexpr (sizeof(T) == 4) r.b = ldg_stream(p + 4);
  return r;
}
template <typename T>
__device__ __forceinline__ void raw_to_f32(const RawV<T>& r, float (&v)[8]) {
  if constexpr (sizeof(T) == 2) {
    const uint32_t w[4] = {r.a.x, r.a.y, r.a.z, r.a.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = Pair<T>::unpack(w[i]);
      v[2 * i] = f.x;
      v[2 * i + 1] = f.y;
    }
  } else {
    v[0] = __uint_as_float(r.a.x); v[1] = __uint_as_float(r.a.y);
    v[2] = __uint_as_float(r.a.z); v[3] = __uint_as_float(r.a.w);
    v[4] = __uint_as_float(r.b.x); v[5] = __uint_as_float(r.b.y);
    v[6] = __uint_as_float(r.b.z); v[7] = __uint_as_float(r.b.w);
  }
}

// accurate silu, out of line (rare path of the approximate form, see prologue)
static __device__ __noinline__ float silu_slow(float x) { return silu_f32<0>(x); }

// gate/up vectors -> the fp32 values to quantise (v) and, for 16-bit T, the same values still
// packed two per word (yw) -- the sign bits of the packed words patch the fp8 codes after an
// exact division (see pack_q8_signed).
//  PRO_SILU_T  : y = T(T(silu(gate)) * up)  -- the product and its rounding are one packed
//                HMUL2 (a product of two 8/11-bit significands is exact before the rounding)
//  PRO_SILU_F32: y = (gate * sigmoid(gate)) * up in fp32
template <typename T, int PRO, int FAST>
__device__ __forceinline__ void prologue(const RawV<T>& g, const RawV<T>& up, float (&v)[8], uint32_t (&yw)[4]) {
  if constexpr (PRO == PRO_NONE) {
    raw_to_f32<T>(g, v);
    if constexpr (sizeof(T) == 2) { yw[0] = g.a.x; yw[1] = g.a.y; yw[2] = g.a.z; yw[3] = g.a.w; }
  } else if constexpr (PRO == PRO_SILU_T && sizeof(T) == 2) {
    const uint32_t gw[4] = {g.a.x, g.a.y, g.a.z, g.a.w};
    const uint32_t uw[4] = {up.a.x, up.a.y, up.a.z, up.a.w};
    // The approximate form flushes 1 / (1 + e^-x) to zero once e^-x passes 2^126 (rcp.approx.ftz), i.e. for
    // gates in (-88.72, -87.34]: three bf16 values (-87.5, -88, -88.5) whose silu is a tiny normal number in
    // the reference.  A vector holding such a gate takes the accurate form (3 packed minima + 1 compare per
    // 8 elements; the branch is never taken on real activations).  Everything else is proven identical by
    // tests/test_silu_exhaustive_gpu.py.
    bool deep = false;
    if constexpr (FAST && std::is_same<T, __nv_bfloat16>::value) {
      using V = typename Pair<T>::V;
      const V m = __hmin2(__hmin2(*reinterpret_cast<const V*>(&gw[0]), *reinterpret_cast<const V*>(&gw[1])),
                          __hmin2(*reinterpret_cast<const V*>(&gw[2]), *reinterpret_cast<const V*>(&gw[3])));
      deep = !__hbgt2(m, __float2bfloat162_rn(-87.25f));       // some gate <= -87.25 (or NaN)
    }
    // (one branch per vector, not one per pair: the guard's BSSY / BRA / BSYNC triples were 5 % of all issued
    // instructions of the silu + 1 x 128 FP8 kernel, profiles/r02b_silu_block.csv)
    float sl[8];
    if constexpr (FAST) {
      if (__builtin_expect(deep, 0)) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 f = Pair<T>::unpack(gw[i]);
          sl[2 * i] = silu_slow(f.x);
          sl[2 * i + 1] = silu_slow(f.y);
        }
      } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 f = Pair<T>::unpack(gw[i]);
          silu2_approx(f.x, f.y, sl[2 * i], sl[2 * i + 1]);
        }
      }
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 f = Pair<T>::unpack(gw[i]);
        sl[2 * i] = silu_f32<0>(f.x);
        sl[2 * i + 1] = silu_f32<0>(f.y);
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float s0 = sl[2 * i], s1 = sl[2 * i + 1];
      const typename Pair<T>::V s2 = Pair<T>::pack(s0, s1);
      const typename Pair<T>::V y2 = __hmul2(s2, *reinterpret_cast<const typename Pair<T>::V*>(&uw[i]));
      yw[i] = *reinterpret_cast<const uint32_t*>(&y2);
      const float2 y = Pair<T>::unpack(yw[i]);
      v[2 * i] = y.x;
      v[2 * i + 1] = y.y;
    }
  } else {
    float u[8];
    raw_to_f32<T>(g, v);
    raw_to_f32<T>(up, u);
#pragma unroll
    for (int j = 0; j < 8; j += 2) {
      if constexpr (PRO == PRO_SILU_T) {
        v[j] = round_to<T>(round_to<T>(silu_f32<FAST>(v[j])) * u[j]);
        v[j + 1] = round_to<T>(round_to<T>(silu_f32<FAST>(v[j + 1])) * u[j + 1]);
      } else if constexpr (FAST) {
        silu2_approx(v[j], v[j + 1], v[j], v[j + 1]);
        mul2(v[j], v[j + 1], u[j], u[j + 1]);
      } else {
        v[j] = (v[j] * __fdiv_rn(1.0f, 1.0f + expf(-v[j]))) * u[j];
        v[j + 1] = (v[j + 1] * __fdiv_rn(1.0f, 1.0f + expf(-v[j + 1]))) * u[j + 1];
      }
    }
  }
}

// IEEE x / s for the vanishing-group case (s subnormal or zero, 1/s overflows).  Out of line and
// scalar on purpose: inlined, ptxas predicates the divide sequence into the hot path; by-reference
// arrays would force the hot path's values through local memory.
static __device__ __noinline__ float slow_div(float x, float s) { return __fdiv_rn(x, s); }

// eight exact quotients x/s (r = RN(1/s)) -> 8 fp8 codes.  Each quotient is the Markstein pair
// q0 = RN(x*r), q = RN(q0 + RN(x - q0*s)*r); a zero quotient can come out with the wrong sign,
// so the sign bits are taken from the packed 16-bit inputs (s > 0 => sign(x/s) = sign(x)):
// one PRMT gathers the four high bytes, one LOP3 merges them under the 0x80808080 mask.
template <int Q>
__device__ __forceinline__ uint2 div_pack_fp8_signed(const float (&v)[8], const uint32_t (&yw)[4], float s, float r) {
  float q[8];
  const float ns = -s;
#pragma unroll
  for (int k = 0; k < 8; k += 2) {
    q[k] = v[k];
    q[k + 1] = v[k + 1];
    mul2(q[k], q[k + 1], r, r);
    float r0, r1;
    fma2_to(r0, r1, q[k], q[k + 1], ns, ns, v[k], v[k + 1]);
    fma2(q[k], q[k + 1], r0, r1, r, r);
  }
  const uint2 c = pack_q8<Q>(q);
  const uint32_t s_lo = __byte_perm(yw[0], yw[1], 0x7531), s_hi = __byte_perm(yw[2], yw[3], 0x7531);
  return make_uint2(c.x ^ ((c.x ^ s_lo) & 0x80808080u), c.y ^ ((c.y ^ s_hi) & 0x80808080u));   // one LOP3 each
}


}  // namespace vxk
