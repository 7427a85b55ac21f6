// Row-resident quantisation kernels (templates).  See quant_row.cu for the op list.
#pragma once
#include "vxk_common.cuh"

#include <cstdlib>

namespace vxk {

enum RowMode { MODE_TOKEN_DIV = 0, MODE_TOKEN_MUL = 1, MODE_GROUP = 2, MODE_MXFP4 = 3, MODE_STATIC = 4 };

struct RowParams {
  void* out;            // [T, H] codes (or [T, H/2] packed fp4)
  const void* in;       // [T, H], row stride in_row_stride
  const void* weight;   // [H] (rms only)
  void* residual;       // [T, H] in/out, nullable
  float* scales;        // per-token [T] or per-group (t*ss0 + g*ss1)
  const float* scale_in;  // scale_ub (token modes) or the static scale
  int64_t H, in_row_stride, ss0, ss1;
  int nvec;             // H / 8
  int group;            // group size (MODE_GROUP / MODE_MXFP4)
  int lanes_log2;       // log2(lanes that share a group) on the vector path
  int gv_log2;          // log2(vectors per group) = log2(group / 8)
  int ue8m0;
  float eps;
  int exact_var;        // rms: variance from an fp64 sum of squares (correctly rounded mean), see row_variance()
  double inv_H;         // 1.0 / H
  uint8_t* sf_packed;   // MODE_MXFP4, nullable: E8M0 bytes also go here in the GEMM's tcgen05.cp order
  int sf_nblk;          // 128-row blocks of that buffer (sf_pack_blocks(T, 0))
};

// Raw (still packed) copy of 8 input elements: 4 registers for bf16/f16, 8 for f32.
template <typename T> struct Raw8 { uint4 a; };
template <> struct Raw8<float> { uint4 a, b; };

template <typename T, bool STREAM>
__device__ __forceinline__ Raw8<T> load_raw(const T* p) {
  Raw8<T> r;
  if constexpr (STREAM) r.a = ldg_stream(p);
  else r.a = *reinterpret_cast<const uint4*>(p);
  if constexpr (sizeof(T) == 4) {
    if constexpr (STREAM) r.b = ldg_stream(p + 4);
    else r.b = *reinterpret_cast<const uint4*>(p + 4);
  }
  return r;
}
template <typename T>
__device__ __forceinline__ void unpack_raw(const Raw8<T>& r, float (&v)[8]) {
  if constexpr (sizeof(T) == 2) {
    const uint32_t w[4] = {r.a.x, r.a.y, r.a.z, r.a.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if constexpr (std::is_same<T, __nv_bfloat16>::value) {
        v[2 * i] = bf16_lo(w[i]);
        v[2 * i + 1] = bf16_hi(w[i]);
      } else {
        const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
        v[2 * i] = f.x;
        v[2 * i + 1] = f.y;
      }
    }
  } else {
    v[0] = __uint_as_float(r.a.x); v[1] = __uint_as_float(r.a.y);
    v[2] = __uint_as_float(r.a.z); v[3] = __uint_as_float(r.a.w);
    v[4] = __uint_as_float(r.b.x); v[5] = __uint_as_float(r.b.y);
    v[6] = __uint_as_float(r.b.z); v[7] = __uint_as_float(r.b.w);
  }
}
// fp32 x8 -> T x8 (RNE), packed
template <typename T>
__device__ __forceinline__ Raw8<T> pack_raw(const float (&v)[8]) {
  Raw8<T> r;
  if constexpr (sizeof(T) == 2) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if constexpr (std::is_same<T, __nv_bfloat16>::value) {
        __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
        w[i] = *reinterpret_cast<uint32_t*>(&h);
      } else {
        __half2 h = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
        w[i] = *reinterpret_cast<uint32_t*>(&h);
      }
    }
    r.a = make_uint4(w[0], w[1], w[2], w[3]);
  } else {
    r.a = make_uint4(__float_as_uint(v[0]), __float_as_uint(v[1]), __float_as_uint(v[2]), __float_as_uint(v[3]));
    r.b = make_uint4(__float_as_uint(v[4]), __float_as_uint(v[5]), __float_as_uint(v[6]), __float_as_uint(v[7]));
  }
  return r;
}
template <typename T>
__device__ __forceinline__ void store_raw(T* p, const Raw8<T>& r) {
  *reinterpret_cast<uint4*>(p) = r.a;
  if constexpr (sizeof(T) == 4) *reinterpret_cast<uint4*>(p + 4) = r.b;
}

// Opaque to the optimiser: stops it from keeping the 8 unpacked fp32 values of every
// vector live between the amax pass and the quantise pass (that would cost 8 registers
// per vector instead of 4 and halve the resident CTAs).
template <typename T>
__device__ __forceinline__ void keep_packed(Raw8<T>& r) {
  asm volatile("" : "+r"(r.a.x), "+r"(r.a.y), "+r"(r.a.z), "+r"(r.a.w));
  if constexpr (sizeof(T) == 4) asm volatile("" : "+r"(r.b.x), "+r"(r.b.y), "+r"(r.b.z), "+r"(r.b.w));
}

__device__ __forceinline__ double block_sum_f64(double v, double* scratch) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if (nwarp == 1) return v;
  __syncthreads();
  if (lane == 0) scratch[warp] = v;
  __syncthreads();
  double r = (lane < nwarp) ? scratch[lane] : 0.0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
  return r;
}

// CTA-wide reduction with ONE barrier: per-warp partials go to one of two scratch rows (the
// caller alternates `slot`), every thread then reads all partials back (broadcast loads) and
// combines them in the same order.  Two slots make the single barrier safe: a warp can only
// overwrite slot k after the barrier of the reduction in between.
template <bool IS_MAX>
__device__ __forceinline__ float block_reduce1(float v, float (*scratch)[32], int slot) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
  v = IS_MAX ? warp_max(v) : warp_sum(v);
  if (nwarp == 1) return v;
  if (lane == 0) scratch[slot][warp] = v;
  __syncthreads();
  // entries past nwarp hold 0 (the identity of both reductions here: sums and maxima of magnitudes)
  const float4* s4 = reinterpret_cast<const float4*>(scratch[slot]);
  auto comb = [](float a, float b) { return IS_MAX ? fmaxf(a, b) : a + b; };
  const float4 a = s4[0];
  float r = comb(comb(a.x, a.y), comb(a.z, a.w));
  if (nwarp > 4) {
    const float4 b = s4[1];
    r = comb(r, comb(comb(b.x, b.y), comb(b.z, b.w)));
    if (nwarp > 8) {
      const float4 c = s4[2], d = s4[3];
      r = comb(r, comb(comb(comb(c.x, c.y), comb(c.z, c.w)), comb(comb(d.x, d.y), comb(d.z, d.w))));
    }
  }
  return r;
}

// One-barrier CTA-wide fp64 sum (same two-slot protocol as block_reduce1).  The terms are squares of
// 8- / 11- / 24-bit significands, so an fp64 sum of <= 64k of them is exact to far below an fp32 ulp and the
// combination order does not matter.
__device__ __forceinline__ double block_sum1_f64(double v, double (*scratch)[32], int slot) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if (nwarp == 1) return v;
  if (lane == 0) scratch[slot][warp] = v;
  __syncthreads();
  const double2* s2 = reinterpret_cast<const double2*>(scratch[slot]);   // entries past nwarp hold 0
  double2 a = s2[0], b = s2[1];
  double r = (a.x + a.y) + (b.x + b.y);
  if (nwarp > 4) {
    a = s2[2]; b = s2[3];
    r += (a.x + a.y) + (b.x + b.y);
    if (nwarp > 8) {
#pragma unroll
      for (int k = 4; k < 8; ++k) { a = s2[k]; r += a.x + a.y; }
    }
  }
  return r;
}

// mean of squares, rounded once to fp32: float(sum / H) with the fp64 quotient taken as sum * (1/H)
// unless that product sits within 2 fp64 ulp of an fp32 rounding boundary (then the true quotient
// decides).  This is the value torch gives for x.double().pow(2).mean().float() -- the reference's
// formula (tests/test_fused_norm_quant.py:41-47) with the reduction carried out exactly instead of in
// whatever order the host's fp32 cascade sum happens to use.
__device__ __forceinline__ float mean_sq_f32(double sum, double inv_H, double H) {
  const double q = sum * inv_H;
  const uint32_t low = static_cast<uint32_t>(__double_as_longlong(q)) & 0x1fffffffu;
  if (__builtin_expect(low - 0x0ffffffeu <= 4u, 0)) return static_cast<float>(sum / H);
  return static_cast<float>(q);
}

// ---- exact sum of squares of 16-bit inputs without a conversion instruction ----------------------------
// The square of a bf16 / f16 value has at most 16 / 22 significant bits, so `x * x` in fp32 is exact (unless
// it leaves the normal fp32 range), and a finite non-negative fp32 bit pattern u, shifted into the fields of
// a double -- high word u >> 3, low word u << 29 -- IS the double  value * 2^-896  (both formats code exponent
// field 0 as "no hidden one", so zero and fp32 subnormals map exactly as well).  One packed FMUL2 per pair, one
// shift and one DADD per element replace F2F.F64.F32 + DFMA: the conversion runs on the quarter-rate XU pipe,
// which these kernels share with the output conversions (ncu, rms_norm + MXFP4: XU 32 %, math-pipe-throttle
// and short-scoreboard stalls on it).  For bf16 the low word is always 0 for a normal square.  It costs half an
// issue slot per element more than the conversions, so only the modes where it measured faster use it (kWidenSq).
// The scaled sum is trusted when it is 0 (every square was 0 or flushed: the mean rounds to 0 either way) or
// lies in [2^-955, 2^-768) = true sum in [2^-59, 2^128): below that, squares that fell into the fp32 subnormal
// range (|x| < 2^-63, error <= 2^-150 each) could matter; at or above it some square was Inf / NaN (those
// patterns widen to >= 2^-768).  Outside the window the caller re-does the row with real conversions.
template <typename T>
__device__ __forceinline__ double widen_sq(float sq) {
  const uint32_t u = __float_as_uint(sq);
  if constexpr (std::is_same<T, __nv_bfloat16>::value) return __hiloint2double(static_cast<int>(u >> 3), 0);
  else return __hiloint2double(static_cast<int>(u >> 3), static_cast<int>(u << 29));
}
__device__ __forceinline__ bool widened_sum_ok(double s) {
  return s == 0.0 || (s >= 0x1p-955 && s < 0x1p-768);
}
// mean_sq_f32 for a sum that carries the 2^-896 scale (power-of-two scaling commutes with every rounding
// here: nothing in the window is subnormal)
__device__ __forceinline__ float mean_sq_f32_widened(double sum, double inv_H, double H) {
  const double q = sum * inv_H;
  const uint32_t low = static_cast<uint32_t>(__double_as_longlong(q)) & 0x1fffffffu;
  if (__builtin_expect(low - 0x0ffffffeu <= 4u, 0)) return static_cast<float>((sum / H) * 0x1p896);
  return static_cast<float>(q * 0x1p896);
}

// Persistent CTAs (grid = min(T, resident CTAs)), each loops over rows.  A thread owns
// VPT vectors of a row (tid + i * blockDim); the row lives in registers as fp32, so every
// element is converted, normalised and scaled exactly once, and the NEXT row's 128-bit loads are
// issued (into the packed registers the current row has just vacated) before the current row's
// arithmetic, so HBM always has a row per CTA in flight.  These kernels are instruction-issue
// bound on B200 (HBM delivers one bf16 element per ~14 issue slots): element-wise multiplies and
// the sum of squares are packed fp32x2 operations, the rms weight is staged once per CTA as fp32
// in shared memory (two planes of float4: conflict-free 128-bit reads, no per-row unpacking), and
// each row reduction costs a single barrier.  FULL: nvec == VPT * blockDim (no tail predicates).
// SAT: clamp the quotient before the conversion (only needed when a scale upper bound can push
// |x / s| past the format maximum).
// VG (group modes): a thread's vectors come in runs of VG vectors of the SAME 1 x G group, so the
// group maximum needs log2(G / 8 / VG) shuffle steps and the scale arithmetic (exact division by
// the format maximum, reciprocal, E8M0 rounding) is done once per VG vectors instead of once per
// vector -- these kernels are issue-bound, and that per-vector scalar work was what kept the
// per-group variants at half the per-token variant's rate.  Vector index of (thread, i):
//   (i / VG) * (threads * VG) + (tid / lpg) * gv + (i % VG) * lpg + tid % lpg,   lpg = gv / VG
// (lpg adjacent lanes read lpg * 16 contiguous bytes: whole 32-byte sectors for lpg >= 2).
template <typename T, int Q, int MODE, bool RMS, bool HAS_RES, int VPT, bool FULL, int VG = 1>
__global__ void __launch_bounds__(512) row_quant_kernel(const RowParams p, const int64_t T_rows) {
  __shared__ __align__(16) float scratch[2][32];
  __shared__ __align__(16) double scratch64[2][32];
  extern __shared__ float4 wsm[];     // RMS: [2][VPT][threads] float4 (elements 0-3 / 4-7 of every vector)
  const int tid = threadIdx.x, nthr = blockDim.x;
  constexpr float kQMax = QTraits<Q>::kMax;
  const bool sat = p.scale_in != nullptr;
  const int nvec = p.nvec;
  int slot = 0;
  // Programmatic dependent launch: the next kernel in the stream may be scheduled as soon as this one's
  // CTAs retire (its launch latency and index set-up overlap our tail); our own global accesses -- inputs
  // AND outputs, which the previous kernel may still be reading -- start only after griddepcontrol.wait.
  pdl_launch_dependents();
  if (tid < 64) {
    scratch[tid >> 5][tid & 31] = 0.0f;
    if constexpr (RMS) scratch64[tid >> 5][tid & 31] = 0.0;
  }
  if (!RMS) __syncthreads();

  constexpr bool kGrouped = MODE == MODE_GROUP || MODE == MODE_MXFP4;
  // widen_sq() variance: measured per mode at 8192 x 4096 / 8192 bf16 (round 2): static 20.1 -> 19.1 / 36.9 -> 34.8 us,
  // MXFP4 27.1 -> 26.6 / 50.4 -> 49.5 us; the per-token and 1 x G FP8 kernels sit exactly at their 128-register bound
  // and lose to the extra live values (20.7 -> 21.3 us, 25.3 -> 27.5 us with spills), so they keep the conversions
  constexpr bool kWidenSq = MODE == MODE_STATIC || MODE == MODE_MXFP4;
  static_assert(VG == 1 || kGrouped, "VG only applies to the per-group modes");
  static_assert(VPT % VG == 0, "whole runs per thread");
  // vector index of (thread, i), see above; VG == 1 without groups degenerates to tid + i * nthr
  int vidx[VPT];
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    if constexpr (kGrouped) {
      const int lpg_log2 = p.lanes_log2;
      vidx[i] = (i / VG) * (nthr * VG) + ((tid >> lpg_log2) << p.gv_log2) + ((i % VG) << lpg_log2) +
                (tid & ((1 << lpg_log2) - 1));
    } else {
      vidx[i] = tid + i * nthr;
    }
  }
  bool ok[VPT];
#pragma unroll
  for (int i = 0; i < VPT; ++i) ok[i] = FULL || (vidx[i] < nvec);

  // per-thread element offsets of its vectors and row pointers that advance by a constant:
  // no per-row 64-bit address arithmetic in the loop
  int voff[VPT];
#pragma unroll
  for (int i = 0; i < VPT; ++i) voff[i] = vidx[i] * kVec;
  const int64_t in_step = static_cast<int64_t>(gridDim.x) * p.in_row_stride;
  const int64_t out_step = static_cast<int64_t>(gridDim.x) * (MODE == MODE_MXFP4 ? p.H / 2 : p.H);
  const T* __restrict__ xnext = static_cast<const T*>(p.in) + static_cast<int64_t>(blockIdx.x) * p.in_row_stride;
  uint8_t* orow = static_cast<uint8_t*>(p.out) + static_cast<int64_t>(blockIdx.x) * (MODE == MODE_MXFP4 ? p.H / 2 : p.H);

  pdl_wait();
  // first row: requested before the weight staging below so that its HBM latency overlaps it
  Raw8<T> raw[VPT];
#pragma unroll
  for (int i = 0; i < VPT; ++i)
    raw[i] = (static_cast<int64_t>(blockIdx.x) < T_rows && ok[i]) ? load_raw<T, true>(xnext + voff[i]) : Raw8<T>{};

  // rms weight, staged once per CTA as fp32 in THREAD order: slot i * nthr + tid holds the weight
  // vector this thread multiplies its i-th vector with (two planes: elements 0-3 / 4-7), so the
  // per-row reads are conflict-free 128-bit loads whatever the vector-to-thread mapping is
  if constexpr (RMS) {
    const T* __restrict__ w = static_cast<const T*>(p.weight);
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      float wv[8] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
      if (ok[i]) load8<T, false>(w + vidx[i] * kVec, wv);
      wsm[i * nthr + tid] = make_float4(wv[0], wv[1], wv[2], wv[3]);
      wsm[(VPT + i) * nthr + tid] = make_float4(wv[4], wv[5], wv[6], wv[7]);
    }
    __syncthreads();
  }

  for (int64_t row = blockIdx.x; row < T_rows; row += gridDim.x, orow += out_step) {
    float v[VPT][8];
    if constexpr (RMS && HAS_RES) {
      T* res = static_cast<T*>(p.residual) + row * p.H;
      Raw8<T> rr[VPT];
#pragma unroll
      for (int i = 0; i < VPT; ++i) rr[i] = ok[i] ? load_raw<T, false>(res + voff[i]) : Raw8<T>{};
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        float b[8];
        unpack_raw<T>(raw[i], v[i]);
        unpack_raw<T>(rr[i], b);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[i][j] = v[i][j] + b[j];
        const Raw8<T> z = pack_raw<T>(v[i]);  // z = T(x + res): new residual and norm input
        if (ok[i]) store_raw<T>(res + voff[i], z);
        unpack_raw<T>(z, v[i]);
      }
    } else {
#pragma unroll
      for (int i = 0; i < VPT; ++i) unpack_raw<T>(raw[i], v[i]);
    }
    xnext += in_step;   // prefetch the next row of this CTA into the packed registers just vacated
    if (row + gridDim.x < T_rows) {
#pragma unroll
      for (int i = 0; i < VPT; ++i)
        if (ok[i]) raw[i] = load_raw<T, true>(xnext + voff[i]);
    }

    if constexpr (RMS) {
      float var;
      if (p.exact_var) {
        bool have = false;
        var = 0.0f;
        if constexpr (sizeof(T) == 2 && kWidenSq) {
          // exact fp32 squares widened with integer operations (see widen_sq)
          double d0 = 0.0, d1 = 0.0;
#pragma unroll
          for (int i = 0; i < VPT; ++i)
#pragma unroll
            for (int j = 0; j < 8; j += 2) {
              float a = v[i][j], b = v[i][j + 1];
              mul2(a, b, a, b);
              d0 += widen_sq<T>(a);
              d1 += widen_sq<T>(b);
            }
          const double s = block_sum1_f64(d0 + d1, scratch64, slot);
          have = widened_sum_ok(s);   // the same value in every thread: a CTA-uniform branch
          if (have) var = mean_sq_f32_widened(s, p.inv_H, static_cast<double>(p.H));
          else slot ^= 1;             // a second reduction follows: it must not reuse this one's scratch row
        }
        if (!have) {
          // fp64 sum of the (exact) squares: two accumulators, 1 conversion + 1 DFMA per element
          double d0 = 0.0, d1 = 0.0;
#pragma unroll
          for (int i = 0; i < VPT; ++i)
#pragma unroll
            for (int j = 0; j < 8; j += 2) {
              const double a = static_cast<double>(v[i][j]), b = static_cast<double>(v[i][j + 1]);
              d0 = fma(a, a, d0);
              d1 = fma(b, b, d1);
            }
          var = mean_sq_f32(block_sum1_f64(d0 + d1, scratch64, slot), p.inv_H, static_cast<double>(p.H));
        }
      } else {
        float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
        for (int i = 0; i < VPT; ++i)
#pragma unroll
          for (int j = 0; j < 8; j += 2) fma2(s0, s1, v[i][j], v[i][j + 1], v[i][j], v[i][j + 1]);
        var = __fdiv_rn(block_reduce1<false>(s0 + s1, scratch, slot), static_cast<float>(p.H));
      }
      slot ^= 1;
      const float inv_rms = __fdiv_rn(1.0f, __fsqrt_rn(var + p.eps));
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        const float4 wa = wsm[i * nthr + tid], wb = wsm[(VPT + i) * nthr + tid];
        mul2(v[i][0], v[i][1], inv_rms, inv_rms);
        mul2(v[i][2], v[i][3], inv_rms, inv_rms);
        mul2(v[i][4], v[i][5], inv_rms, inv_rms);
        mul2(v[i][6], v[i][7], inv_rms, inv_rms);
        mul2(v[i][0], v[i][1], wa.x, wa.y);
        mul2(v[i][2], v[i][3], wa.z, wa.w);
        mul2(v[i][4], v[i][5], wb.x, wb.y);
        mul2(v[i][6], v[i][7], wb.z, wb.w);
      }
    }

    if constexpr (MODE == MODE_TOKEN_DIV || MODE == MODE_TOKEN_MUL) {
      float amax = 0.0f;
#pragma unroll
      for (int i = 0; i < VPT; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) amax = fmaxf(amax, fabsf(v[i][j]));
      amax = block_reduce1<true>(amax, scratch, slot);
      slot ^= 1;
      float s;
      if constexpr (Q == VXK_Q_INT8) {
        s = amax > 0.0f ? div_qmax<Q>(amax) : 1.0f;
      } else {
        if (sat) amax = fminf(amax, __ldg(p.scale_in));
        s = fmaxf(div_qmax<Q>(amax), 1.0f / (kQMax * 512.0f));
      }
      const float r = __frcp_rn(s);
      if (tid == 0) p.scales[row] = s;
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        if constexpr (Q == VXK_Q_INT8 || MODE == MODE_TOKEN_DIV) {
          if (Q != VXK_Q_INT8 && sat) {
#pragma unroll
            for (int j = 0; j < 8; j += 2) div_by2<true>(v[i][j], v[i][j + 1], s, r, kQMax);
          } else {
#pragma unroll
            for (int j = 0; j < 8; j += 2) div_by2<false>(v[i][j], v[i][j + 1], s, r);
          }
        } else {
#pragma unroll
          for (int j = 0; j < 8; j += 2) mul2(v[i][j], v[i][j + 1], r, r);
        }
        if (ok[i]) stg_stream(orow + voff[i], pack_q8<Q>(v[i]));
      }
    } else if constexpr (MODE == MODE_STATIC) {
      const float r = __frcp_rn(__ldg(p.scale_in));
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
#pragma unroll
        for (int j = 0; j < 8; j += 2) mul2(v[i][j], v[i][j + 1], r, r);
        if (ok[i]) stg_stream(orow + voff[i], pack_q8<Q>(v[i]));
      }
    } else {  // MODE_GROUP / MODE_MXFP4: 1 x G groups; VG vectors per thread and 2^lanes_log2 lanes per group
      // MXFP4 groups are 32 wide (4 vectors): the lane count follows from VG at compile time
      const int lanes_log2 = MODE == MODE_MXFP4 ? (VG == 1 ? 2 : (VG == 2 ? 1 : 0)) : p.lanes_log2;
      const bool lead_lane = (tid & ((1 << lanes_log2) - 1)) == 0;
#pragma unroll
      for (int gi = 0; gi < VPT / VG; ++gi) {
        const int i0 = gi * VG;
        const bool valid = ok[i0];              // H % G == 0: a group is valid or invalid as a whole
        float gmax = 0.0f;
#pragma unroll
        for (int i = i0; i < i0 + VG; ++i)
#pragma unroll
          for (int j = 0; j < 8; ++j) gmax = fmaxf(gmax, fabsf(v[i][j]));
        gmax = seg_max_pow2(gmax, lanes_log2);  // all lanes participate (invalid ones carry 0)
        const int g = vidx[i0] >> p.gv_log2;
        const bool leader = valid && lead_lane;
        if constexpr (MODE == MODE_MXFP4) {
          const int e = rceil_e8m0(div_const<6>(gmax));
          const float d = e8m0_descale(e);
          if (leader) {
            p.scales[row * p.ss0 + g * p.ss1] = e8m0_scale(e);
            if (p.sf_packed) p.sf_packed[sf_packed_offset<2>(row, g, p.sf_nblk)] = static_cast<uint8_t>(e);
          }
          uint32_t code[VG];
#pragma unroll
          for (int i = i0; i < i0 + VG; ++i) {
            float q[8];
#pragma unroll
            for (int j = 0; j < 8; j += 2) {
              mul2(v[i][j], v[i][j + 1], d, d);
              // the reference's intermediate rounding to bf16; its clamp to +-6 before that is what
              // the saturating conversion below does anyway (RNE to bf16 is monotonic)
              const __nv_bfloat162 b2 = __floats2bfloat162_rn(v[i][j], v[i][j + 1]);
              const uint32_t wd = *reinterpret_cast<const uint32_t*>(&b2);
              q[j] = bf16_lo(wd);
              q[j + 1] = bf16_hi(wd);
            }
            code[i - i0] = pack_e2m1x8(q);
          }
          if constexpr (VG == 2) {
            // G = 32: a lane pair holds vectors {0, 2} / {1, 3} of the group (interleaved so that the
            // 128-bit loads fill whole sectors); swap one code word so that each lane stores 8
            // contiguous bytes (vectors {0, 1} / {2, 3}) instead of two half-filled sectors
            const bool odd = tid & 1;
            const uint32_t got = __shfl_xor_sync(0xffffffffu, odd ? code[0] : code[1], 1);
            const uint2 o2 = odd ? make_uint2(got, code[1]) : make_uint2(code[0], got);
            if (valid) stg_stream(orow + ((voff[i0] - (odd ? kVec : 0)) >> 1) + (odd ? 8 : 0), o2);
          } else {
#pragma unroll
            for (int i = i0; i < i0 + VG; ++i)
              if (valid) stg_stream(orow + (voff[i] >> 1), code[i - i0]);
          }
        } else {
          float s;
          if constexpr (Q == VXK_Q_INT8) {
            s = gmax > 0.0f ? div_qmax<Q>(gmax) : 1.0f;
          } else {
            s = fmaxf(div_qmax<Q>(gmax), 1.0f / (kQMax * 512.0f));
            if (p.ue8m0) s = exp2_ceil_log2(fmaxf(fabsf(s), 1e-10f));
          }
          const float r = __frcp_rn(s);
          if (leader) p.scales[row * p.ss0 + g * p.ss1] = s;
#pragma unroll
          for (int i = i0; i < i0 + VG; ++i) {
#pragma unroll
            for (int j = 0; j < 8; j += 2) {
              if constexpr (Q == VXK_Q_INT8) div_by2<false>(v[i][j], v[i][j + 1], s, r);
              else mul2(v[i][j], v[i][j + 1], r, r);
            }
            if (valid) stg_stream(orow + voff[i], pack_q8<Q>(v[i]));
          }
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------
// Generic fallback: any H, any alignment.  One CTA per row, scalar accesses,
// the row is re-read from L2 for each pass.  Same arithmetic as above.
// ---------------------------------------------------------------------------
template <typename T, int Q, int MODE, bool RMS>
__global__ void __launch_bounds__(256) row_quant_generic(const RowParams p) {
  __shared__ double scratch64[32];
  float* scratch = reinterpret_cast<float*>(scratch64);
  const int64_t row = blockIdx.x;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const T* __restrict__ xin = static_cast<const T*>(p.in) + row * p.in_row_stride;
  T* res = p.residual ? static_cast<T*>(p.residual) + row * p.H : nullptr;
  const T* __restrict__ w = static_cast<const T*>(p.weight);
  const int H = static_cast<int>(p.H);
  constexpr float kQMax = QTraits<Q>::kMax;

  float inv_rms = 1.0f;
  if constexpr (RMS) {
    float ss = 0.0f;
    double dd = 0.0;
    for (int h = tid; h < H; h += nthr) {
      float x = to_f32<T>(xin[h]);
      if (res) {
        x = round_to<T>(x + to_f32<T>(res[h]));
        res[h] = from_f32<T>(x);
      }
      if (p.exact_var) dd = fma(static_cast<double>(x), static_cast<double>(x), dd);
      else ss = fmaf(x, x, ss);
    }
    float var;
    if (p.exact_var) var = mean_sq_f32(block_sum_f64(dd, scratch64), p.inv_H, static_cast<double>(H));
    else var = __fdiv_rn(block_reduce<false>(ss, scratch), static_cast<float>(H));
    inv_rms = __fdiv_rn(1.0f, __fsqrt_rn(var + p.eps));
  }
  // value to quantise at column h (after the prologue)
  auto value = [&](int h) -> float {
    if constexpr (RMS) {
      const float x = res ? to_f32<T>(res[h]) : to_f32<T>(xin[h]);
      return (x * inv_rms) * to_f32<T>(w[h]);
    } else {
      return to_f32<T>(xin[h]);
    }
  };

  if constexpr (MODE == MODE_TOKEN_DIV || MODE == MODE_TOKEN_MUL) {
    float amax = 0.0f;
    for (int h = tid; h < H; h += nthr) amax = fmaxf(amax, fabsf(value(h)));
    amax = block_reduce<true>(amax, scratch);
    float s;
    if constexpr (Q == VXK_Q_INT8) {
      s = amax > 0.0f ? __fdiv_rn(amax, 127.0f) : 1.0f;
    } else {
      if (p.scale_in != nullptr) amax = fminf(amax, __ldg(p.scale_in));
      s = fmaxf(__fdiv_rn(amax, kQMax), __fdiv_rn(1.0f, kQMax * 512.0f));
    }
    const float r = __fdiv_rn(1.0f, s);
    if (tid == 0) p.scales[row] = s;
    uint8_t* o = static_cast<uint8_t*>(p.out) + row * p.H;
    for (int h = tid; h < H; h += nthr) {
      const float x = value(h);
      const float q = (Q == VXK_Q_INT8) ? div_by<false>(x, s, r) : (MODE == MODE_TOKEN_DIV ? div_by<true>(x, s, r, kQMax) : x * r);
      o[h] = pack_q1<Q>(q);
    }
  } else if constexpr (MODE == MODE_STATIC) {
    const float r = __fdiv_rn(1.0f, __ldg(p.scale_in));
    uint8_t* o = static_cast<uint8_t*>(p.out) + row * p.H;
    for (int h = tid; h < H; h += nthr) o[h] = pack_q1<Q>(value(h) * r);
  } else {
    // one warp per group, round-robin over the row's groups
    const int G = p.group, ngroups = H / G;
    const int lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
    for (int g = warp; g < ngroups; g += nwarp) {
      float gmax = 0.0f;
      for (int h = g * G + lane; h < (g + 1) * G; h += 32) gmax = fmaxf(gmax, fabsf(value(h)));
      gmax = warp_max(gmax);
      if constexpr (MODE == MODE_MXFP4) {
        const int e = rceil_e8m0(__fdiv_rn(gmax, 6.0f));
        const float d = e8m0_descale(e);
        if (lane == 0) {
          p.scales[row * p.ss0 + g * p.ss1] = e8m0_scale(e);
          if (p.sf_packed) p.sf_packed[sf_packed_offset<2>(row, g, p.sf_nblk)] = static_cast<uint8_t>(e);
        }
        uint8_t* o = static_cast<uint8_t*>(p.out) + row * (p.H / 2);
        for (int h2 = g * G / 2 + lane; h2 < (g + 1) * G / 2; h2 += 32) {
          float c0 = fminf(fmaxf(value(2 * h2) * d, -6.0f), 6.0f);
          float c1 = fminf(fmaxf(value(2 * h2 + 1) * d, -6.0f), 6.0f);
          c0 = __bfloat162float(__float2bfloat16_rn(c0));
          c1 = __bfloat162float(__float2bfloat16_rn(c1));
          o[h2] = static_cast<uint8_t>(cvt_e2m1x2(c1, c0));
        }
      } else {
        float s;
        if constexpr (Q == VXK_Q_INT8) {
          s = gmax > 0.0f ? __fdiv_rn(gmax, 127.0f) : 1.0f;
        } else {
          s = fmaxf(__fdiv_rn(gmax, kQMax), __fdiv_rn(1.0f, kQMax * 512.0f));
          if (p.ue8m0) s = exp2_ceil_log2(fmaxf(fabsf(s), 1e-10f));
        }
        const float r = __fdiv_rn(1.0f, s);
        if (lane == 0) p.scales[row * p.ss0 + g * p.ss1] = s;
        uint8_t* o = static_cast<uint8_t*>(p.out) + row * p.H;
        for (int h = g * G + lane; h < (g + 1) * G; h += 32) {
          const float x = value(h);
          o[h] = pack_q1<Q>(Q == VXK_Q_INT8 ? div_by<false>(x, s, r) : x * r);
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------
// host dispatch
// ---------------------------------------------------------------------------
template <typename K>
inline int resident_ctas(K kernel, int threads) {
  int n = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, threads, 0) != cudaSuccess || n < 1) n = 1;
  return n * sm_count();
}

template <typename T, int Q, int MODE, bool RMS, bool HAS_RES, int VPT, bool FULL, int VG>
inline int launch_row_vpt3(const RowParams& p, int64_t T_rows, int threads, cudaStream_t st) {
  auto kernel = row_quant_kernel<T, Q, MODE, RMS, HAS_RES, VPT, FULL, VG>;
  const size_t smem = RMS ? static_cast<size_t>(VPT) * threads * 32 : 0;
  if (smem > 48 * 1024) {
    // per instantiation and per device: raise the limit once to the most this kernel can ever ask for
    static PerDeviceOnce once;
    const PerDeviceOnce::Entry c = once.get([&](PerDeviceOnce::Entry& x) {
      int dev = 0, optin = 0;
      cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
      x.smem = static_cast<size_t>(optin > 2048 ? optin - 2048 : 0);   // static scratch lives in the same budget
      x.err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(x.smem));
      if (x.err != cudaSuccess) cudaGetLastError();
    });
    if (c.err != cudaSuccess || smem > c.smem) return 1;   // row too long for the staged weight: caller falls back
  }
  static thread_local int c_threads = 0, c_per_sm = 0;   // per instantiation; smem follows from nvec = f(threads) mostly
  static thread_local size_t c_smem = 0;
  int per_sm = c_per_sm;
  if (c_threads != threads || c_smem != smem || per_sm < 1) {
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem) != cudaSuccess || per_sm < 1) per_sm = 1;
    c_threads = threads; c_smem = smem; c_per_sm = per_sm;
  }
  const int64_t cap = static_cast<int64_t>(per_sm) * sm_count();
  const unsigned grid = static_cast<unsigned>(T_rows < cap ? T_rows : cap);
  launch_pdl(kernel, dim3(grid), dim3(static_cast<unsigned>(threads)), smem, st, p, T_rows);
  return 0;
}
template <typename T, int Q, int MODE, bool RMS, bool HAS_RES, int VPT, int VG>
inline int launch_row_vpt2(const RowParams& p, int64_t T_rows, int threads, cudaStream_t st) {
  if (p.nvec == VPT * threads) return launch_row_vpt3<T, Q, MODE, RMS, HAS_RES, VPT, true, VG>(p, T_rows, threads, st);
  return launch_row_vpt3<T, Q, MODE, RMS, HAS_RES, VPT, false, VG>(p, T_rows, threads, st);
}
template <typename T, int Q, int MODE, bool RMS, int VPT, int VG>
inline int launch_row_vg(const RowParams& p, int64_t T_rows, int threads, cudaStream_t st) {
  if constexpr (RMS) {
    if (p.residual != nullptr) return launch_row_vpt2<T, Q, MODE, RMS, true, VPT, VG>(p, T_rows, threads, st);
  }
  return launch_row_vpt2<T, Q, MODE, RMS, false, VPT, VG>(p, T_rows, threads, st);
}
// vg = vectors of one group per thread: 1 outside the group modes, else min(VPT, 4, (G / 8) / 2)
template <typename T, int Q, int MODE, bool RMS, int VPT>
inline int launch_row_vpt(const RowParams& p, int64_t T_rows, int threads, int vg, cudaStream_t st) {
  if constexpr ((MODE == MODE_GROUP || MODE == MODE_MXFP4) && VPT >= 2) {
    if (vg == 2) return launch_row_vg<T, Q, MODE, RMS, VPT, 2>(p, T_rows, threads, st);
    if constexpr (VPT >= 4) {
      if (vg == 4) return launch_row_vg<T, Q, MODE, RMS, VPT, 4>(p, T_rows, threads, st);
    }
  }
  return launch_row_vg<T, Q, MODE, RMS, VPT, 1>(p, T_rows, threads, st);
}

template <typename T, int Q, int MODE, bool RMS>
int launch_row(const RowParams& p_in, int64_t T_rows, bool vec_ok, cudaStream_t st) {
  RowParams p = p_in;
  static const int env_exact = [] { const char* e = getenv("VXK_RMS_EXACT_VAR"); return e ? atoi(e) : 1; }();
  p.exact_var = env_exact;
  p.inv_H = p.H > 0 ? 1.0 / static_cast<double>(p.H) : 0.0;
  p.gv_log2 = 0;
  while (p.group > 0 && (kVec << p.gv_log2) < p.group) ++p.gv_log2;
  if (T_rows == 0 || p.H == 0) return VXK_OK;
  const int nvec = p.nvec;
  // per-thread vectors: enough to keep one warp busy; two register copies of the row are live (current +
  // prefetched).  rms kernels take 8 (the per-row scalar work -- reductions, 1/sqrt, scale arithmetic -- is paid
  // per WARP, so fewer, fatter warps per row cut the instruction count: measured 45.9 -> 38.6 us at 8192 x 8192,
  // profiles/r02_row_vpt_sweep.txt; 16 vectors per thread fill the register file and measured slower), the plain
  // per-token kernel 4 (no difference there)
  constexpr int kMaxVpt = 8;
  static const int env_vpt = [] { const char* e = getenv("VXK_ROW_VPT"); return e ? atoi(e) : 0; }();
  int want = env_vpt > 0 ? env_vpt : ((RMS && sizeof(T) == 2) ? 8 : 4);
  if (env_vpt <= 0 && want == 8 && nvec <= 640) {
    // measured at T = 8192 (profiles/r02c_row_vpt_by_hidden.txt): up to H = 5120 the per-group kernels are faster
    // with 4 vectors per thread and twice the warps (H = 2048: 14.9 -> 13.9 us 1x128, 16.2 -> 14.2 us MXFP4; 4096:
    // 25.4 -> 24.2, 26.8 -> 26.1; 5120: 33.7 -> 31.0, 39.5 -> 34.8), and so is every mode when 8 vectors per thread
    // would leave a partly filled last warp but 4 would not (H = 5120: 80 -> 96 threads with tail predicates
    // against 160 full ones: per-token 31.3 -> 30.8, static 29.1 -> 25.6).  Longer rows keep 8.
    const bool grouped = MODE == MODE_GROUP || MODE == MODE_MXFP4;
    if (grouped || (nvec % 256 != 0 && nvec % 128 == 0)) want = 4;
  }
  int vpt = 1;
  while (vpt < want && vpt < kMaxVpt && nvec > 32 * vpt) vpt *= 2;
  while ((nvec + vpt - 1) / vpt > 256 && vpt < 8) vpt *= 2;  // long rows: up to 8 x 512 threads
  if (vpt > kMaxVpt) vpt = kMaxVpt;
  const int threads = ((nvec + vpt - 1) / vpt + 31) / 32 * 32;
  // group modes: runs of vg vectors of one group per thread, 2^lanes_log2 adjacent lanes per group
  int vg = 1;
  if (MODE == MODE_GROUP || MODE == MODE_MXFP4) {
    static const int env_vg = [] { const char* e = getenv("VXK_ROW_VG"); return e ? atoi(e) : 0; }();
    const int gv = 1 << p.gv_log2;
    while (vg * 2 <= vpt && vg * 2 <= 4 && vg * 2 <= gv / 2) vg *= 2;
    if (env_vg == 1 || (env_vg == 2 && vg >= 2)) vg = env_vg;
    int l = 0;
    while ((vg << l) < gv) ++l;
    p.lanes_log2 = l;
  } else {
    p.lanes_log2 = 0;
  }
  int not_launched = 1;
  if (vec_ok && threads <= 512) {
    switch (vpt) {
      case 1: not_launched = launch_row_vpt<T, Q, MODE, RMS, 1>(p, T_rows, threads, vg, st); break;
      case 2: not_launched = launch_row_vpt<T, Q, MODE, RMS, 2>(p, T_rows, threads, vg, st); break;
      case 4: not_launched = launch_row_vpt<T, Q, MODE, RMS, 4>(p, T_rows, threads, vg, st); break;
      default: not_launched = launch_row_vpt<T, Q, MODE, RMS, 8>(p, T_rows, threads, vg, st); break;
    }
  }
  if (not_launched) {
    if (T_rows > 0x7fffffffLL) return fail(VXK_ERR_INVALID, "too many rows (%lld)", (long long)T_rows);
    row_quant_generic<T, Q, MODE, RMS><<<static_cast<unsigned>(T_rows), 256, 0, st>>>(p);
  }
  return check_launch("row_quant");
}


// One entry per (op, quant dtype); instantiated per activation type in quant_row_<type>.cu
enum RowOp { ROW_A1 = 0, ROW_A6 = 1, ROW_A7 = 2, ROW_A8 = 3, ROW_A9 = 4 };

template <typename T>
int row_launch(int op, int q_dtype, const RowParams& p, int64_t T_rows, bool vec_ok, cudaStream_t st) {
  switch (op) {
    case ROW_A1:
      if (q_dtype == VXK_Q_E4M3) return launch_row<T, VXK_Q_E4M3, MODE_TOKEN_DIV, false>(p, T_rows, vec_ok, st);
      return launch_row<T, VXK_Q_E5M2, MODE_TOKEN_DIV, false>(p, T_rows, vec_ok, st);
    case ROW_A6:
      if (q_dtype == VXK_Q_E4M3) return launch_row<T, VXK_Q_E4M3, MODE_TOKEN_MUL, true>(p, T_rows, vec_ok, st);
      return launch_row<T, VXK_Q_INT8, MODE_TOKEN_MUL, true>(p, T_rows, vec_ok, st);
    case ROW_A7:
      if (q_dtype == VXK_Q_E4M3) return launch_row<T, VXK_Q_E4M3, MODE_GROUP, true>(p, T_rows, vec_ok, st);
      return launch_row<T, VXK_Q_INT8, MODE_GROUP, true>(p, T_rows, vec_ok, st);
    case ROW_A8:
      return launch_row<T, VXK_Q_E4M3, MODE_MXFP4, true>(p, T_rows, vec_ok, st);
    case ROW_A9:
      if (q_dtype == VXK_Q_E4M3) return launch_row<T, VXK_Q_E4M3, MODE_STATIC, true>(p, T_rows, vec_ok, st);
      return launch_row<T, VXK_Q_E5M2, MODE_STATIC, true>(p, T_rows, vec_ok, st);
    default: return fail(VXK_ERR_INVALID, "bad row op %d", op);
  }
}

}  // namespace vxk
