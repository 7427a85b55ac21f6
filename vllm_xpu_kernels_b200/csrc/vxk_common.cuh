// Shared device/host helpers for the quantisation kernels (sm_100a only).
//
// Numerics contract (see DESIGN.md "Arithmetic"): this translation unit is
// compiled with -fmad=false, so the only fused multiply-adds are the explicit
// fmaf() calls below; every other float operation is a single IEEE
// round-to-nearest-even op, which is what the torch CPU reference does.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <mutex>

#include "../../include/vxk_b200.h"

namespace vxk {

// ------------------------------------------------------------------ host side
extern thread_local char g_last_error[512];
extern std::atomic<uint64_t> g_launches;
extern std::atomic<int> g_early_weight_loads;

inline int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
  va_end(ap);
  return code;
}

#define VXK_REQUIRE(cond, ...)                                   \
  do {                                                           \
    if (!(cond)) return ::vxk::fail(VXK_ERR_INVALID, __VA_ARGS__); \
  } while (0)

inline int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess)
    return fail(VXK_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return VXK_OK;
}

int sm_count();  // cached cudaDevAttrMultiProcessorCount of the current device

// One-time per-DEVICE configuration of a kernel (cudaFuncSetAttribute is per device: a process that
// calls an op on cuda:0 and then on cuda:1 must configure both).  One static instance per kernel
// instantiation; `init(entry)` runs once for each device ordinal that reaches it.
// MX scale factors in the layout tcgen05.cp.32x128b.warpx4 consumes ("packed", see sf_pack_kernel in gemm_tc.cu):
// 512-byte chunks [32 lanes][4 words]; word (lane, c) of chunk (kb, blk, w) holds the four E8M0 bytes of row
// blk * 128 + c * 32 + lane for the 32-element K groups kb * 4 * SF_WORDS + 4 w .. + 3; chunks ordered [kb][blk][w].
// SF_WORDS = 2 for MXFP4 (a K slice of 256 elements = 8 groups), 1 for MXFP8 (128 elements = 4 groups).
// nblk: 128-row blocks in the buffer = ceil(rows / 128) + 1 for activations, + 2 for weights (sf_pack_blocks).
inline int sf_pack_blocks(int64_t rows, int operand) { return static_cast<int>((rows + 127) / 128) + (operand ? 2 : 1); }
template <int SF_WORDS>
__host__ __device__ __forceinline__ int64_t sf_packed_offset(int64_t row, int g, int nblk) {
  constexpr int kGroups = 4 * SF_WORDS;
  const int kb = g / kGroups, w = (g % kGroups) >> 2, b = g & 3;
  const int64_t chunk = (static_cast<int64_t>(kb) * nblk + (row >> 7)) * SF_WORDS + w;
  return ((chunk * 128 + (row & 31) * 4 + ((row >> 5) & 3)) << 2) + b;
}

// Launch with the programmatic-stream-serialisation attribute (PDL): the kernel may be scheduled while the
// previous kernel in the stream is still draining; it must execute griddepcontrol.wait before its first global
// access that can depend on (or be read by) that kernel.  VXK_PDL=0 turns the attribute off everywhere.
inline bool pdl_enabled() {
  static const int on = [] { const char* e = getenv("VXK_PDL"); return e ? atoi(e) : 1; }();
  return on != 0;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr{};
  attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr.val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = &attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

struct PerDeviceOnce {
  struct Entry { bool done = false; cudaError_t err = cudaSuccess; int aux = 0; size_t smem = 0; };
  std::mutex mu;
  Entry e[64];
  template <typename F>
  Entry get(F&& init) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) dev = 0;
    std::lock_guard<std::mutex> g(mu);
    Entry& x = e[dev & 63];
    if (!x.done) { init(x); x.done = true; }
    return x;
  }
};

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// ---------------------------------------------------------------- device side
constexpr int kVec = 8;  // elements handled per "vector" (16 B of bf16/f16, 32 B of f32)

template <int Q> struct QTraits;
template <> struct QTraits<VXK_Q_E4M3> { static constexpr float kMax = 448.0f; };
template <> struct QTraits<VXK_Q_E5M2> { static constexpr float kMax = 57344.0f; };
template <> struct QTraits<VXK_Q_INT8> { static constexpr float kMax = 127.0f; };

__device__ __forceinline__ uint4 ldg_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void stg_stream(void* p, uint2 v) {
  asm volatile("st.global.L1::no_allocate.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(v.x), "r"(v.y)
               : "memory");
}
__device__ __forceinline__ void stg_stream(void* p, uint32_t v) {
  asm volatile("st.global.L1::no_allocate.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void stg_stream(void* p, uint4 v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

__device__ __forceinline__ float bf16_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }

template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) {
  return __bfloat162float(v);
}
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) {
  return __float2bfloat16_rn(v);
}
// round an fp32 value to T and back (the "stored as orig dtype, read back" steps of the refs)
template <typename T> __device__ __forceinline__ float round_to(float v) {
  return to_f32<T>(from_f32<T>(v));
}

// 8 consecutive elements -> fp32.  STREAM selects the non-coherent streaming
// load (read-only inputs); in-place tensors (residual) use plain loads.
template <typename T, bool STREAM = true>
__device__ __forceinline__ void load8(const T* __restrict__ p, float (&v)[8]) {
  if constexpr (sizeof(T) == 2) {
    uint4 r;
    if constexpr (STREAM) r = ldg_stream(p);
    else r = *reinterpret_cast<const uint4*>(p);
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if constexpr (std::is_same<T, __nv_bfloat16>::value) {
        v[2 * i] = bf16_lo(w[i]);
        v[2 * i + 1] = bf16_hi(w[i]);
      } else {
        const __half2 h = *reinterpret_cast<const __half2*>(&w[i]);
        const float2 f = __half22float2(h);
        v[2 * i] = f.x;
        v[2 * i + 1] = f.y;
      }
    }
  } else {
    uint4 a, b;
    if constexpr (STREAM) {
      a = ldg_stream(p);
      b = ldg_stream(p + 4);
    } else {
      a = *reinterpret_cast<const uint4*>(p);
      b = *reinterpret_cast<const uint4*>(p + 4);
    }
    v[0] = __uint_as_float(a.x); v[1] = __uint_as_float(a.y);
    v[2] = __uint_as_float(a.z); v[3] = __uint_as_float(a.w);
    v[4] = __uint_as_float(b.x); v[5] = __uint_as_float(b.y);
    v[6] = __uint_as_float(b.z); v[7] = __uint_as_float(b.w);
  }
}

// fp32 x8 -> T x8, one 16 B (or 2 x 16 B) store; used for the in-place residual
template <typename T>
__device__ __forceinline__ void store8(T* p, const float (&v)[8]) {
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
    *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
  } else {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
  }
}

// ---- packed saturating conversions (RNE, saturate-to-finite) ----
__device__ __forceinline__ uint32_t cvt_e4m3x2(float hi, float lo) {
  uint16_t r;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ uint32_t cvt_e5m2x2(float hi, float lo) {
  uint16_t r;
  asm("cvt.rn.satfinite.e5m2x2.f32 %0, %1, %2;" : "=h"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ uint32_t cvt_e2m1x2(float hi, float lo) {
  uint16_t r;
  asm("{ .reg .b8 t; cvt.rn.satfinite.e2m1x2.f32 t, %1, %2; cvt.u16.u8 %0, t; }"
      : "=h"(r)
      : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ uint32_t cvt_s8(float v) {
  // round half to even, clamp to [-128, 127]  (torch .round().clamp(-128,127).to(int8))
  int i = __float2int_rn(fminf(fmaxf(v, -128.0f), 127.0f));
  return static_cast<uint32_t>(i) & 0xffu;
}

template <int Q>
__device__ __forceinline__ uint2 pack_q8(const float (&v)[8]) {
  uint32_t lo, hi;
  if constexpr (Q == VXK_Q_E4M3) {
    lo = cvt_e4m3x2(v[1], v[0]) | (cvt_e4m3x2(v[3], v[2]) << 16);
    hi = cvt_e4m3x2(v[5], v[4]) | (cvt_e4m3x2(v[7], v[6]) << 16);
  } else if constexpr (Q == VXK_Q_E5M2) {
    lo = cvt_e5m2x2(v[1], v[0]) | (cvt_e5m2x2(v[3], v[2]) << 16);
    hi = cvt_e5m2x2(v[5], v[4]) | (cvt_e5m2x2(v[7], v[6]) << 16);
  } else {
    lo = cvt_s8(v[0]) | (cvt_s8(v[1]) << 8) | (cvt_s8(v[2]) << 16) | (cvt_s8(v[3]) << 24);
    hi = cvt_s8(v[4]) | (cvt_s8(v[5]) << 8) | (cvt_s8(v[6]) << 16) | (cvt_s8(v[7]) << 24);
  }
  return make_uint2(lo, hi);
}
template <int Q>
__device__ __forceinline__ uint8_t pack_q1(float v) {
  if constexpr (Q == VXK_Q_E4M3) return static_cast<uint8_t>(cvt_e4m3x2(0.0f, v) & 0xff);
  else if constexpr (Q == VXK_Q_E5M2) return static_cast<uint8_t>(cvt_e5m2x2(0.0f, v) & 0xff);
  else return static_cast<uint8_t>(cvt_s8(v));
}
// 8 fp32 -> 4 bytes of packed E2M1 (even element in the low nibble)
__device__ __forceinline__ uint32_t pack_e2m1x8(const float (&v)[8]) {
  return cvt_e2m1x2(v[1], v[0]) | (cvt_e2m1x2(v[3], v[2]) << 8) | (cvt_e2m1x2(v[5], v[4]) << 16) |
         (cvt_e2m1x2(v[7], v[6]) << 24);
}

// ---- packed fp32 pairs (Blackwell FMUL2 / FFMA2: two IEEE RN operations per issue slot) ----
// The quantisation kernels are issue-bound, not FLOP-bound: halving the instruction count of
// the element-wise multiplies is worth more than any scheduling tweak.  Results are bit-identical
// to the scalar mul.rn / fma.rn.
__device__ __forceinline__ void mul2(float& x, float& y, float a, float b) {
  asm("{\n\t.reg .b64 rx, ra;\n\tmov.b64 rx, {%0, %1};\n\tmov.b64 ra, {%2, %3};\n\t"
      "mul.rn.f32x2 rx, rx, ra;\n\tmov.b64 {%0, %1}, rx;\n\t}"
      : "+f"(x), "+f"(y)
      : "f"(a), "f"(b));
}
// (c0, c1) = (a0, a1) * (b0, b1) + (c0, c1)
__device__ __forceinline__ void fma2(float& c0, float& c1, float a0, float a1, float b0, float b1) {
  asm("{\n\t.reg .b64 rc, ra, rb;\n\tmov.b64 rc, {%0, %1};\n\tmov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\t"
      "fma.rn.f32x2 rc, ra, rb, rc;\n\tmov.b64 {%0, %1}, rc;\n\t}"
      : "+f"(c0), "+f"(c1)
      : "f"(a0), "f"(a1), "f"(b0), "f"(b1));
}
__device__ __forceinline__ void add2(float& x, float& y, float a, float b) {
  asm("{\n\t.reg .b64 rx, ra;\n\tmov.b64 rx, {%0, %1};\n\tmov.b64 ra, {%2, %3};\n\t"
      "add.rn.f32x2 rx, rx, ra;\n\tmov.b64 {%0, %1}, rx;\n\t}"
      : "+f"(x), "+f"(y)
      : "f"(a), "f"(b));
}
// (d0, d1) = (a0, a1) * (b0, b1) + (c0, c1)
__device__ __forceinline__ void fma2_to(float& d0, float& d1, float a0, float a1, float b0, float b1, float c0, float c1) {
  asm("{\n\t.reg .b64 rd, ra, rb, rc;\n\tmov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\tmov.b64 rc, {%6, %7};\n\t"
      "fma.rn.f32x2 rd, ra, rb, rc;\n\tmov.b64 {%0, %1}, rd;\n\t}"
      : "=f"(d0), "=f"(d1)
      : "f"(a0), "f"(a1), "f"(b0), "f"(b1), "f"(c0), "f"(c1));
}

// ---- exact division by a row/group constant ----
// r must be RN(1/s), s > 0.  q0 = RN(x*r) is within 2 ulp of x/s; one FMA residual +
// one FMA correction lands on RN(x/s) (Markstein) -- 3 ops instead of the ~10 of
// div.rn.f32.  The sign is re-applied from x (s > 0), which keeps -0 and tiny
// negative quotients that round to zero on the right side; SAT_MAX > 0 also clamps
// the magnitude (fminf drops the NaN that an overflowing x*r would produce, so the
// result is the saturated value the reference's clamp gives).
template <bool SAT>
__device__ __forceinline__ float div_by(float x, float s, float r, float sat_max = 0.0f) {
  const float q0 = x * r;
  const float rem = fmaf(-q0, s, x);
  float q1 = fabsf(fmaf(rem, r, q0));
  if constexpr (SAT) q1 = fminf(q1, sat_max);
  return copysignf(q1, x);
}

// the same Markstein division on a pair of values (3 packed ops + 2 sign patches)
template <bool SAT>
__device__ __forceinline__ void div_by2(float& x0, float& x1, float s, float r, float sat_max = 0.0f) {
  float q0 = x0, q1 = x1;
  mul2(q0, q1, r, r);
  float r0, r1;
  const float ns = -s;                       // x - q0*s as fma(q0, -s, x): the negation is loop-invariant
  fma2_to(r0, r1, q0, q1, ns, ns, x0, x1);
  fma2(q0, q1, r0, r1, r, r);
  q0 = fabsf(q0);
  q1 = fabsf(q1);
  if constexpr (SAT) {
    q0 = fminf(q0, sat_max);
    q1 = fminf(q1, sat_max);
  }
  x0 = copysignf(q0, x0);
  x1 = copysignf(q1, x1);
}

// x / C for a compile-time constant C > 0, correctly rounded (same Markstein step with
// r = RN(1/C) folded at compile time).  Used for amax / fp8_max.
template <int C_NUM>
__device__ __forceinline__ float div_const(float x) {
  constexpr float c = static_cast<float>(C_NUM);
  constexpr float r = 1.0f / c;   // constant-folded by the host compiler: IEEE RN
  const float q0 = x * r;
  const float rem = fmaf(-q0, c, x);
  return fmaf(rem, r, q0);
}
template <int Q>
__device__ __forceinline__ float div_qmax(float x) {
  if constexpr (Q == 0) return div_const<448>(x);
  else if constexpr (Q == 1) return div_const<57344>(x);
  else return div_const<127>(x);
}

// ---- UE8M0 helpers (exact, from the bit pattern) ----
// biased exponent byte of 2^ceil(log2(v)) clamped to [-127,127]; v >= 0
__device__ __forceinline__ int rceil_e8m0(float v) {
  const uint32_t b = __float_as_uint(v);
  const int exp = static_cast<int>((b >> 23) & 0xff);
  const uint32_t man = b & 0x7fffffu;
  int e;
  if (exp == 0) e = (man > 0x400000u) ? 1 : 0;  // subnormal / zero: 2^-126 or 2^-127
  else e = exp + (man != 0 ? 1 : 0);
  return min(e, 254);
}
// 2^(e-127) as fp32 (e = 0 -> the subnormal 2^-127)
__device__ __forceinline__ float e8m0_scale(int e) {
  return __uint_as_float(e == 0 ? 0x00400000u : (static_cast<uint32_t>(e) << 23));
}
// 2^(127-e) as fp32, 1.0 for e == 0 (mx_utils.to_mxfp :327-331)
__device__ __forceinline__ float e8m0_descale(int e) {
  if (e == 0) return 1.0f;
  return __uint_as_float(e == 254 ? 0x00400000u : (static_cast<uint32_t>(254 - e) << 23));
}
// 2^ceil(log2(s)) for positive normal s
__device__ __forceinline__ float exp2_ceil_log2(float s) {
  const uint32_t b = __float_as_uint(s);
  return __uint_as_float(((b >> 23) + ((b & 0x7fffffu) ? 1u : 0u)) << 23);
}

// ---- reductions ----
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// max over `lanes` (power of two <= 32) consecutive lanes
__device__ __forceinline__ float seg_max(float v, int lanes) {
  for (int o = lanes >> 1; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
template <int LANES>
__device__ __forceinline__ float seg_max_c(float v) {
#pragma unroll
  for (int o = LANES >> 1; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// runtime power-of-two lane count, dispatched to the unrolled versions
__device__ __forceinline__ float seg_max_pow2(float v, int lanes_log2) {
  switch (lanes_log2) {
    case 0: return v;
    case 1: return seg_max_c<2>(v);
    case 2: return seg_max_c<4>(v);
    case 3: return seg_max_c<8>(v);
    case 4: return seg_max_c<16>(v);
    default: return seg_max_c<32>(v);
  }
}
// CTA-wide reductions; `scratch` is >= 33 floats of shared memory; all threads get the result.
template <bool IS_MAX>
__device__ __forceinline__ float block_reduce(float v, float* scratch) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
  v = IS_MAX ? warp_max(v) : warp_sum(v);
  if (nwarp == 1) return v;
  __syncthreads();  // scratch may still be read from a previous reduction
  if (lane == 0) scratch[warp] = v;
  __syncthreads();
  float r = (lane < nwarp) ? scratch[lane] : (IS_MAX ? 0.0f : 0.0f);
  r = IS_MAX ? warp_max(r) : warp_sum(r);
  return r;
}

// ---- silu ----
// MODE 0 (accurate): expf (<= 2 ulp) and an IEEE divide, the same formula torch's CPU
// kernels evaluate (x / (1 + exp(-x))).  MODE 1: ex2.approx / rcp.approx (about 2^-22
// relative).  (Sharing one reciprocal between two elements to unload the XU pipe was
// measured slower on B200: the kernels are issue-bound, not XU-bound.)
__device__ __forceinline__ float ex2_approx(float t) {
  float e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(t));
  return e;
}
__device__ __forceinline__ float rcp_approx(float d) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d));
  return r;
}
// silu of two values, approximate mode: the multiplies and the 1 + e add are packed
__device__ __forceinline__ void silu2_approx(float x0, float x1, float& s0, float& s1) {
  float t0 = x0, t1 = x1;
  mul2(t0, t1, -1.4426950408889634f, -1.4426950408889634f);
  float e0 = ex2_approx(t0), e1 = ex2_approx(t1);
  add2(e0, e1, 1.0f, 1.0f);
  s0 = rcp_approx(e0);
  s1 = rcp_approx(e1);
  mul2(s0, s1, x0, x1);
}
template <int MODE>
__device__ __forceinline__ float silu_f32(float x) {
  if constexpr (MODE != 0) {
    return x * rcp_approx(1.0f + ex2_approx(x * -1.4426950408889634f));
  } else {
    return __fdiv_rn(x, 1.0f + expf(-x));
  }
}

}  // namespace vxk
