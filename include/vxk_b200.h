/*
 * vxk_b200.h -- C ABI of libvxk_b200.so, the B200 (sm_100a) implementation of
 * the quantized-linear hot path of vllm-xpu-kernels.
 *
 * One entry point per reference operator on the path.  Every pointer is a
 * DEVICE pointer unless stated otherwise; no torch / C++ types cross this
 * boundary.  All calls are asynchronous on `stream` and allocate nothing
 * (the GEMMs take their output buffer from the caller).  Return value: 0 on
 * success, a negative VXK_ERR_* otherwise; vxk_last_error() returns a
 * thread-local message for the last failure on the calling thread.
 *
 * "ref:" lines cite the reference interface each function replaces
 * (paths relative to the reference repository root).
 */
#ifndef VXK_B200_H_
#define VXK_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VXK_ABI_VERSION 8

typedef struct CUstream_st* vxk_stream_t; /* == cudaStream_t */

/* activation element types (ref: VLLM_DISPATCH_FLOATING_TYPES, csrc/dispatch_utils.h) */
enum { VXK_F32 = 0, VXK_F16 = 1, VXK_BF16 = 2 };
/* quantised element types (ref: VLLM_DISPATCH_FP8_TYPES + int8 of the fused kernels) */
enum { VXK_Q_E4M3 = 0, VXK_Q_E5M2 = 1, VXK_Q_INT8 = 2 };
/* GEMM scale element types */
enum { VXK_S_F32 = 0, VXK_S_E8M0 = 1 };

enum {
  VXK_OK = 0,
  VXK_ERR_INVALID = -1,     /* bad argument (the torch shim turns it into RuntimeError) */
  VXK_ERR_UNSUPPORTED = -2, /* valid in the reference but not implemented here */
  VXK_ERR_CUDA = -3         /* CUDA runtime / driver error */
};

int vxk_abi_version(void);
const char* vxk_last_error(void);
/* number of kernels this library has launched in this process (bench "gpu_launches") */
uint64_t vxk_launch_count(void);
/* Small-M GEMMs (M <= 128 .. 256 of vxk_fp8_gemm / vxk_fp4_gemm / vxk_fp8_gemm_w8a16) are launched with programmatic
 * stream serialisation and request their first WEIGHT tiles before griddepcontrol.wait, i.e. while the previous
 * kernel on the stream is still draining -- the point being that the kernel producing the activations overlaps the
 * weight prefetch.  PRECONDITION (default on): the weight tensor B and its scales must not be written by work that
 * is still in flight on the same stream when the GEMM is enqueued (static model weights satisfy this; quantising a
 * weight and multiplying with it back to back does not).  vxk_set_early_weight_loads(0) (or VXK_SKINNY_PDL=0 in the
 * environment) launches these kernels as ordinary stream-ordered kernels.  Returns the previous setting. */
int vxk_set_early_weight_loads(int on);
/* silu evaluation in the fused silu_and_mul_* kernels.  -1 (default) = auto: the cheapest form that is
 * proven bit-identical to the reference on every input (ex2.approx / rcp.approx for bf16 a10 / a11 and
 * for a12, expf + IEEE divide for f16 a10 / a11); 0 = expf + IEEE divide everywhere; 1 = approximate
 * everywhere.  Returns the previous setting.  Process-wide. */
int vxk_set_silu_fast(int on);
/* Testing aid: out_t[i] = T(silu(in[i])) exactly as the fused silu_and_mul_* kernels evaluate it
 * (mode 0 / 1 as vxk_set_silu_fast; mode 2 = approximate form with caller constants c, beta);
 * out_f32[i] = the fp32 value before the rounding to T.  Either output may be null.  in_dtype F16/BF16. */
int vxk_debug_silu(void* out_t, float* out_f32, const void* in, int64_t n, int in_dtype, int mode,
                   float c, float beta, vxk_stream_t stream);

/* ---------------------------------------------------------------- a1 ----
 * ref: dynamic_per_token_scaled_fp8_quant  csrc/quantization/fp8/fp8_quant.cpp:842-881
 *      schema csrc/torch_bindings.cpp:160-163
 * in [T,H] contiguous, out [T,H] contiguous fp8, scales [T] fp32,
 * scale_ub: nullable pointer to ONE fp32 on the device. */
int vxk_dynamic_per_token_scaled_fp8_quant(void* out, const void* in, float* scales,
                                           const float* scale_ub, int64_t T, int64_t H,
                                           int in_dtype, int q_dtype, vxk_stream_t stream);

/* ---------------------------------------------------------------- a2 ----
 * ref: static_scaled_fp8_quant  fp8_quant.cpp:481-665, schema torch_bindings.cpp:147-149
 * scale element for (row, col) is scale[(row/group_m)*ss_i + (col/group_n)*ss_j]. */
int vxk_static_scaled_fp8_quant(void* out, const void* in, const float* scale, int64_t T,
                                int64_t H, int64_t in_row_stride, int64_t out_row_stride,
                                int64_t group_m, int64_t group_n, int64_t ss_i, int64_t ss_j,
                                int in_dtype, int q_dtype, vxk_stream_t stream);

/* ---------------------------------------------------------------- a3 ----
 * ref: dynamic_scaled_fp8_quant  fp8_quant.cpp:667-722, schema torch_bindings.cpp:153-156
 * scale: ONE fp32, written (amax / fp8_max). */
int vxk_dynamic_scaled_fp8_quant(void* out, const void* in, float* scale, int64_t T, int64_t H,
                                 int64_t in_row_stride, int64_t out_row_stride, int in_dtype,
                                 int q_dtype, vxk_stream_t stream);

/* ---------------------------------------------------------------- a4 ----
 * ref: per_token_group_quant_fp8  fp8_quant.cpp:724-840, schema torch_bindings.cpp:170-177
 * scales element (t, g) at out_s[t*ss0 + g*ss1] (row-major, column-major and
 * TMA-aligned column-major are all stride pairs). */
int vxk_per_token_group_fp8_quant(const void* in, void* out_q, float* out_s, int64_t T, int64_t H,
                                  int64_t group_size, double eps, int scale_ue8m0, int64_t ss0,
                                  int64_t ss1, int in_dtype, int q_dtype, vxk_stream_t stream);

/* ---------------------------------------------------------------- a5 ----
 * ref: per_token_group_quant_mxfp4  csrc/quantization/fp4/mxfp4_quant.cpp:24-126,
 *      schema torch_bindings.cpp:180-182.  out_q [T,H/2] bytes (even element in the
 * low nibble), out_s fp32 power of two 2^(e-127) at out_s[t*ss0 + g*ss1]. */
int vxk_per_token_group_quant_mxfp4(const void* in, void* out_q, float* out_s, int64_t T, int64_t H,
                                    int64_t group_size, double eps, int64_t ss0, int64_t ss1,
                                    int in_dtype, vxk_stream_t stream);

/* ---------------------------------------------------------------- a6 ----
 * ref: rms_norm_dynamic_per_token_quant  csrc/layernorm_quant.cpp:988-1045,
 *      schema torch_bindings.cpp:39-42.  residual (nullable) is updated in place. */
int vxk_rms_norm_dynamic_per_token_quant(void* out, const void* in, const void* weight, float* scales,
                                         double epsilon, const float* scale_ub, void* residual,
                                         int64_t T, int64_t H, int in_dtype, int q_dtype,
                                         vxk_stream_t stream);

/* ---------------------------------------------------------------- a7 ----
 * ref: rms_norm_per_block_quant  layernorm_quant.cpp:1047-1114, schema torch_bindings.cpp:49-53 */
int vxk_rms_norm_per_block_quant(void* out, const void* in, const void* weight, float* scales,
                                 double epsilon, void* residual, int64_t T, int64_t H,
                                 int64_t in_row_stride, int64_t group_size, int64_t ss0, int64_t ss1,
                                 int scale_ue8m0, int in_dtype, int q_dtype, vxk_stream_t stream);

/* ---------------------------------------------------------------- a8 ----
 * ref: rms_norm_mxfp4_quant  layernorm_quant.cpp:1172-1219, schema torch_bindings.cpp:56-59 */
int vxk_rms_norm_mxfp4_quant(void* out, const void* in, const void* weight, float* scales,
                             double epsilon, void* residual, int64_t T, int64_t H,
                             int64_t group_size, int64_t ss0, int64_t ss1, int in_dtype,
                             vxk_stream_t stream);

/* ---------------------------------------------------------------- a9 ----
 * ref: rms_norm_static_fp8_quant / fused_add_rms_norm_static_fp8_quant
 *      layernorm_quant.cpp:1116-1170, schema torch_bindings.cpp:63-73.
 * residual nullable: non-null selects the fused_add variant (updated in place). */
int vxk_rms_norm_static_fp8_quant(void* out, const void* in, const void* weight, const float* scale,
                                  double epsilon, void* residual, int64_t T, int64_t H,
                                  int64_t in_row_stride, int in_dtype, int q_dtype,
                                  vxk_stream_t stream);

/* ---------------------------------------------------------------- a10 ---
 * ref: silu_and_mul_quant  csrc/activation.cpp:552-566, schema torch_bindings.cpp:84-85
 * in [T, 2*d] (gate || up), out [T, d] fp8, scale ONE fp32. */
int vxk_silu_and_mul_quant(void* out, const void* in, const float* scale, int64_t T, int64_t d,
                           int in_dtype, int q_dtype, vxk_stream_t stream);

/* ---------------------------------------------------------------- a11 ---
 * ref: silu_and_mul_per_block_quant  csrc/quantization/fused_kernels/fused_silu_mul_block_quant.cpp:141-239,
 *      schema torch_bindings.cpp:89-97 */
int vxk_silu_and_mul_per_block_quant(void* out, const void* in, float* scales, int64_t T, int64_t H,
                                     int64_t group_size, const float* scale_ub, int64_t ss0,
                                     int64_t ss1, int scale_ue8m0, int in_dtype, int q_dtype,
                                     vxk_stream_t stream);

/* ---------------------------------------------------------------- a12 ---
 * ref: silu_and_mul_mxfp4_quant  fused_silu_mul_mxfp4_quant.cpp:134-209,
 *      schema torch_bindings.cpp:103-106 */
int vxk_silu_and_mul_mxfp4_quant(void* out, const void* in, float* scales, int64_t T, int64_t H,
                                 int64_t group_size, double eps, int64_t ss0, int64_t ss1,
                                 int in_dtype, vxk_stream_t stream);

/* ---------------------------------------------------------------- a13 ---
 * ref: fp8_gemm  csrc/xpu/onednn/onednn_matmul.cpp:54-79 -> dnnl_matmul_w8a8_fp8
 *      csrc/xpu/onednn/fp8_gemm_w8a8.h:12-184, schema csrc/xpu/torch_bindings.cpp:14-16
 *
 * D[M,N] (ldd) = (A o sA) (B o sB) + bias,  fp32 accumulate, D bf16 or f16.
 *   A: [M,K] fp8, row stride lda (K contiguous).
 *   B: element (k,n) at B[k*b_stride_k + n*b_stride_n]; one of the strides is 1
 *      ("nt": b_stride_k == 1, i.e. a transposed view of W[N,K]; "nn": b_stride_n == 1).
 *   a_scale_mode / b_scale_mode: 0 = one scalar, 1 = per row of A ([M]) / per
 *      column of B ([N]), 2 = block: sA[m, k/gk] at sA[m*sa0 + kb*sa1],
 *      sB[k/gk, n/gn] at sB[kb*sb0 + nb*sb1].
 *   scale_dtype: VXK_S_F32 or VXK_S_E8M0 (block mode only).
 *   bias: nullable, [N], bias_dtype VXK_F32/F16/BF16.
 *   workspace: device scratch of vxk_fp8_gemm_workspace_bytes(...) bytes (may be null when 0).
 */
int vxk_fp8_gemm(void* D, const void* A, const void* B, const void* a_scale, const void* b_scale,
                 const void* bias, int64_t M, int64_t N, int64_t K, int64_t lda,
                 int64_t b_stride_k, int64_t b_stride_n, int64_t ldd, int a_scale_mode,
                 int b_scale_mode, int64_t group_k, int64_t group_n, int64_t sa0, int64_t sa1,
                 int64_t sb0, int64_t sb1, int scale_dtype, int fp8_dtype, int out_dtype,
                 int bias_dtype, void* workspace, int64_t workspace_bytes, vxk_stream_t stream);

/* Column-parallel FP8 GEMM fused with the all-gather of its output (the only
 * collective of the path).  Rank r owns W[n0:n0+N, :]; instead of writing a private [M, N]
 * slab that an NCCL all-gather reassembles afterwards, the epilogue stores every finished tile
 * into columns [col_offset, col_offset + N) of ALL ranks' [M, out_ld] outputs through NVLink
 * peer mappings (peer_out[i] = base pointer of rank i's output as mapped into this process,
 * e.g. torch symmetric memory; peer_out[self_rank] is this rank's own buffer: the tile is stored there
 * first and then pushed to the others as whole 512-byte row segments; multicast_out, when
 * non-null, is an NVLS multicast mapping of the same buffers -- one multimem.st per segment
 * then reaches every rank through the NVSwitch, needs M > 256 and N % 256 == 0).  The caller
 * synchronises the ranks afterwards (one barrier) before reading.  Scale arguments as vxk_fp8_gemm
 * (block scales -- DeepSeek-V3 1x128 / 128x128, BASELINE config 5 -- need M > 128).  No reference counterpart: the reference has no parallelism, the
 * all-gather is the caller's (vLLM's) in tensor-parallel column-linear layers. */
int vxk_fp8_gemm_allgather(void* const* peer_out, int n_peers, int self_rank, void* multicast_out,
                           int64_t col_offset, int64_t out_ld, const void* A, const void* B,
                           const void* a_scale, const void* b_scale, const void* bias, int64_t M,
                           int64_t N, int64_t K, int64_t lda, int64_t b_stride_k, int64_t b_stride_n,
                           int a_scale_mode, int b_scale_mode, int64_t group_k, int64_t group_n,
                           int64_t sa0, int64_t sa1, int64_t sb0, int64_t sb1, int scale_dtype,
                           int fp8_dtype, int out_dtype, int bias_dtype, void* workspace,
                           int64_t workspace_bytes, vxk_stream_t stream);

/* Scratch the caller must provide to vxk_fp8_gemm (device memory, any alignment); 0 when
 * none is needed.  ref: the per-call scratchpad of fp8_gemm_w8a8.h:176-179. */
int64_t vxk_fp8_gemm_workspace_bytes(int64_t M, int64_t N, int64_t K, int a_scale_mode, int scale_dtype);

/* How many GEMM calls ran on the tcgen05 kernels / on the CUDA-core fallback so far. */
void vxk_gemm_stats(uint64_t* tensor_core_calls, uint64_t* fallback_calls);
/* Testing aid: route every GEMM to the CUDA-core fallback (also env VXK_GEMM_FORCE_FALLBACK=1).
 * Returns the previous setting. */
int vxk_set_gemm_force_fallback(int on);

/* ---------------------------------------------------------------- a14 ---
 * ref: fp4_gemm  onednn_matmul.cpp:101-123 -> dnnl_matmul_w4a4_fp4
 *      csrc/xpu/onednn/fp4_gemm_w4a4.h:12-167, schema csrc/xpu/torch_bindings.cpp:24-26
 *
 *   A: [M, K/2] packed E2M1 bytes, row stride lda_bytes.
 *   B: W[N, K/2] packed bytes, row stride ldb_bytes (the op receives it as the
 *      transposed view [K/2, N]; only this K-major layout exists for FP4).
 *   sA: [M, K/32] e8m0 bytes at sA[m*sa0 + kb*sa1]; sB: [N, K/32] at sB[n*sb0 + kb*sb1].
 *   workspace: device scratch of vxk_fp4_gemm_workspace_bytes(...) bytes (may be null).
 */
int vxk_fp4_gemm(void* D, const void* A, const void* B, const void* a_scale, const void* b_scale,
                 const void* bias, int64_t M, int64_t N, int64_t K, int64_t lda_bytes,
                 int64_t ldb_bytes, int64_t ldd, int64_t sa0, int64_t sa1, int64_t sb0, int64_t sb1,
                 int out_dtype, int bias_dtype, void* workspace, int64_t workspace_bytes,
                 vxk_stream_t stream);

/* MXFP4 counterpart of vxk_fp8_gemm_allgather (BASELINE config 4: Llama-3-70B MXFP4 GEMM, column-parallel):
 * same contract, operands as vxk_fp4_gemm; needs M > 128 and the vxk_fp4_gemm_workspace_bytes scratch. */
int vxk_fp4_gemm_allgather(void* const* peer_out, int n_peers, int self_rank, int64_t col_offset,
                           int64_t out_ld, const void* A, const void* B, const void* a_scale,
                           const void* b_scale, const void* bias, int64_t M, int64_t N, int64_t K,
                           int64_t lda_bytes, int64_t ldb_bytes, int64_t sa0, int64_t sa1, int64_t sb0,
                           int64_t sb1, int out_dtype, int bias_dtype, void* workspace,
                           int64_t workspace_bytes, vxk_stream_t stream);

/* Scratch the caller must provide to vxk_fp4_gemm (device memory, any alignment; the scale
 * factors are re-packed there in the tensor core's layout); 0 when none is needed.  Without
 * it the call still runs, on the single-CTA kernel.
 * ref: the per-call scratchpad of fp4_gemm_w4a4.h:159-162. */
int64_t vxk_fp4_gemm_workspace_bytes(int64_t M, int64_t N, int64_t K);

/* vxk_fp4_gemm for callers that may not allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same operands, arithmetic and kernels as vxk_fp4_gemm; the scale-factor scratch is a file-scope
 * device array inside the library (vxk_fp4_gemm_static_scratch_bytes() = 4 MiB, held by every GPU that loads the
 * library until the process exits) instead of a caller buffer.  The call allocates nothing (no cudaMalloc /
 * cudaFree, no allocating library routine).
 *   Limit: vxk_fp4_gemm_workspace_bytes(M, N, K) + 256 must fit the array, e.g. M = 2048, N = 8192, K = 8192
 *   (2.7 MB) does, M = 8192 with N = 28672 does not; larger problems return VXK_ERR_INVALID (message names the
 *   sizes) -- use vxk_fp4_gemm with a workspace of your own there.
 *   Sharing: every call on a device uses the same array.  Calls that follow each other on one stream are safe
 *   (each call rewrites the whole region it reads; nothing needs to start at zero); calls that overlap on the GPU
 *   (two streams, two threads) race on it and must not be issued. */
int vxk_fp4_gemm_static_scratch(void* D, const void* A, const void* B, const void* a_scale,
                                const void* b_scale, const void* bias, int64_t M, int64_t N, int64_t K,
                                int64_t lda_bytes, int64_t ldb_bytes, int64_t ldd, int64_t sa0, int64_t sa1,
                                int64_t sb0, int64_t sb1, int out_dtype, int bias_dtype, vxk_stream_t stream);
int64_t vxk_fp4_gemm_static_scratch_bytes(void);

/* ------------------------------------------------------- (f) next #2 ---
 * MX scale factors in the tensor core's own layout ("packed": the 512-byte chunks that
 * tcgen05.cp.32x128b.warpx4 copies into TMEM, DESIGN.md "Scale-factor layouts").  The reference has no
 * counterpart: oneDNN re-orders scales inside the primitive on every call (fp4_gemm_w4a4.h:159-167,
 * scratchpad); here the re-ordering can leave the hot path altogether --
 *   weights:      vxk_mx_sf_pack once after loading, pass the result as b_scale_packed;
 *   activations:  the *_sf variants of the MXFP4 quantisation kernels (a5 / a8 / a12) write the packed
 *                 bytes themselves, next to the fp32 scales the reference API returns.
 *   kind: 0 = FP8 with E8M0 block scales (K slices of 128), 1 = MXFP4 (K slices of 256)
 *   operand: 0 = activations (rows = M), 1 = weights (rows = N)
 * A packed buffer is device memory, 16-byte aligned, vxk_mx_sf_packed_bytes long.  When K is not a
 * multiple of the K slice the caller zero-fills it before a *_sf quantisation kernel writes into it
 * (the kernels only write real groups; vxk_mx_sf_pack writes every byte). */
int64_t vxk_mx_sf_packed_bytes(int64_t rows, int64_t K, int kind, int operand);
/* src: E8M0 bytes of (row r, 32-element K group g) at src[r*s_row + g*s_k]. */
int vxk_mx_sf_pack(void* dst, const void* src, int64_t rows, int64_t K, int64_t s_row, int64_t s_k,
                   int kind, int operand, vxk_stream_t stream);
/* vxk_fp4_gemm with optional packed copies of either scale grid (nullable; the natural-layout
 * pointers stay mandatory: small-M calls and the fallback read those).  With both packed copies no
 * workspace is needed and the call issues no re-packing launch. */
int vxk_fp4_gemm_sf(void* D, const void* A, const void* B, const void* a_scale, const void* b_scale,
                    const void* a_scale_packed, const void* b_scale_packed, const void* bias,
                    int64_t M, int64_t N, int64_t K, int64_t lda_bytes, int64_t ldb_bytes, int64_t ldd,
                    int64_t sa0, int64_t sa1, int64_t sb0, int64_t sb1, int out_dtype, int bias_dtype,
                    void* workspace, int64_t workspace_bytes, vxk_stream_t stream);
/* a5 / a8 / a12 with the extra packed output (sf_packed nullable = the plain op; kind 1, operand 0,
 * rows = T, K = H). */
int vxk_per_token_group_quant_mxfp4_sf(const void* in, void* out_q, float* out_s, void* sf_packed,
                                       int64_t T, int64_t H, int64_t group_size, double eps,
                                       int64_t ss0, int64_t ss1, int in_dtype, vxk_stream_t stream);
int vxk_rms_norm_mxfp4_quant_sf(void* out, const void* in, const void* weight, float* scales,
                                void* sf_packed, double epsilon, void* residual, int64_t T, int64_t H,
                                int64_t group_size, int64_t ss0, int64_t ss1, int in_dtype,
                                vxk_stream_t stream);
int vxk_silu_and_mul_mxfp4_quant_sf(void* out, const void* in, float* scales, void* sf_packed,
                                    int64_t T, int64_t H, int64_t group_size, double eps, int64_t ss0,
                                    int64_t ss1, int in_dtype, vxk_stream_t stream);
/* Scale re-packing launches issued so far by the GEMM entry points (testing aid). */
uint64_t vxk_sf_pack_launches(void);

/* ------------------------------------------------------- (f) next #2 ---
 * Quantisation fused into the GEMM's neighbour (no reference counterpart): the gate_up
 * projection with silu_and_mul and the 1 x 128 block FP8 quantisation in its epilogue.
 *
 *   t      = fp8_gemm(A, W^T) as vxk_fp8_gemm would round it to mid_dtype (bf16 / f16):  [M, 2 * inter],
 *            W [2 * inter, K] K-contiguous (row stride ldb bytes): gate rows first, then up rows
 *   out_q, out_s = silu_and_mul_per_block_quant(t, group 128, e4m3): codes [M, inter] contiguous, scale of
 *            (row m, group g) at out_s[m * ss0 + g * ss1]
 * Bit-identical to the two calls, without writing / re-reading the [M, 2 * inter] intermediate.
 * a_scale_mode / b_scale_mode: 0 scalar, 1 per token [M] / per channel [2 * inter].  bias nullable, [2 * inter].
 * Returns VXK_ERR_UNSUPPORTED when inter % 128 != 0 or an operand is not TMA-expressible (16-byte aligned rows).
 */
int vxk_fp8_gemm_silu_mul_quant(void* out_q, float* out_s, const void* A, const void* B, const float* a_scale,
                                const float* b_scale, const void* bias, int64_t M, int64_t inter, int64_t K,
                                int64_t lda, int64_t ldb, int a_scale_mode, int b_scale_mode, int64_t ss0,
                                int64_t ss1, int fp8_dtype, int mid_dtype, int bias_dtype, vxk_stream_t stream);

/* ------------------------------------------------------- (f) next #4 ---
 * MoE grouped GEMM.
 * ref: cutlass_grouped_gemm_interface  csrc/xpu/grouped_gemm/grouped_gemm_interface.cpp:12-59
 *      -> cutlass_grouped_gemm_xe2_impl  csrc/xpu/grouped_gemm/xe_2/grouped_gemm_xe2_interface.hpp:160-357,
 *      schema csrc/xpu/torch_bindings.cpp:41-53, recipes vllm_xpu_kernels/fused_moe_interface.py:32-45
 *
 *   for every expert e (rows of A / D are grouped by expert, in expert order):
 *     D[r0_e : r0_e + rows_per_expert[e], :] = A[same rows, :] x dequant(B[e]) (+ bias[e, :])
 *   A [total_m, K] and D [total_m, N] contiguous, act_dtype (VXK_BF16 / VXK_F16); fp32 accumulation.
 *   rows_per_expert: int32 [num_experts] ON THE DEVICE (never read by the host); rows past the sum are left
 *   untouched.
 *   weight_kind / B / scales:
 *     VXK_MOE_W16       B [E, K, N] of act_dtype, scales ignored                        (recipe "bf16")
 *     VXK_MOE_FP8_E4M3 / _E5M2   B [E, K, N] FP8 bytes, scales fp32 [E] (per expert)       (recipe "fp8")
 *     VXK_MOE_MXFP4     B [E, N, K/2] packed E2M1 (even k in the low nibble), scales uint8 E8M0 [E, N, K/32];
 *                       a weight is table(code) * 2^(s-127), both factors and the product in act_dtype
 *                       (tests/fused_moe/test_grouped_gemm.py:341-377)                   (recipe "mxfp4")
 *   bias: nullable, [E, N] of act_dtype.  N % 8 == 0 (grouped_gemm_xe2_interface.hpp:178).
 *   The INT4 recipe of the reference op is out of scope (INT4 GEMMs).
 *   force_fallback: testing aid, non-zero routes the call to the CUDA-core kernel.
 */
enum { VXK_MOE_W16 = 0, VXK_MOE_FP8_E4M3 = 1, VXK_MOE_FP8_E5M2 = 2, VXK_MOE_MXFP4 = 3,
       /* extension: the fp8 recipe with the weights stored [E, N, K] (K-contiguous, e.g. B.transpose(1, 2).contiguous()
        * done once for static weights): the widened tile then goes to tensor memory like mxfp4, which the [E, K, N]
        * layout cannot -- about 1.7x faster at decode sizes */
       VXK_MOE_FP8_E4M3_NK = 4, VXK_MOE_FP8_E5M2_NK = 5 };
int vxk_moe_grouped_gemm(void* D, const void* A, const void* B, const void* scales, const void* bias,
                         const int32_t* rows_per_expert, int64_t total_m, int64_t N, int64_t K,
                         int64_t num_experts, int act_dtype, int weight_kind, int force_fallback,
                         vxk_stream_t stream);

/* ------------------------------------------------------- (f) next #1 ---
 * ref: fp8_gemm_w8a16  csrc/xpu/onednn/onednn_matmul.cpp:81-99 -> dnnl_matmul_w8a16_fp8
 *      csrc/xpu/onednn/fp8_gemm_w8a16.h:13-145, schema csrc/xpu/torch_bindings.cpp:19-21
 *
 *   D[M,N] = A[M,K] (bf16 / f16, act_dtype; D has the same type) x dequant(B), + bias.
 *   A: row stride lda (elements).  B(k, n) at B[k*b_stride_k + n*b_stride_n], FP8 bytes
 *      (b_stride_k == 1 is the "nt" layout W[N,K] the tensor-core kernel takes; anything
 *      else runs on the CUDA-core kernel).
 *   b_scale (fp32): b_scale_mode 0 = one scalar, 1 = per output channel [N],
 *      2 = block: sB[k/group_k, n/group_n] at b_scale[kb*sb0 + nb*sb1]
 *      (fp8_gemm_w8a16.h:29-35: 2-D scale with size(0) > 1).
 *   bias: nullable, [N], bias_dtype VXK_F32/F16/BF16.
 *   force_fallback: testing aid, non-zero routes the call to the CUDA-core kernel.
 */
int vxk_fp8_gemm_w8a16(void* D, const void* A, const void* B, const float* b_scale, const void* bias,
                       int64_t M, int64_t N, int64_t K, int64_t lda, int64_t b_stride_k,
                       int64_t b_stride_n, int64_t ldd, int b_scale_mode, int64_t group_k,
                       int64_t group_n, int64_t sb0, int64_t sb1, int act_dtype, int fp8_dtype,
                       int bias_dtype, int force_fallback, vxk_stream_t stream);

/* ------------------------------------------------------- (f) next #3 ---
 * Block-FP8 quantisation into / out of the paged indexer-K cache, and the FP8 cache conversion.
 * Cache block layout (bytes): [block_size * head_dim e4m3 codes][block_size * head_dim / G fp32 scales].
 *
 * ref: indexer_k_quant_and_cache  csrc/cache.cpp:530-612, launcher :1558-1590,
 *      schema csrc/torch_bindings.cpp:298-302 (library _C_cache_ops)
 * k [num_tokens, head_dim] (row stride k_row_stride elements), slot_mapping [num_tokens] int64
 * (negative = skip the token), kv_cache [num_blocks, cache_block_size, cache_stride] bytes with
 * cache_block_stride_bytes between blocks.  s = max(amax, 1e-4) / 448 per 1 x quant_block_size group
 * (use_ue8m0: rounded up to a power of two), q = e4m3(clamp(x / s)). */
int vxk_indexer_k_quant_and_cache(const void* k, void* kv_cache, const int64_t* slot_mapping,
                                  int64_t num_tokens, int64_t head_dim, int64_t k_row_stride,
                                  int64_t quant_block_size, int64_t cache_block_size, int64_t cache_stride,
                                  int64_t cache_block_stride_bytes, int use_ue8m0, int in_dtype,
                                  vxk_stream_t stream);

/* ref: cp_gather_indexer_k_quant_cache  csrc/cache.cpp:616-750, launcher :1625-1680,
 *      schema csrc/torch_bindings.cpp:303-309
 * Gathers the codes and scales of the tokens of `batch_size` sequences (cu_seq_lens [batch+1] int32,
 * block_table [batch, max_blocks_per_seq] int32) from the paged cache into dst_k [num_tokens, head_dim]
 * bytes and dst_scale [num_tokens, 4 * head_dim / quant_block_size] bytes (fp32 scales). */
int vxk_cp_gather_indexer_k_quant_cache(const void* kv_cache, void* dst_k, void* dst_scale,
                                        const int32_t* block_table, const int32_t* cu_seq_lens,
                                        int64_t batch_size, int64_t max_blocks_per_seq, int64_t num_tokens,
                                        int64_t head_dim, int64_t quant_block_size, int64_t cache_block_size,
                                        int64_t cache_block_stride_bytes, int64_t dst_k_stride_bytes,
                                        int64_t dst_scale_stride_bytes, vxk_stream_t stream);

/* ref: convert_fp8  csrc/cache.cpp:1393-1540, schema csrc/torch_bindings.cpp:281-284
 * n elements.  src_kind / dst_kind: VXK_F32 / VXK_F16 / VXK_BF16 or 3 = byte tensor (exactly one side).
 * mode 1 ("fp8" / "fp8_e4m3"): bytes are e4m3 codes, quantise e4m3(clamp(x / scale)) or dequantise
 * float(code) * scale by direction; mode 0 ("auto"): numeric cast of float(src) * scale. */
int vxk_convert_fp8(void* dst, const void* src, double scale, int64_t n, int src_kind, int dst_kind,
                    int mode, vxk_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* VXK_B200_H_ */
