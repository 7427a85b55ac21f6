// Does tcgen05.ld proceed while tcgen05.mma is running?  One thread issues FP8 MMAs (M = 128, N = 256, K = 32
// bytes, single CTA) back to back into TMEM columns 0..255; eight other warps drain columns 256..511 with
// tcgen05.ld.32x32b.x32 (DEPTH loads in flight per warp).  Three runs per configuration: MMA alone, loads alone,
// both.  Prints tensor utilisation (MMA clocks issued / elapsed) and load bandwidth for each.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I vllm_xpu_kernels_b200/csrc -o /tmp/tmem_contention benchmarks/tmem_contention.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#include "gemm_common.cuh"

using namespace vxk;

constexpr int kN = 256;
constexpr int kStageA = 128 * 128, kStageB = kN * 128;   // one 128-byte K slice of A (128 rows) and B (256 rows)

template <int DEPTH>
__global__ void __launch_bounds__(320, 1) probe(int mma_slices, int ld_iters, int mode, unsigned long long* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint32_t tptr;
  __shared__ uint64_t bar;
  __shared__ volatile int mma_done;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < (kStageA + kStageB) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x38383838u;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); mma_done = 0; }
  fence_proxy_async_smem();
  if (warp == 0) tmem_alloc<512>(&tptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t base = tptr;
  unsigned long long t_mma = 0, t_ld = 0;
  const unsigned long long t0 = clock64();
  if (warp == 1) {
    if (mode & 1) {
      const uint32_t idesc = make_idesc_f8(0, 128, kN);
      const uint64_t da = make_kmajor_sw128_desc(smem_u32(smem)), db = make_kmajor_sw128_desc(smem_u32(smem + kStageA));
      for (int s = 0; s < mma_slices; ++s) {
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_f8f6f4(base, da + k * 2, db + k * 2, idesc, (s | k) ? 1u : 0u);
          if ((s & 15) == 15 || s == mma_slices - 1) umma_commit(&bar);
        }
        __syncwarp();
        if ((s & 15) == 15 || s == mma_slices - 1) mbar_wait(&bar, (s >> 4) & 1);   // bounded queue depth: 64 MMAs
      }
      t_mma = clock64() - t0;
      mma_done = 1;
    }
  } else if (warp >= 2) {
    if (mode & 2) {
      const int e = warp - 2, q = warp & 3, h = e >> 2;      // lane quarter = warp % 4, column half
      uint32_t acc = 0;
      int it_at = -1;
      unsigned long long t_at = 0;
      for (int it = 0; it < ld_iters; ++it) {
        if (it_at < 0 && mma_done) { it_at = it; t_at = clock64() - t0; }
        for (int c = 0; c < 128; c += 32 * DEPTH) {
          uint32_t r[DEPTH][32];
#pragma unroll
          for (int d = 0; d < DEPTH; ++d)
            tmem_ld_32x32b_x32(base + (static_cast<uint32_t>(q * 32) << 16) + 256 + h * 128 + c + d * 32, r[d]);
          tmem_ld_wait();
#pragma unroll
          for (int d = 0; d < DEPTH; ++d)
#pragma unroll
            for (int j = 0; j < 32; ++j) acc ^= r[d][j];
        }
      }
      t_ld = clock64() - t0;
      if (it_at >= 0) {          // the MMA stream ended while we were loading: report the overlapped window only
        t_ld = t_at;
        if (lane == 0) out[blockIdx.x * 16 + 10] = it_at;   // (same for all warps to within an iteration)
      }
      if (acc == 0x12345678u) out[4000] = acc;
    }
  }
  if (lane == 0 && warp == 1) out[blockIdx.x * 16] = t_mma;
  if (lane == 0 && warp >= 2) out[blockIdx.x * 16 + warp] = t_ld;
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(base);
}

template <int DEPTH>
void run(int mode, const char* what) {
  unsigned long long* out;
  cudaMalloc(&out, 65536);
  cudaMemset(out, 0, 65536);
  const int smem = kStageA + kStageB + 2048;
  cudaFuncSetAttribute(probe<DEPTH>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int mma_slices = 4096, ld_iters = 16384;
  probe<DEPTH><<<148, 320, smem>>>(mma_slices, ld_iters, mode, out);
  probe<DEPTH><<<148, 320, smem>>>(mma_slices, ld_iters, mode, out);
  cudaError_t err = cudaDeviceSynchronize();
  unsigned long long h[148 * 16];
  cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
  double mma = 0, ld = 0, its = 0;
  for (int b = 0; b < 148; ++b) {
    mma += h[b * 16];
    double m = 0;
    for (int w = 2; w < 10; ++w) m = h[b * 16 + w] > m ? h[b * 16 + w] : m;
    ld += m;
    its += h[b * 16 + 10] ? h[b * 16 + 10] : ld_iters;
  }
  mma /= 148; ld /= 148; its /= 148;
  // one MMA 128 x 256 x 32 B = 128 clk of tensor work at the dense FP8 rate (8192 MAC/clk/SM)
  const double mma_ideal = mma_slices * 4.0 * 128.0;
  const double ld_bytes = 8.0 * its * 32 * 128 * 4;
  printf("depth %d %-10s: ", DEPTH, what);
  if (mode & 1) printf("MMA %9.0f clk = %5.1f %% of the dense FP8 rate   ", mma, 100.0 * mma_ideal / mma);
  if (mode & 2) printf("tcgen05.ld %9.0f clk = %6.1f B/clk/SM", ld, ld_bytes / ld);
  printf("  (%s)\n", cudaGetErrorString(err));
  cudaFree(out);
}

int main() {
  run<1>(1, "mma only");
  run<1>(2, "ld only"); run<2>(2, "ld only"); run<4>(2, "ld only");
  run<1>(3, "both"); run<2>(3, "both"); run<4>(3, "both");
  return 0;
}
