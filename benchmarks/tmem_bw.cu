// TMEM read throughput probe: which tcgen05.ld shape / warp count drains TMEM fastest?
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scratch/tmem_bw scratch/tmem_bw.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#define OUT32(r) "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), \
  "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), \
  "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), \
  "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
#define LIST32 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}"

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

template <int SHAPE>
__device__ __forceinline__ void ld32(uint32_t taddr, uint32_t (&r)[32]) {
  if constexpr (SHAPE == 0)       // 32 lanes x 32 columns
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 " LIST32 ", [%32];" : OUT32(r) : "r"(taddr));
  else if constexpr (SHAPE == 1)  // 16 lanes x 64 columns
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x8.b32 " LIST32 ", [%32];" : OUT32(r) : "r"(taddr));
  else if constexpr (SHAPE == 2)  // 16 lanes x 64 columns
    asm volatile("tcgen05.ld.sync.aligned.16x128b.x16.b32 " LIST32 ", [%32];" : OUT32(r) : "r"(taddr));
  else                            // 16 lanes x 64 columns
    asm volatile("tcgen05.ld.sync.aligned.16x64b.x32.b32 " LIST32 ", [%32];" : OUT32(r) : "r"(taddr));
}

// every warp reads its lane quarter (warp % 4), columns split over warp / 4; ITER passes over the
// CTA's 512 columns; DEPTH loads in flight before a wait
template <int SHAPE, int DEPTH>
__global__ void __launch_bounds__(512, 1) probe(int iters, unsigned long long* clk, uint32_t* sink) {
  __shared__ uint32_t tptr;
  const int warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tptr)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = tptr;
  const int q = warp & 3, cg = warp >> 2, ncg = nwarp >> 2;
  const int cols_per = 512 / ncg;      // columns this warp covers per pass
  uint32_t acc = 0;
  __syncthreads();
  const unsigned long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    // a 32-lane x 32-column block per load for SHAPE 0; 16 lanes x 64 columns for the others
    constexpr int kColsPerLd = SHAPE == 0 ? 32 : 64;
    constexpr int kLaneHalves = SHAPE == 0 ? 1 : 2;
    for (int c = 0; c < cols_per; c += kColsPerLd * DEPTH / kLaneHalves) {
      uint32_t r[DEPTH][32];
#pragma unroll
      for (int d = 0; d < DEPTH; ++d) {
        const int col = cg * cols_per + c + (d / kLaneHalves) * kColsPerLd;
        const int lane0 = q * 32 + (d % kLaneHalves) * 16;
        ld32<SHAPE>(base + (static_cast<uint32_t>(lane0) << 16) + col, r[d]);
      }
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
      for (int d = 0; d < DEPTH; ++d)
#pragma unroll
        for (int j = 0; j < 32; ++j) acc ^= r[d][j];
    }
  }
  __syncthreads();
  const unsigned long long t1 = clock64();
  if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
  if (acc == 0x12345678u) sink[threadIdx.x] = acc;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(base) : "memory");
}

template <int SHAPE, int DEPTH>
void run(const char* name, int warps) {
  unsigned long long* clk;
  uint32_t* sink;
  cudaMalloc(&clk, 148 * 8);
  cudaMalloc(&sink, 4096);
  const int iters = 200;
  probe<SHAPE, DEPTH><<<148, warps * 32>>>(iters, clk, sink);
  probe<SHAPE, DEPTH><<<148, warps * 32>>>(iters, clk, sink);
  cudaError_t e = cudaDeviceSynchronize();
  unsigned long long h[148];
  cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
  double avg = 0;
  for (int i = 0; i < 148; ++i) avg += h[i];
  avg /= 148;
  const double bytes = 128.0 * 512 * 4 * iters;   // whole TMEM per pass
  printf("%-14s depth %d warps %2d : %8.0f clk  %6.1f B/clk/SM  (%s)\n", name, DEPTH, warps, avg, bytes / avg,
         cudaGetErrorString(e));
  cudaFree(clk);
  cudaFree(sink);
}

int main() {
  for (int w : {4, 8, 16}) {
    run<0, 1>("32x32b.x32", w);
    run<0, 2>("32x32b.x32", w);
    run<0, 4>("32x32b.x32", w);
    run<1, 2>("16x256b.x8", w);
    run<1, 4>("16x256b.x8", w);
    run<2, 2>("16x128b.x16", w);
    run<2, 4>("16x128b.x16", w);
    run<3, 2>("16x64b.x32", w);
    run<3, 4>("16x64b.x32", w);
  }
  return 0;
}
