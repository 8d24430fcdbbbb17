// Prints which (TMEM lane, column) each register of tcgen05.ld.16x256b.x2 returns per thread (sm_100a): the lanes are
// first filled with lane*1000+column by tcgen05.st.32x32b.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I ghost_b200/csrc tools/tmem_layout.cu -o tools/bin/tmem_layout
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "ptx.cuh"

using namespace ghost::ptx;

__global__ void __launch_bounds__(128, 1) probe(uint32_t* out) {
  __shared__ uint32_t tmem_base_smem;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) tmem_alloc(&tmem_base_smem, 32);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = tmem_base_smem;
  const uint32_t row = warp * 32 + lane;
  uint32_t v[16];
  for (int c0 = 0; c0 < 32; c0 += 16) {
    for (int i = 0; i < 16; ++i) v[i] = row * 1000 + c0 + i;
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16};" ::"r"(tb + ((warp * 32u) << 16) + c0),
        "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
        "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
        : "memory");
  }
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  // 16x256b.x2: 8 registers; two loads, at lane offset 0 and 16 of this warp's quadrant, columns 8..23
  for (int half = 0; half < 2; ++half) {
    uint32_t r[8];
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(tb + ((warp * 32u + half * 16u) << 16) + 8)
                 : "memory");
    tmem_ld_wait();
    for (int i = 0; i < 8; ++i) out[((warp * 2 + half) * 32 + lane) * 8 + i] = r[i];
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tb, 32);
  }
}

int main() {
  uint32_t* d;
  cudaMalloc(&d, 4 * 2 * 32 * 8 * 4);
  probe<<<1, 128>>>(d);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    printf("failed: %s\n", cudaGetErrorString(e));
    return 1;
  }
  static uint32_t h[4 * 2 * 32 * 8];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  for (int w = 0; w < 2; ++w)
    for (int half = 0; half < 2; ++half) {
      printf("warp %d, lane offset %d: thread -> (lane,col) of r0..r7\n", w, half * 16);
      for (int t = 0; t < 32; ++t) {
        printf("  t%2d:", t);
        for (int i = 0; i < 8; ++i) {
          const uint32_t v = h[((w * 2 + half) * 32 + t) * 8 + i];
          printf(" (%3u,%2u)", v / 1000, v % 1000);
        }
        printf("\n");
      }
    }
  return 0;
}
