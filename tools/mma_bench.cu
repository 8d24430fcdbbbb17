// tcgen05.mma issue / throughput probe (sm_100a): one warp issues batches of four 128xNx16 bf16 MMAs from shared
// memory operands (SS mode, 128B swizzle, the same descriptors igemm_kernel uses), a commit after each batch, no TMA
// traffic.  Prints cycles per MMA for several N, operand ring depths and M, on 1 CTA and on one CTA per SM.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I ghost_b200/csrc tools/mma_bench.cu -o tools/bin/mma_bench
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>
#include "ptx.cuh"

using namespace ghost::ptx;

__device__ __forceinline__ void bulk_load(void* smem, const void* g, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem)),
               "l"(g), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// copy_chunk > 0: warps 2 and 3 each keep 3 bulk copies of copy_chunk bytes (L2 -> shared, into slots behind the operand
// ring) in flight while the MMAs run, like the TMA producers of the conv kernel.
__global__ void __launch_bounds__(128, 1) mma_probe(int M, int N, int iters, int ring, int acc_stages, int wait_each,
                                                    unsigned long long* out, const uint8_t* src, int copy_chunk) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bars[16];
  __shared__ uint64_t done_bar;
  __shared__ uint64_t cbar[8];
  __shared__ int stop_flag;
  __shared__ uint32_t tmem_base_smem;
  __shared__ int abort_flag;
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  const int stage_bytes = 16384 + N * 128;
  for (int i = threadIdx.x; i < ring * stage_bytes / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) {
    abort_flag = 0;
    for (int i = 0; i < 16; ++i) mbar_init(&bars[i], 1);
    for (int i = 0; i < 8; ++i) mbar_init(&cbar[i], 1);
    stop_flag = 0;
    mbar_init(&done_bar, 1);
    fence_mbar_init();
  }
  fence_proxy_async();
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc(&tmem_base_smem, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  if (warp == 1) {
    const uint32_t idesc = make_idesc_bf16(M, N);
    const uint32_t dhi = smem_desc_hi(1024, false);
    const uint32_t ring_u = smem_u32(smem);
    int st = 0, as = 0;
    uint32_t ph = 0;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      const uint32_t sa = ring_u + st * stage_bytes;
      umma_bf16_batch<4>(tmem_base + as * (512 / acc_stages), smem_desc_lo(sa), smem_desc_lo(sa + 16384), dhi, idesc,
                         (it & 15) != 0);
      umma_commit_elect(&bars[st]);
      if (wait_each) {
        mbar_wait_nofail(&bars[st], ph, &abort_flag);
        tc_fence_after();
      }
      if (++st == ring) {
        st = 0;
        ph ^= 1u;
      }
      if ((it & 15) == 15 && ++as == acc_stages) as = 0;
    }
    const long long t1 = clock64();
    umma_commit_elect(&done_bar);
    mbar_wait_nofail(&done_bar, 0, &abort_flag);
    const long long t2 = clock64();
    if ((threadIdx.x & 31) == 0) {
      out[blockIdx.x * 2] = t1 - t0;
      out[blockIdx.x * 2 + 1] = t2 - t0;
      *reinterpret_cast<volatile int*>(&stop_flag) = 1;
    }
  } else if (warp >= 2 && copy_chunk > 0 && (threadIdx.x & 31) == 0) {
    const int w = warp - 2;
    uint8_t* dst = smem + ring * stage_bytes + w * 3 * copy_chunk;
    const uint8_t* g = src + (static_cast<size_t>(blockIdx.x) * 2 + w) * (256u << 10);
    size_t off = 0;
    for (int i = 0; i < 3; ++i) {
      mbar_expect_tx(&cbar[w * 3 + i], copy_chunk);
      bulk_load(dst + i * copy_chunk, g + off, copy_chunk, &cbar[w * 3 + i]);
      off += copy_chunk;
    }
    int sl = 0;
    uint32_t ph = 0;
    unsigned long long copies = 0;
    while (true) {
      mbar_wait_nofail(&cbar[w * 3 + sl], ph, &abort_flag);
      if (*reinterpret_cast<volatile int*>(&stop_flag)) break;
      mbar_expect_tx(&cbar[w * 3 + sl], copy_chunk);
      bulk_load(dst + sl * copy_chunk, g + off, copy_chunk, &cbar[w * 3 + sl]);
      off += copy_chunk;
      if (off + copy_chunk > (256u << 10)) off = 0;
      ++copies;
      if (++sl == 3) {
        sl = 0;
        ph ^= 1u;
      }
    }
    // drain the two copies still in flight
    for (int i = 0; i < 2; ++i) {
      if (++sl == 3) {
        sl = 0;
        ph ^= 1u;
      }
      mbar_wait_nofail(&cbar[w * 3 + sl], ph, &abort_flag);
    }
    out[296 + blockIdx.x * 2 + w] = copies;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

int main() {
  unsigned long long* out;
  cudaMalloc(&out, 148 * 4 * sizeof(unsigned long long));
  uint8_t* src;
  cudaMalloc(&src, 148 * 2 * (256u << 10));
  cudaMemset(src, 0, 148 * 2 * (256u << 10));
  cudaFuncSetAttribute(mma_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int iters = 4096;
  printf("M   N   ring acc chunk ctas | issue cyc/MMA  retire cyc/MMA (median over CTAs)\n");
  const int Ms[2] = {128, 64};
  const int Ns[5] = {16, 64, 128, 192, 256};
  const int chunks[4] = {0, 8192, 16384, 24576};
  for (int mi = 0; mi < 1; ++mi)
    for (int ni = 0; ni < 5; ++ni)
      for (int ci = 0; ci < 4; ++ci)
          for (int ctas = 1; ctas <= 148; ctas += 147) {
            const int M = Ms[mi], N = Ns[ni], ring = 2, wait_each = ci, chunk = chunks[ci];
            const int smem = ring * (16384 + N * 128) + 6 * chunk + 1024;
            cudaMemset(out, 0, 148 * 4 * sizeof(unsigned long long));
            mma_probe<<<ctas, 128, smem>>>(M, N, iters, ring, 2, 0, out, src, chunk);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) {
              printf("launch failed: %s\n", cudaGetErrorString(e));
              return 1;
            }
            static unsigned long long h[592];
            cudaMemcpy(h, out, 592 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
            double copied = 0;
            for (int i = 0; i < 2 * ctas; ++i) copied += double(h[296 + i]) * chunk;
            // median by simple selection
            unsigned long long a[148], b[148];
            for (int i = 0; i < ctas; ++i) a[i] = h[2 * i], b[i] = h[2 * i + 1];
            for (int i = 0; i < ctas; ++i)
              for (int j = i + 1; j < ctas; ++j) {
                if (a[j] < a[i]) { unsigned long long t = a[i]; a[i] = a[j]; a[j] = t; }
                if (b[j] < b[i]) { unsigned long long t = b[i]; b[i] = b[j]; b[j] = t; }
              }
            printf("%3d %3d  %d    2 %5d %3d  | %8.1f %8.1f   copy stream %.1f B/cyc/SM\n", M, N, ring, chunk, ctas,
                   a[ctas / 2] / (4.0 * iters), b[ctas / 2] / (4.0 * iters), copied / ctas / double(b[ctas / 2]));
          }
  return 0;
}
