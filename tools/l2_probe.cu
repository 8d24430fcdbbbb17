// L2 -> shared-memory bulk-copy bandwidth probe (sm_100a): every CTA keeps `slots` cp.async.bulk copies of `chunk`
// bytes in flight from an L2-resident region, either private to the CTA, or shared by all CTAs (every CTA walks the
// same addresses in the same order, like the weight operand of a conv), or a mix.  Prints bytes/cycle/SM and the
// aggregate, for several CTA counts.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I ghost_b200/csrc tools/l2_probe.cu -o tools/bin/l2_probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "ptx.cuh"

using namespace ghost::ptx;

__device__ __forceinline__ void bulk_load(void* smem, const void* g, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem)),
               "l"(g), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// shared_every: 0 = all private, 1 = all shared, k>1 = every k-th chunk private, the rest shared
__global__ void __launch_bounds__(32, 1) probe(const uint8_t* base, size_t priv_bytes, size_t shared_bytes, int chunk,
                                               int slots, int iters, int shared_mode, unsigned long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar[16];
  __shared__ int abort_flag;
  if (threadIdx.x == 0) {
    abort_flag = 0;
    for (int i = 0; i < 16; ++i) mbar_init(&bar[i], 1);
    fence_mbar_init();
  }
  __syncwarp();
  if (threadIdx.x == 0) {
    const uint8_t* priv = base + shared_bytes + static_cast<size_t>(blockIdx.x) * priv_bytes;
    size_t po = 0, so = 0;
    long long c_exp = 0, c_tma = 0, c_wait = 0;
    auto issue = [&](int it, int s) {
      bool sh = shared_mode == 1 || (shared_mode > 1 && (it % shared_mode) != 0);
      const uint8_t* src;
      if (sh) {
        src = base + so;
        so += chunk;
        if (so + chunk > shared_bytes) so = 0;
      } else {
        src = priv + po;
        po += chunk;
        if (po + chunk > priv_bytes) po = 0;
      }
      const long long a = clock64();
      mbar_expect_tx(&bar[s], chunk);
      const long long b = clock64();
      bulk_load(smem + s * chunk, src, chunk, &bar[s]);
      const long long c = clock64();
      c_exp += b - a;
      c_tma += c - b;
    };
    // warm the L2 with one pass, untimed
    for (int it = 0; it < slots; ++it) issue(it, it);
    int s = 0;
    uint32_t ph = 0;
    long long t0 = 0;
    const int warm = 256;
    for (int it = 0; it < iters + warm; ++it) {
      if (it == warm) t0 = clock64();
      const long long w0 = clock64();
      mbar_wait_nofail(&bar[s], ph, &abort_flag);
      c_wait += clock64() - w0;
      issue(it + slots, s);
      if (++s == slots) {
        s = 0;
        ph ^= 1u;
      }
    }
    const long long t1 = clock64();
    for (int i = 0; i < slots; ++i) {
      mbar_wait_nofail(&bar[s], ph, &abort_flag);
      if (++s == slots) {
        s = 0;
        ph ^= 1u;
      }
    }
    out[blockIdx.x] = t1 - t0;
    if (blockIdx.x == 0) {
      out[148] = c_wait;
      out[149] = c_exp;
      out[150] = c_tma;
    }
  }
}

int main() {
  const size_t shared_bytes = 2u << 20, priv_bytes = 256u << 10;
  uint8_t* buf;
  cudaMalloc(&buf, shared_bytes + 148 * priv_bytes);
  cudaMemset(buf, 1, shared_bytes + 148 * priv_bytes);
  unsigned long long* out;
  cudaMalloc(&out, 152 * sizeof(unsigned long long));
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int iters = 4096;
  printf("chunk slots mode ctas | B/cyc/SM  aggregate B/cyc\n");
  const int chunks[3] = {4096, 16384, 32768};
  const int modes[4] = {0, 1, 2, 3};
  const int nm = 2;
  const int ctass[4] = {1, 37, 74, 148};
  for (int ci = 0; ci < 3; ++ci)
    for (int slots = 6; slots <= 12; slots += 6)
      for (int mi = 0; mi < nm; ++mi)
        for (int gi = 0; gi < 4; ++gi) {
          const int chunk = chunks[ci], ctas = ctass[gi];
          if (slots * chunk > 196608) continue;
          probe<<<ctas, 32, slots * chunk>>>(buf, priv_bytes, shared_bytes, chunk, slots, iters, modes[mi], out);
          cudaError_t e = cudaDeviceSynchronize();
          if (e != cudaSuccess) {
            printf("failed: %s\n", cudaGetErrorString(e));
            return 1;
          }
          unsigned long long h[152];
          cudaMemcpy(h, out, 152 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
          double sum = 0;
          for (int i = 0; i < ctas; ++i) sum += h[i];
          const double per = static_cast<double>(chunk) * iters / (sum / ctas);
          printf("%5d %3d   %d   %3d  | %7.1f  %9.0f   per copy: wait %.0f expect_tx %.0f issue %.0f cyc (CTA 0)\n", chunk, slots,
                 modes[mi], ctas, per, per * ctas, h[148] / double(iters + 256), h[149] / double(iters + 256 + slots),
                 h[150] / double(iters + 256 + slots));
        }
  return 0;
}
