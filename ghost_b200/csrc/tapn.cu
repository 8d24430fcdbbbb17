// See tapn.cuh.  Two persistent CTAs per SM (320 threads each): warp 0 = TMA producer, warp 1 = MMA issuer and
// TMEM owner, warps 2-5 and 6-9 = two epilogue groups on alternate tiles.  Per tile: one 16 KB TMA box per 64
// input channels -> four tcgen05.mma
// (M = 128 pixels, N = 32 = 9 taps x 3 channels padded, K = 16 each) -> the accumulator goes through shared memory,
// where every interior pixel collects its nine neighbours' partial sums.
#include "tapn.cuh"

#include <cuda_fp16.h>
#include <string.h>

#include <mutex>

#include "ptx.cuh"

namespace ghost {

namespace {

using namespace ptx;

constexpr int kStages = 4;
constexpr int kTStride = 35;  // odd (lane-strided row writes hit 32 banks) and = 3 mod 32 (the gather reads
                              // pixel-major, channel-minor: consecutive lanes -> consecutive banks)
constexpr int kThreads = 320;
constexpr int kTileW = kTapnTileW, kTileH = kTapnTileH;  // pixels per tile incl. the halo ring
constexpr int kInW = kTileW - 2, kInH = kTileH - 2;  // interior = outputs per tile

struct __align__(1024) TapnSmem {
  uint8_t a[kStages][16384];
  uint8_t w[kTapnMaxChunks][4096];
  float t[2][128 * kTStride];
  uint64_t full[kStages], empty[kStages], wfull, tfull[2], tempty[2];
  uint32_t tmem_base;
  int abort_flag;
};

// bar.sync with a literal barrier id (a register id makes ptxas reserve all 16 barriers of the CTA)
template <int ID>
__device__ __forceinline__ void group_sync() {
  asm volatile("bar.sync %0, 128;" ::"n"(ID) : "memory");
}
__device__ __forceinline__ void group_sync(int g) {
  if (g == 0)
    group_sync<1>();
  else
    group_sync<2>();
}

__device__ __forceinline__ void decode(const TapnParams& p, int tile, int* n, int* gy0, int* gx0) {
  // tiles are walked back to front: the producer of the input wrote it front to back, its tail is still in L2
  tile = p.tiles - 1 - tile;
  const int per = p.tiles_x * p.tiles_y;
  const int nn = tile / per, r = tile - nn * per;
  const int by = r / p.tiles_x, bx = r - by * p.tiles_x;
  *n = nn;
  *gy0 = by * kInH;
  *gx0 = bx * kInW;
}

__global__ void __launch_bounds__(kThreads, 2) tapn_kernel(const __grid_constant__ TapnParams p) {
  extern __shared__ uint8_t tapn_raw[];
  TapnSmem& sm = *reinterpret_cast<TapnSmem*>((reinterpret_cast<uintptr_t>(tapn_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  volatile int* abortp = &sm.abort_flag;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&sm.full[s], 1);
      mbar_init(&sm.empty[s], 1);
    }
    mbar_init(&sm.wfull, 1);
    for (int a = 0; a < 2; ++a) {
      mbar_init(&sm.tfull[a], 1);
      mbar_init(&sm.tempty[a], 4);
    }
    sm.abort_flag = 0;
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(&sm.tmem_base, 64);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;
  const int first = blockIdx.x, step = gridDim.x;

  if (warp == 0) {
    if (lane == 0) {
      prefetch_tmap(&p.tmA);
      prefetch_tmap(&p.tmW);
      mbar_expect_tx(&sm.wfull, p.chunks * 4096);
      for (int c = 0; c < p.chunks; ++c) tma_load_2d(sm.w[c], &p.tmW, &sm.wfull, 0, c * 32);
      int it = 0;
      bool ok = true;
      for (int t = first; t < p.tiles && ok; t += step) {
        int n, gy0, gx0;
        decode(p, t, &n, &gy0, &gx0);
        for (int c = 0; c < p.chunks; ++c, ++it) {
          const int s = it % kStages, ph = (it / kStages) & 1;
          if (!mbar_wait(&sm.empty[s], ph ^ 1, abortp)) {
            ok = false;
            break;
          }
          mbar_expect_tx(&sm.full[s], 16384);
          tma_load_4d(sm.a[s], &p.tmA, &sm.full[s], p.c0 + c * 64, gx0 - 1, gy0 - 1, n);
        }
      }
    }
  } else if (warp == 1) {
    mbar_wait_nofail(&sm.wfull, 0, abortp);
    const uint32_t idesc = make_idesc_f16(128, 32, p.f16 != 0);
    const uint32_t hi = smem_desc_hi(1024, false);
    int it = 0, tl = 0;
    for (int t = first; t < p.tiles; t += step, ++tl) {
      const int acc = tl & 1;
      mbar_wait_nofail(&sm.tempty[acc], ((tl >> 1) & 1) ^ 1, abortp);
      tc_fence_after();
      for (int c = 0; c < p.chunks; ++c, ++it) {
        const int s = it % kStages, ph = (it / kStages) & 1;
        mbar_wait_nofail(&sm.full[s], ph, abortp);
        tc_fence_after();
        umma_bf16_batch<4>(tmem + acc * 32, smem_desc_lo(smem_u32(sm.a[s])), smem_desc_lo(smem_u32(sm.w[c])), hi, idesc,
                           c > 0 ? 1u : 0u);
        umma_commit_elect(&sm.empty[s]);
      }
      umma_commit_elect(&sm.tfull[acc]);
    }
  } else {
    // two epilogue groups of four warps; group g owns accumulator stage g and every second tile of this CTA
    const int g = (warp - 2) >> 2;
    const int e = (threadIdx.x - 64) & 127;  // 0..127 within the group
    const int lg = warp & 3;                 // the TMEM lane group this warp may read
    const int m = lg * 32 + lane;            // tile pixel (row-major over 8 x 16) held by this thread
    float bias[3] = {0.f, 0.f, 0.f};
    if (p.bias) {
#pragma unroll
      for (int c = 0; c < 3; ++c)
        if (c < p.cout) bias[c] = __ldg(p.bias + c);
    }
    float* T = sm.t[g];
    // the (at most two) outputs of this thread inside a tile: pixel-major, channel-minor
    int oy[2], ox[2], co[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int idx = e + 128 * k;
      const int px = idx / 3;
      co[k] = idx - px * 3;
      oy[k] = px / kInW;
      ox[k] = px - oy[k] * kInW;
      if (idx >= kInW * kInH * 3 || co[k] >= p.cout) oy[k] = -1;
    }
    for (int t = first + g * step, tl = g; t < p.tiles; t += 2 * step, tl += 2) {
      int n, gy0, gx0;
      decode(p, t, &n, &gy0, &gx0);
      long long o[2];
      float r[2] = {0.f, 0.f};
      bool on[2];
#pragma unroll
      for (int k = 0; k < 2; ++k) {  // addresses and the residual first: their latency hides behind the MMA wait
        const int gy = gy0 + oy[k], gx = gx0 + ox[k];
        on[k] = oy[k] >= 0 && gy < p.H && gx < p.W;
        o[k] = p.o_off + n * p.o_sn + gy * p.o_sh + gx * p.o_sw + co[k];
        if (on[k] && p.res) {
          const long long ro = p.r_off + n * p.r_sn + gy * p.r_sh + gx * p.r_sw + co[k];
          r[k] = p.out_f32 ? __ldg(reinterpret_cast<const float*>(p.res) + ro)
                 : p.f16    ? __half2float(reinterpret_cast<const __half*>(p.res)[ro])
                            : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.res)[ro]);
        }
      }
      mbar_wait_nofail(&sm.tfull[g], (tl >> 1) & 1, abortp);
      tc_fence_after();
      uint32_t v0[16], v1[16];
      const uint32_t taddr = tmem + (static_cast<uint32_t>(lg * 32) << 16) + g * 32;
      tmem_ld16(taddr, v0);
      tmem_ld16(taddr + 16, v1);
      tmem_ld_wait();
      tc_fence_before();
      if (lane == 0) mbar_arrive(&sm.tempty[g]);
      group_sync(g);  // the group has finished reading the previous tile's T
#pragma unroll
      for (int j = 0; j < 16; ++j) T[m * kTStride + j] = __uint_as_float(v0[j]);
#pragma unroll
      for (int j = 0; j < 11; ++j) T[m * kTStride + 16 + j] = __uint_as_float(v1[j]);
      group_sync(g);
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        if (!on[k]) continue;
        const float* tp = T + (oy[k] * kTileW + ox[k]) * kTStride + co[k];
        float s = 0.f;
#pragma unroll
        for (int kh = 0; kh < 3; ++kh)
#pragma unroll
          for (int kw = 0; kw < 3; ++kw) s += tp[(kh * kTileW + kw) * kTStride + (kh * 3 + kw) * 3];
        s += co[k] == 0 ? bias[0] : (co[k] == 1 ? bias[1] : bias[2]);
        s += r[k];
        if (p.act2 == kAct2Tanh) s = tanhf(s);
        if (p.out_f32)
          reinterpret_cast<float*>(p.out)[o[k]] = s;
        else if (p.f16)
          reinterpret_cast<__half*>(p.out)[o[k]] = __float2half_rn(s);
        else
          reinterpret_cast<__nv_bfloat16*>(p.out)[o[k]] = __float2bfloat16(s);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem, 64);
  }
  if (sm.abort_flag && threadIdx.x == 0 && p.err) *reinterpret_cast<volatile int*>(p.err) = 1;
}

}  // namespace

std::vector<float> tapn_pack_rows(const float* rows_in, int Co, int Ci) {
  const int chunks = (Ci + 63) / 64;
  std::vector<float> rows(static_cast<size_t>(chunks) * 32 * 64, 0.f);
  for (int co = 0; co < Co; ++co)
    for (int t = 0; t < 9; ++t)
      for (int ci = 0; ci < Ci; ++ci)
        rows[(static_cast<size_t>(ci / 64) * 32 + t * 3 + co) * 64 + (ci % 64)] =
            rows_in[(static_cast<size_t>(co) * 9 + t) * Ci + ci];
  return rows;
}

const char* tapn_launch(const TapnParams& p, cudaStream_t s) {
  // the dynamic shared-memory opt-in and the SM count are per device (one process may drive several GPUs)
  static std::mutex mu;
  static int sms_of[64] = {0};
  const int smem = static_cast<int>(sizeof(TapnSmem)) + 1024;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return "tapn: bad device ordinal";
  int sms;
  {
    std::lock_guard<std::mutex> lock(mu);
    if (!sms_of[dev]) {
      cudaError_t e = cudaFuncSetAttribute(tapn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
      int n = 0;
      if (e == cudaSuccess) e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
      if (e != cudaSuccess) {
        cudaGetLastError();
        return cudaGetErrorString(e);
      }
      sms_of[dev] = n > 0 ? n : 148;
    }
    sms = sms_of[dev];
  }
  const int grid = p.tiles < 2 * sms ? p.tiles : 2 * sms;
  tapn_kernel<<<grid, kThreads, smem, s>>>(p);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
}

}  // namespace ghost
