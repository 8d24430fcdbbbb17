#include <type_traits>
#include "simt.cuh"

#include <cuda_fp16.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "sr_kernels.cuh"  // in_stats_launch_ex

#include "igemm.cuh"  // Act1 / Act2 enums
#include "ptx.cuh"    // packed fp32 helpers (fma2 / add2 / sub2 / mul2)

namespace ghost {

namespace {

template <typename T>
__device__ __forceinline__ float ldf(const T* p);
template <>
__device__ __forceinline__ float ldf<float>(const float* p) {
  return __ldg(p);
}
template <>
__device__ __forceinline__ float ldf<__nv_bfloat16>(const __nv_bfloat16* p) {
  return __bfloat162float(*p);
}
template <>
__device__ __forceinline__ float ldf<__half>(const __half* p) {
  return __half2float(*p);
}
template <typename T>
__device__ __forceinline__ void stf(T* p, float v);
template <>
__device__ __forceinline__ void stf<float>(float* p, float v) {
  *p = v;
}
template <>
__device__ __forceinline__ void stf<__nv_bfloat16>(__nv_bfloat16* p, float v) {
  *p = __float2bfloat16(v);
}
template <>
__device__ __forceinline__ void stf<__half>(__half* p, float v) {
  *p = __float2half(v);
}

inline const char* last_err() {
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
}

template <typename T>
__device__ __forceinline__ void load8(const T* p, float (&v)[8]);
template <>
__device__ __forceinline__ void load8<float>(const float* p, float (&v)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p));
  const float4 b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
  v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
template <>
__device__ __forceinline__ void load8<__nv_bfloat16>(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 q = __ldg(reinterpret_cast<const uint4*>(p));
  const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    v[2 * i] = __uint_as_float(w[i] << 16);
    v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
template <>
__device__ __forceinline__ void load8<__half>(const __half* p, float (&v)[8]) {
  const uint4 q = __ldg(reinterpret_cast<const uint4*>(p));
  const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
    v[2 * i] = f.x;
    v[2 * i + 1] = f.y;
  }
}
// the same load in two steps, so that several loads can be in flight with only their raw 16 bytes live
template <typename T>
struct Raw8 {
  uint4 q;
  __device__ __forceinline__ void load(const T* p) { q = __ldg(reinterpret_cast<const uint4*>(p)); }
  __device__ __forceinline__ void get(float (&v)[8]) const {
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (sizeof(T) == 2 && std::is_same<T, __half>::value) {
        const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
        v[2 * i] = f.x;
        v[2 * i + 1] = f.y;
      } else {
        v[2 * i] = __uint_as_float(w[i] << 16);
        v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
      }
    }
  }
};
template <>
struct Raw8<float> {
  float4 a, b;
  __device__ __forceinline__ void load(const float* p) {
    a = __ldg(reinterpret_cast<const float4*>(p));
    b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  }
  __device__ __forceinline__ void get(float (&v)[8]) const {
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
    v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
};
template <typename T>
__device__ __forceinline__ void store8(T* p, const float (&v)[8]);
template <>
__device__ __forceinline__ void store8<float>(float* p, const float (&v)[8]) {
  reinterpret_cast<float4*>(p)[0] = make_float4(v[0], v[1], v[2], v[3]);
  reinterpret_cast<float4*>(p)[1] = make_float4(v[4], v[5], v[6], v[7]);
}
template <>
__device__ __forceinline__ void store8<__nv_bfloat16>(__nv_bfloat16* p, const float (&v)[8]) {
  uint4 q;
  __nv_bfloat162 t;
  t = __floats2bfloat162_rn(v[0], v[1]); q.x = *reinterpret_cast<uint32_t*>(&t);
  t = __floats2bfloat162_rn(v[2], v[3]); q.y = *reinterpret_cast<uint32_t*>(&t);
  t = __floats2bfloat162_rn(v[4], v[5]); q.z = *reinterpret_cast<uint32_t*>(&t);
  t = __floats2bfloat162_rn(v[6], v[7]); q.w = *reinterpret_cast<uint32_t*>(&t);
  *reinterpret_cast<uint4*>(p) = q;
}

template <>
__device__ __forceinline__ void store8<__half>(__half* p, const float (&v)[8]) {
  uint4 q;
  __half2 t;
  t = __floats2half2_rn(v[0], v[1]); q.x = *reinterpret_cast<uint32_t*>(&t);
  t = __floats2half2_rn(v[2], v[3]); q.y = *reinterpret_cast<uint32_t*>(&t);
  t = __floats2half2_rn(v[4], v[5]); q.z = *reinterpret_cast<uint32_t*>(&t);
  t = __floats2half2_rn(v[6], v[7]); q.w = *reinterpret_cast<uint32_t*>(&t);
  *reinterpret_cast<uint4*>(p) = q;
}

// ---------------------------------------------------------------------------
// Direct convolution as a shared-memory tiled GEMM on CUDA cores.
// 64 pixels x 64 output channels per block, 256 threads, 4x4 outputs per thread.
// ---------------------------------------------------------------------------
constexpr int TM = 64, TN = 64, TK = 16;

template <typename T>
__global__ void __launch_bounds__(256) simt_conv_kernel(const SimtConvParams p) {
  __shared__ float As[TK][TM + 4];
  __shared__ float Bs[TK][TN + 4];
  const T* in = reinterpret_cast<const T*>(p.in);
  const int K = p.ntaps * p.Cin;
  const long long M = static_cast<long long>(p.N) * p.Ho * p.Wo;
  const long long m0 = static_cast<long long>(blockIdx.x) * TM;
  const int n0 = blockIdx.y * TN;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;

  // loader mapping: each thread fetches 4 consecutive k for one row
  const int lrow = threadIdx.x >> 2, lk = (threadIdx.x & 3) * 4;
  const long long lm = m0 + lrow;
  int ln = 0, loh = 0, low = 0;
  const bool lvalid = lm < M;
  if (lvalid) {
    low = static_cast<int>(lm % p.Wo);
    long long t = lm / p.Wo;
    loh = static_cast<int>(t % p.Ho);
    ln = static_cast<int>(t / p.Ho);
  }
  const int lco = n0 + lrow;

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < K; k0 += TK) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int k = k0 + lk + q;
      float a = 0.f, b = 0.f;
      if (k < K) {
        const int tap = k / p.Cin, c = k - tap * p.Cin;
        if (lvalid) {
          const int ih = loh * p.stride + p.dh[tap], iw = low * p.stride + p.dw[tap];
          if (ih >= 0 && ih < p.H && iw >= 0 && iw < p.W)
            a = ldf(in + ((static_cast<long long>(ln) * p.H + ih) * p.W + iw) * p.C + p.c0 + c);
        }
        if (lco < p.Cout) b = __ldg(p.w + static_cast<long long>(lco) * K + k);
      }
      As[lk + q][lrow] = a;
      Bs[lk + q][lrow] = b;
    }
    __syncthreads();
    // two-level (blocked) summation: the 16 products of a K tile are summed on their own and then added to the
    // running total, which keeps the fp32 rounding error of K = 9216..18432 term sums near that of a pairwise sum
    float part[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) part[i][j] = 0.f;
#pragma unroll
    for (int kk = 0; kk < TK; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) part[i][j] = fmaf(a[i], b[j], part[i][j]);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] += part[i][j];
    __syncthreads();
  }

  T* out = reinterpret_cast<T*>(p.out);
  const T* res = reinterpret_cast<const T*>(p.res);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long m = m0 + ty * 4 + i;
    if (m >= M) continue;
    const int ow = static_cast<int>(m % p.Wo);
    const long long t = m / p.Wo;
    const int oh = static_cast<int>(t % p.Ho);
    const int n = static_cast<int>(t / p.Ho);
    const long long o_pix = n * p.o_sn + oh * p.o_sh + ow * p.o_sw + p.o_off;
    const long long r_pix = n * p.r_sn + oh * p.r_sh + ow * p.r_sw + p.r_off;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int co = n0 + tx * 4 + j;
      if (co >= p.Cout) continue;
      float v = acc[i][j];
      if (p.bias) v += __ldg(p.bias + co);
      if (p.act1 == kActLeaky)
        v = v > 0.f ? v : v * p.slope;
      else if (p.act1 == kActPRelu)
        v = v > 0.f ? v : v * __ldg(p.slope_c + co);
      else if (p.act1 == kActRelu)
        v = fmaxf(v, 0.f);
      if (res) v += ldf(res + r_pix + co);
      if (p.act2 == kAct2Tanh) v = tanhf(v);
      stf(out + o_pix + co, v);
    }
  }
}

// ---------------------------------------------------------------------------
// Direct convolution for the 3-channel input layers (encoder conv1, iresnet stem): one thread per output
// pixel keeps all COUT accumulators in registers, the [taps*3][COUT] weight matrix sits in shared memory and
// is read as broadcast float4s.  K = taps*3 <= 48, so the layer is FMA-issue bound, not a GEMM-tile problem.
// ---------------------------------------------------------------------------
template <typename T, int COUT>
__global__ void __launch_bounds__(128) rgb_conv_kernel(const SimtConvParams p) {
  __shared__ __align__(16) float ws[48 * COUT];
  __shared__ float bs[COUT], ss[COUT];
  const int K = p.ntaps * 3;
  for (int i = threadIdx.x; i < K * COUT; i += blockDim.x) {
    const int k = i / COUT, co = i - k * COUT;
    ws[i] = __ldg(p.w + static_cast<long long>(co) * K + k);
  }
  for (int i = threadIdx.x; i < COUT; i += blockDim.x) {
    bs[i] = p.bias ? __ldg(p.bias + i) : 0.f;
    ss[i] = p.act1 == kActPRelu ? __ldg(p.slope_c + i) : p.slope;
  }
  __syncthreads();
  const long long M = static_cast<long long>(p.N) * p.Ho * p.Wo;
  // a block walks many groups of 128 pixels: the weight matrix is staged once per block, not once per 128 pixels
  for (long long m = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; m < M;
       m += static_cast<long long>(gridDim.x) * blockDim.x) {
  const int ow = static_cast<int>(m % p.Wo);
  const long long t = m / p.Wo;
  const int oh = static_cast<int>(t % p.Ho);
  const int n = static_cast<int>(t / p.Ho);
  const T* in = reinterpret_cast<const T*>(p.in);
  float acc[COUT];
#pragma unroll
  for (int i = 0; i < COUT; ++i) acc[i] = 0.f;
  for (int tap = 0; tap < p.ntaps; ++tap) {
    const int ih = oh * p.stride + p.dh[tap], iw = ow * p.stride + p.dw[tap];
    if (ih < 0 || ih >= p.H || iw < 0 || iw >= p.W) continue;
    const T* px = in + ((static_cast<long long>(n) * p.H + ih) * p.W + iw) * 3;
    const float x[3] = {ldf(px), ldf(px + 1), ldf(px + 2)};
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float4* wr = reinterpret_cast<const float4*>(ws + (tap * 3 + c) * COUT);
#pragma unroll
      for (int q = 0; q < COUT / 4; ++q) {
        const float4 w = wr[q];
        acc[4 * q] = fmaf(x[c], w.x, acc[4 * q]);
        acc[4 * q + 1] = fmaf(x[c], w.y, acc[4 * q + 1]);
        acc[4 * q + 2] = fmaf(x[c], w.z, acc[4 * q + 2]);
        acc[4 * q + 3] = fmaf(x[c], w.w, acc[4 * q + 3]);
      }
    }
  }
  T* out = reinterpret_cast<T*>(p.out) + n * p.o_sn + oh * p.o_sh + ow * p.o_sw + p.o_off;
#pragma unroll
  for (int q = 0; q < COUT / 8; ++q) {
    float v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float a = acc[8 * q + i] + bs[8 * q + i];
      if (p.act1 == kActLeaky || p.act1 == kActPRelu)
        a = a > 0.f ? a : a * ss[8 * q + i];
      else if (p.act1 == kActRelu)
        a = fmaxf(a, 0.f);
      v[i] = a;
    }
    store8(out + 8 * q, v);
  }
  }
}

// Same layer with the weights in the kernel-parameter constant bank (simt.cuh RgbConvConst).  K x K taps, stride S,
// pad 1; every loop is unrolled so each FFMA reads its weight as c[0][imm].
template <typename T, int COUT, int K, int S>
__global__ void __launch_bounds__(128) rgb_const_kernel(const SimtConvParams p, const __grid_constant__ RgbConvConst cw) {
  const long long M = static_cast<long long>(p.N) * p.Ho * p.Wo;
  const T* in = reinterpret_cast<const T*>(p.in);
  for (long long m = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; m < M;
       m += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int ow = static_cast<int>(m % p.Wo);
    const long long t = m / p.Wo;
    const int oh = static_cast<int>(t % p.Ho);
    const int n = static_cast<int>(t / p.Ho);
    float acc[COUT];
#pragma unroll
    for (int i = 0; i < COUT; ++i) acc[i] = 0.f;
#pragma unroll
    for (int kh = 0; kh < K; ++kh) {
      const int ih = oh * S + kh - 1;
      const bool vh = ih >= 0 && ih < p.H;
#pragma unroll
      for (int kw = 0; kw < K; ++kw) {
        const int iw = ow * S + kw - 1;
        const bool v = vh && iw >= 0 && iw < p.W;
        const T* px = in + ((static_cast<long long>(n) * p.H + (v ? ih : 0)) * p.W + (v ? iw : 0)) * 3;
        float x[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) x[c] = v ? ldf(px + c) : 0.f;
#pragma unroll
        for (int c = 0; c < 3; ++c)
#pragma unroll
          for (int co = 0; co < COUT; ++co) acc[co] = fmaf(x[c], cw.w[((kh * K + kw) * 3 + c) * COUT + co], acc[co]);
      }
    }
    T* out = reinterpret_cast<T*>(p.out) + n * p.o_sn + oh * p.o_sh + ow * p.o_sw + p.o_off;
#pragma unroll
    for (int q = 0; q < COUT / 8; ++q) {
      float v[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float a = acc[8 * q + i] + cw.bias[8 * q + i];
        if (p.act1 == kActLeaky || p.act1 == kActPRelu)
          a = a > 0.f ? a : a * cw.slope[8 * q + i];
        else if (p.act1 == kActRelu)
          a = fmaxf(a, 0.f);
        v[i] = a;
      }
      store8(out + 8 * q, v);
    }
  }
}

template <typename T>
const char* rgb_conv_dispatch(const SimtConvParams& p, cudaStream_t s) {
  const long long M = static_cast<long long>(p.N) * p.Ho * p.Wo;
  long long nb = (M + 127) / 128;
  if (nb > 148 * 8) nb = 148 * 8;
  const unsigned blocks = static_cast<unsigned>(nb);
  if (p.Cout == 32)
    rgb_conv_kernel<T, 32><<<blocks, 128, 0, s>>>(p);
  else
    rgb_conv_kernel<T, 64><<<blocks, 128, 0, s>>>(p);
  return last_err();
}

}  // namespace

namespace {
// 0: not a stem shape; 1: 4x4 s2 p1 -> 32; 2: 3x3 s1 p1 -> 64
int rgb_const_shape(const SimtConvParams& p) {
  const uintptr_t al = reinterpret_cast<uintptr_t>(p.out);
  if (p.Cin != 3 || p.C != 3 || p.c0 != 0 || p.res || p.act2 != kAct2None || (al % 16) || (p.o_sw % 8) ||
      (p.o_sh % 8) || (p.o_sn % 8) || (p.o_off % 8))
    return 0;
  int k = 0, shape = 0;
  if (p.ntaps == 16 && p.stride == 2 && p.Cout == 32) k = 4, shape = 1;
  if (p.ntaps == 9 && p.stride == 1 && p.Cout == 64) k = 3, shape = 2;
  if (!shape) return 0;
  for (int i = 0; i < p.ntaps; ++i)
    if (p.dh[i] != i / k - 1 || p.dw[i] != i % k - 1) return 0;
  return shape;
}
}  // namespace

bool rgb_const_supported(const SimtConvParams& p, int dtype) {
  static const int on = [] {
    const char* e = getenv("GHOST_RGB_CONST");  // A/B switch for profiling: 0 keeps the shared-memory kernel
    return e ? (atoi(e) != 0) : 1;
  }();
  return on && (dtype == kF32 || dtype == kBF16 || dtype == kF16) && rgb_const_shape(p) != 0;
}

const char* rgb_const_prepare(const SimtConvParams& p, RgbConvConst* c) {
  const int K = p.ntaps * 3;
  std::vector<float> w(static_cast<size_t>(p.Cout) * K);
  memset(c, 0, sizeof(*c));
  if (cudaMemcpy(w.data(), p.w, w.size() * sizeof(float), cudaMemcpyDeviceToHost) != cudaSuccess) return "rgb_const: weight copy failed";
  for (int co = 0; co < p.Cout; ++co)
    for (int k = 0; k < K; ++k) c->w[k * p.Cout + co] = w[static_cast<size_t>(co) * K + k];
  if (p.bias && cudaMemcpy(c->bias, p.bias, p.Cout * sizeof(float), cudaMemcpyDeviceToHost) != cudaSuccess)
    return "rgb_const: bias copy failed";
  if (p.act1 == kActPRelu) {
    if (cudaMemcpy(c->slope, p.slope_c, p.Cout * sizeof(float), cudaMemcpyDeviceToHost) != cudaSuccess)
      return "rgb_const: slope copy failed";
  } else {
    for (int i = 0; i < p.Cout; ++i) c->slope[i] = p.slope;
  }
  return nullptr;
}

const char* rgb_const_launch(const SimtConvParams& p, const RgbConvConst& c, int dtype, cudaStream_t s) {
  const long long M = static_cast<long long>(p.N) * p.Ho * p.Wo;
  long long nb = (M + 127) / 128;
  if (nb > 148 * 16) nb = 148 * 16;
  const unsigned blocks = static_cast<unsigned>(nb);
  const int shape = rgb_const_shape(p);
  if (shape == 1 && dtype == kBF16)
    rgb_const_kernel<__nv_bfloat16, 32, 4, 2><<<blocks, 128, 0, s>>>(p, c);
  else if (shape == 1 && dtype == kF32)
    rgb_const_kernel<float, 32, 4, 2><<<blocks, 128, 0, s>>>(p, c);
  else if (shape == 2 && dtype == kBF16)
    rgb_const_kernel<__nv_bfloat16, 64, 3, 1><<<blocks, 128, 0, s>>>(p, c);
  else if (shape == 2 && dtype == kF32)
    rgb_const_kernel<float, 64, 3, 1><<<blocks, 128, 0, s>>>(p, c);
  else if (shape == 1 && dtype == kF16)
    rgb_const_kernel<__half, 32, 4, 2><<<blocks, 128, 0, s>>>(p, c);
  else if (shape == 2 && dtype == kF16)
    rgb_const_kernel<__half, 64, 3, 1><<<blocks, 128, 0, s>>>(p, c);
  else
    return "rgb_const: unsupported shape";
  return last_err();
}

const char* simt_conv_launch(const SimtConvParams& p, int dtype, cudaStream_t s) {
  const long long M = static_cast<long long>(p.N) * p.Ho * p.Wo;
  if (M <= 0 || p.Cout <= 0) return "simt_conv: empty problem";
  const int esz = dtype == kF32 ? 4 : 2;
  const bool out_al16 = (reinterpret_cast<uintptr_t>(p.out) % 16 == 0) && ((p.o_sn * esz) % 16 == 0) &&
                        ((p.o_sh * esz) % 16 == 0) && ((p.o_sw * esz) % 16 == 0) && ((p.o_off * esz) % 16 == 0);
  if (p.Cin == 3 && p.C == 3 && p.c0 == 0 && (p.Cout == 32 || p.Cout == 64) && !p.res && p.act2 == kAct2None &&
      p.ntaps <= 16 && out_al16) {
    if (dtype == kF32) return rgb_conv_dispatch<float>(p, s);
    if (dtype == kBF16) return rgb_conv_dispatch<__nv_bfloat16>(p, s);
    if (dtype == kF16) return rgb_conv_dispatch<__half>(p, s);
  }
  dim3 grid(static_cast<unsigned>((M + TM - 1) / TM), (p.Cout + TN - 1) / TN);
  if (dtype == kF32)
    simt_conv_kernel<float><<<grid, 256, 0, s>>>(p);
  else if (dtype == kBF16)
    simt_conv_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(p);
  else if (dtype == kF16)
    simt_conv_kernel<__half><<<grid, 256, 0, s>>>(p);
  else
    return "simt_conv: unsupported dtype";
  return last_err();
}

// ---------------------------------------------------------------------------
// Instance-norm statistics (nn.InstanceNorm2d, network/AADLayer.py:17,24):
// fp32 sum and sum of squares per (n, c).  One thread owns 8 consecutive
// channels (a 16-byte vector for bf16), a block strides over a slab of pixels,
// partials are combined in shared memory and flushed with one atomic per
// (block, channel).
// ---------------------------------------------------------------------------
namespace {

constexpr int kStatThreads = 256;

// Partial statistics layout for one tensor: per image n a block of (2*P + 1) rows of C floats:
//   row 2p   = sum_p (x - shift)        over the pixels of part p
//   row 2p+1 = sum_p (x - shift)^2
//   row 2P   = shift = x[n, pixel 0, c]  (one common shift per (n,c) keeps the sums linear across parts and
//              removes the E[x^2]-E[x]^2 cancellation: |mean - shift| is of the order of the std)
// Every part is written by exactly one block and the fold kernel adds them in a fixed order, so the
// statistics are deterministic (no atomics, no memset).
//
// s += (v - shift), q += (v - shift)^2 for eight channels, two per packed fp32 instruction.
__device__ __forceinline__ void stats_accumulate8(const float (&v)[8], const float (&sft)[8], float (&s)[8],
                                                  float (&q)[8]) {
#pragma unroll
  for (int i = 0; i < 8; i += 2) {
    const float2 d = ptx::sub2(make_float2(v[i], v[i + 1]), make_float2(sft[i], sft[i + 1]));
    const float2 sn = ptx::add2(make_float2(s[i], s[i + 1]), d);
    const float2 qn = ptx::fma2(d, d, make_float2(q[i], q[i + 1]));
    s[i] = sn.x; s[i + 1] = sn.y;
    q[i] = qn.x; q[i + 1] = qn.y;
  }
}
// (two blocks per SM at 86 registers left the kernel latency-bound at 37 % of the DRAM peak - ncu, round 2)
template <typename T>
__global__ void __launch_bounds__(kStatThreads, 4) in_stats_kernel(const T* __restrict__ x, int HW, int C,
                                                                 int pix_per_block, float* __restrict__ parts,
                                                                 int common_shift) {
  extern __shared__ float sh[];  // [kStatThreads][16]
  // blocks walk the tensor backwards: its producer wrote it front to back, so the tail is what L2 still holds
  const int n = gridDim.y - 1 - blockIdx.y;
  const int bx = gridDim.x - 1 - blockIdx.x;
  const int P = gridDim.x;
  const int vecs = C >> 3;                       // channel vectors per pixel
  const int ppi = kStatThreads / vecs;           // pixels per iteration (vecs <= 256)
  const int cv = threadIdx.x % vecs, pl = threadIdx.x / vecs;
  const int p0 = bx * pix_per_block;
  const int p1 = min(HW, p0 + pix_per_block);
  float s[8], q[8], sft[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) s[i] = q[i] = 0.f;
  const T* base = x + (static_cast<long long>(n) * HW) * C + cv * 8;
  // common_shift: image 0's first pixel for every image, so that the sums add up over the batch (BatchNorm)
  load8(common_shift ? x + cv * 8 : base, sft);
  if (pl < ppi) {
    // four independent 16-byte loads in flight per thread: the kernel is a pure HBM stream
    for (int p = p0 + pl; p < p1; p += 4 * ppi) {
      Raw8<T> raw[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (p + u * ppi < p1) raw[u].load(base + static_cast<long long>(p + u * ppi) * C);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (p + u * ppi < p1) {
          float v[8];
          raw[u].get(v);
          stats_accumulate8(v, sft, s, q);
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    sh[threadIdx.x * 16 + i] = s[i];
    sh[threadIdx.x * 16 + 8 + i] = q[i];
  }
  __syncthreads();
  float* out = parts + static_cast<long long>(n) * (2 * P + 1) * C;
  // thread t < vecs*16 reduces one (channel vector, slot) column over the ppi pixel lanes, in a fixed order
  for (int t = threadIdx.x; t < vecs * 16; t += kStatThreads) {
    const int v = t >> 4, slot = t & 15;
    float a = 0.f;
    for (int l = 0; l < ppi; ++l) a += sh[(l * vecs + v) * 16 + slot];
    const int c = v * 8 + (slot & 7);
    out[(2 * bx + (slot >> 3)) * C + c] = a;
  }
  if (bx == 0 && pl == 0) {
#pragma unroll
    for (int i = 0; i < 8; ++i) out[2 * P * C + cv * 8 + i] = sft[i];
  }
}

// grid (N, C/64): block (n, by) folds 64 channels with 4 part-lanes each.  Every lane adds its parts in a fixed
// order and lane 0 adds the lanes in a fixed order, so the statistics stay deterministic however many parts the
// producers wrote.  The conv_h bias needs a sum over all channels: each block leaves its partial in
// scratch[n][by], the last block of image n to finish (self-resetting counter) adds them in index order.
// blockIdx.z selects the layer: two AADLayers that normalise the SAME tensor (the shortcut layer and the first main
// layer of an AAD block) fold their parameters in one launch.
__global__ void __launch_bounds__(256) aad_fold_kernel(const float* __restrict__ parts, int P, int HW, int C,
                                                        const float* __restrict__ zfc, int ldz, const AadFoldSet fa,
                                                        const AadFoldSet fb) {
  const AadFoldSet& fs = blockIdx.z ? fb : fa;
  const int off_gamma = fs.off_gamma, off_beta = fs.off_beta, ld_out = fs.ld_out;
  const float* __restrict__ w_h = fs.w_h;
  const float b_h = fs.b_h;
  float4* __restrict__ ncpar = fs.ncpar;
  float* __restrict__ wh_fold = fs.wh_fold;
  float* __restrict__ bh_fold = fs.bh_fold;
  float* scratch = fs.scratch;
  __shared__ double shs[256], shq[256];
  __shared__ float red[2];
  __shared__ int is_last;
  const int n = blockIdx.x, by = blockIdx.y, gy = gridDim.y;
  const double inv = 1.0 / static_cast<double>(HW);
  const float* pn = parts + static_cast<long long>(n) * (2 * P + 1) * C;
  const int cl = threadIdx.x & 63, lane = threadIdx.x >> 6;
  const int c = by * 64 + cl;
  double su0 = 0.0, su1 = 0.0, sq0 = 0.0, sq1 = 0.0;
  if (c < C) {
    for (int p = lane; p < P; p += 8) {
      su0 += static_cast<double>(pn[(2 * p) * C + c]);
      sq0 += static_cast<double>(pn[(2 * p + 1) * C + c]);
      if (p + 4 < P) {
        su1 += static_cast<double>(pn[(2 * (p + 4)) * C + c]);
        sq1 += static_cast<double>(pn[(2 * (p + 4) + 1) * C + c]);
      }
    }
  }
  shs[threadIdx.x] = su0 + su1;
  shq[threadIdx.x] = sq0 + sq1;
  __syncthreads();
  float part = 0.f;
  if (lane == 0 && c < C) {
    double su = 0.0, sq = 0.0;
    for (int l = 0; l < 4; ++l) {
      su += shs[l * 64 + cl];
      sq += shq[l * 64 + cl];
    }
    const double d = su * inv;
    const float mu = static_cast<float>(static_cast<double>(pn[2 * P * C + c]) + d);
    const float var = static_cast<float>(fmax(sq * inv - d * d, 0.0));  // biased variance (InstanceNorm2d)
    const float rstd = 1.0f / sqrtf(var + 1e-5f);
    const float g = zfc[static_cast<long long>(n) * ldz + off_gamma + c];
    const float b = zfc[static_cast<long long>(n) * ldz + off_beta + c];
    // laid out per channel pair for the packed-fp32 epilogue: entry c&~1 = {rstd0, rstd1, shift0, shift1},
    // entry (c&~1)+1 = {gamma_id0, gamma_id1, beta_id0, beta_id1}
    float* pp = reinterpret_cast<float*>(ncpar + static_cast<long long>(n) * ld_out + (c & ~1)) + (c & 1);
    pp[0] = rstd;
    pp[2] = -mu * rstd;
    pp[4] = g;
    pp[6] = b;
    const float w = w_h[c] * rstd;
    wh_fold[static_cast<long long>(n) * ld_out + c] = w;
    part = w * mu;
  }
  if (threadIdx.x < 64) {  // warps 0 and 1 hold the 64 channel terms
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = part;
  }
  __syncthreads();
  float* sc = scratch + static_cast<long long>(n) * (gy + 1);
  if (threadIdx.x == 0) {
    sc[1 + by] = red[0] + red[1];
    __threadfence();
    const unsigned done = atomicAdd(reinterpret_cast<unsigned*>(sc), 1u);
    is_last = done == static_cast<unsigned>(gy - 1);
    if (is_last) {
      __threadfence();
      float v = 0.f;
      for (int i = 0; i < gy; ++i) v += *reinterpret_cast<volatile float*>(sc + 1 + i);
      bh_fold[n] = b_h - v;
      *reinterpret_cast<unsigned*>(sc) = 0u;  // ready for the next launch (graph replays included)
    }
  }
}

// One warp per pixel: pass 1 = mask logit (warp reduce), pass 2 = blend + ReLU.
template <typename T>
__global__ void __launch_bounds__(256) aad_apply_kernel(const T* __restrict__ h, const T* __restrict__ gb,
                                                         T* __restrict__ out, long long npix, int HW, int C,
                                                         const float4* __restrict__ ncpar,
                                                         const float* __restrict__ wh_fold,
                                                         const float* __restrict__ bh_fold) {
  const int lane = threadIdx.x & 31;
  const long long warp0 = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  for (long long pix = warp0; pix < npix; pix += nwarps) {
    const int n = static_cast<int>(pix / HW);
    const T* hrow = h + pix * C;
    const T* grow = gb + pix * 2 * C;
    const float* wrow = wh_fold + static_cast<long long>(n) * C;
    float acc = 0.f;
    for (int c = lane; c < C; c += 32) acc = fmaf(ldf(hrow + c), __ldg(wrow + c), acc);
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    const float m = 1.f / (1.f + expf(-(acc + __ldg(bh_fold + n))));
    for (int c = lane; c < C; c += 32) {
      const float* pp = reinterpret_cast<const float*>(ncpar + static_cast<long long>(n) * C + (c & ~1)) + (c & 1);
      const float4 pr = make_float4(__ldg(pp), __ldg(pp + 2), __ldg(pp + 4), __ldg(pp + 6));
      const float hn = fmaf(ldf(hrow + c), pr.x, pr.y);
      const float A = fmaf(ldf(grow + c), hn, ldf(grow + C + c));
      const float I = fmaf(pr.z, hn, pr.w);
      stf(out + pix * C + c, fmaxf((1.f - m) * A + m * I, 0.f));
    }
  }
}

// sigmoid(conv_h(instance_norm(h))) per pixel with the normalisation folded into wh/bh: one warp per pixel,
// 16-byte loads, warp-shuffle reduction over the channels (used when c_x > 128, i.e. several N tiles per pixel).
// The folded weights depend on the image only: every lane keeps its NIT x 8 of them in registers and reloads them
// when its pixel sequence crosses into the next image, so the inner loop is one 16-byte load of h per 8 channels.
// DUAL: a second layer on the same h (the shortcut layer and the first main layer of a block) in the same pass.
template <int NIT, bool DUAL, bool F16>
__global__ void __launch_bounds__(256) aad_mask_kernel(const __nv_bfloat16* __restrict__ h, long long npix, int HW, int C,
                                                        const float* __restrict__ wh_fold,
                                                        const float* __restrict__ bh_fold, float* __restrict__ mask,
                                                        const float* __restrict__ wh_fold2,
                                                        const float* __restrict__ bh_fold2, float* __restrict__ mask2) {
  const int lane = threadIdx.x & 31;
  const long long warp0 = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  float w[NIT][8], u[DUAL ? NIT : 1][8];
  float bh = 0.f, bh2 = 0.f;
  int cur_n = -1;
  // every warp owns a run of consecutive pixels, so it crosses an image boundary at most once or twice
  const long long chunk = (npix + nwarps - 1) / nwarps;
  const long long pend = min(npix, (warp0 + 1) * chunk);
  // the next pixel's channels are requested before the current pixel is reduced: two loads in flight per lane
  uint4 cur[NIT];
  {
    const long long pix = warp0 * chunk;
#pragma unroll
    for (int it = 0; it < NIT; ++it) {
      const int c = it * 256 + lane * 8;
      cur[it] = (pix < pend && c < C) ? __ldg(reinterpret_cast<const uint4*>(h + pix * C + c)) : make_uint4(0u, 0u, 0u, 0u);
    }
  }
  for (long long pix = warp0 * chunk; pix < pend; ++pix) {
    uint4 nxt[NIT];
#pragma unroll
    for (int it = 0; it < NIT; ++it) {
      const int c = it * 256 + lane * 8;
      nxt[it] = (pix + 1 < pend && c < C) ? __ldg(reinterpret_cast<const uint4*>(h + (pix + 1) * C + c))
                                          : make_uint4(0u, 0u, 0u, 0u);
    }
    const int n = static_cast<int>(pix / HW);
    if (n != cur_n) {
      cur_n = n;
      const float* wrow = wh_fold + static_cast<long long>(n) * C;
      bh = __ldg(bh_fold + n);
      if (DUAL) bh2 = __ldg(bh_fold2 + n);
#pragma unroll
      for (int it = 0; it < NIT; ++it) {
        const int c = it * 256 + lane * 8;
        float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0, b0 = a0, b1 = a0;
        if (c < C) {
          a0 = __ldg(reinterpret_cast<const float4*>(wrow + c));
          a1 = __ldg(reinterpret_cast<const float4*>(wrow + c) + 1);
          if (DUAL) {
            const float* urow = wh_fold2 + static_cast<long long>(n) * C;
            b0 = __ldg(reinterpret_cast<const float4*>(urow + c));
            b1 = __ldg(reinterpret_cast<const float4*>(urow + c) + 1);
          }
        }
        w[it][0] = a0.x; w[it][1] = a0.y; w[it][2] = a0.z; w[it][3] = a0.w;
        w[it][4] = a1.x; w[it][5] = a1.y; w[it][6] = a1.z; w[it][7] = a1.w;
        if (DUAL) {
          u[it][0] = b0.x; u[it][1] = b0.y; u[it][2] = b0.z; u[it][3] = b0.w;
          u[it][4] = b1.x; u[it][5] = b1.y; u[it][6] = b1.z; u[it][7] = b1.w;
        }
      }
    }
    float v[NIT][8];
#pragma unroll
    for (int it = 0; it < NIT; ++it) {
      const uint32_t q4[4] = {cur[it].x, cur[it].y, cur[it].z, cur[it].w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if (F16) {
          const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&q4[i]));
          v[it][2 * i] = f.x;
          v[it][2 * i + 1] = f.y;
        } else {
          v[it][2 * i] = __uint_as_float(q4[i] << 16);
          v[it][2 * i + 1] = __uint_as_float(q4[i] & 0xffff0000u);
        }
      }
      cur[it] = nxt[it];
    }
    float acc = 0.f, acc2 = 0.f;
#pragma unroll
    for (int it = 0; it < NIT; ++it) {
      // same association as one expression per 8 channels: products summed left to right, then added to acc
      acc += v[it][0] * w[it][0] + v[it][1] * w[it][1] + v[it][2] * w[it][2] + v[it][3] * w[it][3] +
             v[it][4] * w[it][4] + v[it][5] * w[it][5] + v[it][6] * w[it][6] + v[it][7] * w[it][7];
      if (DUAL)
        acc2 += v[it][0] * u[it][0] + v[it][1] * u[it][1] + v[it][2] * u[it][2] + v[it][3] * u[it][3] +
                v[it][4] * u[it][4] + v[it][5] * u[it][5] + v[it][6] * u[it][6] + v[it][7] * u[it][7];
    }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) mask[pix] = 1.f / (1.f + __expf(-(acc + bh)));
    if (DUAL) {
      for (int o = 16; o > 0; o >>= 1) acc2 += __shfl_xor_sync(0xffffffffu, acc2, o);
      if (lane == 0) mask2[pix] = 1.f / (1.f + __expf(-(acc2 + bh2)));
    }
  }
}

// The same mask with LPP lanes per pixel (32 channels per lane, 32 / LPP pixels per warp and iteration): the 27 % of
// the issue slots that the one-warp-per-pixel kernel spends on the 5-step shuffle reduction, the sigmoid and the
// bookkeeping of every pixel are shared by 4 (c_x = 256) or 2 (c_x = 512) pixels, and every lane has four
// independent 16-byte loads in flight (ncu: the one-warp-per-pixel kernel ran at 61 % issue-slot utilisation and
// 26 % of the DRAM rate on AADBlk6).  C = 32 * LPP; HW and the per-warp runs are multiples of 32 / LPP, so the pixels
// of one iteration share their image.
template <int LPP, bool DUAL, bool F16>
__global__ void __launch_bounds__(256) aad_mask_lpp_kernel(const __nv_bfloat16* __restrict__ h, long long npix, int HW,
                                                            const float* __restrict__ wh_fold,
                                                            const float* __restrict__ bh_fold, float* __restrict__ mask,
                                                            const float* __restrict__ wh_fold2,
                                                            const float* __restrict__ bh_fold2, float* __restrict__ mask2) {
  constexpr int G = 32 / LPP;  // pixels per warp and iteration
  constexpr int C = 32 * LPP;
  const int lane = threadIdx.x & 31;
  const int g = lane / LPP, l = lane % LPP;
  const long long warp0 = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  long long chunk = (npix + nwarps - 1) / nwarps;
  chunk = (chunk + G - 1) / G * G;
  const long long p0 = warp0 * chunk, pend = min(npix, p0 + chunk);
  float w[4][8], u[DUAL ? 4 : 1][8];
  float bh = 0.f, bh2 = 0.f;
  int cur_n = -1;
  for (long long base = p0; base < pend; base += G) {
    const long long pix = base + g;  // (npix is a multiple of G: every lane group has a pixel)
    uint4 q[4];
#pragma unroll
    for (int it = 0; it < 4; ++it) q[it] = __ldg(reinterpret_cast<const uint4*>(h + pix * C + it * (LPP * 8) + l * 8));
    const int n = static_cast<int>(base / HW);
    if (n != cur_n) {
      cur_n = n;
      bh = __ldg(bh_fold + n);
      if (DUAL) bh2 = __ldg(bh_fold2 + n);
#pragma unroll
      for (int it = 0; it < 4; ++it) {
        const int c = it * (LPP * 8) + l * 8;
        const float4 a0 = __ldg(reinterpret_cast<const float4*>(wh_fold + static_cast<long long>(n) * C + c));
        const float4 a1 = __ldg(reinterpret_cast<const float4*>(wh_fold + static_cast<long long>(n) * C + c) + 1);
        w[it][0] = a0.x; w[it][1] = a0.y; w[it][2] = a0.z; w[it][3] = a0.w;
        w[it][4] = a1.x; w[it][5] = a1.y; w[it][6] = a1.z; w[it][7] = a1.w;
        if (DUAL) {
          const float4 b0 = __ldg(reinterpret_cast<const float4*>(wh_fold2 + static_cast<long long>(n) * C + c));
          const float4 b1 = __ldg(reinterpret_cast<const float4*>(wh_fold2 + static_cast<long long>(n) * C + c) + 1);
          u[it][0] = b0.x; u[it][1] = b0.y; u[it][2] = b0.z; u[it][3] = b0.w;
          u[it][4] = b1.x; u[it][5] = b1.y; u[it][6] = b1.z; u[it][7] = b1.w;
        }
      }
    }
    float acc = 0.f, acc2 = 0.f;
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const uint32_t q4[4] = {q[it].x, q[it].y, q[it].z, q[it].w};
      float v[8];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if (F16) {
          const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&q4[i]));
          v[2 * i] = f.x;
          v[2 * i + 1] = f.y;
        } else {
          v[2 * i] = __uint_as_float(q4[i] << 16);
          v[2 * i + 1] = __uint_as_float(q4[i] & 0xffff0000u);
        }
      }
      acc += v[0] * w[it][0] + v[1] * w[it][1] + v[2] * w[it][2] + v[3] * w[it][3] + v[4] * w[it][4] + v[5] * w[it][5] +
             v[6] * w[it][6] + v[7] * w[it][7];
      if (DUAL)
        acc2 += v[0] * u[it][0] + v[1] * u[it][1] + v[2] * u[it][2] + v[3] * u[it][3] + v[4] * u[it][4] + v[5] * u[it][5] +
                v[6] * u[it][6] + v[7] * u[it][7];
    }
#pragma unroll
    for (int o = LPP / 2; o > 0; o >>= 1) {
      acc += __shfl_xor_sync(0xffffffffu, acc, o);
      if (DUAL) acc2 += __shfl_xor_sync(0xffffffffu, acc2, o);
    }
    if (l == 0) {
      mask[pix] = 1.f / (1.f + __expf(-(acc + bh)));
      if (DUAL) mask2[pix] = 1.f / (1.f + __expf(-(acc2 + bh2)));
    }
  }
}

template <int LPP, bool F16>
void aad_mask_lpp_dispatch(unsigned blocks, cudaStream_t s, const __nv_bfloat16* h, long long npix, int HW,
                           const float* wh_fold, const float* bh_fold, float* mask, const float* wh_fold2,
                           const float* bh_fold2, float* mask2) {
  if (wh_fold2)
    aad_mask_lpp_kernel<LPP, true, F16><<<blocks, 256, 0, s>>>(h, npix, HW, wh_fold, bh_fold, mask, wh_fold2, bh_fold2, mask2);
  else
    aad_mask_lpp_kernel<LPP, false, F16><<<blocks, 256, 0, s>>>(h, npix, HW, wh_fold, bh_fold, mask, nullptr, nullptr, nullptr);
}

template <int NIT, bool F16>
void aad_mask_dispatch(unsigned blocks, cudaStream_t s, const __nv_bfloat16* h, long long npix, int HW, int C,
                       const float* wh_fold, const float* bh_fold, float* mask, const float* wh_fold2,
                       const float* bh_fold2, float* mask2) {
  if (wh_fold2)
    aad_mask_kernel<NIT, true, F16><<<blocks, 256, 0, s>>>(h, npix, HW, C, wh_fold, bh_fold, mask, wh_fold2, bh_fold2, mask2);
  else
    aad_mask_kernel<NIT, false, F16><<<blocks, 256, 0, s>>>(h, npix, HW, C, wh_fold, bh_fold, mask, nullptr, nullptr, nullptr);
}
template <int NIT>
void aad_mask_dispatch(unsigned blocks, cudaStream_t s, int dtype, const __nv_bfloat16* h, long long npix, int HW, int C,
                       const float* wh_fold, const float* bh_fold, float* mask, const float* wh_fold2,
                       const float* bh_fold2, float* mask2) {
  if (dtype == kF16)
    aad_mask_dispatch<NIT, true>(blocks, s, h, npix, HW, C, wh_fold, bh_fold, mask, wh_fold2, bh_fold2, mask2);
  else
    aad_mask_dispatch<NIT, false>(blocks, s, h, npix, HW, C, wh_fold, bh_fold, mask, wh_fold2, bh_fold2, mask2);
}

}  // namespace

const char* aad_mask_launch(const void* h, int N, int HW, int C, const float* wh_fold, const float* bh_fold, float* mask,
                            cudaStream_t s, int dtype) {
  return aad_mask2_launch(h, N, HW, C, wh_fold, bh_fold, mask, nullptr, nullptr, nullptr, s, dtype);
}

const char* aad_mask2_launch(const void* h, int N, int HW, int C, const float* wh_fold, const float* bh_fold, float* mask,
                             const float* wh_fold2, const float* bh_fold2, float* mask2, cudaStream_t s, int dtype) {
  if (C % 8) return "aad_mask: C must be a multiple of 8";
  if (dtype != kBF16 && dtype != kF16) return "aad_mask: h must be a 16-bit tensor";
  const long long npix = static_cast<long long>(N) * HW;
  long long blocks = (npix + 7) / 8;
  if (blocks > 148 * 16) blocks = 148 * 16;
  const __nv_bfloat16* hp = reinterpret_cast<const __nv_bfloat16*>(h);
  const unsigned nb = static_cast<unsigned>(blocks);
  static const int lpp_mode = [] {
    const char* e = getenv("GHOST_MASK_LPP");  // A/B switch for profiling
    return e ? atoi(e) : 1;
  }();
  if (lpp_mode && (C == 256 || C == 512) && HW % (C == 256 ? 4 : 2) == 0) {
    // several pixels per warp: a block of 8 warps covers 8 * G pixels per iteration
    const int G = C == 256 ? 4 : 2;
    long long bl = (npix + 8 * G - 1) / (8 * G);
    if (bl > 148 * 16) bl = 148 * 16;
    const unsigned nbl = static_cast<unsigned>(bl);
    if (C == 256) {
      if (dtype == kF16) aad_mask_lpp_dispatch<8, true>(nbl, s, hp, npix, HW, wh_fold, bh_fold, mask, wh_fold2, bh_fold2, mask2);
      else aad_mask_lpp_dispatch<8, false>(nbl, s, hp, npix, HW, wh_fold, bh_fold, mask, wh_fold2, bh_fold2, mask2);
    } else {
      if (dtype == kF16) aad_mask_lpp_dispatch<16, true>(nbl, s, hp, npix, HW, wh_fold, bh_fold, mask, wh_fold2, bh_fold2, mask2);
      else aad_mask_lpp_dispatch<16, false>(nbl, s, hp, npix, HW, wh_fold, bh_fold, mask, wh_fold2, bh_fold2, mask2);
    }
    return last_err();
  }
  if (C <= 256)
    aad_mask_dispatch<1>(nb, s, dtype, hp, npix, HW, C, wh_fold, bh_fold, mask, wh_fold2, bh_fold2, mask2);
  else if (C <= 512)
    aad_mask_dispatch<2>(nb, s, dtype, hp, npix, HW, C, wh_fold, bh_fold, mask, wh_fold2, bh_fold2, mask2);
  else if (C <= 1024)
    aad_mask_dispatch<4>(nb, s, dtype, hp, npix, HW, C, wh_fold, bh_fold, mask, wh_fold2, bh_fold2, mask2);
  else
    return "aad_mask: C must be <= 1024";
  return last_err();
}

namespace {
// pixels per block of the statistics kernel for a tensor of N images x HW pixels x C channels
int stats_pix_per_block(int N, int HW, int C) {
  const int ppi = kStatThreads / (C / 8);
  // aim for ~16 blocks per SM (four waves at four resident blocks): a grid of barely more than one wave loses up
  // to half its time in the tail; but at least 8 iterations per block
  const int blocks_x = (148 * 16 + N - 1) / N;
  int ppb = (HW + blocks_x - 1) / blocks_x;
  if (ppb < ppi * 8) ppb = ppi * 8;
  return ppb;
}
}  // namespace

int in_stats_parts(int N, int HW, int C) {
  const int ppb = stats_pix_per_block(N, HW, C);
  return (HW + ppb - 1) / ppb;
}

const char* in_stats_launch(const void* x, int dtype, int N, int HW, int C, float* parts, cudaStream_t s) {
  return in_stats_launch_ex(x, dtype, N, HW, C, parts, 0, s);
}

const char* in_stats_launch_ex(const void* x, int dtype, int N, int HW, int C, float* parts, int common_shift,
                               cudaStream_t s) {
  if (C % 8 || C > 8 * kStatThreads) return "in_stats: C must be a multiple of 8 and <= 2048";
  const int ppb = stats_pix_per_block(N, HW, C);
  dim3 grid((HW + ppb - 1) / ppb, N);
  const int smem = kStatThreads * 16 * sizeof(float);
  if (dtype == kF32)
    in_stats_kernel<float>
        <<<grid, kStatThreads, smem, s>>>(reinterpret_cast<const float*>(x), HW, C, ppb, parts, common_shift);
  else if (dtype == kBF16)
    in_stats_kernel<__nv_bfloat16>
        <<<grid, kStatThreads, smem, s>>>(reinterpret_cast<const __nv_bfloat16*>(x), HW, C, ppb, parts, common_shift);
  else if (dtype == kF16)
    in_stats_kernel<__half>
        <<<grid, kStatThreads, smem, s>>>(reinterpret_cast<const __half*>(x), HW, C, ppb, parts, common_shift);
  else
    return "in_stats: unsupported dtype";
  return last_err();
}

size_t aad_fold_scratch_floats(int N, int C) { return static_cast<size_t>(N) * ((C + 63) / 64 + 1); }

const char* aad_fold_launch(const float* parts, int P, int N, int HW, int C, const float* zfc, int ldz, int off_gamma,
                            int off_beta, const float* w_h, float b_h, float4* ncpar, float* wh_fold, float* bh_fold,
                            float* scratch, cudaStream_t s, int ld_out) {
  AadFoldSet a;
  a.off_gamma = off_gamma; a.off_beta = off_beta; a.w_h = w_h; a.b_h = b_h; a.ncpar = ncpar; a.wh_fold = wh_fold;
  a.bh_fold = bh_fold; a.scratch = scratch; a.ld_out = ld_out;
  return aad_fold2_launch(parts, P, N, HW, C, zfc, ldz, a, nullptr, s);
}

const char* aad_fold2_launch(const float* parts, int P, int N, int HW, int C, const float* zfc, int ldz,
                             const AadFoldSet& a, const AadFoldSet* b, cudaStream_t s) {
  AadFoldSet fa = a, fb = b ? *b : a;
  if (fa.ld_out <= 0) fa.ld_out = C;
  if (fb.ld_out <= 0) fb.ld_out = C;
  dim3 grid(N, (C + 63) / 64, b ? 2 : 1);
  aad_fold_kernel<<<grid, 256, 0, s>>>(parts, P, HW, C, zfc, ldz, fa, fb);
  return last_err();
}

const char* aad_apply_launch(const void* h, const void* gb, void* out, int dtype, int N, int HW, int C,
                             const float4* ncpar, const float* wh_fold, const float* bh_fold, cudaStream_t s) {
  const long long npix = static_cast<long long>(N) * HW;
  long long blocks = (npix + 7) / 8;
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (dtype == kF32)
    aad_apply_kernel<float><<<static_cast<unsigned>(blocks), 256, 0, s>>>(
        reinterpret_cast<const float*>(h), reinterpret_cast<const float*>(gb), reinterpret_cast<float*>(out), npix, HW,
        C, ncpar, wh_fold, bh_fold);
  else if (dtype == kBF16)
    aad_apply_kernel<__nv_bfloat16><<<static_cast<unsigned>(blocks), 256, 0, s>>>(
        reinterpret_cast<const __nv_bfloat16*>(h), reinterpret_cast<const __nv_bfloat16*>(gb),
        reinterpret_cast<__nv_bfloat16*>(out), npix, HW, C, ncpar, wh_fold, bh_fold);
  else if (dtype == kF16)
    aad_apply_kernel<__half><<<static_cast<unsigned>(blocks), 256, 0, s>>>(
        reinterpret_cast<const __half*>(h), reinterpret_cast<const __half*>(gb),
        reinterpret_cast<__half*>(out), npix, HW, C, ncpar, wh_fold, bh_fold);
  else
    return "aad_apply: unsupported dtype";
  return last_err();
}

// ---------------------------------------------------------------------------
// Bilinear x2 (align_corners=True), layout conversion, small glue kernels.
// ---------------------------------------------------------------------------
namespace {

// ATen upsample_bilinear2d blend of the four neighbours of eight channels (packed fp32):
//   o = (1-lh) * ((1-lw)*a + lw*b) + lh * ((1-lw)*c + lw*d)
__device__ __forceinline__ void bilerp8(const float (&a)[8], const float (&b)[8], const float (&c)[8],
                                        const float (&d)[8], float lw, float lh, float (&o)[8]) {
  const float2 w0 = make_float2(1.f - lw, 1.f - lw), w1 = make_float2(lw, lw);
  const float2 h0 = make_float2(1.f - lh, 1.f - lh), h1 = make_float2(lh, lh);
#pragma unroll
  for (int k = 0; k < 8; k += 2) {
    const float2 t0 = ptx::fma2(w1, make_float2(b[k], b[k + 1]), ptx::mul2(w0, make_float2(a[k], a[k + 1])));
    const float2 t1 = ptx::fma2(w1, make_float2(d[k], d[k + 1]), ptx::mul2(w0, make_float2(c[k], c[k + 1])));
    const float2 r = ptx::fma2(h1, t1, ptx::mul2(h0, t0));
    o[k] = r.x;
    o[k + 1] = r.y;
  }
}

// store eight channels and hand back the values as stored (after rounding to T)
__device__ __forceinline__ void store8_rounded(float* p, const float (&v)[8], float (&r)[8]) {
  store8(p, v);
#pragma unroll
  for (int i = 0; i < 8; ++i) r[i] = v[i];
}
__device__ __forceinline__ void store8_rounded(__nv_bfloat16* p, const float (&v)[8], float (&r)[8]) {
  uint4 q;
  __nv_bfloat162 t;
  t = __floats2bfloat162_rn(v[0], v[1]); q.x = *reinterpret_cast<uint32_t*>(&t);
  t = __floats2bfloat162_rn(v[2], v[3]); q.y = *reinterpret_cast<uint32_t*>(&t);
  t = __floats2bfloat162_rn(v[4], v[5]); q.z = *reinterpret_cast<uint32_t*>(&t);
  t = __floats2bfloat162_rn(v[6], v[7]); q.w = *reinterpret_cast<uint32_t*>(&t);
  *reinterpret_cast<uint4*>(p) = q;
  const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    r[2 * i] = __uint_as_float(w[i] << 16);
    r[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}

__device__ __forceinline__ void store8_rounded(__half* p, const float (&v)[8], float (&r)[8]) {
  uint4 q;
  __half2 t;
  t = __floats2half2_rn(v[0], v[1]); q.x = *reinterpret_cast<uint32_t*>(&t);
  t = __floats2half2_rn(v[2], v[3]); q.y = *reinterpret_cast<uint32_t*>(&t);
  t = __floats2half2_rn(v[4], v[5]); q.z = *reinterpret_cast<uint32_t*>(&t);
  t = __floats2half2_rn(v[6], v[7]); q.w = *reinterpret_cast<uint32_t*>(&t);
  *reinterpret_cast<uint4*>(p) = q;
  const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
    r[2 * i] = f.x;
    r[2 * i + 1] = f.y;
  }
}

template <typename T>
__global__ void __launch_bounds__(256) upsample2x_kernel(const T* __restrict__ x, T* __restrict__ y, int N, int H,
                                                          int W, int C, int oC, int o_c0) {
  // grid: x = chunks of one output row (pixels x channel vectors), y = output row, z = image
  const int Ho = 2 * H, Wo = 2 * W, vecs = C >> 3;
  const int oh = blockIdx.y, n = blockIdx.z;
  const float sh = Ho > 1 ? static_cast<float>(H - 1) / static_cast<float>(Ho - 1) : 0.f;
  const float sw = Wo > 1 ? static_cast<float>(W - 1) / static_cast<float>(Wo - 1) : 0.f;
  // ATen upsample_bilinear2d (align_corners): src = dst * (in-1)/(out-1), lambda in fp32
  const float fh = sh * oh;
  const int h0 = static_cast<int>(fh);
  const int h1 = h0 + (h0 < H - 1 ? 1 : 0);
  const float lh = fh - h0;
  const T* base = x + static_cast<long long>(n) * H * W * C;
  T* orow = y + (static_cast<long long>(n) * Ho + oh) * Wo * oC + o_c0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < Wo * vecs; i += gridDim.x * blockDim.x) {
    const int ow = i / vecs, cv = i - ow * vecs;
    const float fw = sw * ow;
    const int w0 = static_cast<int>(fw);
    const int w1 = w0 + (w0 < W - 1 ? 1 : 0);
    const float lw = fw - w0;
    float a[8], b[8], c[8], d[8], o[8];
    load8(base + (h0 * W + w0) * C + cv * 8, a);
    load8(base + (h0 * W + w1) * C + cv * 8, b);
    load8(base + (h1 * W + w0) * C + cv * 8, c);
    load8(base + (h1 * W + w1) * C + cv * 8, d);
    bilerp8(a, b, c, d, lw, lh, o);
    store8(orow + static_cast<long long>(ow) * oC + cv * 8, o);
  }
}

// Bilinear x2 fused with the InstanceNorm statistics of its own output (the next AAD block normalises exactly
// this tensor, network/AADLayer.py:21): same block / part decomposition and partial-sum layout as
// in_stats_kernel, the sums run over the values as stored (after rounding to T).  The shift row is source
// pixel (0,0), which align_corners copies unchanged to output pixel 0.
template <typename T>
__global__ void __launch_bounds__(kStatThreads, 4) upsample2x_stats_kernel(const T* __restrict__ x, T* __restrict__ y, int H,
                                                                         int W, int C, int pix_per_block,
                                                                         float* __restrict__ parts) {
  extern __shared__ float sh[];  // [kStatThreads][16]
  const int n = blockIdx.y;
  const int P = gridDim.x;
  const int Ho = 2 * H, Wo = 2 * W, HWo = Ho * Wo;
  const int vecs = C >> 3;
  const int ppi = kStatThreads / vecs;
  const int cv = threadIdx.x % vecs, pl = threadIdx.x / vecs;
  const int p0 = blockIdx.x * pix_per_block;
  const int p1 = min(HWo, p0 + pix_per_block);
  const float shh = static_cast<float>(H - 1) / static_cast<float>(Ho - 1);
  const float sww = static_cast<float>(W - 1) / static_cast<float>(Wo - 1);
  float s[8], q[8], sft[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) s[i] = q[i] = 0.f;
  const T* base = x + static_cast<long long>(n) * H * W * C + cv * 8;
  T* obase = y + static_cast<long long>(n) * HWo * C + cv * 8;
  load8(base, sft);
  if (pl < ppi) {
    for (int p = p0 + pl; p < p1; p += ppi) {
      const int oh = p / Wo, ow = p - oh * Wo;
      // ATen upsample_bilinear2d (align_corners): src = dst * (in-1)/(out-1), lambda in fp32
      const float fh = shh * oh, fw = sww * ow;
      const int h0 = static_cast<int>(fh), w0 = static_cast<int>(fw);
      const int h1 = h0 + (h0 < H - 1 ? 1 : 0), w1 = w0 + (w0 < W - 1 ? 1 : 0);
      const float lh = fh - h0, lw = fw - w0;
      float a[8], b[8], c[8], d[8], o[8];
      load8(base + (h0 * W + w0) * C, a);
      load8(base + (h0 * W + w1) * C, b);
      load8(base + (h1 * W + w0) * C, c);
      load8(base + (h1 * W + w1) * C, d);
      bilerp8(a, b, c, d, lw, lh, o);
      float stored[8];
      store8_rounded(obase + static_cast<long long>(p) * C, o, stored);
      stats_accumulate8(stored, sft, s, q);
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    sh[threadIdx.x * 16 + i] = s[i];
    sh[threadIdx.x * 16 + 8 + i] = q[i];
  }
  __syncthreads();
  float* out = parts + static_cast<long long>(n) * (2 * P + 1) * C;
  for (int t = threadIdx.x; t < vecs * 16; t += kStatThreads) {
    const int v = t >> 4, slot = t & 15;
    float acc = 0.f;
    for (int l = 0; l < ppi; ++l) acc += sh[(l * vecs + v) * 16 + slot];
    const int c = v * 8 + (slot & 7);
    out[(2 * blockIdx.x + (slot >> 3)) * C + c] = acc;
  }
  if (blockIdx.x == 0 && pl == 0) {
#pragma unroll
    for (int i = 0; i < 8; ++i) out[2 * P * C + cv * 8 + i] = sft[i];
  }
}

template <typename S, typename D>
__global__ void __launch_bounds__(256) nchw_to_nhwc_kernel(const S* __restrict__ src, D* __restrict__ dst, int C,
                                                            int HW, int dstC, int dst_c0) {
  // tile transpose: 32 pixels x 32 channels through shared memory, both sides coalesced
  __shared__ float tile[32][33];
  const int n = blockIdx.z, p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  for (int j = ty; j < 32; j += 8) {
    const int c = c0 + j, p = p0 + tx;
    tile[j][tx] = (c < C && p < HW) ? ldf(src + (static_cast<long long>(n) * C + c) * HW + p) : 0.f;
  }
  __syncthreads();
  for (int j = ty; j < 32; j += 8) {
    const int p = p0 + j, c = c0 + tx;
    if (c < C && p < HW) stf(dst + (static_cast<long long>(n) * HW + p) * dstC + dst_c0 + c, tile[tx][j]);
  }
}

template <typename S, typename D>
__global__ void __launch_bounds__(256) nhwc_to_nchw_kernel(const S* __restrict__ src, int srcC, int src_c0,
                                                            D* __restrict__ dst, int C, int HW) {
  __shared__ float tile[32][33];
  const int n = blockIdx.z, p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int j = ty; j < 32; j += 8) {
    const int p = p0 + j, c = c0 + tx;
    tile[j][tx] = (c < C && p < HW) ? ldf(src + (static_cast<long long>(n) * HW + p) * srcC + src_c0 + c) : 0.f;
  }
  __syncthreads();
  for (int j = ty; j < 32; j += 8) {
    const int c = c0 + j, p = p0 + tx;
    if (c < C && p < HW) stf(dst + (static_cast<long long>(n) * C + c) * HW + p, tile[tx][j]);
  }
}

// ---------------------------------------------------------------------------
// Bilinear x2, one thread per 2 x 2 OUTPUT block and 8 channels (round 2; ncu of the per-pixel kernels above: issue
// bound at 70-85 % issue slots, 2.6 TB/s).  The four outputs of a block draw on a 3 x 3 source neighbourhood: 9 loads
// and conversions instead of 16, the horizontal blends of the middle source row are shared by both output rows
// (20 instead of 24 packed blends per output), the index arithmetic is paid once per block.  Every output is still
//   o = (1-lh) * ((1-lw)*a + lw*b) + lh * ((1-lw)*c + lw*d)          (ATen upsample_bilinear2d, align_corners)
// on the same four neighbours with the same fp32 lambdas, so the result is bit-identical to the per-pixel kernels;
// blocks whose rows / columns do not form the regular pattern (image borders) take the per-pixel formula.
// STATS: fused InstanceNorm partial statistics of the output (layout of in_stats_kernel), parts per block.
__device__ __forceinline__ void quad_axis(int o, float scale, int n, int* i0, int* i1, float* l) {
  const float f = scale * o;
  *i0 = static_cast<int>(f);
  *i1 = *i0 + (*i0 < n - 1 ? 1 : 0);
  *l = f - *i0;
}
// horizontal blends of one source row for the two output columns: tA from (v0, v1), tB from (v1, v2)
__device__ __forceinline__ void quad_hblend(const float (&v0)[8], const float (&v1)[8], const float (&v2)[8], float lwA,
                                            float lwB, float (&tA)[8], float (&tB)[8]) {
  const float2 a0 = make_float2(1.f - lwA, 1.f - lwA), a1 = make_float2(lwA, lwA);
  const float2 b0 = make_float2(1.f - lwB, 1.f - lwB), b1 = make_float2(lwB, lwB);
#pragma unroll
  for (int k = 0; k < 8; k += 2) {
    const float2 x0 = make_float2(v0[k], v0[k + 1]), x1 = make_float2(v1[k], v1[k + 1]), x2 = make_float2(v2[k], v2[k + 1]);
    const float2 ta = ptx::fma2(a1, x1, ptx::mul2(a0, x0));
    const float2 tb = ptx::fma2(b1, x2, ptx::mul2(b0, x1));
    tA[k] = ta.x; tA[k + 1] = ta.y;
    tB[k] = tb.x; tB[k + 1] = tb.y;
  }
}
__device__ __forceinline__ void quad_vblend(const float (&t0)[8], const float (&t1)[8], float lh, float (&o)[8]) {
  const float2 h0 = make_float2(1.f - lh, 1.f - lh), h1 = make_float2(lh, lh);
#pragma unroll
  for (int k = 0; k < 8; k += 2) {
    const float2 r = ptx::fma2(h1, make_float2(t1[k], t1[k + 1]), ptx::mul2(h0, make_float2(t0[k], t0[k + 1])));
    o[k] = r.x; o[k + 1] = r.y;
  }
}

template <typename T, bool STATS>
__global__ void __launch_bounds__(kStatThreads, 2) upsample2x_quad_kernel(const T* __restrict__ x, T* __restrict__ y, int H,
                                                                        int W, int C, int oC, int o_c0,
                                                                        int quads_per_block, float* __restrict__ parts) {
  extern __shared__ float sh[];  // STATS: [kStatThreads][16]
  const int n = blockIdx.y;
  const int P = gridDim.x;
  const int Ho = 2 * H, Wo = 2 * W, HW = H * W;
  const int vecs = C >> 3;
  const int ppi = kStatThreads / vecs;
  const int cv = threadIdx.x % vecs, pl = threadIdx.x / vecs;
  const int q0 = blockIdx.x * quads_per_block;
  const int q1 = min(HW, q0 + quads_per_block);
  const float shh = Ho > 1 ? static_cast<float>(H - 1) / static_cast<float>(Ho - 1) : 0.f;
  const float sww = Wo > 1 ? static_cast<float>(W - 1) / static_cast<float>(Wo - 1) : 0.f;
  float s[8], q[8], sft[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) s[i] = q[i] = sft[i] = 0.f;
  const T* base = x + static_cast<long long>(n) * HW * C + cv * 8;
  T* obase = y + static_cast<long long>(n) * Ho * Wo * oC + o_c0 + cv * 8;
  if (STATS) load8(base, sft);
  if (pl < ppi) {
    for (int qi = q0 + pl; qi < q1; qi += ppi) {
      const int qy = qi / W, qx = qi - qy * W;
      int ha0, ha1, hb0, hb1, wa0, wa1, wb0, wb1;
      float lha, lhb, lwa, lwb;
      quad_axis(2 * qy, shh, H, &ha0, &ha1, &lha);
      quad_axis(2 * qy + 1, shh, H, &hb0, &hb1, &lhb);
      quad_axis(2 * qx, sww, W, &wa0, &wa1, &lwa);
      quad_axis(2 * qx + 1, sww, W, &wb0, &wb1, &lwb);
      float o[4][8];  // (row a, col A), (a, B), (b, A), (b, B)
      if (ha1 == hb0 && wa1 == wb0) {
        // regular pattern: rows (ha0, hb0, hb1), columns (wa0, wb0, wb1)
        const int rr[3] = {ha0, hb0, hb1}, cc[3] = {wa0, wb0, wb1};
        float tA[3][8], tB[3][8];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
          Raw8<T> raw[3];
#pragma unroll
          for (int c = 0; c < 3; ++c) raw[c].load(base + static_cast<long long>(rr[r] * W + cc[c]) * C);
          float v0[8], v1[8], v2[8];
          raw[0].get(v0);
          raw[1].get(v1);
          raw[2].get(v2);
          quad_hblend(v0, v1, v2, lwa, lwb, tA[r], tB[r]);
        }
        quad_vblend(tA[0], tA[1], lha, o[0]);
        quad_vblend(tB[0], tB[1], lha, o[1]);
        quad_vblend(tA[1], tA[2], lhb, o[2]);
        quad_vblend(tB[1], tB[2], lhb, o[3]);
      } else {
        // borders: the four outputs one by one
        const int hh0[2] = {ha0, hb0}, hh1[2] = {ha1, hb1}, ww0[2] = {wa0, wb0}, ww1[2] = {wa1, wb1};
        const float lh[2] = {lha, lhb}, lw[2] = {lwa, lwb};
#pragma unroll
        for (int r = 0; r < 2; ++r)
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            float a[8], b[8], cc_[8], d[8];
            load8(base + static_cast<long long>(hh0[r] * W + ww0[c]) * C, a);
            load8(base + static_cast<long long>(hh0[r] * W + ww1[c]) * C, b);
            load8(base + static_cast<long long>(hh1[r] * W + ww0[c]) * C, cc_);
            load8(base + static_cast<long long>(hh1[r] * W + ww1[c]) * C, d);
            bilerp8(a, b, cc_, d, lw[c], lh[r], o[r * 2 + c]);
          }
      }
#pragma unroll
      for (int r = 0; r < 2; ++r)
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          T* dst = obase + (static_cast<long long>(2 * qy + r) * Wo + 2 * qx + c) * oC;
          if (STATS) {
            float stored[8];
            store8_rounded(dst, o[r * 2 + c], stored);
            stats_accumulate8(stored, sft, s, q);
          } else {
            store8(dst, o[r * 2 + c]);
          }
        }
    }
  }
  if (STATS) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      sh[threadIdx.x * 16 + i] = s[i];
      sh[threadIdx.x * 16 + 8 + i] = q[i];
    }
    __syncthreads();
    float* out = parts + static_cast<long long>(n) * (2 * P + 1) * C;
    for (int t = threadIdx.x; t < vecs * 16; t += kStatThreads) {
      const int v = t >> 4, slot = t & 15;
      float acc = 0.f;
      for (int l = 0; l < ppi; ++l) acc += sh[(l * vecs + v) * 16 + slot];
      const int c = v * 8 + (slot & 7);
      out[(2 * blockIdx.x + (slot >> 3)) * C + c] = acc;
    }
    if (blockIdx.x == 0 && pl == 0) {
#pragma unroll
      for (int i = 0; i < 8; ++i) out[2 * P * C + cv * 8 + i] = sft[i];
    }
  }
}

template <typename T>
__global__ void zfc_to_act_kernel(const float* __restrict__ zfc, int ldz, int off, int count,
                                  const float* __restrict__ bias, T* __restrict__ dst, int N) {
  const long long total = static_cast<long long>(N) * count;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int n = static_cast<int>(i / count), j = static_cast<int>(i % count);
    stf(dst + i, zfc[static_cast<long long>(n) * ldz + off + j] + (bias ? bias[j] : 0.f));
  }
}

template <typename T>
__global__ void __launch_bounds__(256) channel_affine_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                              long long npix, int C, const float* __restrict__ scale,
                                                              const float* __restrict__ shift,
                                                              const float* __restrict__ prelu) {
  const int vecs = C >> 3;
  const long long total = npix * vecs;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % vecs) * 8;
    float v[8];
    load8(x + i * 8, v);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float t = fmaf(v[k], __ldg(scale + c + k), __ldg(shift + c + k));
      if (prelu) t = t > 0.f ? t : t * __ldg(prelu + c + k);
      v[k] = t;
    }
    store8(y + i * 8, v);
  }
}

inline unsigned grid_for(long long total, int threads) {
  long long b = (total + threads - 1) / threads;
  const long long cap = 148LL * 32;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return static_cast<unsigned>(b);
}

template <typename S>
const char* nchw_to_nhwc_dst(const S* src, void* dst, int dst_dtype, int N, int C, int H, int W, int dstC, int dst_c0,
                             cudaStream_t s) {
  dim3 grid((H * W + 31) / 32, (C + 31) / 32, N);
  if (dst_dtype == kF32)
    nchw_to_nhwc_kernel<S, float><<<grid, 256, 0, s>>>(src, reinterpret_cast<float*>(dst), C, H * W, dstC, dst_c0);
  else if (dst_dtype == kBF16)
    nchw_to_nhwc_kernel<S, __nv_bfloat16>
        <<<grid, 256, 0, s>>>(src, reinterpret_cast<__nv_bfloat16*>(dst), C, H * W, dstC, dst_c0);
  else if (dst_dtype == kF16)
    nchw_to_nhwc_kernel<S, __half>
        <<<grid, 256, 0, s>>>(src, reinterpret_cast<__half*>(dst), C, H * W, dstC, dst_c0);
  else
    return "nchw_to_nhwc: unsupported destination dtype";
  return last_err();
}

template <typename S>
const char* nhwc_to_nchw_dst(const S* src, int srcC, int src_c0, void* dst, int dst_dtype, int N, int C, int H, int W,
                             cudaStream_t s) {
  dim3 grid((H * W + 31) / 32, (C + 31) / 32, N);
  if (dst_dtype == kF32)
    nhwc_to_nchw_kernel<S, float><<<grid, 256, 0, s>>>(src, srcC, src_c0, reinterpret_cast<float*>(dst), C, H * W);
  else if (dst_dtype == kBF16)
    nhwc_to_nchw_kernel<S, __nv_bfloat16>
        <<<grid, 256, 0, s>>>(src, srcC, src_c0, reinterpret_cast<__nv_bfloat16*>(dst), C, H * W);
  else if (dst_dtype == kF16)
    nhwc_to_nchw_kernel<S, __half><<<grid, 256, 0, s>>>(src, srcC, src_c0, reinterpret_cast<__half*>(dst), C, H * W);
  else
    return "nhwc_to_nchw: unsupported destination dtype";
  return last_err();
}

}  // namespace

namespace {
// quads (2 x 2 output blocks = source pixels) per block: ~16 blocks per SM over the batch, at least 8 iterations
int quads_per_block(int N, int HW, int C) {
  const int ppi = kStatThreads / (C / 8);
  const int blocks_x = (148 * 16 + N - 1) / N;
  int qpb = (HW + blocks_x - 1) / blocks_x;
  if (qpb < ppi * 8) qpb = ppi * 8;
  return qpb;
}
static const int g_quad_mode = [] {
  const char* e = getenv("GHOST_UPSAMPLE_QUAD");  // A/B switch for profiling: 0 = the per-pixel kernels
  return e ? atoi(e) : 1;
}();
template <typename T, bool STATS>
const char* upsample_quad_launch(const void* x, void* y, int N, int H, int W, int C, int oC, int o_c0, float* parts,
                                 cudaStream_t s) {
  const int qpb = quads_per_block(N, H * W, C);
  dim3 grid((H * W + qpb - 1) / qpb, N);
  const int smem = STATS ? kStatThreads * 16 * sizeof(float) : 0;
  upsample2x_quad_kernel<T, STATS><<<grid, kStatThreads, smem, s>>>(reinterpret_cast<const T*>(x), reinterpret_cast<T*>(y),
                                                                    H, W, C, oC, o_c0, qpb, parts);
  return last_err();
}
}  // namespace

// number of statistic parts upsample2x_stats_launch writes per image
int upsample2x_stats_parts(int N, int H, int W, int C) {
  if (!g_quad_mode) return in_stats_parts(N, 4 * H * W, C);
  const int qpb = quads_per_block(N, H * W, C);
  return (H * W + qpb - 1) / qpb;
}

const char* upsample2x_launch(const void* x, void* y, int dtype, int N, int H, int W, int C, int oC, int o_c0,
                              cudaStream_t s) {
  if (C % 8 || oC % 8 || o_c0 % 8) return "upsample2x: channels must be multiples of 8";
  if (static_cast<long long>(H) * W * C >= (1LL << 31)) return "upsample2x: image too large";
  if (g_quad_mode && C <= 8 * kStatThreads) {
    if (dtype == kF32) return upsample_quad_launch<float, false>(x, y, N, H, W, C, oC, o_c0, nullptr, s);
    if (dtype == kBF16) return upsample_quad_launch<__nv_bfloat16, false>(x, y, N, H, W, C, oC, o_c0, nullptr, s);
    if (dtype == kF16) return upsample_quad_launch<__half, false>(x, y, N, H, W, C, oC, o_c0, nullptr, s);
    return "upsample2x: unsupported dtype";
  }
  const int row_items = 2 * W * (C / 8);
  dim3 grid((row_items + 255) / 256, 2 * H, N);
  if (grid.x > 64) grid.x = 64;
  if (dtype == kF32)
    upsample2x_kernel<float><<<grid, 256, 0, s>>>(reinterpret_cast<const float*>(x), reinterpret_cast<float*>(y), N, H,
                                                  W, C, oC, o_c0);
  else if (dtype == kBF16)
    upsample2x_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(reinterpret_cast<const __nv_bfloat16*>(x),
                                                          reinterpret_cast<__nv_bfloat16*>(y), N, H, W, C, oC, o_c0);
  else if (dtype == kF16)
    upsample2x_kernel<__half><<<grid, 256, 0, s>>>(reinterpret_cast<const __half*>(x),
                                                          reinterpret_cast<__half*>(y), N, H, W, C, oC, o_c0);
  else
    return "upsample2x: unsupported dtype";
  return last_err();
}

const char* upsample2x_stats_launch(const void* x, void* y, int dtype, int N, int H, int W, int C, float* parts,
                                    cudaStream_t s) {
  if (C % 8 || C > 8 * kStatThreads) return "upsample2x_stats: C must be a multiple of 8 and <= 2048";
  if (static_cast<long long>(H) * W * C >= (1LL << 31)) return "upsample2x_stats: image too large";
  if (g_quad_mode) {
    if (dtype == kF32) return upsample_quad_launch<float, true>(x, y, N, H, W, C, C, 0, parts, s);
    if (dtype == kBF16) return upsample_quad_launch<__nv_bfloat16, true>(x, y, N, H, W, C, C, 0, parts, s);
    if (dtype == kF16) return upsample_quad_launch<__half, true>(x, y, N, H, W, C, C, 0, parts, s);
    return "upsample2x_stats: unsupported dtype";
  }
  const int HWo = 4 * H * W;
  const int ppb = stats_pix_per_block(N, HWo, C);
  dim3 grid((HWo + ppb - 1) / ppb, N);
  const int smem = kStatThreads * 16 * sizeof(float);
  if (dtype == kF32)
    upsample2x_stats_kernel<float><<<grid, kStatThreads, smem, s>>>(reinterpret_cast<const float*>(x),
                                                                    reinterpret_cast<float*>(y), H, W, C, ppb, parts);
  else if (dtype == kBF16)
    upsample2x_stats_kernel<__nv_bfloat16><<<grid, kStatThreads, smem, s>>>(
        reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(y), H, W, C, ppb, parts);
  else if (dtype == kF16)
    upsample2x_stats_kernel<__half><<<grid, kStatThreads, smem, s>>>(
        reinterpret_cast<const __half*>(x), reinterpret_cast<__half*>(y), H, W, C, ppb, parts);
  else
    return "upsample2x_stats: unsupported dtype";
  return last_err();
}

const char* nchw_to_nhwc_launch(const void* src, int src_dtype, void* dst, int dst_dtype, int N, int C, int H, int W,
                                int dstC, int dst_c0, cudaStream_t s) {
  if (src_dtype == kF32)
    return nchw_to_nhwc_dst(reinterpret_cast<const float*>(src), dst, dst_dtype, N, C, H, W, dstC, dst_c0, s);
  if (src_dtype == kBF16)
    return nchw_to_nhwc_dst(reinterpret_cast<const __nv_bfloat16*>(src), dst, dst_dtype, N, C, H, W, dstC, dst_c0, s);
  if (src_dtype == kF16)
    return nchw_to_nhwc_dst(reinterpret_cast<const __half*>(src), dst, dst_dtype, N, C, H, W, dstC, dst_c0, s);
  return "nchw_to_nhwc: unsupported source dtype";
}

const char* nhwc_to_nchw_launch(const void* src, int src_dtype, int srcC, int src_c0, void* dst, int dst_dtype, int N,
                                int C, int H, int W, cudaStream_t s) {
  if (src_dtype == kF32)
    return nhwc_to_nchw_dst(reinterpret_cast<const float*>(src), srcC, src_c0, dst, dst_dtype, N, C, H, W, s);
  if (src_dtype == kBF16)
    return nhwc_to_nchw_dst(reinterpret_cast<const __nv_bfloat16*>(src), srcC, src_c0, dst, dst_dtype, N, C, H, W, s);
  if (src_dtype == kF16)
    return nhwc_to_nchw_dst(reinterpret_cast<const __half*>(src), srcC, src_c0, dst, dst_dtype, N, C, H, W, s);
  return "nhwc_to_nchw: unsupported source dtype";
}

const char* zfc_to_act_launch(const float* zfc, int ldz, int off, int count, const float* bias, void* dst, int dtype,
                              int N, cudaStream_t s) {
  const long long total = static_cast<long long>(N) * count;
  if (dtype == kF32)
    zfc_to_act_kernel<float>
        <<<grid_for(total, 256), 256, 0, s>>>(zfc, ldz, off, count, bias, reinterpret_cast<float*>(dst), N);
  else if (dtype == kBF16)
    zfc_to_act_kernel<__nv_bfloat16>
        <<<grid_for(total, 256), 256, 0, s>>>(zfc, ldz, off, count, bias, reinterpret_cast<__nv_bfloat16*>(dst), N);
  else if (dtype == kF16)
    zfc_to_act_kernel<__half>
        <<<grid_for(total, 256), 256, 0, s>>>(zfc, ldz, off, count, bias, reinterpret_cast<__half*>(dst), N);
  else
    return "zfc_to_act: unsupported dtype";
  return last_err();
}

const char* channel_affine_launch(const void* x, void* y, int dtype, long long npix, int C, const float* scale,
                                  const float* shift, const float* prelu, cudaStream_t s) {
  if (C % 8) return "channel_affine: C must be a multiple of 8";
  const long long total = npix * (C / 8);
  if (dtype == kF32)
    channel_affine_kernel<float><<<grid_for(total, 256), 256, 0, s>>>(
        reinterpret_cast<const float*>(x), reinterpret_cast<float*>(y), npix, C, scale, shift, prelu);
  else if (dtype == kBF16)
    channel_affine_kernel<__nv_bfloat16><<<grid_for(total, 256), 256, 0, s>>>(
        reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(y), npix, C, scale, shift, prelu);
  else if (dtype == kF16)
    channel_affine_kernel<__half><<<grid_for(total, 256), 256, 0, s>>>(
        reinterpret_cast<const __half*>(x), reinterpret_cast<__half*>(y), npix, C, scale, shift, prelu);
  else
    return "channel_affine: unsupported dtype";
  return last_err();
}

}  // namespace ghost
