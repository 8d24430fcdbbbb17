// K8 / K10 / K11: u8 <-> activation conversion and the fused paste-back.
//
// Exact mode reproduces utils/inference/image_processing.py:51-77 (get_final_image)
// bit for bit: cv2.resize 256->224 (11-bit fixed point), cv2.invertAffineTransform,
// cv2.warpAffine uint8 (10+5-bit fixed-point coordinates, 15-bit weights, replicate
// border) for the swap, cv2.warpAffine float64 (zero border) for the mask, float64
// blend across faces, truncating uint8 store.  All double arithmetic uses explicit
// round-to-nearest intrinsics so nvcc cannot contract it into FMAs.
// Float mode follows utils/inference/video_processing.py:212-233 (kornia warp_affine
// == affine_grid + grid_sample, fp32, zero padding, faces compose through uint8).
#include "paste.cuh"

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <math.h>

namespace ghost {

namespace {

inline const char* last_err() {
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
}

struct CropSrc {
  const void* base;
  int kind;
  // uint8 BGR value of crop pixel (y, x), channel c (0=B,1=G,2=R) of the 256x256 generator output
  __device__ __forceinline__ int px(int crop, int y, int x, int c) const {
    const long long pix = (static_cast<long long>(crop) * 256 + y) * 256 + x;
    if (kind == kCropU8BGR) return reinterpret_cast<const uint8_t*>(base)[pix * 3 + c];
    float v;
    if (kind == kCropYF16) {
      // fp16 generator (inference.py:30 `G.half()`): the reference evaluates (Y*0.5 + 0.5)*255 in fp16, one rounding
      // per operation (faceshifter_run.py:20), then truncates
      const __half h = reinterpret_cast<const __half*>(base)[pix * 3 + (2 - c)];
      const __half t = __hmul(__hadd(__hmul(h, __float2half_rn(0.5f)), __float2half_rn(0.5f)), __float2half_rn(255.f));
      v = fminf(fmaxf(__half2float(t), 0.f), 255.f);
      return static_cast<int>(v);
    }
    if (kind == kCropYBF16)
      v = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(base)[pix * 3 + (2 - c)]);
    else
      v = __ldg(reinterpret_cast<const float*>(base) + pix * 3 + (2 - c));
    // (Y*0.5 + 0.5)*255 then truncating cast (faceshifter_run.py:20-21), separate roundings
    v = __fmul_rn(__fadd_rn(__fmul_rn(v, 0.5f), 0.5f), 255.f);
    v = fminf(fmaxf(v, 0.f), 255.f);
    return static_cast<int>(v);
  }
};

// cv2.resize(256->224, INTER_LINEAR, uint8) of every crop, once per batch: rz[crop][y][x] = {B, G, R, 0}.
// The compositing kernel then samples 4 resized pixels per destination pixel instead of 16 crop pixels.
__global__ void __launch_bounds__(256) resize_crops_kernel(const PasteParams p) {
  const CropSrc cs{p.crops, p.crop_kind};
  const long long total = static_cast<long long>(p.n_crops) * 224 * 224;
  for (long long i = blockIdx.x * 256LL + threadIdx.x; i < total; i += gridDim.x * 256LL) {
    const int x = static_cast<int>(i % 224), y = static_cast<int>((i / 224) % 224), crop = static_cast<int>(i / (224 * 224));
    const int sy = __ldg(p.rs_idx + y), sx = __ldg(p.rs_idx + x);
    const int ay0 = __ldg(p.rs_a0 + y), ay1 = __ldg(p.rs_a1 + y);
    const int ax0 = __ldg(p.rs_a0 + x), ax1 = __ldg(p.rs_a1 + x);
    const int sy1 = min(sy + 1, 255), sx1 = min(sx + 1, 255);
    uint32_t out = 0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const int r0 = cs.px(crop, sy, sx, c) * ax0 + cs.px(crop, sy, sx1, c) * ax1;
      const int r1 = cs.px(crop, sy1, sx, c) * ax0 + cs.px(crop, sy1, sx1, c) * ax1;
      const int v = (((ay0 * (r0 >> 4)) >> 16) + ((ay1 * (r1 >> 4)) >> 16) + 2) >> 2;
      out |= static_cast<uint32_t>(v & 255) << (8 * c);
    }
    reinterpret_cast<uint32_t*>(p.rz)[i] = out;
  }
}

// packed BGR0 of resized pixel (y, x) of one crop
__device__ __forceinline__ uint32_t rz_px(const PasteParams& p, int crop, int y, int x) {
  return __ldg(reinterpret_cast<const uint32_t*>(p.rz) + (static_cast<long long>(crop) * 224 + y) * 224 + x);
}

__device__ __forceinline__ double mask_at(const PasteParams& p, int mi, int y, int x) {
  if (y < 0 || y >= 224 || x < 0 || x >= 224) return 0.0;
  const long long i = (static_cast<long long>(mi) * 224 + y) * 224 + x;
  return p.mask_f64 ? __ldg(reinterpret_cast<const double*>(p.masks) + i)
                    : static_cast<double>(__ldg(reinterpret_cast<const float*>(p.masks) + i));
}

// One destination pixel, exact mode.  fin[3] is the float64 running composite; it is loaded from the frame (offset
// `off`) the first time a face's warped mask is non-zero at this pixel (`touched`).
__device__ __forceinline__ void paste_exact_px(const PasteParams& p, const PasteFace& f, int x, int y, double (&fin)[3],
                                               long long off, bool& touched) {
  // cv2.warpAffine fixed-point coordinates (AB_BITS = 10, INTER_BITS = 5)
  const long long adelta = __double2ll_rn(__dmul_rn(__dmul_rn(f.M[0], static_cast<double>(x)), 1024.0));
  const long long bdelta = __double2ll_rn(__dmul_rn(__dmul_rn(f.M[3], static_cast<double>(x)), 1024.0));
  const long long X0 =
      __double2ll_rn(__dmul_rn(__dadd_rn(__dmul_rn(f.M[1], static_cast<double>(y)), f.M[2]), 1024.0)) + 16;
  const long long Y0 =
      __double2ll_rn(__dmul_rn(__dadd_rn(__dmul_rn(f.M[4], static_cast<double>(y)), f.M[5]), 1024.0)) + 16;
  const long long X = (X0 + adelta) >> 5, Y = (Y0 + bdelta) >> 5;
  long long sxl = X >> 5, syl = Y >> 5;
  sxl = sxl < -32768 ? -32768 : (sxl > 32767 ? 32767 : sxl);
  syl = syl < -32768 ? -32768 : (syl > 32767 ? 32767 : syl);
  const int sx = static_cast<int>(sxl), sy = static_cast<int>(syl);
  const int ax = static_cast<int>(X & 31), ay = static_cast<int>(Y & 31);
  // bilinear weights: (32-ay)(32-ax)/1024 ... exact in fp32, 32x the same integers in the int16 table
  const int w00 = (32 - ay) * (32 - ax), w01 = (32 - ay) * ax, w10 = ay * (32 - ax), w11 = ay * ax;
  const double k = 1.0 / 1024.0;
  double m = __dmul_rn(mask_at(p, f.mask_index, sy, sx), w00 * k);
  m = __dadd_rn(m, __dmul_rn(mask_at(p, f.mask_index, sy, sx + 1), w01 * k));
  m = __dadd_rn(m, __dmul_rn(mask_at(p, f.mask_index, sy + 1, sx), w10 * k));
  m = __dadd_rn(m, __dmul_rn(mask_at(p, f.mask_index, sy + 1, sx + 1), w11 * k));
  if (m == 0.0) return;  // 0*swap + (1-0)*final == final exactly
  if (!touched) {
    fin[0] = static_cast<double>(p.frames_in[off]);
    fin[1] = static_cast<double>(p.frames_in[off + 1]);
    fin[2] = static_cast<double>(p.frames_in[off + 2]);
    touched = true;
  }
  const int x0 = min(max(sx, 0), 223), x1 = min(max(sx + 1, 0), 223);
  const int y0 = min(max(sy, 0), 223), y1 = min(max(sy + 1, 0), 223);
  const uint32_t q00 = rz_px(p, f.crop_index, y0, x0), q01 = rz_px(p, f.crop_index, y0, x1);
  const uint32_t q10 = rz_px(p, f.crop_index, y1, x0), q11 = rz_px(p, f.crop_index, y1, x1);
  const double om = __dsub_rn(1.0, m);
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const int sh = 8 * c;
    const int acc = static_cast<int>((q00 >> sh) & 255) * w00 + static_cast<int>((q01 >> sh) & 255) * w01 +
                    static_cast<int>((q10 >> sh) & 255) * w10 + static_cast<int>((q11 >> sh) & 255) * w11;
    const int sw = (acc + 512) >> 10;  // == (sum(src * 32*w) + 16384) >> 15
    fin[c] = __dadd_rn(__dmul_rn(m, static_cast<double>(sw)), __dmul_rn(om, fin[c]));
  }
}

// Float mode: one destination pixel; fin[] holds the uint8 composite as float.
__device__ __forceinline__ void paste_float_px(const PasteParams& p, const PasteFace& f, int x, int y, float (&fin)[3]) {
  const float fx = __fadd_rn(__fadd_rn(__fmul_rn(f.A[0], x), __fmul_rn(f.A[1], y)), f.A[2]);
  const float fy = __fadd_rn(__fadd_rn(__fmul_rn(f.A[3], x), __fmul_rn(f.A[4], y)), f.A[5]);
  const float x0f = floorf(fx), y0f = floorf(fy);
  const int sx = static_cast<int>(x0f), sy = static_cast<int>(y0f);
  if (sx < -1 || sx >= 224 || sy < -1 || sy >= 224) return;
  const float tx = fx - x0f, ty = fy - y0f;
  const float w00 = (1.f - ty) * (1.f - tx), w01 = (1.f - ty) * tx, w10 = ty * (1.f - tx), w11 = ty * tx;
  const bool vx0 = sx >= 0, vx1 = sx + 1 < 224, vy0 = sy >= 0, vy1 = sy + 1 < 224;
  float m = 0.f;
  uint32_t q00 = 0, q01 = 0, q10 = 0, q11 = 0;
  if (vy0 && vx0) { m += static_cast<float>(mask_at(p, f.mask_index, sy, sx)) * w00; q00 = rz_px(p, f.crop_index, sy, sx); }
  if (vy0 && vx1) { m += static_cast<float>(mask_at(p, f.mask_index, sy, sx + 1)) * w01; q01 = rz_px(p, f.crop_index, sy, sx + 1); }
  if (vy1 && vx0) { m += static_cast<float>(mask_at(p, f.mask_index, sy + 1, sx)) * w10; q10 = rz_px(p, f.crop_index, sy + 1, sx); }
  if (vy1 && vx1) { m += static_cast<float>(mask_at(p, f.mask_index, sy + 1, sx + 1)) * w11; q11 = rz_px(p, f.crop_index, sy + 1, sx + 1); }
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const int sh = 8 * c;
    float sw = 0.f;
    if (vy0 && vx0) sw += static_cast<float>((q00 >> sh) & 255) * w00;
    if (vy0 && vx1) sw += static_cast<float>((q01 >> sh) & 255) * w01;
    if (vy1 && vx0) sw += static_cast<float>((q10 >> sh) & 255) * w10;
    if (vy1 && vx1) sw += static_cast<float>((q11 >> sh) & 255) * w11;
    const float v = __fadd_rn(__fmul_rn(m, sw), __fmul_rn(__fsub_rn(1.f, m), fin[c]));
    // faces compose through uint8 (video_processing.py:233,235)
    fin[c] = static_cast<float>(static_cast<int>(fminf(fmaxf(v, 0.f), 255.f)));
  }
}

// One thread per destination pixel of one face box (blockIdx.z = face).  Frames are copied in -> out beforehand;
// this kernel rewrites only the pixels inside face boxes.  A pixel covered by several boxes of its frame is owned
// by the first of them, and its owner composites every face of the frame in order, so the float64 running blend
// matches the reference's per-frame loop exactly.
__global__ void __launch_bounds__(256) paste_faces_kernel(const PasteParams p) {
  const int j = blockIdx.z;
  const PasteFace& me = p.faces[j];
  const int x = me.x0 + blockIdx.x * 32 + (threadIdx.x & 31);
  const int y = me.y0 + blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= me.x1 || y >= me.y1) return;
  const int f = me.frame;
  const int fs = __ldg(p.face_start + f), fe = __ldg(p.face_start + f + 1);
  for (int k = fs; k < j; ++k) {
    const PasteFace& o = p.faces[k];
    if (x >= o.x0 && x < o.x1 && y >= o.y0 && y < o.y1) return;  // an earlier box owns this pixel
  }
  const long long off = ((static_cast<long long>(f) * p.H + y) * p.W + x) * 3;
  if (p.mode == kPasteExact) {
    // Most pixels of a face box lie outside the blend mask (the box bounds the rotated 224 x 224 crop, the mask is
    // an ellipse inside it): there the composite IS the frame pixel, which frames_out already holds (copied, or the
    // same buffer), so such a pixel costs the two coordinate maps and four mask taps and nothing else.
    double fd[3] = {0.0, 0.0, 0.0};
    bool touched = false;
    for (int k = j; k < fe; ++k) {
      const PasteFace& face = p.faces[k];
      if (x < face.x0 || x >= face.x1 || y < face.y0 || y >= face.y1) continue;
      paste_exact_px(p, face, x, y, fd, off, touched);
    }
    if (touched) {
#pragma unroll
      for (int c = 0; c < 3; ++c) p.frames_out[off + c] = static_cast<uint8_t>(static_cast<int>(fd[c]));
    }
  } else {
    const int b0 = p.frames_in[off], b1 = p.frames_in[off + 1], b2 = p.frames_in[off + 2];
    float ff[3] = {static_cast<float>(b0), static_cast<float>(b1), static_cast<float>(b2)};
    for (int k = j; k < fe; ++k) {
      const PasteFace& face = p.faces[k];
      if (x < face.x0 || x >= face.x1 || y < face.y0 || y >= face.y1) continue;
      paste_float_px(p, face, x, y, ff);
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) p.frames_out[off + c] = static_cast<uint8_t>(static_cast<int>(ff[c]));
  }
}

template <typename T>
__device__ __forceinline__ void st_act(T* p, float v);
template <>
__device__ __forceinline__ void st_act<float>(float* p, float v) {
  *p = v;
}
template <>
__device__ __forceinline__ void st_act<__nv_bfloat16>(__nv_bfloat16* p, float v) {
  *p = __float2bfloat16(v);
}
template <>
__device__ __forceinline__ void st_act<__half>(__half* p, float v) {
  *p = __float2half_rn(v);
}
template <typename T>
struct YKind { static constexpr int value = kCropYF32; };
template <>
struct YKind<__nv_bfloat16> { static constexpr int value = kCropYBF16; };
template <>
struct YKind<__half> { static constexpr int value = kCropYF16; };

template <typename T>
__global__ void preproc_u8_kernel(const uint8_t* __restrict__ src, T* __restrict__ dst, long long npix,
                                  int half_rounding) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < npix;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float v = __fdiv_rn(static_cast<float>(src[i * 3 + (2 - c)]), 255.f);  // BGR -> RGB, /255.
      if (half_rounding) {
        // reference: .half() first, then (x - 0.5)/0.5 evaluated in fp16 (core.py:20-23)
        v = __half2float(__float2half_rn(v));
        v = __half2float(__float2half_rn(__fsub_rn(v, 0.5f)));
        v = __half2float(__float2half_rn(__fdiv_rn(v, 0.5f)));
      } else {
        v = __fdiv_rn(__fsub_rn(v, 0.5f), 0.5f);
      }
      st_act(dst + i * 3 + c, v);
    }
  }
}

template <typename T>
__global__ void postproc_u8_kernel(const T* __restrict__ y, uint8_t* __restrict__ dst, long long npix) {
  CropSrc cs{y, YKind<T>::value};
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < npix;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int crop = static_cast<int>(i >> 16), rem = static_cast<int>(i & 65535);
#pragma unroll
    for (int c = 0; c < 3; ++c) dst[i * 3 + c] = static_cast<uint8_t>(cs.px(crop, rem >> 8, rem & 255, c));
  }
}

template <typename T>
__global__ void arcface_preproc_kernel(const void* __restrict__ src, int is_u8, T* __restrict__ dst, int N) {
  const long long total = static_cast<long long>(N) * 112 * 112;
  const float sc = 223.f / 111.f;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int ow = static_cast<int>(i % 112), oh = static_cast<int>((i / 112) % 112), n = static_cast<int>(i / 12544);
    const float fh = sc * oh, fw = sc * ow;
    const int h0 = static_cast<int>(fh), w0 = static_cast<int>(fw);
    const int h1 = h0 + (h0 < 223 ? 1 : 0), w1 = w0 + (w0 < 223 ? 1 : 0);
    const float lh = fh - h0, lw = fw - w0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float t[4];
      const int hs[4] = {h0, h0, h1, h1}, ws[4] = {w0, w1, w0, w1};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const long long idx = ((static_cast<long long>(n) * 224 + hs[k]) * 224 + ws[k]) * 3 + c;
        float v = is_u8 ? __fdiv_rn(static_cast<float>(reinterpret_cast<const uint8_t*>(src)[idx]), 255.f)
                        : reinterpret_cast<const float*>(src)[idx];
        t[k] = __fdiv_rn(__fsub_rn(v, 0.5f), 0.5f);
      }
      const float v = (1.f - lh) * ((1.f - lw) * t[0] + lw * t[1]) + lh * ((1.f - lw) * t[2] + lw * t[3]);
      st_act(dst + i * 3 + c, v);
    }
  }
}

inline unsigned grid_for(long long total, int threads) {
  long long b = (total + threads - 1) / threads;
  if (b > 148LL * 32) b = 148LL * 32;
  if (b < 1) b = 1;
  return static_cast<unsigned>(b);
}

}  // namespace

void paste_resize_tables(short* idx, short* a0, short* a1) {
  // cv2.resize INTER_LINEAR coefficient generation for 256 -> 224 (float32 arithmetic, 11-bit coefficients)
  const double scale = 256.0 / 224.0;
  for (int d = 0; d < 224; ++d) {
    float f = static_cast<float>((d + 0.5) * scale - 0.5);
    int s = static_cast<int>(floorf(f));
    f -= static_cast<float>(s);
    if (s < 0) { s = 0; f = 0.f; }
    if (s >= 255) { s = 255; f = 0.f; }
    idx[d] = static_cast<short>(s);
    a0[d] = static_cast<short>(lrintf((1.f - f) * 2048.f));
    a1[d] = static_cast<short>(lrintf(f * 2048.f));
  }
}

void paste_prepare_face(PasteFace* f, const double* t, int H, int W, int crop_index, int mask_index) {
  // cv2.invertAffineTransform(tfm) (image_processing.py:69)
  double D = t[0] * t[4] - t[1] * t[3];
  D = D != 0.0 ? 1.0 / D : 0.0;
  double r[6];
  r[0] = t[4] * D;
  r[4] = t[0] * D;
  r[1] = -t[1] * D;
  r[3] = -t[3] * D;
  r[2] = -r[0] * t[2] - r[1] * t[5];
  r[5] = -r[3] * t[2] - r[4] * t[5];
  // cv2.warpAffine inverts the matrix it is given once more, in double
  double M[6] = {r[0], r[1], r[2], r[3], r[4], r[5]};
  double D2 = M[0] * M[4] - M[1] * M[3];
  D2 = D2 != 0.0 ? 1.0 / D2 : 0.0;
  const double A11 = M[4] * D2, A22 = M[0] * D2;
  M[0] = A11;
  M[1] *= -D2;
  M[3] *= -D2;
  M[4] = A22;
  const double b1 = -M[0] * M[2] - M[1] * M[5];
  const double b2 = -M[3] * M[2] - M[4] * M[5];
  M[2] = b1;
  M[5] = b2;
  for (int i = 0; i < 6; ++i) {
    f->M[i] = M[i];
    f->A[i] = static_cast<float>(M[i]);
  }
  // bounding box of the crop square [-1.5, 224.5]^2 mapped into the frame by mat_rev (crop -> frame)
  double xmin = 1e30, xmax = -1e30, ymin = 1e30, ymax = -1e30;
  const double cs[2] = {-1.5, 224.5};
  for (int a = 0; a < 2; ++a)
    for (int b = 0; b < 2; ++b) {
      const double x = r[0] * cs[a] + r[1] * cs[b] + r[2];
      const double y = r[3] * cs[a] + r[4] * cs[b] + r[5];
      xmin = x < xmin ? x : xmin; xmax = x > xmax ? x : xmax;
      ymin = y < ymin ? y : ymin; ymax = y > ymax ? y : ymax;
    }
  auto clampi = [](double v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : static_cast<int>(v)); };
  f->x0 = clampi(floor(xmin) - 2, 0, W);
  f->y0 = clampi(floor(ymin) - 2, 0, H);
  f->x1 = clampi(ceil(xmax) + 3, 0, W);
  f->y1 = clampi(ceil(ymax) + 3, 0, H);
  f->crop_index = crop_index;
  f->mask_index = mask_index;
}

const char* paste_back_launch(PasteParams p, const PasteFace* host_faces, cudaStream_t s) {
  if (p.F <= 0 || p.H <= 0 || p.W <= 0) return "paste_back: empty frame batch";
  if (p.frames_out != p.frames_in) {
    const size_t bytes = static_cast<size_t>(p.F) * p.H * p.W * 3;
    cudaError_t e = cudaMemcpyAsync(p.frames_out, p.frames_in, bytes, cudaMemcpyDeviceToDevice, s);
    if (e != cudaSuccess) return cudaGetErrorString(e);
  }
  if (p.n_faces <= 0) return last_err();
  p.max_bw = p.max_bh = 0;
  for (int i = 0; i < p.n_faces; ++i) {
    const int bw = host_faces[i].x1 - host_faces[i].x0, bh = host_faces[i].y1 - host_faces[i].y0;
    p.max_bw = bw > p.max_bw ? bw : p.max_bw;
    p.max_bh = bh > p.max_bh ? bh : p.max_bh;
  }
  if (p.max_bw <= 0 || p.max_bh <= 0) return last_err();  // every face lies outside its frame
  resize_crops_kernel<<<grid_for(static_cast<long long>(p.n_crops) * 224 * 224, 256), 256, 0, s>>>(p);
  dim3 grid((p.max_bw + 31) / 32, (p.max_bh + 7) / 8, p.n_faces);
  paste_faces_kernel<<<grid, 256, 0, s>>>(p);
  return last_err();
}

const char* preproc_u8_launch(const uint8_t* src, void* dst, int dst_dtype, long long npix, int half_rounding,
                              cudaStream_t s) {
  if (dst_dtype == kF32)
    preproc_u8_kernel<float><<<grid_for(npix, 256), 256, 0, s>>>(src, reinterpret_cast<float*>(dst), npix, half_rounding);
  else if (dst_dtype == kBF16)
    preproc_u8_kernel<__nv_bfloat16>
        <<<grid_for(npix, 256), 256, 0, s>>>(src, reinterpret_cast<__nv_bfloat16*>(dst), npix, half_rounding);
  else if (dst_dtype == kF16)  // fp16 net: the reference's own rounding order (core.py:20-23), whatever half_rounding says
    preproc_u8_kernel<__half><<<grid_for(npix, 256), 256, 0, s>>>(src, reinterpret_cast<__half*>(dst), npix, 1);
  else
    return "preproc_u8: unsupported dtype";
  return last_err();
}

const char* postproc_u8_launch(const void* y, int y_dtype, uint8_t* dst, long long npix, cudaStream_t s) {
  if (y_dtype == kF32)
    postproc_u8_kernel<float><<<grid_for(npix, 256), 256, 0, s>>>(reinterpret_cast<const float*>(y), dst, npix);
  else if (y_dtype == kBF16)
    postproc_u8_kernel<__nv_bfloat16>
        <<<grid_for(npix, 256), 256, 0, s>>>(reinterpret_cast<const __nv_bfloat16*>(y), dst, npix);
  else if (y_dtype == kF16)
    postproc_u8_kernel<__half><<<grid_for(npix, 256), 256, 0, s>>>(reinterpret_cast<const __half*>(y), dst, npix);
  else
    return "postproc_u8: unsupported dtype";
  return last_err();
}

const char* arcface_preproc_launch(const void* src, int is_u8, void* dst, int dst_dtype, int N, cudaStream_t s) {
  const long long total = static_cast<long long>(N) * 112 * 112;
  if (dst_dtype == kF32)
    arcface_preproc_kernel<float><<<grid_for(total, 256), 256, 0, s>>>(src, is_u8, reinterpret_cast<float*>(dst), N);
  else if (dst_dtype == kBF16)
    arcface_preproc_kernel<__nv_bfloat16>
        <<<grid_for(total, 256), 256, 0, s>>>(src, is_u8, reinterpret_cast<__nv_bfloat16*>(dst), N);
  else if (dst_dtype == kF16)
    arcface_preproc_kernel<__half><<<grid_for(total, 256), 256, 0, s>>>(src, is_u8, reinterpret_cast<__half*>(dst), N);
  else
    return "arcface_preproc: unsupported dtype";
  return last_err();
}

}  // namespace ghost
