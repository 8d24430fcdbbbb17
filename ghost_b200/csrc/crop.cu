// Crop extraction in front of the generator, bit-exact against the reference's OpenCV calls:
//   utils/inference/video_processing.py:159-168  cv2.warpAffine(frame, M, (224, 224), borderValue=0.0)
//   utils/inference/video_processing.py:174-188  cv2.resize(crop, (256, 256))
// warpAffine: 10+5-bit fixed-point source coordinates, 15-bit bilinear weights, BORDER_CONSTANT 0 (same coordinate
// model as the paste-back, paste.cu).  resize 224 -> 256: 11-bit coefficients; horizontally index and fraction are
// clamped at the borders, vertically only the row indices (the fraction is kept) - oracle/crop_ref.py.
#include "crop.cuh"

#include <math.h>

namespace ghost {

namespace {

inline const char* last_err() {
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
}

// one thread per pixel of one 224x224 crop (blockIdx.y = face)
__global__ void __launch_bounds__(256) crop_warp_kernel(const uint8_t* __restrict__ frames, int H, int W,
                                                         const CropFace* __restrict__ faces,
                                                         uint8_t* __restrict__ out224) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= 224 * 224) return;
  const CropFace& f = faces[blockIdx.y];
  const int y = i / 224, x = i - y * 224;
  // cv2.warpAffine fixed-point coordinates (AB_BITS = 10, INTER_BITS = 5); explicit roundings, no FMA contraction
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
  const int w00 = (32 - ay) * (32 - ax), w01 = (32 - ay) * ax, w10 = ay * (32 - ax), w11 = ay * ax;
  const uint8_t* fr = frames + static_cast<long long>(f.frame) * H * W * 3;
  const bool vx0 = sx >= 0 && sx < W, vx1 = sx + 1 >= 0 && sx + 1 < W;
  const bool vy0 = sy >= 0 && sy < H, vy1 = sy + 1 >= 0 && sy + 1 < H;
  uint8_t* o = out224 + (static_cast<long long>(blockIdx.y) * 224 * 224 + i) * 3;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    int acc = 0;  // taps outside the frame carry the border value 0
    if (vy0 && vx0) acc += fr[(static_cast<long long>(sy) * W + sx) * 3 + c] * w00;
    if (vy0 && vx1) acc += fr[(static_cast<long long>(sy) * W + sx + 1) * 3 + c] * w01;
    if (vy1 && vx0) acc += fr[(static_cast<long long>(sy + 1) * W + sx) * 3 + c] * w10;
    if (vy1 && vx1) acc += fr[(static_cast<long long>(sy + 1) * W + sx + 1) * 3 + c] * w11;
    o[c] = static_cast<uint8_t>((acc + 512) >> 10);  // == (sum(src * 32*w) + 16384) >> 15
  }
}

// cv2.resize 224 -> 256; tables t[7][256]: x index, x a0, x a1, row0, row1, b0, b1
__global__ void __launch_bounds__(256) crop_resize_kernel(const uint8_t* __restrict__ in224, const short* __restrict__ t,
                                                           uint8_t* __restrict__ out256, long long total) {
  for (long long i = blockIdx.x * 256LL + threadIdx.x; i < total; i += gridDim.x * 256LL) {
    const int x = static_cast<int>(i & 255), y = static_cast<int>((i >> 8) & 255);
    const long long face = i >> 16;
    const int sx = t[x], a0 = t[256 + x], a1 = t[512 + x];
    const int r0 = t[768 + y], r1 = t[1024 + y], b0 = t[1280 + y], b1 = t[1536 + y];
    const int sx1 = min(sx + 1, 223);
    const uint8_t* src = in224 + face * 224 * 224 * 3;
    uint8_t* o = out256 + i * 3;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const int h0 = src[(r0 * 224 + sx) * 3 + c] * a0 + src[(r0 * 224 + sx1) * 3 + c] * a1;
      const int h1 = src[(r1 * 224 + sx) * 3 + c] * a0 + src[(r1 * 224 + sx1) * 3 + c] * a1;
      o[c] = static_cast<uint8_t>((((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2);
    }
  }
}

}  // namespace

void crop_prepare_face(CropFace* f, const double* t, int frame) {
  // cv2.warpAffine (no WARP_INVERSE_MAP) inverts the frame -> crop matrix it is given, in double
  double M[6] = {t[0], t[1], t[2], t[3], t[4], t[5]};
  double D = M[0] * M[4] - M[1] * M[3];
  D = D != 0.0 ? 1.0 / D : 0.0;
  const double A11 = M[4] * D, A22 = M[0] * D;
  M[0] = A11;
  M[1] *= -D;
  M[3] *= -D;
  M[4] = A22;
  const double b1 = -M[0] * M[2] - M[1] * M[5];
  const double b2 = -M[3] * M[2] - M[4] * M[5];
  M[2] = b1;
  M[5] = b2;
  for (int i = 0; i < 6; ++i) f->M[i] = M[i];
  f->frame = frame;
  f->pad_ = 0;
}

void crop_resize_tables(short* t) {
  const double scale = 224.0 / 256.0;
  for (int d = 0; d < 256; ++d) {
    float f = static_cast<float>((d + 0.5) * scale - 0.5);
    int s = static_cast<int>(floorf(f));
    f -= static_cast<float>(s);
    t[768 + d] = static_cast<short>(s < 0 ? 0 : (s > 223 ? 223 : s));
    t[1024 + d] = static_cast<short>(s + 1 < 0 ? 0 : (s + 1 > 223 ? 223 : s + 1));
    t[1280 + d] = static_cast<short>(lrintf((1.f - f) * 2048.f));
    t[1536 + d] = static_cast<short>(lrintf(f * 2048.f));
    if (s < 0) { s = 0; f = 0.f; }
    if (s >= 223) { s = 223; f = 0.f; }
    t[d] = static_cast<short>(s);
    t[256 + d] = static_cast<short>(lrintf((1.f - f) * 2048.f));
    t[512 + d] = static_cast<short>(lrintf(f * 2048.f));
  }
}

const char* crop_faces_launch(const uint8_t* frames, int H, int W, const CropFace* faces_dev, int n,
                              const short* tables_dev, uint8_t* crops224, uint8_t* crops256, cudaStream_t s) {
  if (n <= 0) return nullptr;
  dim3 grid((224 * 224 + 255) / 256, n);
  crop_warp_kernel<<<grid, 256, 0, s>>>(frames, H, W, faces_dev, crops224);
  if (crops256) {
    const long long total = static_cast<long long>(n) * 65536;
    long long blocks = (total + 255) / 256;
    if (blocks > 148LL * 32) blocks = 148LL * 32;
    crop_resize_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(crops224, tables_dev, crops256, total);
  }
  return last_err();
}

}  // namespace ghost
