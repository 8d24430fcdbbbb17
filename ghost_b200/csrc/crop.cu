// Crop extraction in front of the generator, bit-exact against the reference's OpenCV calls:
//   utils/inference/video_processing.py:159-168  cv2.warpAffine(frame, M, (224, 224), borderValue=0.0)
//   utils/inference/video_processing.py:174-188  cv2.resize(crop, (256, 256))
// warpAffine: 10+5-bit fixed-point source coordinates, 15-bit bilinear weights, BORDER_CONSTANT 0 (same coordinate
// model as the paste-back, paste.cu).  resize 224 -> 256: 11-bit coefficients; horizontally index and fraction are
// clamped at the borders, vertically only the row indices (the fraction is kept) - oracle/crop_ref.py.
#include "crop.cuh"

#include <math.h>
#include <stdio.h>

#include <mutex>

#include "../../include/ghost_b200.h"

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

// ------------------------------------------------------------------------------------------------------------------
// ghost_crop_faces_noalloc: the same two kernels with every byte of scratch (face table, resize tables, the 224x224
// intermediate when the caller does not want it) in ONE file-scope device array instead of a stream-ordered
// allocation - for callers that cannot call cudaMalloc (CUDA graph capture, serving loops, harnesses that forbid
// it).  No counterpart in the reference.
//
// Layout of the array (per GPU that loads the library, held until exit):
//   [0, 2048)        CropFace[GHOST_CROP_NOALLOC_MAX_FACES]   rewritten by every call
//   [2048, 6144)     7 x 256 shorts of cv2.resize 224 -> 256   written once per device at set-up
//   [6144, ...)      MAX_FACES x 224 x 224 x 3 bytes           intermediate crops when crops224 == NULL
// Nothing in it has to start at zero: every byte a kernel reads was written earlier in the same call (or at set-up).
// Every call shares it, so calls that overlap on the GPU (different streams, different threads) race.
namespace {

constexpr int kNaMaxFaces = GHOST_CROP_NOALLOC_MAX_FACES;
constexpr size_t kNaTablesOff = 2048, kNa224Off = 6144;
constexpr size_t kNaBytes = kNa224Off + static_cast<size_t>(kNaMaxFaces) * 224 * 224 * 3;
static_assert(sizeof(ghost::CropFace) * kNaMaxFaces <= kNaTablesOff, "face table overruns its slot");
static_assert(7 * 256 * sizeof(short) <= kNa224Off - kNaTablesOff, "resize tables overrun their slot");

static __device__ __align__(256) unsigned char g_crop_noalloc_scratch[kNaBytes];

constexpr int kNaMaxDevices = 64;
unsigned char* g_na_ptr[kNaMaxDevices];   // device address of the array, per device (NOT the host-side symbol)
std::mutex g_na_mu;

// once per device, with that device current: resolve the device address of the array and write the constant tables
const char* noalloc_setup(unsigned char** out, char* msg, size_t cap) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    snprintf(msg, cap, "ghost_crop_faces_noalloc: cudaGetDevice: %s", cudaGetErrorString(e));
    return msg;
  }
  if (dev < 0 || dev >= kNaMaxDevices) {
    snprintf(msg, cap, "ghost_crop_faces_noalloc: device index %d out of range", dev);
    return msg;
  }
  std::lock_guard<std::mutex> lk(g_na_mu);
  if (!g_na_ptr[dev]) {
    unsigned char* p = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_crop_noalloc_scratch);
    if (e != cudaSuccess || !p) {
      snprintf(msg, cap, "ghost_crop_faces_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return msg;
    }
    short tables[7 * 256];
    ghost::crop_resize_tables(tables);
    e = cudaMemcpy(p + kNaTablesOff, tables, sizeof(tables), cudaMemcpyHostToDevice);   // synchronous, 3.5 KB, once
    if (e != cudaSuccess) {
      snprintf(msg, cap, "ghost_crop_faces_noalloc: table upload: %s", cudaGetErrorString(e));
      return msg;
    }
    g_na_ptr[dev] = p;
  }
  *out = g_na_ptr[dev];
  return nullptr;
}

}  // namespace

extern "C" int ghost_crop_faces_noalloc(const uint8_t* frames, int n_frames, int H, int W, const double* tfm,
                                        const int* face_frame, int n, uint8_t* crops224, uint8_t* crops256,
                                        void* stream) {
  char msg[256];
  if (!(frames && tfm && face_frame && n > 0 && n_frames > 0 && H > 0 && W > 0)) {
    ghost::set_last_error("ghost_crop_faces_noalloc: bad argument");
    return GHOST_ERR_INVALID;
  }
  if (!(crops224 || crops256)) {
    ghost::set_last_error("ghost_crop_faces_noalloc: no output buffer");
    return GHOST_ERR_INVALID;
  }
  if (n > kNaMaxFaces) {
    snprintf(msg, sizeof(msg), "ghost_crop_faces_noalloc: n = %d exceeds the static scratch limit of %d faces per call "
                               "(split the batch or use ghost_crop_faces)", n, kNaMaxFaces);
    ghost::set_last_error(msg);
    return GHOST_ERR_INVALID;
  }
  ghost::CropFace faces[kNaMaxFaces];
  for (int i = 0; i < n; ++i) {
    if (face_frame[i] < 0 || face_frame[i] >= n_frames) {
      ghost::set_last_error("ghost_crop_faces_noalloc: face_frame out of range");
      return GHOST_ERR_INVALID;
    }
    ghost::crop_prepare_face(&faces[i], tfm + 6 * i, face_frame[i]);
  }
  unsigned char* scratch = nullptr;
  if (const char* err = noalloc_setup(&scratch, msg, sizeof(msg))) {
    ghost::set_last_error(err);
    return GHOST_ERR_CUDA;
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  // pageable source: staged by the time the call returns, ordered on `s` behind the previous call's kernels
  cudaError_t e = cudaMemcpyAsync(scratch, faces, sizeof(ghost::CropFace) * n, cudaMemcpyHostToDevice, s);
  if (e != cudaSuccess) {
    snprintf(msg, sizeof(msg), "ghost_crop_faces_noalloc: face table upload: %s", cudaGetErrorString(e));
    ghost::set_last_error(msg);
    return GHOST_ERR_CUDA;
  }
  uint8_t* c224 = crops224 ? crops224 : scratch + kNa224Off;
  const char* err = ghost::crop_faces_launch(frames, H, W, reinterpret_cast<const ghost::CropFace*>(scratch), n,
                                             reinterpret_cast<const short*>(scratch + kNaTablesOff), c224, crops256, s);
  if (err) {
    snprintf(msg, sizeof(msg), "ghost_crop_faces_noalloc: %s", err);
    ghost::set_last_error(msg);
    return GHOST_ERR_CUDA;
  }
  return GHOST_OK;
}

