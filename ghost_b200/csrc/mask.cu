// Blend-mask synthesis from 106-point landmarks, bit-exact against the OpenCV calls of
// utils/inference/masks.py:5-108 (expand_eyebrows, get_mask, face_mask_static, erode_and_blur):
//   host  : eyebrow lift, convex hull, cv::FillConvexPoly rasterisation (outline by cv::Line / clipLine + scan-line
//           fill with 16.16 fixed-point edges) into a 224x224 bitmap, bit-exact Gaussian kernels (8 fractional bits,
//           error diffusion) - a few hundred integer operations per face;
//   device: one CTA per face, everything in shared memory: bitmap -> uint8 mask, k x k erode / dilate (separable
//           min / max, border ignored), border clip, fixed-point separable Gaussian (8.8 horizontal, 16.16 vertical,
//           BORDER_REFLECT_101, round half up), /255 -> float64 / float32 mask in the layout the paste-back reads.
// The CPU restatement this was written against is oracle/mask_ref.py (pinned to cv2 and to the reference).
#include "mask.cuh"

#include <math.h>
#include <string.h>

#include <algorithm>
#include <vector>

namespace ghost {

namespace {

inline const char* last_err() {
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
}

constexpr int S = kMaskSize;
constexpr int kXYShift = 16;
constexpr long long kXYOne = 1LL << kXYShift;

struct Pt {
  long long x, y;
};

inline void set_px(uint32_t* bm, long long x, long long y) {
  if (x >= 0 && x < S && y >= 0 && y < S) bm[y * kMaskWords + (x >> 5)] |= 1u << (x & 31);
}

// cv::clipLine (imgproc/src/drawing.cpp)
bool clip_line(Pt& p1, Pt& p2) {
  const long long right = S - 1, bottom = S - 1;
  long long &x1 = p1.x, &y1 = p1.y, &x2 = p2.x, &y2 = p2.y;
  int c1 = (x1 < 0) + (x1 > right) * 2 + (y1 < 0) * 4 + (y1 > bottom) * 8;
  int c2 = (x2 < 0) + (x2 > right) * 2 + (y2 < 0) * 4 + (y2 > bottom) * 8;
  if ((c1 & c2) == 0 && (c1 | c2) != 0) {
    long long a;
    if (c1 & 12) {
      a = c1 < 8 ? 0 : bottom;
      x1 += static_cast<long long>(static_cast<double>(a - y1) * (x2 - x1) / (y2 - y1));
      y1 = a;
      c1 = (x1 < 0) + (x1 > right) * 2;
    }
    if (c2 & 12) {
      a = c2 < 8 ? 0 : bottom;
      x2 += static_cast<long long>(static_cast<double>(a - y2) * (x2 - x1) / (y2 - y1));
      y2 = a;
      c2 = (x2 < 0) + (x2 > right) * 2;
    }
    if ((c1 & c2) == 0 && (c1 | c2) != 0) {
      if (c1) {
        a = c1 == 1 ? 0 : right;
        y1 += static_cast<long long>(static_cast<double>(a - x1) * (y2 - y1) / (x2 - x1));
        x1 = a;
        c1 = 0;
      }
      if (c2) {
        a = c2 == 1 ? 0 : right;
        y2 += static_cast<long long>(static_cast<double>(a - x2) * (y2 - y1) / (x2 - x1));
        x2 = a;
        c2 = 0;
      }
    }
  }
  return (c1 | c2) == 0;
}

// cv::Line, connectivity 8 (LineIterator, leftToRight) after clipLine
void draw_line(uint32_t* bm, Pt p1, Pt p2) {
  if (!clip_line(p1, p2)) return;
  long long dx = p2.x - p1.x, dy = p2.y - p1.y;
  long long x = p1.x, y = p1.y;
  if (dx < 0) {
    x = p2.x;
    y = p2.y;
    dx = -dx;
    dy = -dy;
  }
  const long long sy = dy >= 0 ? 1 : -1;
  if (dy < 0) dy = -dy;
  if (dy > dx) {
    long long err = dy - 2 * dx;
    for (long long i = 0; i <= dy; ++i) {
      set_px(bm, x, y);
      const bool m = err < 0;
      err += -2 * dx + (m ? 2 * dy : 0);
      y += sy;
      if (m) x += 1;
    }
  } else {
    long long err = dx - 2 * dy;
    for (long long i = 0; i <= dx; ++i) {
      set_px(bm, x, y);
      const bool m = err < 0;
      err += -2 * dy + (m ? 2 * dx : 0);
      x += 1;
      if (m) y += sy;
    }
  }
}

// cv::FillConvexPoly (line_type 8, shift 0) onto a zero bitmap
void fill_convex_poly(uint32_t* bm, const std::vector<Pt>& v) {
  const int n = static_cast<int>(v.size());
  if (n == 0) return;
  for (int i = 0; i < n; ++i) draw_line(bm, v[(i + n - 1) % n], v[i]);
  long long ymin = v[0].y, ymax = v[0].y, xmin = v[0].x, xmax = v[0].x;
  int imin = 0;
  for (int i = 0; i < n; ++i) {
    if (v[i].y < ymin) {
      ymin = v[i].y;
      imin = i;
    }
    ymax = std::max(ymax, v[i].y);
    xmax = std::max(xmax, v[i].x);
    xmin = std::min(xmin, v[i].x);
  }
  if (n < 3 || xmax < 0 || ymax < 0 || xmin >= S || ymin >= S) return;
  ymax = std::min<long long>(ymax, S - 1);
  int edges = n;
  struct Edge {
    int idx, di;
    long long x, dx;
    long long ye;
  } edge[2];
  edge[0].idx = edge[1].idx = imin;
  edge[0].ye = edge[1].ye = ymin;
  edge[0].di = 1;
  edge[1].di = n - 1;
  edge[0].x = edge[1].x = -kXYOne;
  edge[0].dx = edge[1].dx = 0;
  long long y = ymin;
  const long long delta = kXYOne >> 1;
  do {
    for (int i = 0; i < 2; ++i) {
      if (y >= edge[i].ye) {
        int idx0 = edge[i].idx, di = edge[i].di;
        int idx = idx0 + di;
        if (idx >= n) idx -= n;
        for (; edges-- > 0;) {
          const long long ty = v[idx].y;
          if (ty > y) {
            const long long xs = v[idx0].x << kXYShift, xe = v[idx].x << kXYShift;
            edge[i].ye = ty;
            edge[i].dx = ((xe - xs) * 2 + (ty - y)) / (2 * (ty - y));
            edge[i].x = xs;
            edge[i].idx = idx;
            break;
          }
          idx0 = idx;
          idx += di;
          if (idx >= n) idx -= n;
        }
      }
    }
    if (edges < 0) break;
    if (y >= 0) {
      int left = 0, right = 1;
      if (edge[0].x > edge[1].x) left = 1, right = 0;
      long long xx1 = (edge[left].x + delta) >> kXYShift;
      long long xx2 = (edge[right].x + delta) >> kXYShift;
      if (xx2 >= 0 && xx1 < S) {
        if (xx1 < 0) xx1 = 0;
        if (xx2 >= S) xx2 = S - 1;
        for (long long x = xx1; x <= xx2; ++x) set_px(bm, x, y);
      }
    }
    edge[0].x += edge[0].dx;
    edge[1].x += edge[1].dx;
  } while (++y <= ymax);
}

// convex hull without collinear vertices (monotone chain, exact integer arithmetic); any exact hull gives the same
// rasterisation as cv2.convexHull's (checked against cv2 in tests/test_oracle_mask.py)
std::vector<Pt> convex_hull(std::vector<Pt> pts) {
  std::sort(pts.begin(), pts.end(), [](const Pt& a, const Pt& b) { return a.x < b.x || (a.x == b.x && a.y < b.y); });
  pts.erase(std::unique(pts.begin(), pts.end(), [](const Pt& a, const Pt& b) { return a.x == b.x && a.y == b.y; }),
            pts.end());
  const int n = static_cast<int>(pts.size());
  if (n <= 2) return pts;
  auto cross = [](const Pt& o, const Pt& a, const Pt& b) {
    return (a.x - o.x) * (b.y - o.y) - (a.y - o.y) * (b.x - o.x);
  };
  std::vector<Pt> lower, upper;
  for (int i = 0; i < n; ++i) {
    while (lower.size() >= 2 && cross(lower[lower.size() - 2], lower.back(), pts[i]) <= 0) lower.pop_back();
    lower.push_back(pts[i]);
  }
  for (int i = n - 1; i >= 0; --i) {
    while (upper.size() >= 2 && cross(upper[upper.size() - 2], upper.back(), pts[i]) <= 0) upper.pop_back();
    upper.push_back(pts[i]);
  }
  lower.pop_back();
  upper.pop_back();
  lower.insert(lower.end(), upper.begin(), upper.end());
  return lower;
}

// getGaussianKernelBitExact + getGaussianKernelFixedPoint_ED (imgproc/src/smooth.dispatch.cpp), 8 fractional bits
int gaussian_kernel_fixed(double sigma, uint16_t* out) {
  const int n = static_cast<int>(nearbyint(sigma * 3 * 2 + 1)) | 1;
  if (n > kMaskMaxTaps) return -1;
  const int n2 = (n - 1) / 2;
  const double scale2x = -0.125 / (sigma * sigma);
  std::vector<double> vals(n2);
  double s = 0.0;
  long long x = 1 - n;
  for (int i = 0; i < n2; ++i, x += 2) {
    const double t = exp(static_cast<double>(x * x) * scale2x);
    vals[i] = t;
    s += t;
  }
  s = s * 2.0 + 1.0;
  const double mul1 = 1.0 / s;
  double err = 0.0;
  long long tot = 0;
  for (int i = 0; i < n2; ++i) {
    const double adj = vals[i] * mul1 * 256.0 + err;
    const long long v0 = static_cast<long long>(nearbyint(adj));
    err = adj - static_cast<double>(v0);
    out[i] = out[n - 1 - i] = static_cast<uint16_t>(v0);
    tot += v0;
  }
  out[n2] = static_cast<uint16_t>(256 - 2 * tot);
  return n;
}

__device__ __forceinline__ int reflect101(int i, int n) {
  while (i < 0 || i >= n) i = i < 0 ? -i : 2 * (n - 1) - i;
  return i;
}

// One CTA per face.  Shared memory: A, B uint8 [224][224], H uint16 [224][224].
template <typename T>
__global__ void __launch_bounds__(1024) face_mask_kernel(const MaskJob* __restrict__ jobs, T* __restrict__ out) {
  extern __shared__ __align__(16) uint8_t smem[];
  uint8_t* A = smem;
  uint8_t* B = smem + S * S;
  uint16_t* H = reinterpret_cast<uint16_t*>(smem + 2 * S * S);
  __shared__ uint16_t kx[kMaskMaxTaps], ky[kMaskMaxTaps];
  const MaskJob& job = jobs[blockIdx.x];
  const int tid = threadIdx.x, nt = blockDim.x;
  for (int i = tid; i < kMaskMaxTaps; i += nt) {
    kx[i] = job.kx[i];
    ky[i] = job.ky[i];
  }
  // 1. bitmap -> 0 / 255
  for (int i = tid; i < S * S; i += nt) {
    const int y = i / S, x = i - y * S;
    A[i] = (job.bitmap[y * kMaskWords + (x >> 5)] >> (x & 31)) & 1u ? 255 : 0;
  }
  __syncthreads();
  // 2. k x k erode (min) / dilate (max): rows then columns; window [p - k/2, p - k/2 + k - 1], outside ignored
  const int k = job.erode > 0 ? job.erode : -job.erode;
  const bool er = job.erode > 0;
  const int a = k / 2;
  for (int i = tid; i < S * S; i += nt) {
    const int y = i / S, x = i - y * S;
    const int x0 = max(0, x - a), x1 = min(S - 1, x - a + k - 1);
    int v = er ? 255 : 0;
    for (int xx = x0; xx <= x1; ++xx) v = er ? min(v, static_cast<int>(A[y * S + xx])) : max(v, static_cast<int>(A[y * S + xx]));
    B[i] = static_cast<uint8_t>(v);
  }
  __syncthreads();
  const int c = job.clip;
  for (int i = tid; i < S * S; i += nt) {
    const int y = i / S, x = i - y * S;
    const int y0 = max(0, y - a), y1 = min(S - 1, y - a + k - 1);
    int v = er ? 255 : 0;
    for (int yy = y0; yy <= y1; ++yy) v = er ? min(v, static_cast<int>(B[yy * S + x])) : max(v, static_cast<int>(B[yy * S + x]));
    // fade_to_border (masks.py:100-105): numpy slices [:c] / [-c:] (c == 0 would clear everything, as in numpy)
    const bool cleared = c == 0 || y < c || y >= S - c || x < c || x >= S - c;
    A[i] = cleared ? 0 : static_cast<uint8_t>(v);
  }
  __syncthreads();
  // 3. horizontal pass, ufixedpoint16 (8.8): sum <= 255 * 256
  const int nx = job.nx, ny = job.ny, rx = nx / 2, ry = ny / 2;
  for (int i = tid; i < S * S; i += nt) {
    const int y = i / S, x = i - y * S;
    uint32_t acc = 0;
    for (int t = 0; t < nx; ++t) acc += static_cast<uint32_t>(kx[t]) * A[y * S + reflect101(x + t - rx, S)];
    H[i] = static_cast<uint16_t>(acc);
  }
  __syncthreads();
  // 4. vertical pass, ufixedpoint32 (16.16), rounding cast to uint8, then mask / 255 (numpy true division)
  T* o = out + static_cast<long long>(blockIdx.x) * S * S;
  for (int i = tid; i < S * S; i += nt) {
    const int y = i / S, x = i - y * S;
    uint32_t acc = 0;
    for (int t = 0; t < ny; ++t) acc += static_cast<uint32_t>(ky[t]) * H[reflect101(y + t - ry, S) * S + x];
    const uint32_t u8 = (acc + 32768u) >> 16;
    o[i] = static_cast<T>(__ddiv_rn(static_cast<double>(u8), 255.0));
  }
}

}  // namespace

const char* mask_prepare_job(MaskJob* job, const float* lm, int erode, int sigma_x, int sigma_y) {
  if (erode == 0 || erode > 64 || erode < -64) return "face mask: erode must be a non-zero integer in [-64, 64]";
  if (sigma_x <= 0 || sigma_y <= 0) return "face mask: sigmaX / sigmaY must be positive";
  memset(job, 0, sizeof(*job));
  // expand_eyebrows (masks.py:5-20): int32 truncation, float64 arithmetic, truncation again
  static const int brow_l[5] = {43, 48, 49, 51, 50}, eye_l[5] = {35, 41, 40, 42, 39};
  static const int brow_r[5] = {102, 103, 104, 105, 101}, eye_r[5] = {89, 95, 94, 96, 93};
  const double mod = erode == 15 ? 2.7 : (erode == -5 ? 0.5 : 2.0);  // masks.py:71-76
  int pts[106][2];
  for (int i = 0; i < 106; ++i) {
    pts[i][0] = static_cast<int>(lm[2 * i]);
    pts[i][1] = static_cast<int>(lm[2 * i + 1]);
  }
  const double f = mod * 0.5;
  int lifted[10][2];
  for (int s = 0; s < 2; ++s) {
    const int* brow = s == 0 ? brow_l : brow_r;
    const int* eye = s == 0 ? eye_l : eye_r;
    for (int i = 0; i < 5; ++i)
      for (int d = 0; d < 2; ++d) {
        const double top = pts[brow[i]][d], diff = pts[brow[i]][d] - pts[eye[i]][d];
        const double prod = f * diff;
        lifted[s * 5 + i][d] = static_cast<int>(top + prod);
      }
  }
  for (int s = 0; s < 2; ++s) {
    const int* brow = s == 0 ? brow_l : brow_r;
    for (int i = 0; i < 5; ++i) {
      pts[brow[i]][0] = lifted[s * 5 + i][0];
      pts[brow[i]][1] = lifted[s * 5 + i][1];
    }
  }
  std::vector<Pt> v(106);
  for (int i = 0; i < 106; ++i) v[i] = Pt{pts[i][0], pts[i][1]};
  fill_convex_poly(job->bitmap, convex_hull(v));
  job->erode = erode;
  job->clip = sigma_y * 2;
  job->nx = gaussian_kernel_fixed(static_cast<double>(sigma_x), job->kx);
  job->ny = gaussian_kernel_fixed(static_cast<double>(sigma_y), job->ky);
  if (job->nx < 0 || job->ny < 0) return "face mask: sigma too large (kernel longer than 127 taps)";
  return nullptr;
}

const char* face_masks_launch(const MaskJob* jobs_dev, int n, void* out, int f64, cudaStream_t s) {
  if (n <= 0) return nullptr;
  const int smem = 2 * S * S + 2 * S * S;
  {
    cudaError_t e = f64 ? cudaFuncSetAttribute(face_mask_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)
                        : cudaFuncSetAttribute(face_mask_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) {
      cudaGetLastError();
      return cudaGetErrorString(e);
    }
  }
  if (f64)
    face_mask_kernel<double><<<n, 1024, smem, s>>>(jobs_dev, reinterpret_cast<double*>(out));
  else
    face_mask_kernel<float><<<n, 1024, smem, s>>>(jobs_dev, reinterpret_cast<float*>(out));
  return last_err();
}

}  // namespace ghost
