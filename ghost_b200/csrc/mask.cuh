// Blend-mask synthesis from landmarks (utils/inference/masks.py:5-108): host-side polygon
// rasterisation + per-face device kernel (erode / dilate, border clip, fixed-point Gaussian).  See mask.cu.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ghost {

constexpr int kMaskSize = 224;                 // the aligned crop the landmarks refer to (image_processing.py:63-68)
constexpr int kMaskWords = kMaskSize / 32;     // bitmap words per row
constexpr int kMaskMaxTaps = 128;

// Everything the device needs for one face (6.8 KB): the filled hull as a bitmap, the morphology size, the border
// clip and the two fixed-point Gaussian kernels.
struct MaskJob {
  uint32_t bitmap[kMaskSize * kMaskWords];
  int erode;  // > 0: erode with a k x k box, < 0: dilate with |k| x |k|
  int clip;   // 2 * sigmaY rows / columns cleared at every border
  int nx, ny;
  uint16_t kx[kMaskMaxTaps], ky[kMaskMaxTaps];
};

// lm: 106 x 2 float32 landmarks in crop coordinates (what Handler.get_without_detection_without_transform returns).
// Returns nullptr or an error message.
const char* mask_prepare_job(MaskJob* job, const float* lm, int erode, int sigma_x, int sigma_y);
// out: [n][224][224] float64 (f64 = 1) or float32 on the device
const char* face_masks_launch(const MaskJob* jobs_dev, int n, void* out, int f64, cudaStream_t s);

}  // namespace ghost
