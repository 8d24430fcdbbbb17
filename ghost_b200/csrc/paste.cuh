// Pre/post conversion (K8/K10) and the fused paste-back (K11).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "simt.cuh"

namespace ghost {

enum PasteMode : int { kPasteExact = 0, kPasteFloat = 1 };
enum CropKind : int { kCropU8BGR = 0, kCropYBF16 = 1, kCropYF32 = 2, kCropYF16 = 3 };

// One (frame, face) pair.  Filled on the host by paste_prepare_face().
struct PasteFace {
  double M[6];     // exact mode: src = M * dst, the matrix cv2.warpAffine iterates with (double)
  float A[6];      // float mode: src = A * dst in fp32 (kornia warp_affine / grid_sample, align_corners=True)
  int x0, y0, x1, y1;  // destination bounding box of the 224x224 crop square (exclusive upper bounds)
  int crop_index;  // which 256x256 crop of the batch
  int mask_index;  // which 224x224 mask
  int frame;       // destination frame
  int pad_;
};

struct PasteParams {
  const void* crops;  // u8 BGR [n,256,256,3], or generator output Y NHWC-3 (RGB, [-1,1]) bf16 / fp32 / fp16
  int crop_kind;
  const void* masks;  // [n,224,224] float64 or float32
  int mask_f64;
  const uint8_t* frames_in;  // [F,H,W,3] u8 BGR
  uint8_t* frames_out;       // [F,H,W,3]
  int F, H, W;
  const PasteFace* faces;       // faces of frame f are faces[face_start[f] .. face_start[f+1])
  const int* face_start;        // [F+1]
  const short* rs_idx;          // [224] cv2.resize 256->224 source index
  const short* rs_a0;           // [224] 11-bit coefficient of the left/top sample
  const short* rs_a1;           // [224] 11-bit coefficient of the right/bottom sample
  int mode;
  // filled by the launcher
  int n_faces, n_crops;
  int max_bw, max_bh;  // largest face box of the batch (grid extent of the compositing kernel)
  uint8_t* rz;         // scratch [n_crops][224][224][4]: crops after cv2.resize 256->224, BGR0 per pixel
};

// tfm: 2x3 float64 frame->crop(224) affine (row-major).  Computes M / A and the bounding box.
void paste_prepare_face(PasteFace* f, const double* tfm, int H, int W, int crop_index, int mask_index);
// 224-entry cv2.resize tables (host): idx, a0, a1
void paste_resize_tables(short* idx, short* a0, short* a1);

// bytes of scratch paste_back_launch needs behind p.rz
inline size_t paste_rz_bytes(int n_crops) { return static_cast<size_t>(n_crops) * 224 * 224 * 4; }
// host_faces: the same PasteFace records p.faces points at (device), for the launch geometry
const char* paste_back_launch(PasteParams p, const PasteFace* host_faces, cudaStream_t s);

// F1: uint8 BGR NHWC 256x256 crops -> normalised RGB NHWC-3 activations (utils/inference/core.py:13-26)
const char* preproc_u8_launch(const uint8_t* src, void* dst, int dst_dtype, long long npix, int half_rounding,
                              cudaStream_t s);
// F2: generator output NHWC-3 RGB in [-1,1] -> uint8 BGR NHWC (utils/inference/faceshifter_run.py:20-22)
const char* postproc_u8_launch(const void* y, int y_dtype, uint8_t* dst, long long npix, cudaStream_t s);
// A0: 224x224 RGB HWC (uint8 or float32, /255 when is_u8) -> (x-.5)/.5 -> bilinear 1/2 (align_corners=True)
//     -> NHWC-3 112x112 (image_processing.py:23-48 + core.py:45,54)
const char* arcface_preproc_launch(const void* src, int is_u8, void* dst, int dst_dtype, int N, cudaStream_t s);

}  // namespace ghost
