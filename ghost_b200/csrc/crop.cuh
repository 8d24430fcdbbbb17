// Crop extraction: cv2.warpAffine(frame, M, (224,224), borderValue=0) + cv2.resize(., (256,256)).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ghost {

struct CropFace {
  double M[6];  // src = M * dst: the matrix cv2.warpAffine iterates with (the inverse of the frame -> crop affine)
  int frame;    // source frame of this face
  int pad_;
};

// tfm: 2x3 float64 frame -> 224-crop affine (row-major), as tfm_array holds them (video_processing.py:162-166)
void crop_prepare_face(CropFace* f, const double* tfm, int frame);
// 7 x 256 shorts: x index, x a0, x a1, row0, row1, b0, b1 of cv2.resize 224 -> 256
void crop_resize_tables(short* t);
// frames [F,H,W,3] u8; crops224 [n,224,224,3] (required, also the intermediate), crops256 [n,256,256,3] or null
const char* crop_faces_launch(const uint8_t* frames, int H, int W, const CropFace* faces_dev, int n,
                              const short* tables_dev, uint8_t* crops224, uint8_t* crops256, cudaStream_t s);

// message returned by ghost_last_error() on the calling thread (defined in c_api.cu); for entry points that live
// next to their kernels instead of in c_api.cu
void set_last_error(const char* msg);

}  // namespace ghost
