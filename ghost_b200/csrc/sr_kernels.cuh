// CUDA-core kernels of the pix2pix super-resolution generator: everything around the
// tcgen05 convolutions of LIPSPADEGenerator - normalisation folds, the LIP pooling, the SPADE modulation,
// nearest x2, the spectral-norm power iteration and the uint8 conversions of face_enhancement.
// All activation tensors are NHWC with a channel count that is a multiple of 8 (16-byte vectors).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "simt.cuh"

namespace ghost {

enum SrAct : int { kSrNone = 0, kSrRelu = 1, kSrLeaky02 = 2 };

// 3x3 neighbourhoods of the nearest-downsampled conditioning image (SPADE.forward, normalization.py:99-100:
// F.interpolate(segmap, size, 'nearest') then a 3x3 pad-1 conv): patches[n,i,j, tap*3+c] = img[n,(i+dh)*f,(j+dw)*f,c]
// (zero outside the S x S grid, channels 27..31 zero), f = 256/S.  Every 3->nh conv on the image then is a 1x1
// GEMM with K = 32 on the tensor cores.
const char* seg_patches_launch(const void* img_nhwc3, void* patches, int dtype, int N, int S, cudaStream_t s);

// Statistics fold.  parts as written by in_stats_launch(_ex).  pool_n = 0: InstanceNorm2d, one (mean, var) per
// (n, c); pool_n = 1: BatchNorm2d in training mode, one per channel over (N, H, W) - the parts must then have been
// produced with common_shift = 1.  ss[n*ss_stride + c] = {scale, shift} with scale = weight[c] * rstd,
// shift = bias[c] - mean * scale (weight / bias may be null = 1 / 0); biased variance, eps inside the sqrt.
const char* norm_fold_launch(const float* parts, int P, int N, int HW, int C, int pool_n, const float* weight,
                             const float* bias, float eps, float2* ss, cudaStream_t s);
// BatchNorm (training) fold written as the parameter table of the fused tcgen05 SPADE epilogue (igemm.cuh `ncpar`):
// [N][C] float4, per channel pair {scale0, scale1, shift0, shift1}, {0, 0, 0, 0}; the table must start zeroed.
const char* norm_fold_ncpar_launch(const float* parts, int P, int N, int HW, int C, float eps, float4* ncpar,
                                   cudaStream_t s);
// in_stats_launch with one shift per channel shared by all images (sums become additive over the batch)
const char* in_stats_launch_ex(const void* x, int dtype, int N, int HW, int C, float* parts, int common_shift,
                               cudaStream_t s);

// y = act(x * scale[n,c] + shift[n,c]);  ss may be null (identity); ss_stride_n = C (per image) or 0 (per channel)
const char* norm_act_launch(const void* x, void* y, int dtype, int N, int HW, int C, const float2* ss,
                            int ss_stride_n, int act, cudaStream_t s);

// Local importance pooling (generator.py:311-351): w = exp(12 * sigmoid(logit * scale + shift)) with the
// InstanceNorm2d(affine) fold in ss[n][c]; out = sum_3x3,s2,p1 (x * w) / sum (w).
const char* lip_pool_launch(const void* x, const void* logit, void* out, int dtype, int N, int H, int W, int C,
                            const float2* ss, cudaStream_t s);

// SPADE modulation (normalization.py:94-107) + optional LeakyReLU(0.2) (architecture.py:69-70):
//   out[n,h,w,c] = act((x[n, h>>up, w>>up, c] * scale[c] + shift[c]) * gamma + beta) * (post ? *post : 1)
// gb = [N,H,W,2C] = (gamma | beta); up = 1 reads x through nn.Upsample(scale_factor=2) (nearest, generator.py:63);
// post = 1/sigma of the spectral-normalised conv that consumes `out` (conv(x, W/sigma) = conv(x/sigma, W)).
const char* spade_apply_launch(const void* x, const void* gb, void* out, int dtype, int N, int H, int W, int C, int up,
                               const float2* ss, int act, const float* post, cudaStream_t s);

const char* nearest_up2x_launch(const void* x, void* y, int dtype, int N, int H, int W, int C, cudaStream_t s);

// One power iteration of torch.nn.utils.spectral_norm (training mode) on W [R][K] fp32 (weight_orig.view(R,-1)):
//   v = normalize(W^T u); u = normalize(W v); sigma = u . (W v); inv_sigma[0] = 1 / sigma   (eps 1e-12)
// u [R] and v [K] are updated in place; scratch holds sn_scratch_floats(R, K) floats.
size_t sn_scratch_floats(int R, int K);
const char* sn_power_iter_launch(const float* W, int R, int K, float* u, float* v, float* inv_sigma, float* scratch,
                                 cudaStream_t s);

// The same iteration for a table of convs in four launches (blockIdx.z = conv); descs_dev is a device array.
struct SnDesc {
  const float* W;
  float *u, *v, *inv_sigma, *scratch;
  int R, K;
};
const char* sn_power_iter_batched_launch(const SnDesc* descs_dev, int n, int max_R, int max_K, cudaStream_t s);

// face_enhancement pre/post (video_processing.py:253-256, 277-282)
const char* sr_preproc_launch(const uint8_t* crops_bgr, void* dst_nhwc3, int dtype, int N, int HW, cudaStream_t s);
const char* sr_postproc_launch(const void* y_nhwc3, int dtype, uint8_t* dst_bgr, int N, int HW, cudaStream_t s);

}  // namespace ghost
