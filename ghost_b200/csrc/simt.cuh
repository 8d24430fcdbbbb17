// CUDA-core kernels: the fp32 "correctness mode" convolution, the thin layers
// that do not fit a tensor-core tile (Cin = 3), and every bandwidth-bound
// elementwise / reduction step of the hot path.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace ghost {

enum DType : int { kF32 = 0, kBF16 = 1, kF16 = 2, kU8 = 3 };

struct SimtConvParams {
  const void* in;  // NHWC, element type = act dtype
  int N, H, W, C;  // C = pixel stride in elements
  int c0, Cin;     // channel slice read by the conv
  int stride;      // input pos = stride * o + d
  int ntaps;
  short dh[16], dw[16];
  int Ho, Wo;      // output pixel grid
  const float* w;  // [Cout][ntaps*Cin]
  int Cout;
  void* out;
  long long o_sn, o_sh, o_sw, o_off;
  const float* bias;
  int act1;
  float slope;
  const float* slope_c;
  const void* res;
  long long r_sn, r_sh, r_sw, r_off;
  int act2;
};

// dtype: kF32 or kBF16 activations (weights and accumulation are always fp32)
const char* simt_conv_launch(const SimtConvParams& p, int dtype, cudaStream_t s);

// The Cin = 3 stems (AEI_Net encoder conv1: 4x4 s2 p1 -> 32; iresnet conv1: 3x3 s1 p1 -> 64) with the whole weight
// matrix, bias and slopes passed BY VALUE as a kernel parameter (13 KB; CUDA >= 12.1 allows 32 KB): the FFMAs take
// their weight operand straight from the constant bank, so the loop is pure FMA issue.  (Weights in shared memory
// cost one LDS.128 per four FFMA, and a broadcast LDS.128 still occupies the SM's shared-memory pipe for four cycles
// - the layer ran at 29 % of the FMA rate.)
struct RgbConvConst {
  float w[48 * 64];  // [tap*3 + c][Cout]
  float bias[64];
  float slope[64];   // LeakyReLU / PReLU slope per channel (act1), unused otherwise
};
// true when (p, taps) is one of the two stem shapes rgb_const_launch handles
bool rgb_const_supported(const SimtConvParams& p, int dtype);
// fills `c` from the device-side weights / bias / slopes of p (synchronous copies: plan-build time only)
const char* rgb_const_prepare(const SimtConvParams& p, RgbConvConst* c);
const char* rgb_const_launch(const SimtConvParams& p, const RgbConvConst& c, int dtype, cudaStream_t s);

// Instance-norm statistics, NHWC, as deterministic per-block partial sums about a per-(n,c) shift.
// parts: [N][2*P + 1][C] fp32 with P = in_stats_parts(N, HW, C) (layout in simt.cu); fully overwritten by the call.
int in_stats_parts(int N, int HW, int C);
const char* in_stats_launch(const void* x, int dtype, int N, int HW, int C, float* parts, cudaStream_t s);

// Per-AAD-layer fold of the statistics (fixed-order sum of the P parts, in double) with the identity FCs and conv_h.
//   ncpar[n][c] = per channel pair: {rstd0, rstd1, -mu0*rstd0, -mu1*rstd1}, {gamma_id0, gamma_id1, beta_id0, beta_id1}; wh_fold[n][c] = w_h[c]*rstd; bh_fold[n] = b_h - sum_c w_h*mu*rstd
// scratch: aad_fold_scratch_floats(N, C) floats, zero before the first launch (the kernel leaves it zeroed)
size_t aad_fold_scratch_floats(int N, int C);
// the per-layer half of a fold (see aad_fold_launch); two layers on the same statistics share one launch
struct AadFoldSet {
  int off_gamma = 0, off_beta = 0;
  const float* w_h = nullptr;
  float b_h = 0.f;
  float4* ncpar = nullptr;
  float* wh_fold = nullptr;
  float* bh_fold = nullptr;
  float* scratch = nullptr;  // own scratch per layer
  int ld_out = 0;
};
const char* aad_fold2_launch(const float* parts, int P, int N, int HW, int C, const float* zfc, int ldz,
                             const AadFoldSet& a, const AadFoldSet* b, cudaStream_t s);
const char* aad_fold_launch(const float* parts, int P, int N, int HW, int C, const float* zfc, int ldz, int off_gamma,
                            int off_beta, const float* w_h, float b_h, float4* ncpar, float* wh_fold, float* bh_fold,
                            float* scratch, cudaStream_t s, int ld_out = 0);  // ld_out: row stride of ncpar / wh_fold (0 = C)

// bf16 mode, c_x > 128: the per-pixel sigmoid mask of an AADLayer, mask[N*HW] fp32.
const char* aad_mask_launch(const void* h, int N, int HW, int C, const float* wh_fold, const float* bh_fold, float* mask,
                            cudaStream_t s, int dtype = kBF16);  // dtype of h: kBF16 or kF16
// the masks of two AADLayers that read the same h (shortcut layer + first main layer of a block) in one pass over h
const char* aad_mask2_launch(const void* h, int N, int HW, int C, const float* wh_fold, const float* bh_fold, float* mask,
                             const float* wh_fold2, const float* bh_fold2, float* mask2, cudaStream_t s,
                             int dtype = kBF16);

// Unfused AAD blend (+ReLU) used by the fp32 mode: gb = [N,H,W,2C] (gamma | beta, bias included).
const char* aad_apply_launch(const void* h, const void* gb, void* out, int dtype, int N, int HW, int C,
                             const float4* ncpar, const float* wh_fold, const float* bh_fold, cudaStream_t s);

// Bilinear x2, align_corners=True, NHWC (F.interpolate call sites network/AEI_Net.py:94,125-137).
// The output may be a channel slice of a wider tensor (oC = pixel stride, o_c0 = first channel).
const char* upsample2x_launch(const void* x, void* y, int dtype, int N, int H, int W, int C, int oC, int o_c0,
                              cudaStream_t s);
// Same, fused with in_stats of the (dense, oC == C) output: parts in the layout of in_stats_launch with
// P = upsample2x_stats_parts(N, H, W, C) parts per image.
int upsample2x_stats_parts(int N, int H, int W, int C);
const char* upsample2x_stats_launch(const void* x, void* y, int dtype, int N, int H, int W, int C, float* parts,
                                    cudaStream_t s);

// Layout / dtype conversion at the operator boundary.
const char* nchw_to_nhwc_launch(const void* src, int src_dtype, void* dst, int dst_dtype, int N, int C, int H, int W,
                                int dstC, int dst_c0, cudaStream_t s);
const char* nhwc_to_nchw_launch(const void* src, int src_dtype, int srcC, int src_c0, void* dst, int dst_dtype, int N,
                                int C, int H, int W, cudaStream_t s);

// up1 (ConvTranspose2d 512->1024 k2 on a 1x1 map) is part of the z_id GEMM; this scatters its
// fp32 result (+bias) into the NHWC activation m1[N,2,2,1024].
const char* zfc_to_act_launch(const float* zfc, int ldz, int off, int count, const float* bias, void* dst, int dtype,
                              int N, cudaStream_t s);

// y[n,p,c] = x[n,p,c]*scale[c] + shift[c]  (iresnet BatchNorm that precedes a zero-padded conv)
const char* channel_affine_launch(const void* x, void* y, int dtype, long long npix, int C, const float* scale,
                                  const float* shift, const float* prelu, cudaStream_t s);

}  // namespace ghost
