// Host-side convolution descriptors: one ConvOp describes a conv / conv-transpose
// phase / linear layer and is lowered either to the tcgen05 implicit-GEMM kernel
// (bf16 mode) or to the CUDA-core kernel (fp32 mode, Cin not a multiple of 32).
#pragma once
#include <memory>
#include <string>
#include <vector>

#include "igemm.cuh"
#include "simt.cuh"
#include "tapn.cuh"

namespace ghost {

// 16-bit storage types of the tensor-core path: bf16, or IEEE fp16 (the reference's own inference dtype, inference.py:30)
inline bool is16(int dtype) { return dtype == kBF16 || dtype == kF16; }

// Dense NHWC activation tensor (C = pixel stride in elements).
struct Act {
  void* p = nullptr;
  int N = 0, H = 0, W = 0, C = 0;
  int dtype = kBF16;
  size_t elems() const { return static_cast<size_t>(N) * H * W * C; }
  size_t bytes() const { return elems() * (dtype == kF32 ? 4 : 2); }
};

// Geometry of the sliding window: input position = stride * o + (dh, dw).
struct Taps {
  int n = 0;
  int stride = 1;
  short dh[kIgemmMaxTaps], dw[kIgemmMaxTaps];
  static Taps square(int k, int pad, int stride);
  static Taps deconv_phase(int r, int s);  // ConvTranspose2d k4 s2 p1, output phase (r, s)
  static Taps point();
};

// Weights of one op, already in the packed [Cout][tap][Cin] order, resident on the device.
struct PackedW {
  const __nv_bfloat16* wb = nullptr;  // [rows_pad][K] bf16 or fp16 bits, the net's 16-bit storage type (igemm)
  const float* wf = nullptr;          // [Cout][K] fp32 (CUDA-core path)
  const float* bias = nullptr;        // [rows_pad] fp32 or null
  const __nv_bfloat16* wt = nullptr;  // Cout <= 3, K = 9*Cin: the "taps in N" rows of tapn.cuh, or null
  int Cout = 0, rows_pad = 0, K = 0;
};

struct Epilogue {
  // tensor-core path: bias per border class of the output pixel, [9][bias9_ld] (igemm.cuh bias9), instead of PackedW::bias
  const float* bias9 = nullptr;
  int bias9_ld = 0;
  int act1 = kActNone;
  float slope = 0.f;
  const float* slope_c = nullptr;
  const void* res = nullptr;  // residual, same dtype as the output
  long long r_sn = 0, r_sh = 0, r_sw = 0, r_off = 0;
  int act2 = kAct2None;
  // optional second output (tensor-core path, bf16, same strides as the first): out2 = out * scale2[c] + shift2[c]
  // computed from the value as stored - iresnet's bn1 of the next block (arcface iresnet.py IBasicBlock.forward)
  void* out2 = nullptr;
  const float* scale2 = nullptr;
  const float* shift2 = nullptr;
  // tensor-core path: also produce the InstanceNorm partial statistics of the output (igemm.cuh st_mode); after
  // lowering, LaunchOp::st_floats says how large the parts buffer must be (set_stats()), LaunchOp::ig.st_P is the P the
  // fold kernel reads it with
  bool want_stats = false;
  // tensor-core path (CTA-pair kernel): a second input tensor whose x2_cin channels (a multiple of 64) follow the conv's
  // own K columns in the packed weight matrix, read at the conv's stride with a 1x1 tap: out = conv(in) + pointwise(x2)
  // in one accumulator (igemm.cuh x_kb).  x2 has the spatial size of `in`.  lower_conv fails when the CTA-pair kernel
  // does not apply (the caller keeps the two-launch form for that shape).
  Act x2;
  int x2_cin = 0;
};

struct OutView {
  void* p = nullptr;
  long long sn = 0, sh = 0, sw = 0, off = 0;  // element strides over (n, oh, ow) + constant offset
  int dtype = kBF16;                          // kBF16 / kF32
  static OutView dense(const Act& a, int c_off = 0) {
    OutView o;
    o.p = a.p;
    o.sn = static_cast<long long>(a.H) * a.W * a.C;
    o.sh = static_cast<long long>(a.W) * a.C;
    o.sw = a.C;
    o.off = c_off;
    o.dtype = a.dtype;
    return o;
  }
  // ConvTranspose2d output phase (r, s): pixel (2q+r, 2p+s) of a tensor twice the grid size
  static OutView deconv_phase(const Act& a, int r, int s, int c_off = 0) {
    OutView o;
    o.p = a.p;
    o.sn = static_cast<long long>(a.H) * a.W * a.C;
    o.sh = 2LL * a.W * a.C;
    o.sw = 2LL * a.C;
    o.off = (static_cast<long long>(r) * a.W + s) * a.C + c_off;
    o.dtype = a.dtype;
    return o;
  }
};

struct AadArgs {  // only for the fused AAD epilogue
  const void* h = nullptr;
  int c_x = 0;
  const float4* ncpar = nullptr;
  const float* wh_fold = nullptr;
  const float* bh_fold = nullptr;
  const float* mask = nullptr;  // precomputed sigmoid mask [N*H*W]; required when c_x > 128
  // SPADE variant (igemm.cuh): no mask, out = max(v, v*slope) * post[0]
  int spade = 0;
  float slope = 0.f;
  const float* post = nullptr;
  int no_h = 0;  // SPADE variant without an h tensor (h = nullptr): out = act(beta half + bias) * post
  int pair = 0;  // two layers on the same h / z_attr as the two N tiles (igemm.cuh aad_pair); c_x = 2 * channels of h
};

// The four output phases of a ConvTranspose2d(k4, s2, p1) as one launch (igemm.cuh `nphase`): `w` passed to lower_conv
// holds the four phase matrices stacked along the rows (Cout = 4 * Cout_phase, phase-major, bias stacked likewise),
// `taps` is phase 0's, `out` / `epi.res` the views of phase (0, 0); here: the other phases' taps and view offsets.
struct DeconvPhases {
  Taps taps[4];
  long long o_off[4] = {0, 0, 0, 0};  // OutView::off of each phase
  long long r_off[4] = {0, 0, 0, 0};  // residual view offset of each phase (ignored without a residual)
};

// A lowered, launch-ready op.
struct LaunchOp {
  enum Kind { kIgemm, kSimt, kTapn } kind = kIgemm;
  IgemmParams ig;
  TapnParams tp;
  SimtConvParams sc;
  int simt_dtype = kBF16;
  std::shared_ptr<RgbConvConst> rgb;  // Cin = 3 stems: weights by value in the kernel parameters (simt.cuh)
  std::string name;
  double flops = 0.0;
  // split-K (ig.splits > 1): bytes of fp32 workspace the caller must provide through set_workspace() before launch
  size_t ws_bytes = 0;
  void set_workspace(void* ws) { ig.ws = static_cast<float*>(ws); }
  // fused output statistics (Epilogue::want_stats): floats of the parts buffer [N][2P+1][C]; zeroed at every launch
  size_t st_floats = 0;
  void set_stats(float* parts) { ig.st_parts = parts; }
  const char* launch(cudaStream_t s) const {
    if (kind == kSimt && rgb) return rgb_const_launch(sc, *rgb, simt_dtype, s);
    if (kind == kIgemm && ig.st_mode) {
      if (!ig.st_parts) return "conv with fused statistics launched without its parts buffer";
      if (cudaMemsetAsync(ig.st_parts, 0, st_floats * sizeof(float), s) != cudaSuccess) return "stats parts memset failed";
    }
    if (kind == kIgemm && ig.splits > 1) {
      if (!ig.ws) return "split-K op launched without its workspace";
      if (const char* e = igemm_launch(ig, s)) return e;
      return splitk_reduce_launch(ig, s);
    }
    return kind == kIgemm ? igemm_launch(ig, s) : kind == kTapn ? tapn_launch(tp, s) : simt_conv_launch(sc, simt_dtype, s);
  }
};

// Lower a convolution.  `in` is read on channels [c0, c0+Cin); the output grid is Ho x Wo per image.
// force_simt selects the CUDA-core kernel (fp32 mode).  Returns nullptr on success.
const char* lower_conv(LaunchOp* op, const Act& in, int c0, int Cin, const Taps& taps, int Ho, int Wo,
                       const PackedW& w, const OutView& out, const Epilogue& epi, bool force_simt, int* err_word,
                       const AadArgs* aad = nullptr, int bn_hint = 0, const DeconvPhases* phases = nullptr);

// Conv chains (igemm.cuh `chain`): consecutive lowered convs that feed one another, all on the CTA-pair kernel with
// identical geometry, as ONE launch.  chain_compatible() says whether `next` may follow `first` in a chain;
// make_chain() turns steps (>= 2) into one launch-ready op - `dev_params` is device memory for steps.size() parameter
// blocks (chain_param_bytes(), 128-byte aligned), filled here with a synchronous copy.  Returns nullptr on success and
// sets *fused = false (leaving `out` untouched) when the chain would cost more rounds of the persistent grid than the
// separate launches (groups of whole images per cluster do not always divide the tile count as well).
bool chain_compatible(const LaunchOp& first, const LaunchOp& next);
inline size_t chain_param_bytes(size_t steps) { return steps * sizeof(IgemmParams); }
const char* make_chain(LaunchOp* out, const std::vector<LaunchOp>& steps, void* dev_params, bool* fused);

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time libcuda dependency).
const char* encode_tmap(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                        const uint32_t* box, int swizzle_bytes);

}  // namespace ghost
