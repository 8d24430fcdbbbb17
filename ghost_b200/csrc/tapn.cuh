// 3x3 stride-1 pad-1 convolution with at most 3 output channels on the tensor cores, "taps in N":
//   AADBlk8's two 64 -> 3 convs (network/AADLayer.py:69-79 as used by network/AEI_Net.py:113,139) and the SR
//   generator's to_rgb conv (models/networks/generator.py:133-135).
// The implicit GEMM of igemm.cu spends one K block per (tap, 64 channels) and a tcgen05.mma costs the same whether
// N is 16 or 256, so a Cout = 3 conv costs as much as a Cout = 64 one (0.37 ms at 64 x 256 x 256).  Here the nine
// taps become N instead of K:  T[pixel][tap*3 + co] = sum_ci in[pixel][ci] * w[co][ci][tap]  is ONE K block per 64
// input channels over an 8 x 16 pixel tile (halo included), and the epilogue gathers
//   out[y][x][co] = sum_tap T[y + dh(tap)][x + dw(tap)][tap*3 + co]
// from shared memory for the 6 x 14 interior pixels.  The tile's halo comes from the TMA box (zero-filled outside
// the image = the conv padding).  9x fewer MMAs; the kernel is bound by reading the input once.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <vector>

#include "igemm.cuh"
#include "simt.cuh"

namespace ghost {

constexpr int kTapnMaxChunks = 2;  // Cin <= 128

struct TapnParams {
  CUtensorMap tmA;  // input NHWC bf16, box {64 channels, 16, 8, 1}, SWIZZLE_128B
  CUtensorMap tmW;  // weights [chunks*32][64] bf16 (row = tap*3 + co, rows 27..31 zero), box {64, 32}, SWIZZLE_128B
  const float* bias;  // [3] or null
  int chunks, c0, cout;
  int N, H, W;
  int tiles_x, tiles_y, tiles;
  void* out;
  int out_f32;
  int f16;  // 16-bit storage type: 0 = bf16, 1 = fp16 (input, weights, residual and a 16-bit output)
  long long o_sn, o_sh, o_sw, o_off;
  const void* res;  // residual in the output's dtype, or null
  long long r_sn, r_sh, r_sw, r_off;
  int act2;  // kAct2Tanh / kAct2None
  int* err;
};

// rows [chunks*32][64] fp32 (for DevPool::upload_bf16) from the igemm row layout [Co <= 3][9 taps][Ci]
std::vector<float> tapn_pack_rows(const float* rows_co_tap_ci, int Co, int Ci);

// tile geometry: 8 x 16 pixels per tile with the halo ring, 6 x 14 outputs (lower_conv fills TapnParams)
constexpr int kTapnTileW = 16, kTapnTileH = 8;
const char* tapn_launch(const TapnParams& p, cudaStream_t s);

}  // namespace ghost
