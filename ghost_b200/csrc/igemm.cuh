// Parameter block shared by host and device for the tcgen05 implicit-GEMM
// convolution kernel (K1/K2/K4/K6/K9 in DESIGN.md).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace ghost {

constexpr int kIgemmMaxTaps = 16;
constexpr int kIgemmEpiWarps = 16;                           // four warps per TMEM lane group, splitting the columns
constexpr int kIgemmMmaWarps = 2;                            // MMA issuers: warp 1 takes even tiles, warp 2 odd tiles
constexpr int kIgemmEpiWarp0 = 4;                            // first epilogue warp (a multiple of 4: TMEM lane groups)
// warp 0: TMA producer of the activation operand (+ AAD h tiles, resident weights), warps 1-2: MMA issuers,
// warp 3: TMA producer of the weight operand, warps 4-19: epilogue
constexpr int kIgemmThreads = 32 * (kIgemmEpiWarp0 + kIgemmEpiWarps);

enum IgemmMode : int { kModeLinear = 0, kModeAAD = 1 };
enum Act1 : int { kActNone = 0, kActLeaky = 1, kActPRelu = 2, kActRelu = 3 };
enum Act2 : int { kAct2None = 0, kAct2Tanh = 1 };

struct alignas(64) IgemmParams {
  CUtensorMap tmA;  // activations, NHWC (4-D) or parity-split NHWC (5-D, stride-2 convs)
  CUtensorMap tmB;  // packed weights [Cout_pad][taps*Cin], K-major
  CUtensorMap tmH;  // AAD mode: block input h, NHWC, boxes of 64 channels x (bw, bh, bn), SWIZZLE_128B
  CUtensorMap tmO;  // AAD mode: output tensor, boxes of 64 channels x the 32 tile rows one epilogue warp owns (TMA store)

  // ---- M tiling: one tile = bn images x bh rows x bw cols (<= 128 GEMM rows)
  int a_rank;  // 4 or 5
  int bw, bh, bn;
  int tiles_w, tiles_h, tiles_nb, tiles_n;
  int Nimg, Ho, Wo;
  // ---- N / K
  int BN;       // UMMA N (multiple of 16, <= 256)
  int n_valid;  // real Cout (columns >= n_valid are never stored)
  int kc;       // K elements per stage: 64 (128B swizzle) or 32 (64B swizzle)
  int ntaps, kc_per_tap;
  int a_c0;                       // first input channel inside the inner dimension
  int b_k0;                       // first K column inside the packed weight matrix
  int splits;                     // split-K factor (grid.y); > 1: CTA row y handles K blocks [y, y+1) * nkb / splits and
                                  // stores its fp32 partial tile into ws (below); splitk_reduce_launch finishes the op
  short tap_c[kIgemmMaxTaps];     // added to the channel coordinate (parity * C for 5-D maps)
  short tap_w[kIgemmMaxTaps];     // added to the w coordinate
  short tap_p[kIgemmMaxTaps];     // h-parity coordinate (5-D maps only)
  short tap_h[kIgemmMaxTaps];     // added to the h coordinate
  int stages, tmem_cols;
  int sets;    // TMEM accumulator stages (tmem_cols / sets columns each): 2, or 4 for AAD tiles of 128 columns
  int h_bufs;  // AAD mode: shared-memory h tiles (sets or 2*sets); the blended tile overwrites its h tile in place
  // Row-window mode (3x3 stride-1 convs on wide images, bh == bn == 1, kc == 64): one TMA box of bw+2 pixels per
  // (kernel row, 64-channel chunk) serves the three horizontal taps as row-shifted UMMA views, so the A operand
  // crosses L2->shared 3x instead of 9x.  b_res: the whole packed weight slab stays resident in shared memory.
  int dual;    // 1: two MMA issuer warps (even / odd tiles), each with its own half of the operand ring
  int a_win;   // 0 off, 1 row window (the UMMA descriptor start address carries the row shift), 2 tile window:
               // CTA-pair kernel, narrow images - the (bh + 2) x (bw + 2) pixel patch of a tile of bh whole image rows
               // is loaded once per 64-channel chunk (a_slot bytes per ring slot) and all NINE taps are shifted views
               // of it; GEMM row m = pixel (m / (bw + 2), m % (bw + 2)), two junk rows per image row; weights resident
  int a_slot;
  int b_res;
  // Linear mode, bf16 output, BN 128 / 256: 1 = for the last tile of a CTA every epilogue warp stages its 32 rows x
  // BN/4 columns in the (then idle) operand ring and moves whole row segments to / from global memory (residual in,
  // out and out2 out): the one epilogue that no main loop hides.
  int stg;

  // ---- epilogue (linear):  v = act2(act1(acc + bias) + res)
  int mode;
  void* out;
  long long o_sn, o_sh, o_sw, o_off;  // element strides of the output pixel grid
  int out_f32;                        // 1: fp32 output, else bf16
  const float* bias;                  // [tiles_n*BN] or null
  // Bias per border class of the output pixel, [9][bias9_ld] with class = 3 * {first row, inner, last row} + {first
  // column, inner, last column}: what a per-channel affine (BatchNorm) in front of a zero-padded 3x3 conv turns into
  // when it is folded into the conv - the shift only reaches the output through the taps that lie inside the image
  // (iresnet bn1 -> conv1).  Replaces `bias` (which must be null); linear mode, no split-K.
  const float* bias9;
  int bias9_ld;
  int act1;
  float slope;
  const float* slope_c;  // per-channel PReLU slopes
  const __nv_bfloat16* res;
  long long r_sn, r_sh, r_sw, r_off;
  int act2;
  void* out2;            // optional second bf16 output, same strides: bf16(out) * scale2[c] + shift2[c]
  const float* scale2;
  const float* shift2;

  // ---- epilogue (AAD blend): columns [0,cn) = gamma_attr, [cn,2cn) = beta_attr, cn = BN/2
  // The h tile is TMA-loaded into shared memory, blended there in place and TMA-stored (nothing is read row by
  // row from global memory).  Every epilogue warp owns a 32-row x 64-channel piece of a tile and never syncs with
  // the others; the 16 warps form `sets` groups that work on different tiles at the same time.  When c_x > 128
  // (several N tiles per pixel) the sigmoid mask comes precomputed in `mask`.
  const __nv_bfloat16* h;  // raw block input, NHWC, pixel stride c_x (kept for reference / debugging)
  int c_x;
  const float4* ncpar;  // [Nimg][c_x] per channel pair: {rstd0, rstd1, -mu0*rstd0, -mu1*rstd1}, {gamma_id0, gamma_id1, beta_id0, beta_id1}
  const float* wh_fold;     // [Nimg][c_x] conv_h weight * rstd
  const float* bh_fold;     // [Nimg]      conv_h bias - sum(w*mu*rstd)
  const float* mask;    // optional precomputed sigmoid mask per pixel (else computed in-kernel)
  // SPADE variant of the same epilogue (sr.cu; models/networks/normalization.py:94-107 + architecture.py:69-70):
  // no attention mask (M = 0, the identity half of ncpar is ignored), out = max(v, v*slope) * post[0] with
  // v = gamma * (h*scale + shift) + beta: slope 0 = ReLU, 0.2 = LeakyReLU(0.2), 1 = no activation; post (device
  // scalar, may be null) is 1/sigma of the spectral-normalised conv that reads the result.
  int aad_spade;
  float aad_slope;
  const float* aad_post;
  // aad_no_h (SPADE variant only): there is no h tensor - nothing is loaded, h counts as 0, so with a zero ncpar
  // table the epilogue is out = act(beta-half of the accumulator + bias) * post, written through the same
  // shared-memory staging + TMA store.  This is how the thin-K convs of sr.cu (K = 32 .. 128 per pixel, where the
  // row-per-thread stores of the linear epilogue are the whole cost) leave the SM.
  int aad_no_h;
  // aad_pair: two AAD layers that read the SAME h and z_attr (the shortcut layer and the first main layer of an
  // AAD block, AADLayer.py / AEI_Net.py AAD_ResBlk) as the two N tiles of one launch: weights / bias / ncpar /
  // wh_fold rows of layer L at offset L * cn, bh_fold at L * Nimg, output channels [L*cn, (L+1)*cn) of a tensor of
  // c_x = 2*cn channels; h has cn channels and each layer's mask comes from its own tile.  Tiles walk the N index
  // fastest and the grid is even, so CTA pairs (2i, 2i+1) work on the same pixels at the same time: h and z_attr
  // cross HBM once for both layers (the second reader hits L2), and every CTA keeps one layer's weights resident.
  int aad_pair;

  // CTA-pair mode (igemm2_kernel, tcgen05 cta_group::2): clusters of two CTAs work on two neighbouring M tiles of the
  // same N tile as ONE 256 x BN MMA - each CTA loads its own 128 A rows and HALF of the weight tile (tmB boxes of
  // BN/2 rows), so a K block brings 16 KB + BN*64 B instead of 16 KB + BN*128 B into each SM.  Linear mode, 64-channel
  // chunks, streaming weights only.
  int cta2;

  // Output phases (the four phases of a ConvTranspose2d k4 s2 p1 as ONE launch): the N-tile index runs over
  // nphase * tpn tiles, tile nt belongs to phase nt / tpn.  The phase weight matrices (and bias vectors) are stacked
  // along the rows, so weight rows and bias columns follow nt * BN as usual; phase ph uses taps [ph*ntaps,
  // (ph+1)*ntaps) of the tap tables and adds ph_o[ph] / ph_r[ph] to the output / residual pixel offsets (these
  // already carry -ph * Cout, so "pixel offset + stacked column" addresses the right channel).  nphase = 1: off.
  int nphase, tpn;
  long long ph_o[4], ph_r[4];

  // Deterministic split-K for layers with a handful of output tiles and a long K (2x2 .. 8x8 feature maps, the
  // ArcFace FC): partial accumulators [splits][tiles_m * 128 rows][ws_cols] fp32, plain stores; the reduce kernel adds
  // them in split order and applies the whole linear epilogue (bias, activation, residual, second output).
  float* ws;
  int ws_cols;  // tiles_n * BN

  // InstanceNorm statistics of the OUTPUT in the epilogue (the next AADLayer normalises exactly this tensor,
  // AADLayer.py:21): per-column sum and sum of squares of the values as stored, in the partial-sum layout of
  // in_stats_kernel (simt.cu) with a zero shift row: st_parts [Nimg][2*st_P + 1][st_C].  Tiles lie inside one image.
  //   st_mode 1: one part per tile of an image (st_P = tiles per image), every part written by exactly one CTA;
  //   st_mode 2: one part per CTA (st_P >= grid size, buffer zeroed before the launch): a CTA adds up its
  //              consecutive tiles of an image in registers and flushes when the image changes (one N tile only).
  // Either way the fold kernel adds the parts in index order: deterministic.
  float* st_parts;
  int st_mode, st_P, st_C;

  // TMA-store epilogue of thin-K linear tiles (tst = 1): the row-per-lane 16-byte stores of the plain epilogue cost 32
  // LSU wavefronts per instruction (32 lanes, 32 different 128-byte lines) and bound the 3x3 convs of AADBlk7-8
  // (measured: the MMA issuers wait half of their time for a free accumulator stage).  Instead the four warps of a
  // TMEM lane group write their 32 rows x BN columns into a double-buffered shared-memory piece (128B swizzle, boxes
  // of 64 channels) and one lane sends it out through tmO (box 64 channels x 32 pixels).  BN 64 / 128, bw % 32 == 0,
  // 16-bit output with regular NHWC strides.
  int tst;
  int tst_c0;  // first output channel (view offset) inside the tensor tmO describes
  // Whole-tile TMA epilogue of the staged last tile (tte = 1, with stg): tmH / tmO / tmO2 describe the residual, the
  // output and the second output as NHWC tensors with boxes of 64 channels x (bw, bh, bn) pixels (128B swizzle); the
  // residual box lands in the idle operand ring, every lane adds its row in place, one TMA store per 64-column group
  // (and output) sends the tile out.  tst_c0 / tte_r0: channel offsets of the output / residual views.
  int tte;
  int tte_r0;
  CUtensorMap tmO2;

  int f16;   // 16-bit storage type of activations / weights / outputs: 0 = bf16, 1 = IEEE fp16 (same MMA rate)

  // Conv chain (CTA-pair kernel only): `chain_n` convs of IDENTICAL geometry (tiling, N tile, taps, ring) that feed
  // one another run as ONE launch.  A 3x3 conv only reads pixels of its own image, so when every cluster owns whole
  // images - `chain_spg` consecutive super tiles (pairs of M tiles x N tiles) form one group of images, cluster c works
  // on groups c, c + clusters, ... in every step - step k+1 of a group depends on step k of the SAME cluster only: a
  // cluster barrier (plus the generic -> async proxy fence for the TMA loads that follow plain stores) replaces the
  // kernel boundary, and with it the launch gap, the prologue and the teardown (~4 us of every ~20 us iresnet conv).
  // `chain` points at the device copies of all steps' parameter blocks (tensor maps are read from there); the block
  // passed to the launch is step 0's.  0 = off.
  const IgemmParams* chain;
  int chain_n, chain_spg;

  // Second activation tensor as extra K blocks (CTA-pair kernel, plain tap-shifted loads): after the ntaps * kc_per_tap
  // blocks of the conv, x_kb more blocks load 64-channel chunks of ANOTHER tensor through tmH (same box, same rank as
  // tmA, tap coordinates x_tap_*) against the K columns that follow in the packed weight matrix: out = conv(A) + 1x1
  // conv(X) in one accumulator - iresnet's stride-2 blocks, conv2(u) + downsample(x), without the downsample launch
  // and without a residual read.
  int x_kb, x_c0;
  short x_tap_c, x_tap_w, x_tap_p, x_tap_h;

  int* err;  // device word set non-zero when a pipeline wait timed out
};

// Host side (igemm.cu)
const char* igemm_launch(const IgemmParams& p, cudaStream_t stream);
// second pass of a split-K op: `p` is the op's parameter block (real output, epilogue, ws, splits)
const char* splitk_reduce_launch(const IgemmParams& p, cudaStream_t stream);
int igemm_smem_bytes(const IgemmParams& p);
// bytes behind the operand ring of a chain launch: two parameter blocks + two epilogue parameter tables
constexpr int kChainSmemBytes = 2 * static_cast<int>(sizeof(IgemmParams)) + 2 * 4096;

}  // namespace ghost
